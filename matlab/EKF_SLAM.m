classdef EKF_SLAM < EKFSLAM_B200_Base
    % Known-correspondence EKF-SLAM filter with the reference's interface (EKF_SLAM.m), B200-native body.
    %   s = EKF_SLAM();  s.predict(u);  s.measure(laserData, u, landmark_list);  s.x, s.P, s.s
    % Defaults follow EKF_SLAM.m:12-16 (C = 0.2, Rc = [.01 5], s_cost = 1e-11, s_thresh = 1e9); the
    % idx = ii slot rule of EKF_SLAM.m:123 is reproduced inside the library (variant 0).

    properties (Dependent)
        s_cost;
        s_thresh;
    end

    methods
        function h = EKF_SLAM(opts)
            if nargin < 1
                opts = struct();
            end
            h@EKFSLAM_B200_Base(0, opts);
        end
        function v = get.s_cost(h)
            v = h.s_cost_;
        end
        function set.s_cost(h, v)
            h.s_cost_ = v;
            h.pushNoise();
        end
        function v = get.s_thresh(h)
            v = h.s_thresh_;
        end
        function set.s_thresh(h, v)
            h.s_thresh_ = v;
            h.pushNoise();
        end
    end
end
