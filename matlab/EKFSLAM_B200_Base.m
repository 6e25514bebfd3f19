classdef EKFSLAM_B200_Base < handle
    % Shared body of the B200-native EKF_SLAM / EKF_SLAM_UC classes in this directory.
    %
    % The public surface is the reference's (SamShue/EKF_SLAM EKF_SLAM.m:5-22, :40, :56, :67, :100):
    % properties x, P, Q, s, C, Rc, landmark_list, observed and methods predict(u), f(x,u),
    % append(u,R,landmarkPos,signature), measure(laserData,u,landmark_list), plot(landmark_list).
    % All filter arithmetic runs on the GPU inside libekfslam_b200.so; this class only marshals
    % arguments through ekfslam_mex (one MEX call per method).  x, P and s are Dependent: reading
    % them copies device -> host (P is 3.2 GB at 10k landmarks, so read it only when needed), assigning
    % one of them (h.x = v; h.P = M; h.s = v) uploads that property at the current size and keeps the
    % other two; setState uploads all three and may change the number of landmarks.

    properties (Dependent)
        x;   % 1 x n row vector [x y theta_deg l1x l1y ...]
        P;   % n x n covariance
        s;   % N x 1 landmark signatures
    end

    properties
        Q;               % process-noise matrix of the last predict (host-built on demand, zeros(n) + 3x3 corner)
        landmark_list;
        observed;
        actions;         % per observation of the last measure: +idx = update of slot idx, -idx = append with signature idx
    end

    properties (Access = protected)
        hnd = uint64(0);         % ekf_handle* as uint64
        C_ = 0.2;
        Rc_ = [.01, 5];
        s_cost_ = 1e-11;
        s_thresh_ = 1e9;
        last_u = [];
        last_theta = [];
    end

    properties (Dependent)
        C;
        Rc;
    end

    methods
        function h = EKFSLAM_B200_Base(variant, opts)
            % variant 0 = EKF_SLAM, 1 = EKF_SLAM_UC; opts: struct with optional fields
            % capacity_landmarks (default 1024), storage (0 full / 1 sym), assoc_mode, device, defer_m,
            % devices ([0 1 2 3]: ONE filter in column slabs over several GPUs, still one MATLAB process).
            if nargin < 2 || isempty(opts)
                opts = struct();
            end
            if ~isfield(opts, 'capacity_landmarks')
                opts.capacity_landmarks = 1024;
            end
            h.hnd = ekfslam_mex('create', variant, opts);
            if variant == 1
                h.Rc_ = [.1, 5];
            end
        end

        function delete(h)
            if h.hnd ~= 0
                ekfslam_mex('destroy', h.hnd);
                h.hnd = uint64(0);
            end
        end

        % ---- state (device resident) -------------------------------------
        function v = get.x(h)
            v = ekfslam_mex('get_x', h.hnd);
        end
        function v = get.P(h)
            v = ekfslam_mex('get_P', h.hnd);
        end
        function v = get.s(h)
            v = ekfslam_mex('get_s', h.hnd);
        end
        % x, P, s are public read/write properties upstream (EKF_SLAM.m:6-9): assigning one of them keeps the other two.
        % The size must match the current state (setState changes the number of landmarks).
        function set.x(h, v)
            ekfslam_mex('set_x', h.hnd, v(:)');
        end
        function set.P(h, M)
            ekfslam_mex('set_P', h.hnd, M);
        end
        function set.s(h, v)
            ekfslam_mex('set_s', h.hnd, v(:));
        end
        function setState(h, x, P, s)
            ekfslam_mex('set_state', h.hnd, x(:)', P, s(:));
        end
        function B = Pblock(h, r1, c1, nr, nc)
            B = ekfslam_mex('get_P_block', h.hnd, r1, c1, nr, nc);
        end
        function N = numLandmarks(h)
            N = ekfslam_mex('n_landmarks', h.hnd);
        end

        % ---- noise constants (public and mutable in the reference) -------
        function v = get.C(h)
            v = h.C_;
        end
        function set.C(h, v)
            h.C_ = v;
            h.pushNoise();
        end
        function v = get.Rc(h)
            v = h.Rc_;
        end
        function set.Rc(h, v)
            h.Rc_ = v;
            h.pushNoise();
        end

        % ---- filter methods ----------------------------------------------
        function predict(h, u)
            % EKF_SLAM.m:40-51: W, Q, f, P = F*P*F' + Q, theta = wrapTo360(theta)  -> K1 on the device
            ekfslam_mex('predict', h.hnd, u(1:2));
            h.last_u = u(1:2);
        end

        function [x_new, F] = f(~, x, u)
            % EKF_SLAM.m:56-65, pure host helper (the device never materialises the dense F)
            x_new = x;
            x_new(1) = x(1) + u(1) * cosd(x(3) + u(2));
            x_new(2) = x(2) + u(1) * sind(x(3) + u(2));
            x_new(3) = x(3) + u(2);
            F = eye(numel(x));
            F(1, 3) = -u(1) * sind(x(3));
            F(2, 3) = u(1) * cosd(x(3));
        end

        function append(h, u, R, landmarkPos, signature)
            % EKF_SLAM.m:67-98 -> K5 (in place at capacity; the reference reallocates P)
            ekfslam_mex('append', h.hnd, u(1:2), R, landmarkPos(1:2), signature);
        end

        function measure(h, laserData, u, landmark_list)
            % EKF_SLAM.m:100-151 / EKF_SLAM_UC.m:102-152: the whole per-observation loop (R build,
            % append / associate / update decision, idx=ii quirk of the known-correspondence class)
            % runs inside ekf_measure_batch; this method only flattens the landmark table.
            observed_LL = landmark_list.getLandmark(laserData, h.x);
            h.observed = observed_LL;
            if isempty(observed_LL)
                h.actions = [];
                return;
            end
            lm = landmark_list.landmarkObj.landmark;
            if isempty(lm)
                idx = zeros(0, 1);
                loc = zeros(0, 2);
            else
                idx = reshape([lm.index], [], 1);
                loc = reshape([lm.loc], 2, [])';
            end
            h.actions = ekfslam_mex('measure', h.hnd, observed_LL, u(1:2), loc, idx);
        end

        function plot(~, ~)
            % EKF_SLAM.m:154-234 draws figures; headless build: no-op (see covarianceEllipses).
        end

        function E = covarianceEllipses(h)
            % rows [cx, cy, semi-axis along the major eigenvector, semi-axis across, arc_rad] of the ellipses
            % plot() draws upstream (EKF_SLAM.m:180-229); row 1 = robot, row 1+b = landmark b.
            E = ekfslam_mex('get_ellipses', h.hnd);
        end

        function flush(h)
            ekfslam_mex('flush', h.hnd);
        end
    end

    methods (Access = protected)
        function pushNoise(h)
            if h.hnd ~= 0
                ekfslam_mex('set_noise', h.hnd, h.C_, h.Rc_, h.s_cost_, h.s_thresh_);
            end
        end
    end
end
