classdef EKF_SLAM_UC < EKFSLAM_B200_Base
    % Unknown-correspondence EKF-SLAM filter with the reference's interface (EKF_SLAM_UC.m), B200-native body.
    % Owns a Correspondence(1e-11, 1e9, 'EKF_SLAM_UC') like EKF_SLAM_UC.m:16; measure() associates every
    % observation on the device (K4 gating + first-argmin) before updating or appending (variant 1).

    properties
        correspondence;
    end

    methods
        function h = EKF_SLAM_UC(opts)
            if nargin < 1
                opts = struct();
            end
            h@EKFSLAM_B200_Base(1, opts);
            h.correspondence = Correspondence(1e-11, 1e9, 'EKF_SLAM_UC');
            h.correspondence = h.correspondence.bind(h);
        end

        function syncCorrespondence(h)
            % pick up edits made to h.correspondence.s_cost / s_thresh since the last call
            if h.correspondence.s_cost ~= h.s_cost_ || h.correspondence.s_thresh ~= h.s_thresh_
                h.s_cost_ = h.correspondence.s_cost;
                h.s_thresh_ = h.correspondence.s_thresh;
                h.pushNoise();
            end
        end

        function measure(h, laserData, u, landmark_list)
            h.syncCorrespondence();
            measure@EKFSLAM_B200_Base(h, laserData, u, landmark_list);
        end

        function [newLL, index] = associateResident(h, z, R)
            % estimateCorrespondence against the device-resident x, P, s (no copies)
            h.syncCorrespondence();
            [newLL, index] = ekfslam_mex('associate', h.hnd, z(1:3), R);
        end
    end
end
