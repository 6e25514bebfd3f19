classdef EKFSLAM_B200_Batch < handle
    % n_filters independent EKF_SLAM / EKF_SLAM_UC filters stepped in lock step on one GPU (Monte-Carlo runs: BASELINE
    % config 5, 4096 filters x 200 landmarks).  Replaces a MATLAB loop over n_filters reference objects, each doing
    %     slam.predict(u); slam.measure(laserData, u, LM)            (SLAM.m:110,116)
    % One CUDA block owns one filter and runs the whole iteration on the device (association, append / update
    % decision, deferred covariance downdate); filters never communicate.  Thin marshalling over the ekfb_* C ABI
    % through ekfslam_mex('batch_*', ...); no filter arithmetic lives here.
    %
    %   b = EKFSLAM_B200_Batch('EKF_SLAM_UC', 4096, struct('capacity_landmarks', 200, 'max_obs', 5, 'storage', 1));
    %   b.setStateAll(x0, P0, s0);
    %   actions = b.step(U, observed, counts, LM);     % U: 2 x nf, observed: max_obs x 3 x nf, counts: nf x 1
    %   poses = b.poses();                             % 3 x nf, x(1:3) of every filter
    %   xf = b.x(f);  Pf = b.P(f);  sf = b.s(f);       % one filter's state, with the reference's shapes

    properties (SetAccess = private)
        hnd;            % uint64 handle of the gateway's batch record
        n_filters;
        max_obs;
        max_table;
    end

    methods
        function b = EKFSLAM_B200_Batch(cls, n_filters, opts)
            if nargin < 3
                opts = struct();
            end
            variant = double(strcmp(cls, 'EKF_SLAM_UC'));
            cap = EKFSLAM_B200_Batch.opt(opts, 'capacity_landmarks', 200);
            b.n_filters = n_filters;
            b.max_obs = EKFSLAM_B200_Batch.opt(opts, 'max_obs', 8);
            b.max_table = EKFSLAM_B200_Batch.opt(opts, 'max_table', cap);
            if ~isfield(opts, 'capacity_landmarks'), opts.capacity_landmarks = cap; end
            if ~isfield(opts, 'storage'), opts.storage = 1; end
            if ~isfield(opts, 'defer_m'), opts.defer_m = 8; end
            b.hnd = ekfslam_mex('batch_create', variant, n_filters, b.max_obs, b.max_table, opts);
        end

        function delete(b)
            if ~isempty(b.hnd)
                ekfslam_mex('batch_destroy', b.hnd);
                b.hnd = [];
            end
        end

        function actions = step(b, U, observed, counts, landmark_list, do_predict, do_measure)
            % One SLAM iteration of every filter.  landmark_list: a Landmark-like object whose
            % landmarkObj.landmark(:) entries carry .loc and .index (RANSAC.m:238-241), shared by all filters.
            if nargin < 6, do_predict = true; end
            if nargin < 7, do_measure = true; end
            lm = landmark_list.landmarkObj.landmark;
            T = numel(lm);
            loc = zeros(T, 2); index = zeros(T, 1);
            for i = 1:T
                loc(i, :) = lm(i).loc; index(i) = lm(i).index;
            end
            actions = ekfslam_mex('batch_step', b.hnd, U, observed, counts, loc, index, double(do_predict), double(do_measure));
        end

        function N = nLandmarks(b)
            N = ekfslam_mex('batch_n_landmarks', b.hnd);
        end

        function v = x(b, f), v = ekfslam_mex('batch_get_x', b.hnd, f); end     % 1 x n, EKF_SLAM.m:28
        function v = s(b, f), v = ekfslam_mex('batch_get_s', b.hnd, f); end     % N x 1
        function v = P(b, f), v = ekfslam_mex('batch_get_P', b.hnd, f); end     % n x n
        function v = poses(b), v = ekfslam_mex('batch_poses', b.hnd); end       % 3 x nf

        function setState(b, f, x, P, s)
            ekfslam_mex('batch_set_state', b.hnd, f, x, P, s);
        end

        function setStateAll(b, x, P, s)
            ekfslam_mex('batch_set_state', b.hnd, 0, x, P, s);
        end

        function sync(b)
            ekfslam_mex('batch_sync', b.hnd);
        end
    end

    methods (Static, Access = private)
        function v = opt(o, name, dflt)
            if isfield(o, name), v = o.(name); else, v = dflt; end
        end
    end
end
