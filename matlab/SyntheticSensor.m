classdef SyntheticSensor < handle
    % Deterministic range-bearing "extractor" that stands in for RANSAC behind Landmark.getLandmark:
    % a seeded world of N point landmarks, a true pose advanced by step(), and getLandmark(~, x)
    % returning observed_LL rows [range, bearing_deg, index] plus the .landmark(i).{index,loc} table
    % the filters read (EKF_SLAM.m:111,120).  Same world rules as ekf_slam_b200/synth.py (SURVEY 8d):
    % square of side 10*sqrt(N/20) m around the centre of a circular trajectory of radius 0.35*L that starts at the
    % filter's initial state [0 0 0], k nearest landmarks, observations within 20 deg of the 0/360 bearing seam
    % dropped (the innovation is not wrapped).  Landmarks are numbered in order of first sighting, as RANSAC.m:262
    % numbers them: EKF_SLAM_UC looks a new landmark up under index N+1 (EKF_SLAM_UC.m:123) and EKF_SLAM appends when
    % the observed index exceeds N (EKF_SLAM.m:118), so the table index must be the sighting order, not the world id.

    properties
        landmark;        % struct array: index, loc, observe, fresh (RANSAC.m:238-241 field names)
        truth;           % N x 2 true landmark positions
        assigned;        % N x 1: index given to world landmark i at first sight, 0 = not seen yet
        pose = [0 0 0];  % true robot pose [x y theta_deg]
        k = 5;
        prefix = false;  % true: emit landmarks 1..m in index order (needed by EKF_SLAM's idx = ii rule)
        L;
        dD;
    end

    methods
        function h = SyntheticSensor(N, seed)
            rng(seed, 'twister');
            h.L = 10 * sqrt(N / 20);
            h.truth = [(rand(N, 1) - 0.5) * h.L, 0.35 * h.L + (rand(N, 1) - 0.5) * h.L];
            h.assigned = zeros(N, 1);
            h.dD = 2 * pi * 0.35 * h.L / 360;
            h.landmark = struct('index', {}, 'loc', {}, 'observe', {}, 'fresh', {});
        end

        function u = step(h)
            % advance the true pose one step along the circle and return noisy odometry [dD, dtheta_deg]
            th = h.pose(3) + 1;
            h.pose = [h.pose(1) + h.dD * cosd(th), h.pose(2) + h.dD * sind(th), th];
            u = [h.dD * (1 + 0.01 * randn), 1 + 0.05 * randn];
        end

        function observed_LL = getLandmark(h, ~, ~)
            d = h.truth - h.pose(1:2);
            rngs = sqrt(sum(d .^ 2, 2));
            brg = mod(atan2d(d(:, 2), d(:, 1)) - h.pose(3), 360);
            ok = brg > 20 & brg < 340;
            if h.prefix
                m = find(~ok, 1) - 1;
                if isempty(m), m = numel(ok); end
                sel = (1:min(m, numel(h.landmark) + 1))';
            else
                [~, order] = sort(rngs);
                order = order(1:min(h.k, numel(order)));    % the k nearest ...
                sel = order(ok(order));                     % ... minus those in the seam zone
            end
            observed_LL = zeros(numel(sel), 3);
            for r = 1:numel(sel)
                i = sel(r);
                if h.assigned(i) == 0
                    h.assigned(i) = numel(h.landmark) + 1;
                    h.landmark(end + 1) = struct('index', h.assigned(i), 'loc', h.truth(i, :) + 0.05 * randn(1, 2), ...
                                                 'observe', 1, 'fresh', 1);
                end
                observed_LL(r, :) = [rngs(i) + 0.01 * randn, mod(brg(i) + 0.1 * randn, 360), h.assigned(i)];
            end
            [~, o] = sort(observed_LL(:, 3));               % rows in index order
            observed_LL = observed_LL(o, :);
        end
    end
end
