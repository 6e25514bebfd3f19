classdef RefFixtureStub < handle
    % Duck-typed stand-in for the reference's `Landmark` object (Landmark.m:25-33), used by run_reference_fixture.m to
    % drive the UNMODIFIED reference classes with recorded inputs: getLandmark(laserData, x) returns the step's
    % observed_LL rows; landmarkObj.landmark(k).index / .loc is the landmark table the filter reads
    % (EKF_SLAM.m:111,120; EKF_SLAM_UC.m:113,123; fields as in RANSAC.m:238-241).
    properties
        observed_LL = zeros(0, 3);
        landmarkObj = struct('landmark', []);
    end
    methods
        function h = RefFixtureStub(rows, table_index, table_loc)
            h.observed_LL = rows;
            lm = struct('index', {}, 'loc', {}, 'observe', {}, 'fresh', {});
            for k = 1:numel(table_index)
                lm(k).index = table_index(k);
                lm(k).loc = table_loc(k, :);
                lm(k).observe = 1;
                lm(k).fresh = 0;
            end
            h.landmarkObj = struct('landmark', lm);
        end
        function observed_LL = getLandmark(h, ~, ~)
            observed_LL = h.observed_LL;
        end
    end
end
