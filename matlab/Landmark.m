classdef Landmark < handle
    % Front-end seam with the reference's shape (Landmark.m:12,25): .method, .landmarkObj and
    % getLandmark(laserdata, x) -> observed_LL rows [range, bearing_deg, index].  The reference's only
    % extractor (RANSAC over a ROS LaserScan) is out of scope; 'SYNTH' wraps any object that offers
    % getLandmark(laserdata, x) and a .landmark struct array with fields .index and .loc, e.g.
    % SyntheticSensor in this directory.

    properties
        landmarkObj;
        method;
    end

    methods
        function h = Landmark(method, obj)
            h.method = method;
            if nargin > 1
                h.landmarkObj = obj;
            elseif strcmp(method, 'SYNTH')
                h.landmarkObj = SyntheticSensor(20, 20261104);
            else
                error('ekfslam:landmark', 'Only the SYNTH front-end ships with the B200 filter core.');
            end
        end

        function observed_LL = getLandmark(h, laserdata, x)
            observed_LL = h.landmarkObj.getLandmark(laserdata, x);
        end
    end
end
