classdef Correspondence
    % Data association with the reference's interface (Correspondence.m:12,28), scored on the GPU.
    %   c = Correspondence(cost, thresh, 'EKF_SLAM_UC');
    %   [newLL, index] = c.estimateCorrespondence(z, R, x, P, s)
    % Value class like upstream.  With explicit x, P, s the state is uploaded into a scratch filter
    % and scored by kernel K4; when the object belongs to an EKF_SLAM_UC (bind) and x is omitted or
    % empty, the owner's device-resident state is scored without any copy.

    properties
        s_cost;
        s_thresh;
        method;
    end

    properties (Hidden)
        owner = [];
    end

    methods
        function h = Correspondence(cost, thresh, method)
            h.s_cost = cost;
            h.s_thresh = thresh;
            if strcmp(method, 'EKF_SLAM_UC')
                h.method = method;
            else
                warning('ekfslam:method', 'Unknown correspondence method, falling back to ML.');
                h.method = 'ML';
            end
        end

        function h = bind(h, filter)
            h.owner = filter;
        end

        function [newLL, index] = estimateCorrespondence(h, z, R, x, P, s)
            if (nargin < 4 || isempty(x)) && ~isempty(h.owner)
                [newLL, index] = h.owner.associateResident(z, R);
                return;
            end
            N = (numel(x) - 3) / 2;
            scratch = EKF_SLAM_UC(struct('capacity_landmarks', max(N, 1)));
            scratch.correspondence.s_cost = h.s_cost;
            scratch.correspondence.s_thresh = h.s_thresh;
            scratch.setState(x, P, s);
            [newLL, index] = scratch.associateResident(z, R);
            delete(scratch);
        end
    end
end
