% example_sim.m -- a runnable stand-in for the reference's example.m / test_SLAM.m (both need names or a
% robot that do not exist: example.m:7,9,10,20 read undefined h, Q, s, observed; test_SLAM.m:5-18 needs ROS).
% Same call pattern as SLAM.m:110,116: construct, then loop predict(u); measure(laserData, u, landmark_list).
% ~20 landmarks, simulated 2-D range-bearing sensor, filter arithmetic on the GPU through ekfslam_mex.

N = 20; T = 1000;
sensor = SyntheticSensor(N, 20261104);
LM = Landmark('SYNTH', sensor);
slam = EKF_SLAM_UC(struct('capacity_landmarks', N, 'storage', 1, 'defer_m', 8));

for t = 1:T
    u = sensor.step();
    slam.predict(u);
    slam.measure([], u, LM);
end

x = slam.x;
fprintf('%d landmarks mapped, pose [%.3f %.3f %.2f deg], true pose [%.3f %.3f %.2f deg]\n', ...
        slam.numLandmarks(), x(1), x(2), x(3), sensor.pose(1), sensor.pose(2), mod(sensor.pose(3), 360));
fprintf('trace(P_rr) = %.3e\n', trace(slam.Pblock(1, 1, 3, 3)));
delete(slam);
