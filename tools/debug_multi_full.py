"""Debug: 2-device FULL mode-B loop vs single GPU: first step where x or P differ, and what that step did."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import ekf_slam_b200 as E
from ekf_slam_b200.synth import make_world
for mode in ('position', 'signature'):
    w = make_world(N=16, T=120, k=4, mode='nearest', seed=5)
    fs = [E.EKF_SLAM_UC(capacity_landmarks=16, storage='full', assoc_mode=mode, defer_m=8, devices=[0, 1]),
          E.EKF_SLAM_UC(capacity_landmarks=16, storage='full', assoc_mode=mode, defer_m=8, device=0)]
    for t in range(w.T):
        n0 = len(fs[0].chosen)
        for f in fs:
            f.predict(w.u[t])
        xa, xb = fs[0].x, fs[1].x
        if not np.array_equal(xa, xb):
            print(mode, 'x differs after predict', t, np.abs(xa - xb).max()); break
        for f in fs:
            f.measure(None, w.u[t], w.stub(t))
        xa, xb = fs[0].x, fs[1].x
        Pa, Pb = fs[0].P, fs[1].P
        if not np.array_equal(xa, xb) or not np.array_equal(Pa, Pb):
            d = np.abs(Pa - Pb)
            print(mode, 'differs after measure', t, 'actions', fs[0].chosen[n0:], 'max dx', np.abs(xa - xb).max(), 'max dP', d.max(),
                  'where', np.argwhere(d > 0)[:12].tolist(), 'n', xa.size, 'slabs', [fs[0].shard_info()])
            break
    else:
        print(mode, 'identical over', w.T, 'steps')
    for f in fs:
        f.close()
