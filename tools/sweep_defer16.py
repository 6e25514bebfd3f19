"""K3 tile-shape sweep for the rank-32 flush of the cross-step deferral (defer_m = 16) at 10k landmarks, one process:
    python tools/sweep_defer16.py [variant ...]        (EKF_DOWNDATE_VARIANT values; '' = default shape)
Prints the average K3 launch time (CUDA events inside the library) and the step time per variant."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as E
from ekf_slam_b200 import _lib as L
from bench import make_inputs

N, k, W, K = 10000, 8, 4, 40
storage = os.environ.get('SWEEP_STORAGE', 'sym')
x, A, d, s, lms, us, obs = make_inputs(N, k, W + K)
variants = sys.argv[1:] or ['', '9', '11', '14', '16', '17', '18', '20']
for v in variants:
    if v:
        os.environ['EKF_DOWNDATE_VARIANT'] = v
    else:
        os.environ.pop('EKF_DOWNDATE_VARIANT', None)
    f = E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=16)
    f.set_state_lowrank(x, s, A, d)
    lib, h, Rc = f._lib, f._h, f.Rc
    zs = [np.asfortranarray(o[:, :2]) for o in obs]
    Rds = [np.asfortranarray(np.column_stack([o[:, 0] * Rc[0], o[:, 1] * Rc[1]])) for o in obs]
    ids = [np.ascontiguousarray(o[:, 2].astype(np.int32)) for o in obs]

    def step(t):
        L.check(h, lib.ekf_predict(h, L.dptr(us[t])))
        L.check(h, lib.ekf_update_batch(h, k, L.dptr(zs[t]), L.dptr(Rds[t]), L.iptr(ids[t])))

    for t in range(W):
        step(t)
    f.flush(); f.sync()
    f.profile_enable(True)
    t0 = time.perf_counter()
    for t in range(W, W + K):
        step(t)
    f.flush(); f.sync()
    wall = (time.perf_counter() - t0) * 1e3
    ms, nd, nl = f.profile_read()
    print('variant %-3s  K3 avg %.4f ms over %d launches   step %.4f ms   %.0f updates/s' %
          (v or '-', ms / max(nd, 1), nd, wall / K, K * k / (wall * 1e-3)), flush=True)
    f.close()
