"""K3 kernel sweep at 10k landmarks, one process:
    python tools/sweep_flush.py [--defer 16] [--storage sym] [--steps 40] [--out gpurun_out/sweep.jsonl] [variant ...]
variant = EKF_DOWNDATE_VARIANT value ('-' = library default).  Per variant: average K3 launch time (CUDA events inside
the library), algorithmic GB/s, and the resident step time; one JSON line each."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as E
from ekf_slam_b200 import _lib as L
from bench import make_inputs

ap = argparse.ArgumentParser()
ap.add_argument('--defer', type=int, default=16)
ap.add_argument('--storage', default='sym')
ap.add_argument('--landmarks', type=int, default=10000)
ap.add_argument('--steps', type=int, default=40)
ap.add_argument('--out', default=None)
ap.add_argument('variants', nargs='*')
a = ap.parse_args()
N, k, W, K = a.landmarks, 8, 4, a.steps
x, A, d, s, lms, us, obs = make_inputs(N, k, W + K)
peak = 6538.9
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    pass
out = open(a.out, 'a') if a.out else None
for v in a.variants or ['-']:
    if v != '-':
        os.environ['EKF_DOWNDATE_VARIANT'] = v
    else:
        os.environ.pop('EKF_DOWNDATE_VARIANT', None)
    f = E.EKF_SLAM(capacity_landmarks=N, storage=a.storage, defer_m=a.defer)
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
    b = f.downdate_bytes()
    assert np.all(np.isfinite(f.x[:3]))
    rec = {'variant': v, 'defer': a.defer, 'storage': a.storage, 'landmarks': N, 'k3_ms': ms / max(nd, 1), 'k3_launches': nd,
           'gbs': b / (ms / max(nd, 1) * 1e-3) / 1e9, 'frac': b / (ms / max(nd, 1) * 1e-3) / 1e9 / peak,
           'step_ms': wall / K, 'updates_per_s': K * k / (wall * 1e-3)}
    print(json.dumps(rec), flush=True)
    if out:
        out.write(json.dumps(rec) + '\n'); out.flush()
    f.close()
