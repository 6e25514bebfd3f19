#!/usr/bin/env python
"""Generate the golden vectors in this directory.  Run from the repo root:  python tests/golden/make_golden.py

The reference (SamShue/EKF_SLAM, pure MATLAB) ships no golden vectors, tests or fixtures and cannot be executed in
this image (no MATLAB / Octave), so these vectors are produced by the dense-literal oracle (oracle/dense_literal.py,
a line-by-line restatement of EKF_SLAM.m / EKF_SLAM_UC.m / Correspondence.m) on seeded synthetic worlds
(ekf_slam_b200/synth.py).  They pin the oracle against regressions and give the GPU tests fixed
byte-identical inputs/outputs; they do NOT pin the oracle to the reference -- make_reference_run.py (refrun_*.npz: the
reference's unmodified source executed by oracle/mini_matlab.py) does that for its statements, DESIGN.md section 6.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from ekf_slam_b200.synth import make_world            # noqa: E402
from oracle import dense_literal as O                 # noqa: E402

CASES = {
    # name: (oracle class, world kwargs)
    'known_prefix_N20': ('EKF_SLAM', dict(N=20, T=300, k=5, mode='prefix', seed=20261104)),
    'known_quirk_N20': ('EKF_SLAM', dict(N=20, T=50, k=5, mode='nearest', seed=20261104)),      # idx = ii quirk (F8)
    'uc_nearest_N20': ('EKF_SLAM_UC', dict(N=20, T=300, k=5, mode='nearest', seed=20261104)),
    'uc_nearest_N8_seed7': ('EKF_SLAM_UC', dict(N=8, T=120, k=3, mode='nearest', seed=7)),
}


def pack_world(w):
    """Flatten a World into plain arrays (ragged observation lists -> rows + per-step counts)."""
    counts = np.array([len(o) for o in w.observed], dtype=np.int32)
    rows = np.concatenate([np.asarray(o, dtype=np.float64).reshape(-1, 3) for o in w.observed] + [np.zeros((0, 3))])
    tab_idx = np.array([e.index for e in w.table], dtype=np.int32)
    tab_loc = np.array([e.loc for e in w.table], dtype=np.float64).reshape(-1, 2)
    return dict(u=np.asarray(w.u), obs_rows=rows, obs_counts=counts, table_index=tab_idx, table_loc=tab_loc,
                n_table=np.asarray(w.n_table, dtype=np.int32))


def run_case(cls_name, kw, every=50):
    w = make_world(**kw)
    o = getattr(O, cls_name)()
    xs, steps = [], []
    for t in range(w.T):
        o.predict(w.u[t])
        o.measure(None, w.u[t], w.stub(t))
        if (t + 1) % every == 0 or t == w.T - 1:
            steps.append(t + 1)
            xs.append(np.pad(o.x, (0, 3 + 2 * kw['N'] - o.x.size), constant_values=np.nan))
    act = np.array([(1 if a == 'update' else -1) * i for a, i in o.chosen], dtype=np.int32)
    out = pack_world(w)
    out.update(x_final=o.x, P_final=o.P, s_final=np.asarray(o.s, dtype=np.float64), actions=act,
               x_checkpoints=np.array(xs), checkpoint_steps=np.array(steps, dtype=np.int32))
    return out


def association_cases(seed=5):
    """Mode A (Correspondence.m:75 as shipped) and mode B (:74) decisions on one fixed state, incl. ties and gating."""
    rng = np.random.default_rng(seed)
    N = 12
    n = 3 + 2 * N
    A = rng.normal(size=(n, n))
    P = A @ A.T / n + np.diag(rng.uniform(0.01, 0.1, n))
    x = np.concatenate([[0.3, -0.2, 40.0], rng.uniform(-8, 8, 2 * N)])
    s = np.arange(1, N + 1, dtype=np.float64)
    s[4] = 3.0                       # duplicate signature: landmark 5 ties with landmark 3 -> first min wins (:81)
    zs, Rs, outA, outB = [], [], [], []
    for q in range(40):
        kk = int(rng.integers(0, N))
        d = x[3 + 2 * kk:5 + 2 * kk] - x[:2]
        z = np.array([np.hypot(*d) + rng.normal(0, 0.05),
                      (np.degrees(np.arctan2(d[1], d[0])) - x[2] + rng.normal(0, 0.5)) % 360.0,
                      float(kk + 1) if q % 5 else float(N + 3)])     # every 5th: unseen signature -> newLL
        R = np.diag([z[0] * 0.1, z[1] * 5.0])
        a = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC').estimateCorrespondence(z, R, x, P, s)
        sB = np.ones(N)              # all signatures equal -> the Mahalanobis term decides
        zB = z.copy(); zB[2] = 1.0
        b = O.Correspondence(1e-11, 9.21, 'EKF_SLAM_UC', position_cost_in_score=True).estimateCorrespondence(zB, R, x, P, sB)
        zs.append(z); Rs.append(np.diag(R)); outA.append([int(a[0]), a[1]]); outB.append([int(b[0]), b[1]])
    return dict(x=x, P=P, s=s, z=np.array(zs), Rdiag=np.array(Rs), modeA=np.array(outA, dtype=np.int32),
                modeB=np.array(outB, dtype=np.int32))


def main():
    for name, (cls_name, kw) in CASES.items():
        np.savez_compressed(os.path.join(HERE, name + '.npz'), **run_case(cls_name, kw))
        print('wrote', name)
    np.savez_compressed(os.path.join(HERE, 'association_N12.npz'), **association_cases())
    print('wrote association_N12')


if __name__ == '__main__':
    main()
