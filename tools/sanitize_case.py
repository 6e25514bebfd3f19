"""Small end-to-end case for compute-sanitizer (memcheck / racecheck / synccheck): both filter classes, both storages,
deferred TMA flush, batched association, decoupled state chain, Monte-Carlo batch with early pose read-back.
    compute-sanitizer --tool memcheck python tools/sanitize_case.py
(compute-sanitizer is closed on the pool this round was measured on -- it answers "closed on this pool" -- so only the
plain run of this script was made there: it passes.)
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ekf_slam_b200 as E
from ekf_slam_b200.synth import make_world

T = int(os.environ.get('SAN_STEPS', '12'))
for cls, mode in ((E.EKF_SLAM, 'prefix'), (E.EKF_SLAM_UC, 'nearest')):
    for storage, defer in (('sym', 8), ('full', 0), ('sym', 2)):
        w = make_world(N=12, T=T, k=4, mode=mode)
        g = cls(capacity_landmarks=16, storage=storage, defer_m=defer)
        for t in range(w.T):
            g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t)); x = g.x
        P = g.P; el = g.covariance_ellipses()
        assert np.all(np.isfinite(x)) and np.all(np.isfinite(P)) and np.all(np.isfinite(el))
        g.close()
        print('ok', cls.__name__, storage, defer, flush=True)
# a larger map so that K3 runs full 128-row tiles (N = 200 -> n = 403) from a synthetic SPD state
from bench import make_inputs
N, k = 200, 8
x, A, d, s, lms, us, obs = make_inputs(N, k, 3)
for storage in ('sym', 'full'):
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, defer_m=8)
    g.set_state_lowrank(x, s, A, d)
    for t in range(3):
        g.predict(us[t])
        r = obs[t]
        g.update_batch(r[:, :2], np.column_stack([r[:, 0] * g.Rc[0], r[:, 1] * g.Rc[1]]), r[:, 2].astype(np.int32))
        assert np.all(np.isfinite(g.x))
    assert np.all(np.isfinite(g.P_block(0, 0, 1, 3 + 2 * N)))
    g.close()
    print('ok 200 landmarks', storage, flush=True)
nf = 6
ws = [make_world(N=10, T=T, k=4, mode='nearest', seed=100 + f) for f in range(nf)]
for storage in ('sym', 'full'):
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=10, max_obs=4, max_table=10, storage=storage, defer_m=4)
    for t in range(T):
        b.step(np.array([w.u[t] for w in ws]), [w.observed[t] for w in ws], [list(w.table[:w.n_table[t]]) for w in ws])
        p = b.poses()
    b.sync()
    assert np.all(np.isfinite(p)) and np.all(np.isfinite(b[0].P))
    b.close()
    print('ok batch', storage, flush=True)
print('SANITIZE_CASE_OK')
