"""Debug aid: continuous flush (EKF_SHADOW=1) against the in-place cross-step deferral (EKF_SHADOW=0) at a given size,
step by step (x after every step, blocks of P at the end), through update_batch or through the class-API measure()."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from types import SimpleNamespace


def rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300))


def run(N, T, mode, flag, sync_each):
    os.environ['EKF_SHADOW'] = flag
    import ekf_slam_b200 as E
    from bench import make_inputs
    from tests.golden_util import Stub
    k = 8
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=16)
    g.set_state_lowrank(x, s, A, d)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    xs = []
    for t in range(T):
        g.predict(us[t])
        if mode == 'measure':
            g.measure(None, us[t], Stub(obs[t], table))
        else:
            o = np.asarray(obs[t])
            Rd = np.column_stack([o[:, 0] * 0.1, o[:, 1] * 5.0])
            g.update_batch(o[:, :2], Rd, o[:, 2].astype(np.int32))
        if sync_each:
            g._ck(g._lib.ekf_sync(g._h))
        xs.append(g.x)
    n = 3 + 2 * N
    blocks = {'top': g.P_block(0, 0, 8, n), 'diag': g.P_block(n - 300, n - 300, 300, 300), 'mid': g.P_block(5000 % n, 7000 % n, 200, 200),
              'corner': g.P_block(0, 0, 8, 8)}
    g.close()
    return xs, blocks


if __name__ == '__main__':
    N = int(sys.argv[1]); T = int(sys.argv[2])
    modes = sys.argv[3].split(',') if len(sys.argv) > 3 else ('batch', 'measure')
    for mode in modes:
        for sync_each in ((0,) if len(sys.argv) > 3 else (0, 1)):
            a = run(N, T, mode, '0', sync_each)
            b = run(N, T, mode, '1', sync_each)
            ex = [rel(p, q) for p, q in zip(b[0], a[0])]
            print('N=%d mode=%s sync_each=%d  x rel per step: %s' % (N, mode, sync_each, ' '.join('%.1e' % e for e in ex)))
            for key in a[1]:
                dd = np.abs(b[1][key] - a[1][key])
                print('    P %-6s rel %.2e  max abs %.2e at %s' % (key, rel(b[1][key], a[1][key]), dd.max(), np.unravel_index(dd.argmax(), dd.shape)))
