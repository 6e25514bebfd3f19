"""Where the end-to-end step of bench.py spends its wall time: per-call host timings of the class-API loop
(predict / measure incl. its read of x / read x) at N landmarks, continuous flush on or off."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from types import SimpleNamespace


def run(N, T, k=8, mode='measure'):
    import ekf_slam_b200 as E
    from bench import make_inputs
    from ekf_slam_b200.synth import LandmarkStub
    x, A, d, s, lms, us, obs = make_inputs(N, k, T + 10)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=16)
    g.set_state_lowrank(x, s, A, d)
    for t in range(10):
        g.predict(us[t]); g.measure(None, us[t], LandmarkStub(obs[t], table)); g.x
    g.flush(); g.sync()
    g.profile_enable(True)
    acc = np.zeros(5)
    raw = []
    t00 = time.perf_counter()
    for t in range(10, 10 + T):
        t0 = time.perf_counter()
        g.predict(us[t])
        t1 = time.perf_counter()
        stub = LandmarkStub(obs[t], table)
        if mode in ('measure', 'batch'):
            xx = g.x                               # what measure() does first (getLandmark(laserData, h.x))
        t2 = time.perf_counter()
        if mode == 'measure':
            g.measure(None, us[t], stub, pass_x=False)
        else:
            o = np.asarray(obs[t])
            g.update_batch(o[:, :2], np.column_stack([o[:, 0] * 0.1, o[:, 1] * 5.0]), o[:, 2].astype(np.int32))
        t3 = time.perf_counter()
        if mode != 'batch_nox2':
            xx = g.x
        t4 = time.perf_counter()
        acc += [t1 - t0, t2 - t1, t3 - t2, t4 - t3, t4 - t0]
        raw.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
    g.flush(); g.sync()
    tot = time.perf_counter() - t00
    dd_ms, dd_n, _ = g.profile_read()
    print('   K3: %d passes, avg %.4f ms; loop %.4f ms per step' % (dd_n, dd_ms / max(dd_n, 1), tot / T * 1e3))
    g.close()
    r = np.array(raw) * 1e6
    print('   last 8 steps (predict, x1, measure, x2) us:', ' | '.join(' '.join('%.0f' % v for v in row) for row in r[-8:]))
    print('   even/odd step means of x2: %.0f / %.0f   of measure: %.0f / %.0f' % (r[0::2, 3].mean(), r[1::2, 3].mean(), r[0::2, 2].mean(), r[1::2, 2].mean()))
    return acc / T * 1e6, tot / T * 1e6


if __name__ == '__main__':
    N = int(sys.argv[1]); T = int(sys.argv[2])
    mode = sys.argv[3] if len(sys.argv) > 3 else 'measure'
    for flag in (('1',) if len(sys.argv) > 3 else ('1', '0')):
        os.environ['EKF_SHADOW'] = flag
        a, tot = run(N, T, mode=mode)
        print('N=%d EKF_SHADOW=%s  us per step: predict %.1f | x read #1 %.1f | measure %.1f | x read #2 %.1f | sum %.1f | loop %.1f'
              % (N, flag, a[0], a[1], a[2], a[3], a[4], tot))
