#!/usr/bin/env python
"""bench.py -- EKF updates/sec at 10k landmarks on B200 (BASELINE.json metric), with the P-downdate roofline.

A "step" is one pass of the hot path over one batch of synthetic input: predict(u) followed by the
measurement update of k=8 range-bearing observations on a fully populated 10 000-landmark map
(P 20003^2 fp64).  `value` = updates/s with all inputs resident (direct C-ABI calls, scalars only);
`e2e` = the same through the reference-facing class API (EKF_SLAM_UC.predict / .measure with a
landmark_list stub: association + update per observation, host buffers, x read back twice per step), wall clock
between two device-wide synchronisations.  The reads of x return behind the gains of the step, not behind its flush
(DESIGN.md "Decoupled state chain"), so that loop keeps the device busy; EKF_ASYNC_X=0 shows the difference.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'EKF updates/sec at 10k landmarks'
UNIT = 'updates/s'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='10k', choices=['10k', 'mc'],
                    help="10k: one 10 000-landmark filter (BASELINE metric, default); mc: 4096 independent EKF_SLAM_UC filters x 200 landmarks")
    ap.add_argument('--filters', type=int, default=4096, help='mc workload: total filters over all GPUs')
    ap.add_argument('--landmarks', type=int, default=None)
    ap.add_argument('--k', type=int, default=8)
    ap.add_argument('--storage', default='sym', choices=['sym', 'full'])
    ap.add_argument('--defer', type=int, default=16,
                    help='deferred downdate: one pass over P per this many observations (1 = every observation; <= 8: within one '
                         'measure, bit-identical to the immediate path; 9..16: terms also stay pending across predict, single GPU)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-streaming-variant', action='store_true', help='skip the secondary defer_m = 8 line')
    ap.add_argument('--sustained', type=int, default=200, help='extra resident steps timed after the main region (0 = off)')
    ap.add_argument('--parity-steps', type=int, default=6, help='steps of the parity block (class API vs the matrix-free oracle; 0 = off)')
    ap.add_argument('--also-80k', type=int, default=1, help='N >= 2 only: add the 80 000-landmark block (BASELINE config 4) to the line')
    ap.add_argument('--steps-80k', type=int, default=20)
    ap.add_argument('--kernel-name', default='k_downdate_tma/k_downdate_hyb', help=argparse.SUPPRESS)
    ap.add_argument('--cpu-budget', type=float, default=12.0, help='seconds of CPU work for the cpu_baseline sample')
    return ap.parse_args()


def workload_config(N, k):
    """`config` of the JSON line: names the WORKLOAD only, so that the repo arm and the `--impl reference` arm print the
    same object (how each arm executes it is in `implementation` / `cpu_baseline`)."""
    n = 3 + 2 * N
    return {'workload': ('10k landmarks (P 20003^2 fp64), predict + %d updates per step' % k) if N == 10000 else
                        '%d landmarks, predict + %d updates per step' % (N, k),
            'landmarks': N, 'k': k, 'n_state': n,
            'l2': 'no flush needed: every update touches all of P (%.2f GB fp64 as a full square), far beyond the 126 MB L2 '
                  'and any CPU cache' % (8.0 * n * n / 1e9)}


def use_all_host_threads():
    """The CPU arms use every host core whatever the launcher left in the environment (torch.distributed.run exports
    OMP_NUM_THREADS=1, which made the round-1 N >= 2 reference numbers 25x slower than the N = 1 one)."""
    cores = os.cpu_count() or 1
    try:
        import threadpoolctl
        threadpoolctl.threadpool_limits(limits=cores)
        blas = sorted({(d.get('internal_api'), d.get('num_threads')) for d in threadpoolctl.threadpool_info()})
    except Exception as e:                                   # pragma: no cover
        blas = [('threadpoolctl unavailable: %s' % e, None)]
    return cores, blas


def structured_port(N, k, x, P, us, obs):
    """Like-for-like CPU number: the same O(n^2) structured algebra the GPU runs, in C on the host threads
    (oracle/structured.c, pthreads): 1 predict + k updates in full.  NOT the reference's algorithm (that is the dense
    port); reported beside it in both arms."""
    try:
        from oracle import structured as S
        so = S.EKF_SLAM_UC(capacity_landmarks=N)
        so.set_state(x, P, np.arange(1, N + 1, dtype=np.float64))
        r = obs[0][0]
        so.predict(us[0]); so.update(r[:2], [r[0] * 0.1, r[1] * 5.0], int(r[2]))            # warm the pages
        t0 = time.perf_counter()
        so.predict(us[1 % len(us)])
        for q in range(k):
            rq = obs[1 % len(obs)][q]
            so.update(rq[:2], [rq[0] * 0.1, rq[1] * 5.0], int(rq[2]))
        dt = time.perf_counter() - t0
        return {'value': k / dt, 'unit': UNIT, 'cores': min(16, os.cpu_count() or 1), 'measured_wall_s': dt,
                'what': 'oracle/structured.c (5-column O(n^2) form, pthreads): 1 predict + %d updates in full' % k}
    except Exception as e:                      # test infrastructure missing a compiler must not break the bench line
        return {'unavailable': str(e)[:120]}


def make_inputs(N, k, steps, seed=20261104):
    """Populated synthetic map + per-step odometry and k observations (host, generated once)."""
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    x, A, d, s, lms = synthetic_spd_state(N, seed=seed)
    rng = np.random.default_rng(seed + 1)
    L = 10.0 * np.sqrt(N / 20.0)
    Rr = 0.35 * L
    dD = 2.0 * np.pi * Rr / 360.0
    pose = np.zeros(3)
    us, obs = [], []
    for t in range(steps):
        th = pose[2] + 1.0
        pose = np.array([pose[0] + dD * np.cos(np.radians(th)), pose[1] + dD * np.sin(np.radians(th)), th])
        us.append(np.array([dD + rng.normal(0, 0.01 * dD), 1.0 + rng.normal(0, 0.05)]))
        dist = np.hypot(lms[:, 0] - pose[0], lms[:, 1] - pose[1])
        cand = np.argsort(dist)[:4 * k]
        rows = observe_from_truth(lms, pose, cand, rng)[:k]
        rows = rows[np.argsort(rows[:, 2])]
        assert rows.shape[0] == k, 'not enough observations outside the seam zone'
        obs.append(rows)
    return x, A, d, s, lms, us, obs


class ClockSampler:
    """SM clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe).  NVML in a thread at
    ~2 ms (a 20-step timed region at 10k is only ~12 ms long); falls back to `nvidia-smi -lms 100`."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index
        self.nv, self.stop_flag, self.samples = None, False, []

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            if not uuid.startswith('GPU-'):
                uuid = 'GPU-' + uuid
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def start(self):
        try:
            self.nv, self.h = self._nvml_handle()
            self.max_mhz = float(self.nv.nvmlDeviceGetMaxClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                self.samples.append((mhz, rs, time.perf_counter()))
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self, window=None):
        """window = (t0, t1) in time.perf_counter() seconds: keep only the samples taken inside the timed region."""
        if self.nv is not None:
            self.stop_flag = True
            self.th.join(timeout=1.0)
            if window is not None:
                inside = [q for q in self.samples if window[0] <= q[2] <= window[1]]
                self.samples = inside if inside else self.samples[-1:]
            nv = self.nv
            masks = {'hw_slowdown': getattr(nv, 'nvmlClocksEventReasonHwSlowdown', 0x8),
                     'hw_thermal_slowdown': getattr(nv, 'nvmlClocksEventReasonHwThermalSlowdown', 0x40),
                     'sw_thermal_slowdown': getattr(nv, 'nvmlClocksEventReasonSwThermalSlowdown', 0x20),
                     'sw_power_cap': getattr(nv, 'nvmlClocksEventReasonSwPowerCap', 0x4)}
            sm = [q[0] for q in self.samples]
            reasons = sorted(nm for nm, bit in masks.items() if any(q[1] & bit for q in self.samples))
            return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': self.max_mhz, 'samples': len(sm),
                    'reasons': reasons, 'source': 'nvml @2ms'}
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith('active'):
                        reasons.add(nm)
            except Exception:
                pass
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons), 'source': 'nvidia-smi -lms 100'}


def cpu_baseline(N, k, x, A, d, us, obs, budget_s):
    """Reference's dense algorithm (oracle port, NumPy/OpenBLAS, all host threads) on a bounded sample."""
    from oracle import cpu_baseline as cb
    from ekf_slam_b200.synth import lowrank_to_dense
    cores, blas = use_all_host_threads()
    n = 3 + 2 * N
    rate = cb.gemm_rate()
    P = lowrank_to_dense(A, d)
    m = cb.pick_rows(n, budget_s / 3.0, rate)
    t0 = time.perf_counter()
    t_pred = cb.sampled_predict_seconds(x, P, us[0], 0.2, m)
    r = obs[0][0]
    t_upd, _, _ = cb.sampled_update_seconds(x, P, r[:2], [r[0] * 0.1, r[1] * 5.0], int(r[2]), m)
    wall = time.perf_counter() - t0
    t_step = t_pred + k * t_upd
    sp = structured_port(N, k, x, P, us, obs)
    return {'value': k / t_step, 'unit': UNIT, 'cores': cores, 'blas_threads': blas, 'kind': 'port', 'extrapolated': m < n,
            'measured_wall_s': wall, 'structured_port': sp,
            'sample': 'dense reference algebra (EKF_SLAM.m:47,141-145) at n=%d via NumPy/OpenBLAS: 1 predict + 1 update, '
                      'n^3 products on the first %d of %d output rows scaled by n/m (O(n^2) parts in full); '
                      'step = predict + %d updates; gemm %.0f GFLOP/s' % (n, m, n, k, rate / 1e9),
            's_per_update': t_upd, 's_per_predict': t_pred}


def run_reference(args):
    """--impl reference: the reference's own CPU algorithm (oracle port; MATLAB/Octave are absent) on host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from oracle import cpu_baseline as cb
    from ekf_slam_b200.synth import lowrank_to_dense
    cores, blas = use_all_host_threads()
    N, k = args.landmarks, args.k
    n = 3 + 2 * N
    x, A, d, s, lms, us, obs = make_inputs(N, k, max(2, args.steps + args.warmup))
    P = lowrank_to_dense(A, d)
    rate = cb.gemm_rate()
    per_step_budget = min(6.0, 150.0 / max(1, args.steps + args.warmup))
    m = cb.pick_rows(n, per_step_budget / 3.0, rate)
    times = []
    w0 = time.perf_counter()
    for t in range(args.warmup + args.steps):
        r = obs[t][0]
        tp = cb.sampled_predict_seconds(x, P, us[t], 0.2, m)
        tu, _, _ = cb.sampled_update_seconds(x, P, r[:2], [r[0] * 0.1, r[1] * 5.0], int(r[2]), m)
        if t >= args.warmup:
            times.append(tp + k * tu)
    wall = time.perf_counter() - w0
    t_step = float(np.mean(times))
    val = k / t_step
    sample = ('each step: 1 dense predict + 1 dense update at n=%d, n^3 products on the first %d output rows scaled by '
              'n/m, update cost x%d observations; %d BLAS threads' % (n, m, k, cores))
    sp = structured_port(N, k, x, P, us, obs)
    line = {'impl': 'reference', 'metric': METRIC if N == 10000 else 'EKF updates/sec at %d landmarks' % N, 'value': val, 'unit': UNIT,
            'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': t_step * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(N, k),
            # what the value is: an n/m extrapolation of row-sampled dense products, not a region timed in full
            'extrapolated': m < n, 'measured_wall_s': wall, 'claimed_s': t_step * args.steps, 'cores': cores, 'blas_threads': blas,
            'cpu_baseline': {'value': val, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample, 'extrapolated': m < n,
                             'measured_wall_s': wall, 'structured_port': sp},
            'cpu_structured': sp,
            'e2e': {'value': val, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


def measure_filter(E, L, dist, world, rank, local, stream, barrier, N, k, K, W, storage, defer, clocks=None, want_e2e=True,
                   sustained_steps=0, parity_steps=6):
    """One large filter of N landmarks (single GPU, or ONE filter in `world` column slabs): resident arm (direct C-ABI calls,
    inputs resident), end-to-end arm (class API, host buffers, x read back), parity block (class API vs the matrix-free
    oracle).  Returns the raw measurements; every time is CUDA-event or wall time on this rank (reduced by the caller)."""
    import torch
    from ekf_slam_b200.synth import LandmarkStub
    from types import SimpleNamespace
    n = 3 + 2 * N
    x, A, d, s, lms, us, obs = make_inputs(N, k, 2 * (K + W) + sustained_steps)

    def make(cls):
        if world > 1:      # ONE filter sharded over all ranks (column slabs of P, peer-memory column exchange per batch)
            return E.shard.create_sharded(cls, dist, local, storage=storage, capacity_landmarks=N, stream=stream, defer_m=defer)
        return cls(capacity_landmarks=N, storage=storage, device=local, stream=stream, defer_m=defer)

    # ---------------- device-resident arm: direct C-ABI, known correspondence (idx given) ----------------
    f = make(E.EKF_SLAM)
    f.set_state_lowrank(x, s, A, d)
    lib, h = f._lib, f._h
    Rc = f.Rc
    zs = [np.asfortranarray(o[:, :2]) for o in obs]
    Rds = [np.asfortranarray(np.column_stack([o[:, 0] * Rc[0], o[:, 1] * Rc[1]])) for o in obs]
    ids = [np.ascontiguousarray(o[:, 2].astype(np.int32)) for o in obs]

    def step_resident(t):
        L.check(h, lib.ekf_predict(h, L.dptr(us[t])))
        L.check(h, lib.ekf_update_batch(h, k, L.dptr(zs[t]), L.dptr(Rds[t]), L.iptr(ids[t])))

    for t in range(W):
        step_resident(t)
    f.flush()
    f.profile_enable(True)
    if clocks is not None:
        clocks.start()                  # before the barrier: NVML start-up must not make rank 0 a straggler inside the timed region
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    th0 = time.perf_counter()
    for t in range(W, W + K):
        step_resident(t)
    f.flush()                                            # cross-step deferral: the last pending terms belong to the timed region
    host_ms = (time.perf_counter() - th0) * 1e3          # host time to ENQUEUE the K steps (no sync inside)
    e1.record()
    barrier()
    th1 = time.perf_counter()
    out = {'ms': e0.elapsed_time(e1), 'host_ms': host_ms, 'window': (th0, th1)}
    out['clk'] = clocks.stop((th0, th1)) if clocks is not None else None
    dd_ms, dd_n, launches = f.profile_read()
    xch_ms, xch_n, xch_p2p = f.profile_exchange() if world > 1 else (0.0, 0, False)
    out.update(dd_ms=dd_ms, dd_n=dd_n, launches=launches, xch_ms=xch_ms, xch_n=xch_n, xch_p2p=xch_p2p,
               dd_bytes=f.downdate_bytes(), slab=f.shard_info(), pend_cap=f.pend_cap(), continuous=f.flush_info()[0])
    if sustained_steps > 0:             # same loop, longer: what the rate settles at once the power cap has had time to act
        f.profile_enable(True)
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for t in range(W + K, W + K + sustained_steps):
            step_resident(t)
        f.flush()
        s1.record()
        barrier()
        sd_ms, sd_n, _ = f.profile_read()
        out['sustained'] = {'steps': sustained_steps, 'ms': s0.elapsed_time(s1), 'k3_avg_launch_ms': sd_ms / max(sd_n, 1)}
    f.profile_enable(False)
    assert np.all(np.isfinite(f.x[:3])), 'filter diverged'
    f.close()

    # ---------------- end-to-end arm: reference-facing class API, host buffers ----------------
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    if want_e2e:
        g = make(E.EKF_SLAM_UC)
        g.set_state_lowrank(x, s, A, d)

        def step_e2e(t):
            g.predict(us[t])
            g.measure(None, us[t], LandmarkStub(obs[t], table))
            return g.x

        for t in range(W):
            step_e2e(t)
        g.flush()
        barrier()
        t0 = time.perf_counter()
        for t in range(W, W + K):
            step_e2e(t)
        g.flush()
        barrier()
        out['e2e_s'] = time.perf_counter() - t0
        assert all(a[0] == 'update' for a in g.chosen), 'e2e arm appended unexpectedly'
        g.close()

    # ---------------- parity block: the same class-API loop on a fresh filter vs the matrix-free oracle ----------------
    # (oracle/lowrank_lazy.py: every row of the current P in O(n) per accumulated outer product, so it also covers the
    # 80 000-landmark slabs).  Every rank steps the oracle on identical inputs and checks x and sampled rows of ITS slab.
    if parity_steps > 0:
        from oracle.lowrank_lazy import LazyLowRankFilter
        g = make(E.EKF_SLAM_UC)
        g.set_state_lowrank(x, s, A, d)
        o = LazyLowRankFilter(x, A, d, s, Rc=g.Rc)
        want = []
        for t in range(parity_steps):
            g.predict(us[t]); g.measure(None, us[t], LandmarkStub(obs[t], table))
            o.predict(us[t]); want += o.measure_rows(obs[t])
        _, _, c_lo, c_hi = g.shard_info()
        c_hi = min(c_hi, n)
        ex = float(np.linalg.norm(g.x - o.x) / np.linalg.norm(o.x))
        rng = np.random.default_rng(5 + rank)
        rows = sorted(set([0, 2, 3 + 2 * (int(obs[0][0][2]) - 1), c_lo, max(c_lo, c_hi - 1)] + [int(v) for v in rng.integers(0, c_hi, 8)]))
        worst, checked = 0.0, 0
        for i in rows:
            j0 = max(c_lo, i) if storage == 'sym' else c_lo      # SYM slab: rows <= column are live
            if j0 >= c_hi:
                continue
            got = g.P_block(i, j0, 1, c_hi - j0)
            ref = o.block([i], j0, c_hi)
            worst = max(worst, float(np.linalg.norm(got - ref) / np.linalg.norm(ref)))
            checked += 1
        out['parity'] = {'steps': parity_steps, 'updates': len(want), 'actions_identical': bool(g.chosen == want), 'max_rel_x': ex,
                         'max_rel_P': worst, 'rows_checked': checked}
        g.close()
    out['inputs'] = (x, A, d, us, obs)
    return out


def reduce_measurements(torch, dist, world, m, K, k):
    """MAX over ranks of the timed quantities; the roofline is that of the slowest rank's slab; parity = worst rank."""
    vals = [m['ms'], m.get('e2e_s', 0.0) * 1e3, m['dd_ms'] / max(m['dd_n'], 1), m.get('sustained', {}).get('ms', 0.0)]
    tt = torch.tensor(vals, dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    m['ms'], m['e2e_ms'], m['dd_avg_ms'] = float(tt[0]), float(tt[1]), float(tt[2])
    if 'sustained' in m:
        m['sustained']['ms'] = float(tt[3])
    if world > 1:
        allv = [None] * world
        dist.all_gather_object(allv, (m['dd_ms'] / max(m['dd_n'], 1), m['dd_bytes'], m['slab'], m.get('parity'), m['xch_ms'] / max(m['xch_n'], 1)))
        slow = max(allv, key=lambda v: v[0])
        m['dd_one_ms'], m['dd_bytes'], m['slab'] = slow[0], slow[1], slow[2]
        m['per_rank'] = [{'k3_avg_launch_ms': v[0], 'k3_bytes': v[1], 'cols': [v[2][2], v[2][3]], 'exchange_avg_ms': v[4]} for v in allv]
        ps = [v[3] for v in allv if v[3]]
        if ps:
            m['parity'] = {'steps': ps[0]['steps'], 'updates': ps[0]['updates'],
                           'actions_identical': all(p['actions_identical'] for p in ps),
                           'max_rel_x': max(p['max_rel_x'] for p in ps), 'max_rel_P': max(p['max_rel_P'] for p in ps),
                           'rows_checked': sum(p['rows_checked'] for p in ps), 'ranks': len(ps)}
    else:
        m['dd_one_ms'] = m['dd_ms'] / max(m['dd_n'], 1)
    return m


# ncu --set full captures of the shipped K3 kernels (profiles/): dram__bytes_read.sum + dram__bytes_write.sum per launch.
# Keyed by (landmarks, storage, world, terms per launch); anything else prints null rather than a number from another kernel.
NCU_TRAFFIC = {
    (10000, 'sym', 1, 16): (1.700704e9 + 1.564460e9, 'profiles/ncu_downdate_r02_shadow_rank32_raw.csv: k_downdate_hyb<8,8,4,1,3> out of place (P -> P2, the continuous flush), ncu --set full of the shipped kernel (505.9 us under ncu, where kernels are serialised)'),
    (10000, 'sym', 1, 8): (1.629e9 + 1.551e9, 'profiles/ncu_downdate_r01_final_rank16_raw.csv: k_downdate_tma<4,128,8,4,1> (kernel unchanged since round 1)'),
}


def main():
    args = parse()
    if args.impl == 'ours':
        args.warmup = max(args.warmup, 3)       # timing rule: at least 3 untimed warm-up steps (the line reports the count used)
    if args.landmarks is None:
        args.landmarks = 200 if args.workload == 'mc' else 10000
    if args.workload == 'mc':
        if args.k == 8:
            args.k = 5                      # BASELINE config 5: k = 5 observations per step
        return run_mc_reference(args) if args.impl == 'reference' else run_mc(args)
    if args.impl == 'reference':
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import ekf_slam_b200 as E
    from ekf_slam_b200 import _lib as L

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    N, k, K, W = args.landmarks, args.k, args.steps, args.warmup
    n = 3 + 2 * N
    tstream = torch.cuda.Stream(priority=-1)   # non-default stream: torch events and the library share it; above the lowest priority so that the
                                              # state chain overlaps the library's K3 stream (continuous flush)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local) if rank == 0 else None
    m = measure_filter(E, L, dist, world, rank, local, stream, barrier, N, k, K, W, args.storage, args.defer, clocks=clocks,
                       sustained_steps=args.sustained, parity_steps=args.parity_steps)
    x, A, d, us, obs = m.pop('inputs')
    m = reduce_measurements(torch, dist, world, m, K, k)

    # ---------------- secondary line (single GPU, cross-step deferral only): the same workload with one pass over P per
    # batch of 8 (defer_m = 8), where K3 is a pure streaming kernel -- its HBM fraction is reported beside the main roofline
    streaming = None
    in_place = None
    shadow_on = bool(m.get('continuous'))            # what the handle reports (ekf_flush_info)
    if shadow_on and world == 1 and not args.no_streaming_variant:
        # the same workload with the K3 pass in place on the handle's stream (EKF_SHADOW=0): what the rank-32 kernel does when it
        # has the GPU to itself, beside the main roofline, which times it while the state chain runs next to it
        os.environ['EKF_SHADOW'] = '0'
        try:
            mi = measure_filter(E, L, dist, world, rank, local, stream, barrier, N, k, min(K, 100), W, args.storage, args.defer,
                                want_e2e=False, parity_steps=0)
        finally:
            os.environ.pop('EKF_SHADOW', None)
        mi.pop('inputs')
        Ki = min(K, 100)
        di = mi['dd_ms'] / max(mi['dd_n'], 1)
        in_place = {'steps': Ki, 'updates_per_s': Ki * k / (mi['ms'] * 1e-3), 'ms_per_step': mi['ms'] / Ki, 'k3_avg_launch_ms': di,
                    'k3_bytes_per_launch': mi['dd_bytes'], 'achieved': mi['dd_bytes'] / (di * 1e-3) / 1e9, 'unit': 'GB/s',
                    'what': 'EKF_SHADOW=0: the same rank-32 kernel in place on the one stream, nothing running beside it'}
    if world == 1 and args.defer > 8 and not args.no_streaming_variant:
        m8 = measure_filter(E, L, dist, world, rank, local, stream, barrier, N, k, min(K, 100), W, args.storage, 8,
                            want_e2e=False, parity_steps=0)
        m8.pop('inputs')
        K8 = min(K, 100)
        d8 = m8['dd_ms'] / max(m8['dd_n'], 1)
        streaming = {'defer_m': 8, 'steps': K8, 'updates_per_s': K8 * k / (m8['ms'] * 1e-3), 'ms_per_step': m8['ms'] / K8,
                     'k3_avg_launch_ms': d8, 'k3_bytes_per_launch': m8['dd_bytes'], 'achieved': m8['dd_bytes'] / (d8 * 1e-3) / 1e9,
                     'unit': 'GB/s', 'what': 'rank-16 flush, one pass over P per measure(): bit-identical to the immediate path'}

    # ---------------- secondary block at N >= 2: BASELINE config 4, the 80 000-landmark filter (P 205 GB dense, 102 GB live
    # in SYM slabs), the size the north star's multi-GPU claim is about
    big = None
    if world > 1 and N == 10000 and args.also_80k:
        try:
            K80 = args.steps_80k
            b = measure_filter(E, L, dist, world, rank, local, stream, barrier, 80000, k, K80, 3, 'sym', args.defer,
                               want_e2e=False, parity_steps=4)
            b.pop('inputs')
            b = reduce_measurements(torch, dist, world, b, K80, k)
            big = b
        except Exception as e:                      # must not cost the main line
            big = {'error': str(e)[:200]}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'MEASURED_PEAKS.json hbm_gbs (of measured)' if peaks else 'fallback 6650 (of fallback)'

    def roofline_of(mm, KK, NN, storage):
        per_pass = int(round(KK * k / max(mm['dd_n'], 1)))           # observations one K3 launch applied, as executed
        ach = mm['dd_bytes'] / (mm['dd_one_ms'] * 1e-3) / 1e9
        key = (NN, storage, world, per_pass)
        tr = NCU_TRAFFIC.get(key)
        return {'bound': 'hbm', 'kernel': '%s (rank-%d flush of %d observations in one pass over P)' % (args.kernel_name, 2 * per_pass, per_pass),
                'achieved': ach, 'peak': peak, 'unit': 'GB/s', 'frac': ach / peak, 'peak_source': peak_src,
                'frac_of_8TBps_nominal': ach / 8000.0,
                'traffic': tr[0] if tr else None, 'traffic_source': tr[1] if tr else None,
                'algorithmic_bytes_per_launch': mm['dd_bytes'], 'avg_launch_ms': mm['dd_one_ms'], 'launches': mm['dd_n'],
                'share_of_step': mm['dd_one_ms'] * mm['dd_n'] / mm['ms'], 'observations_per_launch': per_pass,
                # fp64 tensor-core side of the same launch: 2 flop x rank per live element; measured DMMA peak 37.2 TFLOP/s
                'fp64_tflops': (mm['dd_bytes'] / 16.0) * 4.0 * per_pass / (mm['dd_one_ms'] * 1e-3) / 1e12, 'fp64_dmma_peak_tflops': 37.2}

    roof = roofline_of(m, K, N, args.storage)
    per_pass = roof['observations_per_launch']
    line = {
        'metric': METRIC if N == 10000 else 'EKF updates/sec at %d landmarks' % N, 'value': K * k / (m['ms'] * 1e-3), 'unit': UNIT,
        'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': m['ms'] / K, 'higher_is_better': True,
        'scaling': 'strong' if world > 1 else 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(N, k),
        'implementation': {
            'storage': args.storage, 'defer_m': m['pend_cap'],
            'downdate': ('one rank-%d pass over P per %d observations%s' % (2 * per_pass, per_pass, ' (terms stay pending across predict: %d steps per pass)' % (per_pass // k) if per_pass > k else ''))
                        + ('; continuous flush: the pass runs out of place (P -> P2) on its own stream, back to back, while predict / gather / '
                           'linearise / gain of the next batches run beside it from P + the robot strip + the window of pending terms' if shadow_on else ''),
            'parallelism': ('one filter, P in %d column slabs (%s storage), per batch one exchange of the 3+2k current columns '
                            '(peer-memory scatter; EKF_SHARD_EXCHANGE=nccl: NCCL %s); slowest slab cols [%d,%d)'
                            % (world, args.storage, 'all-reduce' if args.storage == 'sym' else 'broadcast group', m['slab'][2], m['slab'][3])) if world > 1 else 'single GPU',
            'serial_per_step': ('nothing on the K3 stream (passes are queued one ahead); beside it: predict on the strip, k_gather_cols_strip, '
                                'k_batch_lin (k dependent linearisations), k_batch_gain, and per pass strip fix + window rotation') if shadow_on else
                               'predict (P0 rows/cols 0-1 + pending slabs), exchange (sharded), k_batch_lin (k dependent linearisations), k_batch_gain'},
        'roofline': roof,
        'exchange': ({'avg_ms': m['xch_ms'] / max(m['xch_n'], 1), 'per_step': m['xch_n'] / K,
                      'path': 'peer-memory scatter (cudaIpc, NVLink stores + flag words)' if m['xch_p2p'] else 'NCCL'} if world > 1 else None),
        'e2e': {'value': K * k / (m['e2e_ms'] * 1e-3), 'unit': UNIT,
                'h2d_bytes_per_step': int(16 + obs[0].nbytes + k * (24 + 16)), 'd2h_bytes_per_step': int(2 * 8 * n + k * 24),
                'api': 'EKF_SLAM_UC.predict(u); .measure(laserData,u,landmark_list); read x'},
        'parity': m.get('parity'),
        'gpu_launches': int(m['launches']), 'clocks': m['clk'], 'host_enqueue_ms_per_step': m['host_ms'] / K,
    }
    if world > 1:
        line['per_rank'] = m.get('per_rank')
    if 'sustained' in m:
        sm = m['sustained']
        line['sustained'] = {'steps': sm['steps'], 'value': sm['steps'] * k / (sm['ms'] * 1e-3), 'unit': UNIT,
                             'ms_per_step': sm['ms'] / sm['steps'], 'k3_avg_launch_ms': sm['k3_avg_launch_ms'],
                             'what': 'the same resident loop continued for %d more steps (what the rate settles at under the power cap)' % sm['steps']}
    if in_place is not None:
        in_place['frac'] = in_place['achieved'] / peak
        line['roofline_in_place_variant'] = in_place
        line['roofline']['measured'] = 'while the state chain of the next batches runs beside the pass (see roofline_in_place_variant for the kernel alone)'
    if streaming is not None:
        streaming['frac'] = streaming['achieved'] / peak
        line['roofline_streaming_variant'] = streaming
    if big is not None:
        if 'error' in big:
            line['scaling_80k'] = big
        else:
            K80 = args.steps_80k
            r80 = roofline_of(big, K80, 80000, 'sym')
            line['scaling_80k'] = {'workload': workload_config(80000, k), 'storage': 'sym', 'steps': K80, 'warmup': 3,
                                   'value': K80 * k / (big['ms'] * 1e-3), 'unit': UNIT, 'ms_per_step': big['ms'] / K80,
                                   'roofline': r80, 'exchange_avg_ms': big['xch_ms'] / max(big['xch_n'], 1),
                                   'per_rank': big.get('per_rank'), 'parity': big.get('parity')}
    if not args.no_cpu_baseline:
        line['cpu_baseline'] = cpu_baseline(N, k, x, A, d, us, obs, args.cpu_budget)
        line['cpu_structured'] = line['cpu_baseline'].get('structured_port')
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


# ======================================================================================================================
# Monte-Carlo workload (BASELINE config 5): --workload mc
# ======================================================================================================================
METRIC_MC = 'EKF updates/sec, 4096 independent EKF_SLAM_UC filters x 200 landmarks'


def make_mc_inputs(N, k, nf, steps, seed=20261104):
    """One populated N-landmark map shared by all filters; per step every filter observes the same k landmarks
    (those nearest the true pose, outside the seam zone) with its own measurement noise and its own odometry noise."""
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense
    x, A, d, s, lms = synthetic_spd_state(N, seed=seed)
    P = lowrank_to_dense(A, d)
    rng = np.random.default_rng(seed + 7)
    L = 10.0 * np.sqrt(N / 20.0)
    dD = 2.0 * np.pi * 0.35 * L / 360.0
    pose = np.zeros(3)
    U = np.zeros((steps, nf, 2))
    OBS = np.zeros((steps, nf, 3, k))
    for t in range(steps):
        th = pose[2] + 1.0
        pose = np.array([pose[0] + dD * np.cos(np.radians(th)), pose[1] + dD * np.sin(np.radians(th)), th])
        U[t, :, 0] = dD + rng.normal(0, 0.01 * dD, nf)
        U[t, :, 1] = 1.0 + rng.normal(0, 0.05, nf)
        dx, dy = lms[:, 0] - pose[0], lms[:, 1] - pose[1]
        brg = np.mod(np.degrees(np.arctan2(dy, dx)) - pose[2], 360.0)
        ok = np.where((brg > 25.0) & (brg < 335.0))[0]
        ids = np.sort(ok[np.argsort(np.hypot(dx, dy)[ok])[:k]])
        assert ids.size == k
        OBS[t, :, 0, :] = np.hypot(dx, dy)[ids][None, :] + rng.normal(0, 0.01, (nf, k))
        OBS[t, :, 1, :] = np.mod(brg[ids][None, :] + rng.normal(0, 0.1, (nf, k)), 360.0)
        OBS[t, :, 2, :] = (ids + 1)[None, :]
    loc = np.ascontiguousarray(lms.T.reshape(1, 2, N))
    idx = np.arange(1, N + 1, dtype=np.int32).reshape(1, N)
    return x, P, s, lms, U, OBS, loc, idx


def run_mc(args):
    import torch
    import torch.distributed as dist
    import ekf_slam_b200 as E
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    N, k, K, W = args.landmarks, args.k, args.steps, args.warmup
    mine = E.shard_filters(args.filters, rank, world)          # filter f lives on rank f mod world; no collective
    nf = len(mine)
    x, P, s, lms, U, OBS, loc, idx = make_mc_inputs(N, k, args.filters, 2 * (K + W))
    U, OBS = np.ascontiguousarray(U[:, mine]), np.ascontiguousarray(OBS[:, mine])
    counts = np.full(nf, k, dtype=np.int32)
    ntab = np.array([N], dtype=np.int32)
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def make():
        b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=N, max_obs=k, max_table=N, storage=args.storage,
                          defer_m=args.defer, device=local, stream=tstream.cuda_stream)
        b.set_state_all(x, P, s)
        return b

    # ---------------- device-resident arm: inputs already in HBM ----------------
    b = make()
    dU, dO = torch.from_numpy(U).cuda(), torch.from_numpy(OBS).cuda()
    dC, dL, dI, dT = (torch.from_numpy(a).cuda() for a in (counts, loc, idx, ntab))

    def step_resident(t):
        b.step_device(dU[t].data_ptr(), dO[t].data_ptr(), dC.data_ptr(), dL.data_ptr(), dI.data_ptr(), dT.data_ptr(), True)

    for t in range(W):
        step_resident(t)
    b.profile(True)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    th0 = time.perf_counter()
    for t in range(W, W + K):
        step_resident(t)
    e1.record()
    barrier()
    th1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop((th0, th1)) if rank == 0 else None
    kern_ms, kern_n, flush_ms = b.profile(False)
    dd_bytes = b.downdate_bytes()
    b.sync()
    assert np.all(b.n_landmarks() == N), 'a filter appended: the timed steps must be updates only'
    assert np.all(np.isfinite(b.poses())), 'a filter diverged'
    b.close()

    # ---------------- end-to-end arm: host buffers in, poses out, every step ----------------
    g = make()
    hU, hO = torch.from_numpy(U).pin_memory(), torch.from_numpy(OBS).pin_memory()
    nU, nO = hU.numpy(), hO.numpy()

    def step_e2e(t):
        g.step(nU[t], packed=(nO[t], counts, loc, idx, ntab, True))
        return g.poses()

    for t in range(W):
        step_e2e(t)
    barrier()
    t0 = time.perf_counter()
    for t in range(W, W + K):
        poses = step_e2e(t)
    barrier()
    e2e_s = time.perf_counter() - t0
    assert np.all(np.isfinite(poses))
    g.close()

    tt = torch.tensor([ms, e2e_s * 1e3, kern_ms / max(kern_n, 1), flush_ms / max(kern_n, 1)], dtype=torch.float64, device='cuda')
    tot = torch.tensor([float(nf), dd_bytes], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms, e2e_ms, kern_avg_ms, flush_avg_ms = float(tt[0]), float(tt[1]), float(tt[2]), float(tt[3])
    nf_total = int(tot[0])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    achieved = dd_bytes / (flush_avg_ms * 1e-3) / 1e9           # rank 0's filters / slowest rank's flush kernel time
    n = 3 + 2 * N
    line = {
        'metric': METRIC_MC if (args.filters, N) == (4096, 200) else 'EKF updates/sec, %d filters x %d landmarks' % (args.filters, N),
        'value': K * k * nf_total / (ms * 1e-3), 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W,
        'ms_per_step': ms / K, 'higher_is_better': True, 'scaling': 'strong' if world > 1 else 'weak', 'vs_baseline': None,
        'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': '%d independent EKF_SLAM_UC filters x %d landmarks (n=%d), predict + %d associated updates per filter per step'
                               % (args.filters, N, n, k), 'filters': args.filters, 'landmarks': N, 'k': k, 'storage': args.storage,
                   'defer_m': args.defer, 'parallelism': 'filter f on rank f mod %d, no communication (replicas only)' % world,
                   'l2': 'no flush needed: each step streams %.2f GB of covariances per GPU, far beyond the 126 MB L2' % (dd_bytes / 1e9)},
        'roofline': {'bound': 'hbm', 'kernel': 'k_mc_flush_w (one deferred rank-%d pass over every filter\'s P; k_mc_front does predict + %d associate/linearise/gain per filter before it)' % (2 * min(k, args.defer), k),
                     'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'peak_source': 'MEASURED_PEAKS.json hbm_gbs (of measured)' if peaks else 'fallback 6650 (of fallback)',
                     'frac_of_8TBps_nominal': achieved / 8000.0,
                     # from profiles/ncu_mcflush_r01_final_raw.csv (this configuration only): 3.044 GB read + 2.642 GB written
                     'traffic': 5.686e9 if (N == 200 and k == 5 and args.filters == 4096 and args.storage == 'sym' and world == 1) else None,
                     'algorithmic_bytes_per_launch': dd_bytes,
                     'avg_launch_ms': flush_avg_ms, 'launches': kern_n, 'share_of_step': flush_avg_ms * K / ms,
                     'front_kernel_avg_ms': kern_avg_ms - flush_avg_ms},
        'e2e': {'value': K * k * nf_total / (e2e_ms * 1e-3), 'unit': UNIT,
                'h2d_bytes_per_step': int(nU[0].nbytes + nO[0].nbytes + counts.nbytes + loc.nbytes + idx.nbytes + ntab.nbytes),
                'd2h_bytes_per_step': int(nf * 24 + 8), 'api': 'FilterBatch.step(u, observed_LL, tables); .poses()'},
        'gpu_launches': int(2 * kern_n), 'clocks': clk,
    }
    if not args.no_cpu_baseline:
        line['cpu_baseline'] = mc_cpu_baseline(N, k, x, P, s, lms, U, OBS, args.cpu_budget)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def mc_cpu_baseline(N, k, x, P, s, lms, U, OBS, budget_s):
    """The reference's algorithm (dense-literal oracle port) on ONE filter for as many steps as fit the budget;
    independent filters run one after another on the CPU, so filters x steps / time is the whole-batch rate."""
    from oracle import dense_literal as O
    from types import SimpleNamespace
    from tests.golden_util import Stub
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    o = O.EKF_SLAM_UC()
    o.x, o.P, o.s = x.copy(), P.copy(), list(s)
    t0 = time.perf_counter()
    done = 0
    for t in range(U.shape[0]):
        o.predict(U[t, 0])
        o.measure(None, U[t, 0], Stub(OBS[t, 0].T.copy(), table))
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return {'value': done * k / dt, 'unit': UNIT, 'cores': os.cpu_count(), 'kind': 'port',
            'sample': 'dense-literal oracle (EKF_SLAM_UC.m:42-152, Correspondence.m:28-88 incl. its N dense H P H\' products), '
                      'one filter, %d steps of predict + %d associated updates at n=%d, NumPy/OpenBLAS' % (done, k, 3 + 2 * N)}


def run_mc_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    N, k = args.landmarks, args.k
    x, P, s, lms, U, OBS, loc, idx = make_mc_inputs(N, k, 1, max(4, args.steps + args.warmup) * 4)
    per = min(8.0, 150.0 / max(1, args.steps + args.warmup))
    vals = []
    for i in range(args.warmup + args.steps):
        cb = mc_cpu_baseline(N, k, x, P, s, lms, U, OBS, per)
        if i >= args.warmup:
            vals.append(cb['value'])
    val = float(np.mean(vals))
    line = {'impl': 'reference', 'metric': METRIC_MC, 'value': val, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1e3 * k * args.filters / val, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': '%d independent EKF_SLAM_UC filters x %d landmarks' % (args.filters, N), 'filters': args.filters,
                       'landmarks': N, 'k': k},
            'cpu_baseline': {'value': val, 'unit': UNIT, 'cores': os.cpu_count(), 'kind': 'port', 'sample': cb['sample']},
            'e2e': {'value': val, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}, 'gpu_launches': 0}
    print(json.dumps(line))


if __name__ == '__main__':
    main()
