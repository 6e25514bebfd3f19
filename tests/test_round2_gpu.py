"""Round-2 GPU parity tests: the benchmarked configurations at their full size AND over a long horizon, mode-B
association at 10 000 landmarks with engineered near-ties, and the boundary items added this round (property writes,
EKF_ENONFINITE, in-place mutated landmark tables, state after a failed measure).

Tolerance (BASELINE.json north_star): association indices / actions identical; x and P <= 1e-9 relative (Frobenius)
at every checkpoint.  The oracles are oracle/structured.c (O(n^2) per update, what makes 10k x 100 steps and
1k x 1000 steps affordable on the host) and oracle/dense_literal.py (the reference's dense algebra)."""
import time
from types import SimpleNamespace

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300))


def _blocked_rel(Pg, Po, step=1024):
    num = den = 0.0
    for j in range(0, Po.shape[1], step):
        a, b = Pg[:, j:j + step], Po[:, j:j + step]
        num += float(np.sum((a - b) ** 2)); den += float(np.sum(b * b))
    return (num / den) ** 0.5


# ---------------------------------------------------------------- config 3 (the bench's): 10 000 landmarks x 100 steps
def test_config3_10k_landmarks_100_steps_defer16_and_defer8():
    """BASELINE config 3 at full size over a long horizon: N = 10 000 (n = 20 003, P 3.2 GB), the populated map and
    the observation stream of bench.py, 100 steps of predict + 8 associated updates through the class API
    (EKF_SLAM_UC.predict / .measure) = 800 updates.  Two filters run side by side against ONE pass of
    oracle/structured.c over the same inputs: defer_m = 16 (the bench default: terms stay pending across predict, one
    rank-32 flush per two steps, >= 48 of them here) and defer_m = 8 (one rank-16 flush per step, bit-identical to the
    immediate path).  Actions identical; x and P <= 1e-9 relative at steps 25, 50, 75 and 100."""
    import ekf_slam_b200 as E
    from oracle import structured as S
    from bench import make_inputs
    from tests.golden_util import Stub
    N, k, T = 10000, 8, 100
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    gs = {m: E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=m) for m in (16, 8)}
    for g in gs.values():
        g.set_state_lowrank(x, s, A, d)
    o = S.EKF_SLAM_UC(capacity_landmarks=N)
    o.set_state(x, gs[16].P, s)                      # the oracle starts from the very bits the device built
    worst = {m: [0.0, 0.0] for m in gs}
    t0 = time.perf_counter()
    for t in range(T):
        stub = Stub(obs[t], table)
        for g in gs.values():
            g.predict(us[t]); g.measure(None, us[t], stub)
        o.predict(us[t]); o.measure(None, us[t], stub)
        if (t + 1) % 25 == 0:
            for m, g in gs.items():
                assert g.chosen == o.chosen, 'defer %d: actions differ by step %d' % (m, t)
                ex = _rel(g.x, o.x)
                eP = _blocked_rel(g.P, o._P)
                worst[m] = [max(worst[m][0], ex), max(worst[m][1], eP)]
                assert ex <= TOL and eP <= TOL, (m, t, ex, eP)
    assert len(o.chosen) == T * k and all(a == 'update' for a, _ in o.chosen)
    print('10k x %d steps (%.0f s): defer 16 worst rel err x %.2e P %.2e; defer 8 x %.2e P %.2e' %
          (T, time.perf_counter() - t0, worst[16][0], worst[16][1], worst[8][0], worst[8][1]))
    for g in gs.values():
        g.close()


# ---------------------------------------------------------------- config 2: 1 000 landmarks x 1000 steps, full loop
@pytest.fixture(scope='module')
def config2_oracle():
    """One pass of oracle/structured.c over SURVEY 8(d) config 2 (N = 1000, k = 8, T = 1000): x and P at every 100th
    step, plus the dense-literal oracle (the reference's O(n^3) algebra) over the first 10 steps."""
    from oracle import structured as S
    from oracle import dense_literal as O
    from bench import make_inputs
    from ekf_slam_b200.synth import lowrank_to_dense
    from tests.golden_util import Stub
    N, k, T = 1000, 8, 1000
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    P0 = lowrank_to_dense(A, d)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    o = S.EKF_SLAM_UC(capacity_landmarks=N)
    o.set_state(x, P0, s)
    cps = {}
    for t in range(T):
        o.predict(us[t]); o.measure(None, us[t], Stub(obs[t], table))
        if (t + 1) % 100 == 0 or t + 1 == 10:
            cps[t + 1] = (o.x, o.P)
    dl = O.EKF_SLAM_UC()
    dl.x, dl.P, dl.s = x.copy(), P0.copy(), list(s)
    for t in range(10):                              # slots from the structured oracle's association (mode A = signature);
        dl.predict(us[t])                            # the dense N x (2 x n)(n x n) association products are skipped
        for r in obs[t]:
            dl._update(r, np.diag([r[0] * dl.Rc[0], r[1] * dl.Rc[1]]), int(r[2]))
    assert _rel(cps[10][0], dl.x) <= 1e-12 and _rel(cps[10][1], dl.P) <= 1e-12     # the two oracles agree
    return dict(N=N, k=k, T=T, x=x, P0=P0, s=s, lms=lms, us=us, obs=obs, table=table, cps=cps, chosen=list(o.chosen),
                dense10=(dl.x.copy(), dl.P.copy()))


@pytest.mark.parametrize('storage', ['sym', 'full'])
@pytest.mark.parametrize('defer_m', [0, 8, 16])
def test_config2_1k_landmarks_1000_steps_full_loop(config2_oracle, storage, defer_m):
    """SURVEY 8(d) config 2 as stated: N = 1000, n = 2003, k = 8, T = 1000, the whole predict / measure loop through the
    class API, both storages, immediate / per-measure deferred / cross-step deferred downdate: actions identical, x and
    P <= 1e-9 against oracle/structured.c at every 100th step and against the dense-literal oracle after 10 steps."""
    import ekf_slam_b200 as E
    from tests.golden_util import Stub
    c = config2_oracle
    g = E.EKF_SLAM_UC(capacity_landmarks=c['N'], storage=storage, defer_m=defer_m)
    g.set_state(c['x'], c['P0'], c['s'])
    worst = [0.0, 0.0]
    for t in range(c['T']):
        g.predict(c['us'][t]); g.measure(None, c['us'][t], Stub(c['obs'][t], c['table']))
        if t + 1 == 10:
            ex, eP = _rel(g.x, c['dense10'][0]), _rel(g.P, c['dense10'][1])
            assert ex <= TOL and eP <= TOL, ('dense', ex, eP)
        if (t + 1) in c['cps']:
            xo, Po = c['cps'][t + 1]
            ex, eP = _rel(g.x, xo), _rel(g.P, Po)
            worst = [max(worst[0], ex), max(worst[1], eP)]
            assert ex <= TOL and eP <= TOL, (t, ex, eP)
    assert g.chosen == c['chosen']
    print('1k x 1000 steps %s/defer %d: worst rel err x %.2e P %.2e' % (storage, defer_m, worst[0], worst[1]))
    g.close()


# ---------------------------------------------------------------- mode-B association at 10 000 landmarks
@pytest.mark.parametrize('storage', ['sym', 'full'])
def test_mode_b_association_10k_landmarks_with_near_ties(storage):
    """Correspondence.m:49-87 with the Mahalanobis term in the score (:74, commented out upstream) at N = 10 000:
    1200 queries, GPU index == oracle/structured.c index for every one of them.  120 of the queries aim at engineered
    pairs: exact duplicates (same position, same covariance blocks: costs tie, the first index must win, :81) and
    twins whose position differs by one ulp (costs a few ulp apart: only bit-identical arithmetic on both sides picks
    the same one; the GPU and the C oracle share every operation incl. atan2, ekf_math_shared.h)."""
    import ekf_slam_b200 as E
    from oracle import structured as S
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense, observe_from_truth
    N = 10000
    x, A, d, s, lms = synthetic_spd_state(N, seed=5)
    P = lowrank_to_dense(A, d)
    rng = np.random.default_rng(8)
    base = rng.choice(np.arange(50, N - 50), 60, replace=False)
    pairs = []
    for q, a in enumerate(base):                     # slot b = a + 17 becomes a copy of landmark a
        b = int(a) + 17
        ia, ib = 3 + 2 * int(a), 3 + 2 * b
        x[ib:ib + 2] = x[ia:ia + 2]
        lms[b] = lms[a]
        P[:3, ib:ib + 2] = P[:3, ia:ia + 2]; P[ib:ib + 2, :3] = P[ia:ia + 2, :3]
        P[ib:ib + 2, ib:ib + 2] = P[ia:ia + 2, ia:ia + 2]
        if q % 2:                                    # near-tie: one ulp apart in one coordinate
            x[ib + (q // 2) % 2] = np.nextafter(x[ib + (q // 2) % 2], np.inf if q % 4 == 1 else -np.inf)
        pairs.append((int(a), b))
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, assoc_mode='position')
    g.correspondence.s_thresh = 9.21                 # chi-square(2) 99 % gate
    g._sync_corr()
    ones = np.ones(N)
    g.set_state(x, P, ones)
    o = S.EKF_SLAM_UC(capacity_landmarks=N, assoc_mode=1)
    o.s_thresh = 9.21
    o.set_state(x, P, ones)
    del P
    ids = np.concatenate([[a for a, _ in pairs], [b for _, b in pairs], rng.choice(N, 1080, replace=False)])
    rows = observe_from_truth(lms, x[:3], ids, rng, seam_deg=0.0)
    assert rows.shape[0] == 1200
    n_new = n_pair_hits = 0
    pair_slots = {a + 1 for a, _ in pairs} | {b + 1 for _, b in pairs}
    t_gpu = 0.0
    for r in rows:
        z = np.array([r[0], r[1], 1.0])
        Rd = np.array([z[0] * 0.1, z[1] * 5.0])
        t0 = time.perf_counter()
        got = g.correspondence.estimateCorrespondence(z, np.diag(Rd))
        t_gpu += time.perf_counter() - t0
        want = o.associate(z, Rd)
        assert got == want, (z, got, want)
        n_new += int(want[0]); n_pair_hits += int(want[1] in pair_slots)
    assert n_pair_hits >= 60 and n_new < 600
    # 25 doubles of P per landmark, one 32-byte sector each in the worst case, + x and s: SURVEY 2.2's 25*32*N B
    print('mode B at 10k/%s: 1200 queries identical (%d on engineered pairs, %d new); %.1f us per synchronous call '
          '(launch + D2H), %.1f GB/s of sector traffic' % (storage, n_pair_hits, n_new, t_gpu / 1200 * 1e6,
                                                           25 * 32 * N / (t_gpu / 1200) / 1e9))
    g.close()


# ---------------------------------------------------------------- hybrid flush kernels (variants 30+)
@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('defer_m', [3, 8, 11, 16])
@pytest.mark.parametrize('variant', ['30', '31', '32', '33', '34', '35', '36', '37', '38'])
def test_hybrid_flush_is_bit_identical_to_the_tma_flush(storage, defer_m, variant, monkeypatch):
    """k_downdate_hyb (bulk-async in, registers out; B operand in the stage for BST variants) applies the same terms in
    the same order as k_downdate_tma: x and P bit-identical, at ranks below and above 16, odd term counts included,
    with n not a multiple of the 128-row tile."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 333
    x, A, d, s, lms = synthetic_spd_state(N, seed=9)
    fs = []
    for var in (None, variant):
        if var is None:
            monkeypatch.delenv('EKF_DOWNDATE_VARIANT', raising=False)
        else:
            monkeypatch.setenv('EKF_DOWNDATE_VARIANT', var)
        fs.append(E.EKF_SLAM(capacity_landmarks=N + 5, storage=storage, defer_m=defer_m))
    for f in fs:
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(4)
    for step in range(5):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 16, replace=False), rng)[:7 + step % 2]
        Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
        for f in fs:
            f.predict([0.2, 1.0])
            f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
    assert np.array_equal(fs[0].x, fs[1].x)
    assert np.array_equal(fs[0].P, fs[1].P), _rel(fs[1].P, fs[0].P)
    for f in fs:
        f.close()


# ---------------------------------------------------------------- boundary items
@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_public_properties_x_P_s_are_writable(storage):
    """x, P, s are public read/write properties upstream (EKF_SLAM.m:6-9): assigning one keeps the other two; the
    filter continues from the assigned value exactly as if it had been uploaded with set_state."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense, observe_from_truth
    N = 40
    x, A, d, s, lms = synthetic_spd_state(N, seed=2)
    P = lowrank_to_dense(A, d)
    g = E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=8)
    g.set_state(x, P, s)
    x2 = x + 0.01
    g.x = x2
    assert np.array_equal(g.x, x2) and np.array_equal(g.P, P) and np.array_equal(g.s, s)
    P2 = P * 1.5
    g.P = P2
    assert np.array_equal(g.P, P2) and np.array_equal(g.x, x2)
    s2 = s[::-1].copy()
    g.s = s2
    assert np.array_equal(g.s, s2) and np.array_equal(g.P, P2)
    ref = E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=8)
    ref.set_state(x2, P2, s2)
    rows = observe_from_truth(lms, x2[:3], np.arange(0, N, 5), np.random.default_rng(1))
    for f in (g, ref):
        f.predict([0.3, 2.0])
        for r in rows:
            f.update(r[:2], [r[0] * 0.01, r[1] * 5.0], int(r[2]))
    g.x = g.x                                        # assignment with terms pending (flushes first)
    assert np.array_equal(g.x, ref.x) and np.array_equal(g.P, ref.P)
    with pytest.raises(E.EkfError):
        g.x = x[:-2]                                 # wrong length: changing N is set_state's job
    with pytest.raises(ValueError):
        g.P = P[:-2, :-2]
    g.close(); ref.close()


def test_nonfinite_innovation_is_reported_once():
    """EKF_SLAM.m:137 divides by q = |landmark - robot|^2: a robot standing on the observed landmark makes H, phi and K
    non-finite and poisons x and P.  The reference carries the NaNs silently; the library raises EKF_ENONFINITE at
    the next synchronising call (once), on the per-observation, the fused-batch and the cross-step deferred path."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state
    N = 30
    x, A, d, s, lms = synthetic_spd_state(N, seed=4)
    for defer_m, batch in ((0, False), (8, True), (16, True)):
        g = E.EKF_SLAM(capacity_landmarks=N, storage='sym', defer_m=defer_m)
        xr = x.copy()
        xr[0:2] = xr[3 + 2 * 6: 5 + 2 * 6]           # robot exactly on landmark 7
        g.set_state_lowrank(xr, s, A, d)
        z, Rd, idx = np.array([[1.0, 90.0], [2.0, 100.0]]), np.array([[0.01, 450.0], [0.02, 500.0]]), np.array([7, 9], dtype=np.int32)
        if batch:
            g.update_batch(z, Rd, idx)
        else:
            g.update(z[0], Rd[0], 7)
        with pytest.raises(E.EkfError) as ei:
            g.x
        assert ei.value.code == -5 and 'non-finite' in str(ei.value)
        assert not np.all(np.isfinite(g.x))          # reported once; the state stays poisoned like the reference's
        g.set_state_lowrank(x, s, A, d)              # a new state clears it
        g.update(z[0], Rd[0], 7)
        assert np.all(np.isfinite(g.x))
        g.close()


def test_landmark_table_mutated_in_place_between_steps():
    """The reference's landmark database changes entries in place without changing the list or its length
    (RANSAC.m:262 promotes index 0 to max+1, :268 rewrites .loc).  measure() must read the table as it is NOW."""
    import ekf_slam_b200 as E
    from oracle import dense_literal as O
    from tests.golden_util import Stub
    table = [SimpleNamespace(index=1, loc=np.array([2.0, 1.0])), SimpleNamespace(index=0, loc=np.array([-3.0, 2.0])),
             SimpleNamespace(index=0, loc=np.array([1.0, -4.0]))]
    g, o = E.EKF_SLAM_UC(capacity_landmarks=8), O.EKF_SLAM_UC()
    u = np.array([0.1, 2.0])

    def both(rows):
        for f in (g, o):
            f.predict(u); f.measure(None, u, Stub(np.array(rows, dtype=np.float64), table))

    both([[2.2, 30.0, 1.0]])                         # first landmark: appended with signature 1 at table[0].loc
    table[1].index = 2; table[1].loc = np.array([-3.5, 2.5])      # promoted and moved IN PLACE
    both([[2.2, 31.0, 1.0], [4.0, 140.0, 2.0]])
    assert np.array_equal(g.x[5:7], [-3.5, 2.5])     # the moved location, not the one the table held at the first call
    table[2].index = 3
    table[0].loc = np.array([9.0, 9.0])              # no effect: landmark 1 is already in the state
    both([[2.1, 32.0, 1.0], [4.1, 141.0, 2.0], [4.2, 280.0, 3.0]])
    assert g.chosen == o.chosen == [('append', 1), ('update', 1), ('append', 2), ('update', 1), ('update', 2), ('append', 3)]
    assert _rel(g.x, o.x) <= TOL and _rel(g.P, o.P) <= TOL
    g.close()


@pytest.mark.parametrize('defer_m', [0, 8, 16])
def test_failed_measure_leaves_the_rows_before_the_error_applied(defer_m):
    """A table miss in the middle of observed_LL raises (MATLAB would, at EKF_SLAM_UC.m:123); every row before it has
    been applied, as in the reference, on the immediate and on the queued (deferred) path alike."""
    import ekf_slam_b200 as E
    from oracle import dense_literal as O
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense, observe_from_truth
    from tests.golden_util import Stub
    N = 25
    x, A, d, s, lms = synthetic_spd_state(N, seed=6)
    P = lowrank_to_dense(A, d)
    g = E.EKF_SLAM_UC(capacity_landmarks=N + 4, storage='sym', defer_m=defer_m)
    g.set_state(x, P, s)
    o = O.EKF_SLAM_UC()
    o.x, o.P, o.s = x.copy(), P.copy(), list(s)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    rows = observe_from_truth(lms, x[:3], [2, 5, 9, 14], np.random.default_rng(3), seam_deg=0.0)
    rows = np.vstack([rows[:3], [[3.0, 100.0, 77.0]], rows[3:]])   # row 4: unseen signature 77, not in the table
    u = np.array([0.2, 1.0])
    g.predict(u); o.predict(u)
    with pytest.raises(E.EkfError):
        g.measure(None, u, Stub(rows, table))
    for r in rows[:3]:
        o._update(r, np.diag([r[0] * 0.1, r[1] * 5.0]), int(r[2]))
    assert _rel(g.x, o.x) <= TOL and _rel(g.P, o.P) <= TOL
    g.close()


@pytest.mark.parametrize('defer_m', [16, 11])
def test_lazy_flush_matches_flush_then_linearise(defer_m, monkeypatch):
    """Lazy flush (cross-step deferral, SYM): a full set of pending terms is flushed when the NEXT batch arrives -- the K3
    pass runs on the main stream while that batch is linearised on a side stream from copies of its slot columns + the
    pending terms, gains into the other pair of slabs.  The linearisation itself sees the same operands in the same
    order as flush-first; what differs from EKF_LAZY_FLUSH=0 is that a full set now also crosses the predict in between
    (transformed there, as every pending term is), i.e. rounding in rows / columns 0-1 only: x at every step and the
    final P agree to 1e-12, with ragged batches, single updates, getters and an append in between."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 700
    x, A, d, s, lms = synthetic_spd_state(N, seed=13)
    runs = []
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_LAZY_FLUSH', flag)
        f = E.EKF_SLAM(capacity_landmarks=N + 4, storage='sym', defer_m=defer_m)
        f.set_state_lowrank(x, s, A, d)
        rng = np.random.default_rng(31)
        xs = []
        for step in range(14):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:(8, 8, 8, 11, 5, 8, 1, 8, 8, 3, 8, 8, 8, 8)[step]]
            Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
            f.predict([0.3, 1.0])
            if step == 6:
                f.update(rows[0, :2], Rd[0], int(rows[0, 2]))
            else:
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
            xs.append(f.x)
            if step == 8:
                xs.append(f.P_block(0, 0, 5, 5))
            if step == 10:
                f.append([0.3, 1.0], np.diag([0.02, 100.0]), [1.0, 2.0], N + 1)
        runs.append((xs, f.P))
        f.close()
    assert max(_rel(a, b) for a, b in zip(runs[0][0], runs[1][0])) <= 1e-12
    assert _rel(runs[0][1], runs[1][1]) <= 1e-12


@pytest.mark.parametrize('defer_m', [16, 11])
def test_continuous_flush_matches_in_place_deferral(defer_m, monkeypatch):
    """Continuous flush (single GPU, SYM, cross-step deferral; EKF_SHADOW=0 disables): the K3 pass of a full set runs out
    of place on its own stream while predict and the next batches carry on from the untouched matrix + the robot strip +
    the whole window of pending terms.  Same estimator as the in-place deferral; what differs is when a term reaches
    the stored matrix relative to the predicts around it, i.e. rounding in rows / columns 0-1: x at every step, blocks of P
    read mid-run (which settle everything) and the final P agree to 1e-12 over 60 steps (>= 25 passes) with ragged
    batches, single updates, an append, a property write and getters in between."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 700
    x, A, d, s, lms = synthetic_spd_state(N, seed=17)
    sizes = (8, 8, 8, 11, 5, 8, 1, 8, 8, 3, 8, 8, 8, 8, 16, 8, 2, 8, 8, 8)
    runs = []
    monkeypatch.setenv('EKF_SHADOW_MIN_N', '0')      # the mode is for large maps (default: n >= 8000); force it at this size
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_SHADOW', flag)
        f = E.EKF_SLAM(capacity_landmarks=N + 4, storage='sym', defer_m=defer_m)
        f.set_state_lowrank(x, s, A, d)
        rng = np.random.default_rng(41)
        xs = []
        for step in range(60):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:sizes[step % len(sizes)]]
            Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
            f.predict([0.3, 1.0])
            if step % 20 == 6:
                f.update(rows[0, :2], Rd[0], int(rows[0, 2]))
            else:
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
            xs.append(f.x)
            if step in (8, 33):
                xs.append(f.P_block(0, 0, 7, 7))
                xs.append(f.P_block(3, 100, 40, 300))
            if step == 10:
                f.append([0.3, 1.0], np.diag([0.02, 100.0]), [1.0, 2.0], N + 1)
            if step == 21:
                f.predict([0.1, -2.0])          # two predicts in a row while a pass is in flight
            if step == 40:
                xv = f.x
                xv[0] += 0.25
                f.x = xv                          # property write: pending terms are applied first
        runs.append((xs, f.P))
        f.close()
    assert max(_rel(a, b) for a, b in zip(runs[0][0], runs[1][0])) <= 1e-12
    assert _rel(runs[0][1], runs[1][1]) <= 1e-12
    assert np.array_equal(runs[1][1], runs[1][1].T)


def test_continuous_flush_settles_for_everything_that_needs_P(monkeypatch):
    """Entry points that need the whole matrix while passes are in flight: the ellipses, a mode-B association (reads P
    blocks), an append, replacing the state (set_state_lowrank: everything pending is dropped) and closing the handle
    with passes outstanding.  Continuous flush (forced at this size) against the in-place deferral, 1e-12."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 500
    x, A, d, s, lms = synthetic_spd_state(N, seed=29)
    monkeypatch.setenv('EKF_SHADOW_MIN_N', '0')
    runs = []
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_SHADOW', flag)
        f = E.EKF_SLAM_UC(capacity_landmarks=N + 2, storage='sym', defer_m=16, assoc_mode='position')
        f.correspondence.s_thresh = 9.21
        f._sync_corr()
        f.set_state_lowrank(x, np.ones(N), A, d)
        assert f.flush_info()[0] == (flag == '1')
        rng = np.random.default_rng(3)
        out = []
        for step in range(30):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:8]
            Rd = np.column_stack([rows[:, 0] * f.Rc[0], rows[:, 1] * f.Rc[1]])
            f.predict([0.3, 1.0])
            f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
            if step == 7:
                out.append(f.covariance_ellipses())
            if step == 11:
                z = np.array([rows[0, 0], rows[0, 1], 1.0])
                out.append(np.array(f.correspondence.estimateCorrespondence(z, np.diag([z[0] * 0.1, z[1] * 5.0])), dtype=float))
            if step == 15:
                f.append([0.3, 1.0], np.diag([0.02, 100.0]), [1.0, 2.0], 1.0)
            if step == 20:                                   # replace the state with passes in flight
                f.set_state_lowrank(x, np.ones(N), A, d)
            out.append(f.x)
        out.append(f.P)
        runs.append(out)
        g = E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=16)      # closed with passes outstanding
        g.set_state_lowrank(x, s, A, d)
        for step in range(5):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:8]
            g.predict([0.3, 1.0])
            g.update_batch(rows[:, :2], np.column_stack([rows[:, 0] * g.Rc[0], rows[:, 1] * g.Rc[1]]), rows[:, 2].astype(np.int32))
        g.close(); f.close()
    assert len(runs[0]) == len(runs[1])
    assert max(_rel(a, b) for a, b in zip(runs[0], runs[1])) <= 1e-12


@pytest.mark.parametrize('defer_m', [16, 11])
def test_batch_front_kernel_is_bit_identical_to_gather_lin_gain(defer_m, monkeypatch):
    """k_batch_front runs the linearisation chain (block 0) beside the gather + pending-term correction of every row's slot
    entries, and k_batch_gain then starts with nothing pending: the same terms in the same order on the same operands as
    gather -> k_batch_lin -> k_batch_gain(m0) (EKF_BATCH_FRONT=0).  x at every step and the final P must be the same bits,
    with windows of up to 32 pending terms, ragged batches, single updates and the device-decided measure()."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth, LandmarkStub
    N = 600
    x, A, d, s, lms = synthetic_spd_state(N, seed=23)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    sizes = (8, 8, 8, 11, 5, 8, 1, 8, 8, 3, 8, 16)
    monkeypatch.setenv('EKF_SHADOW_MIN_N', '0')
    runs = []
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_BATCH_FRONT', flag)
        f = E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=defer_m)
        f.set_state_lowrank(x, s, A, d)
        assert f.flush_info()[0]
        rng = np.random.default_rng(7)
        xs = []
        for step in range(36):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:sizes[step % len(sizes)]]
            Rd = np.column_stack([rows[:, 0] * f.Rc[0], rows[:, 1] * f.Rc[1]])
            f.predict([0.3, 1.0])
            if step % 12 == 6:
                f.update(rows[0, :2], Rd[0], int(rows[0, 2]))
            elif step % 3 == 0:
                f.measure(None, [0.3, 1.0], LandmarkStub(rows, table))          # decided on the device
            else:
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
            xs.append(f.x)
        runs.append((xs, f.P, list(f.chosen)))
        f.close()
    assert runs[0][2] == runs[1][2]
    assert all(np.array_equal(a, b) for a, b in zip(runs[0][0], runs[1][0]))
    assert np.array_equal(runs[0][1], runs[1][1])


@pytest.mark.parametrize('defer_m', [0, 8, 16])
def test_x_read_back_after_predict_is_the_full_state(defer_m):
    """ekf_get_x keeps the last x it returned in pinned memory and, when only predict has run since (it moves x(1:3),
    EKF_SLAM.m:57-60), copies just the pose.  A filter whose x is read after EVERY call must return the bits of a twin
    that reads x only after a measure (always a full copy), through updates, appends, property writes and double predicts."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    w = make_world(N=16, T=60, k=4, mode='nearest', seed=5)
    a = E.EKF_SLAM_UC(capacity_landmarks=16, storage='sym', defer_m=defer_m)
    b = E.EKF_SLAM_UC(capacity_landmarks=16, storage='sym', defer_m=defer_m)
    for t in range(w.T):
        a.predict(w.u[t]); xa = a.x
        b.predict(w.u[t])
        if t % 5 == 2:
            a.predict(w.u[t]); xa2 = a.x; b.predict(w.u[t])
            assert np.array_equal(xa2[3:], xa[3:])
        if t == 30:
            for f in (a, b):
                v = f.x; v[1] -= 0.5; f.x = v
            assert np.array_equal(a.x, b.x)
        a.measure(None, w.u[t], w.stub(t)); b.measure(None, w.u[t], w.stub(t))
        xa, xb = a.x, b.x
        assert np.array_equal(xa, xb), t
        assert np.array_equal(a.x, xa)                  # a second read with nothing in between
    assert a.chosen == b.chosen and np.array_equal(a.P, b.P)
    a.close(); b.close()


@pytest.mark.parametrize('storage,defer_m', [('sym', 16), ('sym', 8), ('full', 8), ('full', 3)])
def test_measure_decided_on_device_is_bit_identical_to_the_host_loop(storage, defer_m, monkeypatch):
    """ekf_measure_batch_async (EKF_SLAM_UC, signature association): rows whose signatures are all in the map are
    associated by k_measure_decide and the update kernels read the slots from device memory -- no device-to-host copy
    inside the call; rows with an unseen signature (appends) take the host loop.  Same kernels, same slots: actions, x
    at every step and the final P are bit-identical to EKF_ASYNC_MEASURE=0, over an exploration phase (appends mixed in)
    and a populated-map phase with batches of 1..11 rows (several fused chunks)."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world, synthetic_spd_state, observe_from_truth, LandmarkStub
    w = make_world(N=24, T=160, k=6, mode='nearest', seed=11)
    N2 = 300
    x2, A2, d2, s2, lms2 = synthetic_spd_state(N2, seed=19)
    table2 = [SimpleNamespace(index=i + 1, loc=lms2[i]) for i in range(N2)]
    runs = {}
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_ASYNC_MEASURE', flag)
        g = E.EKF_SLAM_UC(capacity_landmarks=24, storage=storage, defer_m=defer_m)
        xs = []
        for t in range(w.T):
            g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t), pass_x=(t % 2 == 0))
            if t % 3 == 0:
                xs.append(g.x)
        out = [xs, list(g.chosen), g.P, g.s]
        g.close()
        g = E.EKF_SLAM_UC(capacity_landmarks=N2, storage=storage, defer_m=defer_m)
        g.set_state_lowrank(x2, s2, A2, d2)
        rng = np.random.default_rng(3)
        for step in range(12):
            rows = observe_from_truth(lms2, x2[:3], rng.choice(N2, 30, replace=False), rng)[:(8, 11, 1, 5, 8, 2, 9, 8, 3, 8, 8, 7)[step]]
            g.predict([0.3, 1.0]); g.measure(None, [0.3, 1.0], LandmarkStub(rows, table2), pass_x=False)
            if step % 4 == 1:
                xs.append(g.x)
            if step == 6:
                g.s = g.s                                # property write in between (rebuilds the signature mirror)
        out += [list(g.chosen), g.x, g.P]
        g.close()
        runs[flag] = out
    a, b = runs['0'], runs['1']
    assert a[1] == b[1] and a[4] == b[4], 'actions differ'
    assert len(a[0]) == len(b[0]) and all(np.array_equal(p, q) for p, q in zip(a[0], b[0]))
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3]) and np.array_equal(a[5], b[5]) and np.array_equal(a[6], b[6])
    assert all(act == 'update' for act, _ in b[4]) and len(b[4]) == 78
