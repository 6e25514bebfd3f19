"""GPU parity tests proper: the CUDA path (through the C ABI / class mirror) vs the dense-literal oracle
on identical seeded inputs.  Tolerance (BASELINE.json north_star): association indices / actions identical,
state and covariance <= 1e-9 relative Frobenius over 1000 steps."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-9


def _rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _run_pair(gpu, ora, w, check_every=100):
    worst = (0.0, 0.0)
    for t in range(w.T):
        gpu.predict(w.u[t]); gpu.measure(None, w.u[t], w.stub(t))
        ora.predict(w.u[t]); ora.measure(None, w.u[t], w.stub(t))
        if (t + 1) % check_every == 0 or t == w.T - 1:
            assert gpu.chosen == ora.chosen, 'actions differ by step %d' % t
            ex, eP = _rel(gpu.x, ora.x), _rel(gpu.P, ora.P)
            worst = (max(worst[0], ex), max(worst[1], eP))
            assert ex <= TOL and eP <= TOL, (t, ex, eP)
    return worst


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('variant,mode,T', [('known', 'prefix', 1000), ('known', 'nearest', 50), ('uc', 'nearest', 1000)])
def test_config1_example_style_1000_steps(storage, variant, mode, T):
    """BASELINE config 1 (N=20): 1a prefix rule, 1b idx=ii quirk run (EKF_SLAM.m:123), UC k-nearest."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=20, T=T, k=5, mode=mode)
    gpu = (E.EKF_SLAM if variant == 'known' else E.EKF_SLAM_UC)(capacity_landmarks=20, storage=storage)
    ora = (O.EKF_SLAM if variant == 'known' else O.EKF_SLAM_UC)()
    ex, eP = _run_pair(gpu, ora, w)
    assert np.array_equal(gpu.s, ora.s)
    assert gpu.N == (len(ora.x) - 3) // 2
    print('config1 %s/%s/%s: worst rel err x %.2e P %.2e' % (variant, mode, storage, ex, eP))
    gpu.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_uc_200_landmarks(storage):
    """One filter of BASELINE config 5 (N=200, EKF_SLAM_UC, k=5), 300 steps."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=200, T=300, k=5, mode='nearest', seed=7)
    gpu = E.EKF_SLAM_UC(capacity_landmarks=200, storage=storage)
    ora = O.EKF_SLAM_UC()
    _run_pair(gpu, ora, w, check_every=100)
    gpu.close()


def _populated(N, storage, cls_name='EKF_SLAM', seed=3):
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense
    x, A, d, s, lms = synthetic_spd_state(N, seed=seed)
    g = getattr(E, cls_name)(capacity_landmarks=N, storage=storage)
    g.set_state_lowrank(x, s, A, d)
    return g, x, lowrank_to_dense(A, d), s, lms


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_lowrank_init_and_roundtrip(storage):
    g, x, P, s, _ = _populated(150, storage)
    assert _rel(g.P, P) < 1e-14
    assert np.array_equal(g.x, x) and np.array_equal(g.s, s)
    g.set_state(x, P, s)
    assert np.array_equal(g.P, P)
    blk = g.P_block(5, 17, 40, 33)
    assert np.array_equal(blk, P[5:45, 17:50])
    g.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_config2_1k_landmarks_updates_vs_dense_literal(storage):
    """BASELINE config 2 (N=1000, n=2003): populated map, predict + 8 updates per step, 5 steps, vs the dense
    O(n^3) oracle."""
    from ekf_slam_b200.synth import observe_from_truth
    from oracle import dense_literal as O
    N = 1000
    g, x, P, s, lms = _populated(N, storage)
    o = O.EKF_SLAM()
    o.x, o.P, o.s = x.copy(), P.copy(), s.copy()
    rng = np.random.default_rng(11)
    for step in range(5):
        u = np.array([0.3, 1.0])
        g.predict(u); o.predict(u)
        ids = rng.choice(N, 12, replace=False)
        rows = observe_from_truth(lms, o.x[:3], ids, rng)[:8]
        for r in rows:
            R = np.diag([r[0] * 0.01, r[1] * 5.0])
            g.update(r[:2], R, int(r[2]))
            o._update(r, R, int(r[2]))
    ex, eP = _rel(g.x, o.x), _rel(g.P, o.P)
    print('1k/%s: rel err x %.2e P %.2e' % (storage, ex, eP))
    assert ex <= TOL and eP <= TOL
    g.close()


def test_predict_leaves_landmark_block_bit_identical():
    g, x, _, s, _ = _populated(64, 'full')
    P = g.P
    g.predict([0.4, 2.0])
    P2 = g.P
    assert np.array_equal(P2[3:, 3:], P[3:, 3:])
    assert not np.array_equal(P2[:3, :], P[:3, :])
    g.close()


def test_append_keeps_old_block_bit_identical_and_capacity_error():
    import ekf_slam_b200 as E
    g, x, P, s, _ = _populated(10, 'full')
    with pytest.raises(E.EkfError) as ei:
        g.append([0.1, 1.0], np.diag([0.02, 100.0]), [1.0, 2.0], 11)
    assert ei.value.code == -2
    g.close()
    g = E.EKF_SLAM(capacity_landmarks=12)
    g.set_state(x, P, s)
    P = g.P
    g.append([0.1, 1.0], np.diag([0.02, 100.0]), [1.0, 2.0], 11)
    P2 = g.P
    assert P2.shape == (25, 25) and np.array_equal(P2[:23, :23], P)
    assert np.array_equal(P2[23:, :23], P2[:23, 23:].T)
    assert g.x[-2:].tolist() == [1.0, 2.0] and g.s[-1] == 11
    g.close()


def test_update_index_out_of_range_is_an_error():
    import ekf_slam_b200 as E
    g, *_ = _populated(10, 'full')
    with pytest.raises(E.EkfError):
        g.update([1.0, 90.0], np.diag([0.01, 450.0]), 11)
    g.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('mode', ['signature', 'position'])
def test_association_indices_bit_exact(mode, storage):
    """Correspondence.m:28-88: mode A as shipped (:75) and mode B (:74) with all signatures equal and a
    chi-square gate, incl. the first-minimum tie rule (:81) on duplicated landmarks."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense, observe_from_truth
    from oracle import dense_literal as O
    N = 300
    x, A, d, s, lms = synthetic_spd_state(N, seed=5)
    P = lowrank_to_dense(A, d)
    if mode == 'position':
        s = np.ones(N)
        x[3 + 2 * 40: 5 + 2 * 40] = x[3 + 2 * 7: 5 + 2 * 7]          # duplicate landmark 8 into slot 41
        P[:, 3 + 2 * 40: 5 + 2 * 40] = P[:, 3 + 2 * 7: 5 + 2 * 7]
        P[3 + 2 * 40: 5 + 2 * 40, :] = P[3 + 2 * 7: 5 + 2 * 7, :]
        P[3 + 2 * 40: 5 + 2 * 40, 3 + 2 * 40: 5 + 2 * 40] = P[3 + 2 * 7: 5 + 2 * 7, 3 + 2 * 7: 5 + 2 * 7]
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, assoc_mode=mode)
    oc = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC', position_cost_in_score=(mode == 'position'))
    if mode == 'position':
        g.correspondence.s_thresh = 9.21
        oc.s_thresh = 9.21
        g._sync_corr()
    g.set_state(x, P, s)
    rng = np.random.default_rng(2)
    ids = np.concatenate([[7, 40], rng.choice(N, 60, replace=False)])
    rows = observe_from_truth(lms, x[:3], ids, rng, seam_deg=0.0)
    n_new = 0
    for r in rows:
        z = r.copy()
        if mode == 'position':
            z[2] = 1.0
        elif rng.random() < 0.2:
            z[2] = N + 5.0                                            # unseen signature -> new landmark
        R = np.diag([z[0] * 0.1, z[1] * 5.0])
        got = g.correspondence.estimateCorrespondence(z, R)
        want = oc.estimateCorrespondence(z, R, x, P, s)
        assert got == want, (z, got, want)
        n_new += int(want[0])
    assert 0 < n_new < len(rows) or mode == 'position'
    # value-class form with explicit arrays (Correspondence.m:28 signature)
    c2 = E.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC')
    z = rows[0]
    assert c2.estimateCorrespondence(z, np.diag([z[0] * 0.1, z[1] * 5.0]), x, P, np.arange(1, N + 1.0)) == \
        O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC').estimateCorrespondence(z, np.diag([z[0] * 0.1, z[1] * 5.0]), x, P, np.arange(1, N + 1.0))
    g.close()


def test_empty_observation_list_is_a_noop():
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import LandmarkStub
    g = E.EKF_SLAM(capacity_landmarks=4)
    g.measure(None, [0.1, 1.0], LandmarkStub(np.zeros((0, 3)), []))
    assert g.N == 0 and np.array_equal(g.P, 0.1 * np.eye(3))
    g.close()


def test_smoke_entry():
    import __graft_entry__ as ge
    ge.smoke()


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('defer_m', [2, 5, 8])
def test_deferred_flush_is_bit_identical_to_immediate(storage, defer_m):
    """Deferred rank-2m downdate: same sequence of linearisations, one pass over P per m
    observations.  All consumers share rank2_sub, so x and P must equal the immediate path bit for bit."""
    from ekf_slam_b200.synth import observe_from_truth
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state
    N = 257
    x, A, d, s, lms = synthetic_spd_state(N, seed=9)
    fs = [E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=m) for m in (0, defer_m)]
    for f in fs:
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(4)
    for step in range(4):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 16, replace=False), rng)[:11]
        for f in fs:
            f.predict([0.2, 1.0])
            for r in rows:
                f.update(r[:2], np.diag([r[0] * 0.01, r[1] * 5.0]), int(r[2]))
    assert np.array_equal(fs[0].x, fs[1].x)
    assert np.array_equal(fs[0].P, fs[1].P)
    for f in fs:
        f.close()


@pytest.mark.parametrize('variant', ['known', 'uc'])
def test_deferred_full_loop_vs_oracle(variant):
    """1000-step config-1 loop with defer_m=8 (appends, associations and predicts force flushes) vs the oracle."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=20, T=1000, k=5, mode='prefix' if variant == 'known' else 'nearest')
    gpu = (E.EKF_SLAM if variant == 'known' else E.EKF_SLAM_UC)(capacity_landmarks=20, storage='sym', defer_m=8)
    ora = (O.EKF_SLAM if variant == 'known' else O.EKF_SLAM_UC)()
    _run_pair(gpu, ora, w)
    gpu.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('defer_m', [2, 3, 8])
@pytest.mark.parametrize('variant', ['0', '4', '6', '8', '9', '11', '13', '14', '15', '16', '17', '18'])
def test_tensor_core_flush_matches_immediate(storage, defer_m, variant, monkeypatch):
    """K3b (mma.sync.m8n8k4.f64 flush) vs the immediate CUDA-core path: same terms in the same order."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 300
    x, A, d, s, lms = synthetic_spd_state(N, seed=9)
    fs = []
    for m, var in ((0, '0'), (defer_m, variant)):
        monkeypatch.setenv('EKF_DOWNDATE_VARIANT', var)
        fs.append(E.EKF_SLAM(capacity_landmarks=N + 3, storage=storage, defer_m=m))
    for f in fs:
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(4)
    for step in range(3):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 16, replace=False), rng)[:11]
        for f in fs:
            f.predict([0.2, 1.0])
            for r in rows:
                f.update(r[:2], np.diag([r[0] * 0.01, r[1] * 5.0]), int(r[2]))
    P0, P1 = fs[0].P, fs[1].P
    print('mma flush v%s m=%d %s: bit-identical=%s rel=%.2e' % (variant, defer_m, storage, np.array_equal(P0, P1), _rel(P1, P0)))
    assert _rel(P1, P0) < 1e-13 and _rel(fs[1].x, fs[0].x) < 1e-13
    for f in fs:
        f.close()


# ---------------------------------------------------------------- committed golden vectors (tests/golden/)
@pytest.mark.parametrize('storage,defer_m', [('full', 0), ('sym', 0), ('sym', 8), ('full', 4)])
@pytest.mark.parametrize('name', ['known_prefix_N20', 'known_quirk_N20', 'uc_nearest_N20', 'uc_nearest_N8_seed7'])
def test_gpu_reproduces_golden_vectors(name, storage, defer_m):
    """Same inputs as the committed fixtures: actions identical, x at every checkpoint and final x, P within 1e-9."""
    import ekf_slam_b200 as E
    from tests import golden_util as GU
    g = GU.load(name)
    f = getattr(E, GU.CASES[name])(capacity_landmarks=20, storage=storage, defer_m=defer_m)
    errs = []

    def cp(i, flt):
        xg = g['x_checkpoints'][i]
        xg = xg[~np.isnan(xg)]
        errs.append(GU.rel(flt.x, xg))

    acts = GU.replay(f, g, cp)
    assert np.array_equal(acts, g['actions'])
    assert len(errs) == len(g['checkpoint_steps']) and max(errs) <= 1e-9, errs
    assert GU.rel(f.x, g['x_final']) <= 1e-9 and GU.rel(f.P, g['P_final']) <= 1e-9
    assert np.array_equal(f.s, g['s_final'])
    f.close()


@pytest.mark.parametrize('storage,defer_m', [('full', 0), ('sym', 0), ('sym', 8), ('sym', 16)])
@pytest.mark.parametrize('name', ['known_prefix_N20', 'known_quirk_N20', 'uc_nearest_N20', 'uc_nearest_N8_seed7'])
def test_gpu_matches_the_reference_text_run(name, storage, defer_m):
    """tests/golden/refrun_*.npz: what the UNMODIFIED EKF_SLAM.m / EKF_SLAM_UC.m / Correspondence.m computed on these
    inputs, executed from their own source by oracle/mini_matlab.py (tests/golden/make_reference_run.py).  The CUDA path
    through the class API: landmark count after every step and s exactly, x at every checkpoint, final x and P <= 1e-9."""
    import os
    import ekf_slam_b200 as E
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_%s.npz' % name)))
    f = getattr(E, GU.CASES[name])(capacity_landmarks=20, storage=storage, defer_m=defer_m)
    cps = {int(s): i for i, s in enumerate(g['checkpoint_steps'])}
    for t, u, stub in GU.steps(g):
        f.predict(u)
        f.measure(None, u, stub)
        assert f.N == int(g['n_landmarks'][t]), (name, t)
        if (t + 1) in cps:
            want = g['x_checkpoints'][cps[t + 1]]
            want = want[~np.isnan(want)]
            assert GU.rel(f.x, want) <= 1e-9, (name, t)
    assert np.array_equal(f.s, g['s_final'].ravel())
    assert GU.rel(f.x, g['x_final']) <= 1e-9 and GU.rel(f.P, g['P_final']) <= 1e-9
    f.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_gpu_golden_association_cases(storage):
    import ekf_slam_b200 as E
    from tests import golden_util as GU
    g = GU.load('association_N12')
    N = len(g['s'])
    fa = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage)
    fa.set_state(g['x'], g['P'], g['s'])
    fb = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, assoc_mode='position')
    fb.correspondence.s_thresh = 9.21
    fb.set_state(g['x'], g['P'], np.ones(N))
    for q in range(len(g['z'])):
        z, R = g['z'][q], np.diag(g['Rdiag'][q])
        a = fa.correspondence.estimateCorrespondence(z, R)
        assert [int(a[0]), a[1]] == list(g['modeA'][q]), q
        zB = z.copy(); zB[2] = 1.0
        b = fb.correspondence.estimateCorrespondence(zB, R)
        assert [int(b[0]), b[1]] == list(g['modeB'][q]), q
    fa.close(); fb.close()


# ---------------------------------------------------------------- fused batch path (k_batch_lin + k_batch_gain)
@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('k,defer_m', [(2, 8), (8, 8), (9, 8), (11, 8), (7, 3), (5, 2)])
def test_fused_batch_is_bit_identical(storage, k, defer_m):
    """ekf_update_batch (2 launches per chunk) == k sequential ekf_update calls on the immediate path, bit for bit,
    including a landmark observed twice in one batch (duplicate slots)."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import synthetic_spd_state, observe_from_truth
    N = 300
    x, A, d, s, lms = synthetic_spd_state(N, seed=21)
    seq = E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=0)
    bat = E.EKF_SLAM(capacity_landmarks=N, storage=storage, defer_m=defer_m)
    for f in (seq, bat):
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(k * 10 + defer_m)
    for step in range(3):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 3 * k + 8, replace=False), rng)[:k]
        assert rows.shape[0] == k
        if k >= 3:
            rows[k - 1, 2] = rows[0, 2]                       # same landmark again within the batch
        Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
        for f in (seq, bat):
            f.predict([0.2, 1.0])
        for r, rd in zip(rows, Rd):
            seq.update(r[:2], rd, int(r[2]))
        bat.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
    assert np.array_equal(seq.x, bat.x), _rel(bat.x, seq.x)
    assert np.array_equal(seq.P, bat.P), _rel(bat.P, seq.P)
    with pytest.raises(E.EkfError):
        bat.update_batch(rows[:2, :2], Rd[:2], np.array([1, N + 1], dtype=np.int32))
    seq.close(); bat.close()


def test_measure_batches_associations_and_handles_appends_mid_batch():
    """EKF_SLAM_UC.measure with several new landmarks inside one observed_LL (association results computed ahead must
    be refreshed after every append) against the oracle, with deferral on."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=30, T=120, k=9, mode='nearest', seed=77)
    g = E.EKF_SLAM_UC(capacity_landmarks=30, storage='sym', defer_m=8)
    o = O.EKF_SLAM_UC()
    worst = _run_pair(g, o, w, check_every=40)
    multi = sum(1 for t in range(w.T) if sum(1 for a in o.chosen if a[0] == 'append') > 0)
    assert multi > 0 and g.N == (o.x.size - 3) // 2
    print('UC k=9 with mid-batch appends: worst rel err x %.2e P %.2e' % worst)
    g.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_covariance_ellipses_match_reference_plot_math(storage):
    """K6: the ellipses plot() would draw (EKF_SLAM.m:180-229) vs the literal eig-based restatement."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=20, T=150, k=5, mode='nearest')
    g = E.EKF_SLAM_UC(capacity_landmarks=20, storage=storage, defer_m=4)
    o = O.EKF_SLAM_UC()
    for t in range(w.T):
        g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t))
        o.predict(w.u[t]); o.measure(None, w.u[t], w.stub(t))
    got, want = g.covariance_ellipses(), O.ellipse_params(o.x, o.P)
    assert got.shape == want.shape == (g.N + 1, 5)
    assert np.allclose(got[:, :4], want[:, :4], rtol=1e-9, atol=0)
    d = np.abs(got[:, 4] - want[:, 4])
    d = np.minimum(d, np.abs(d - np.pi))                               # atan branch: +-pi/2 is the same axis
    assert d.max() < 1e-7, d.max()
    # exact diagonal blocks (fresh filter: P = 0.1*eye(3)): equal eigenvalues -> eig returns I, column 2 -> pi/2
    f = E.EKF_SLAM(capacity_landmarks=2, storage=storage)
    e0 = f.covariance_ellipses()
    assert e0.shape == (1, 5) and abs(e0[0, 4] - np.pi / 2) < 1e-15
    assert np.allclose(e0[0, 2:4], 0.25 * 2 * np.sqrt(2.2788 * 0.1))
    f.close(); g.close()


# ---------------------------------------------------------------- BASELINE's full size: 10 000 landmarks vs the structured C oracle
def _blocked_rel(Pg, Po, step=1024):
    """Relative Frobenius error of two 3.2 GB matrices without a third one."""
    num = den = 0.0
    for j in range(0, Po.shape[1], step):
        a, b = Pg[:, j:j + step], Po[:, j:j + step]
        num += float(np.sum((a - b) ** 2)); den += float(np.sum(b * b))
    return (num / den) ** 0.5


@pytest.mark.parametrize('storage,defer_m', [('sym', 8), ('full', 1), ('sym', 16)])
def test_config3_10k_landmarks_vs_structured_oracle(storage, defer_m):
    """BASELINE config 3 at full size (N = 10 000, n = 20 003, P 3.2 GB): the populated synthetic map of the bench,
    2 steps of predict + 8 associated updates through the class API (EKF_SLAM_UC.measure: batched association, fused
    deferred update), against oracle/structured.c stepping the same P on the host.  Association actions identical;
    x and P within 1e-9 relative (Frobenius)."""
    import ekf_slam_b200 as E
    from oracle import structured as S
    from bench import make_inputs
    from tests.golden_util import Stub
    from types import SimpleNamespace
    N, k, T = 10000, 8, 2
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, defer_m=defer_m)
    g.set_state_lowrank(x, s, A, d)
    o = S.EKF_SLAM_UC(capacity_landmarks=N)
    o.set_state(x, g.P, s)                       # the oracle starts from the very bits the device built
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    for t in range(T):
        stub = Stub(obs[t], table)
        g.predict(us[t]); g.measure(None, us[t], stub)
        o.predict(us[t]); o.measure(None, us[t], stub)
    assert g.chosen == o.chosen and len(g.chosen) == T * k and all(a == 'update' for a, _ in g.chosen)
    ex = np.linalg.norm(g.x - o.x) / np.linalg.norm(o.x)
    eP = _blocked_rel(g.P, o._P)
    print('10k %s/defer %d vs structured oracle: rel err x %.2e P %.2e' % (storage, defer_m, ex, eP))
    assert ex <= 1e-9 and eP <= 1e-9
    g.close()


@pytest.mark.parametrize('storage', ['sym', 'full'])
def test_config3_10k_sampled_rows_vs_matrix_free_oracle(storage):
    """Same configuration, 3 steps, against the matrix-free oracle (oracle/lowrank_lazy.py) that also checks the
    80 000-landmark slabs: here it additionally covers the device-side construction of the synthetic map."""
    import ekf_slam_b200 as E
    from oracle.lowrank_lazy import LazyLowRankFilter
    from bench import make_inputs
    N, k, T = 10000, 8, 3
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, defer_m=8)
    g.set_state_lowrank(x, s, A, d)
    o = LazyLowRankFilter(x, A, d, s, Rc=g.Rc)
    for t in range(T):
        g.predict(us[t]); o.predict(us[t])
        rows = obs[t]
        g.update_batch(rows[:, :2], np.column_stack([rows[:, 0] * g.Rc[0], rows[:, 1] * g.Rc[1]]), rows[:, 2].astype(np.int32))
        o.measure_rows(rows)
    n = 3 + 2 * N
    assert np.linalg.norm(g.x - o.x) / np.linalg.norm(o.x) <= 1e-9
    for i in [0, 1, 2, 3 + 2 * (int(obs[0][0][2]) - 1), 4711, n - 1]:
        got, want = g.P_block(i, 0, 1, n), o.block([i], 0, n)
        assert np.linalg.norm(got - want) / np.linalg.norm(want) <= 1e-9, i
    g.close()


@pytest.mark.parametrize('storage,defer_m', [('sym', 8), ('full', 0), ('sym', 3), ('sym', 16), ('full', 13)])
@pytest.mark.parametrize('variant', ['known', 'uc'])
def test_decoupled_state_chain_is_bit_identical(variant, storage, defer_m, monkeypatch):
    """EKF_ASYNC_X (default on): the state half of predict, the batched signature association and the x read-back run
    on a side stream, ordered by events against the covariance work.  Same kernels, same arithmetic: x read at every
    point of the loop, the actions and the final P must be bit-identical to the single-stream path (EKF_ASYNC_X=0),
    whatever the interleaving of predict / read x / measure / read x / append / double predict."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    cls = E.EKF_SLAM if variant == 'known' else E.EKF_SLAM_UC
    w = make_world(N=24, T=150, k=6, mode='prefix' if variant == 'known' else 'nearest', seed=11)
    runs = {}
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_ASYNC_X', flag)
        g = cls(capacity_landmarks=24, storage=storage, defer_m=defer_m)
        xs = []
        for t in range(w.T):
            g.predict(w.u[t])
            if t % 3 == 0:
                xs.append(g.x)                       # x after the motion step, before the covariance half has run
            if t % 7 == 3:
                g.predict(w.u[t])                    # two predicts in a row (xprev must not be overwritten early)
            g.measure(None, w.u[t], w.stub(t), pass_x=(t % 2 == 0))
            xs.append(g.x)                           # x after measure, possibly while the flush still streams P
            if t % 25 == 24:
                xs.append(g.covariance_ellipses())   # main-stream reader of x and P
                xs.append(g.P_block(0, 0, 3, 3))
        runs[flag] = (xs, list(g.chosen), g.P, g.s)
        g.close()
    a, b = runs['0'], runs['1']
    assert a[1] == b[1], 'actions differ'
    assert len(a[0]) == len(b[0]) and all(np.array_equal(p, q, equal_nan=True) for p, q in zip(a[0], b[0])), 'x differs somewhere in the loop'
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])


def test_decoupled_state_chain_large_map_overlaps_the_flush(monkeypatch):
    """Same check where it matters: 5 000 landmarks (P 0.8 GB), so that the flush of one measure() is still streaming
    P when x is read and the next predict / association are enqueued.  x after every predict and every measure, the
    actions and sampled rows of P must be bit-identical with and without the side stream."""
    import ekf_slam_b200 as E
    from bench import make_inputs
    from tests.golden_util import Stub
    from types import SimpleNamespace
    N, k, T = 5000, 8, 24
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    n = 3 + 2 * N
    runs = {}
    for flag in ('0', '1'):
        monkeypatch.setenv('EKF_ASYNC_X', flag)
        g = E.EKF_SLAM_UC(capacity_landmarks=N, storage='sym', defer_m=8)
        g.set_state_lowrank(x, s, A, d)
        xs = []
        for t in range(T):
            g.predict(us[t])
            if t % 5 == 4:
                g.predict(us[t])
            xs.append(g.x)
            g.measure(None, us[t], Stub(obs[t], table))
            xs.append(g.x)
        rows = [g.P_block(i, 0, 1, n) for i in (0, 2, 3 + 2 * (int(obs[0][0][2]) - 1), 4711, n - 1)]
        runs[flag] = (xs, list(g.chosen), rows)
        g.close()
    a, b = runs['0'], runs['1']
    assert a[1] == b[1] and len(a[1]) == T * k
    assert all(np.array_equal(p, q) for p, q in zip(a[0], b[0])), 'x differs somewhere in the loop'
    assert all(np.array_equal(p, q) for p, q in zip(a[2], b[2])), 'P differs'


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('defer_m', [16, 11])
@pytest.mark.parametrize('variant,mode,T', [('known', 'prefix', 1000), ('uc', 'nearest', 1000)])
def test_cross_step_deferral_1000_steps_vs_oracle(variant, mode, T, defer_m, storage):
    """defer_m = 9..16: rank-2 terms stay pending across measure() and predict() (predict transforms the pending slabs,
    K_q <- F K_q, B_q <- B_q F', instead of flushing) and one rank-<=32 pass over P applies up to 16 observations.  Same
    estimator, different rounding than the immediate path: the bar is the north star's (actions identical, x and P
    <= 1e-9 relative over 1000 steps, checked every 100), P read back mid-run included (that read flushes)."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=20, T=T, k=5, mode=mode)
    g = (E.EKF_SLAM if variant == 'known' else E.EKF_SLAM_UC)(capacity_landmarks=24, storage=storage, defer_m=defer_m)
    o = (O.EKF_SLAM if variant == 'known' else O.EKF_SLAM_UC)()
    worst = _run_pair(g, o, w)
    print('cross-step defer %d %s/%s: worst rel err x %.2e P %.2e' % (defer_m, variant, storage, worst[0], worst[1]))
    g.close()


@pytest.mark.parametrize('storage', ['full', 'sym'])
def test_cross_step_deferral_200_landmarks_ragged_batches(storage):
    """200 landmarks, batches of 1..9 observations per step (chunks of every size meet every pending count), explicit
    slots through update_batch and single updates mixed, against the dense-literal oracle."""
    import ekf_slam_b200 as E
    from bench import make_inputs
    from ekf_slam_b200.synth import lowrank_to_dense
    from oracle import dense_literal as O
    N, T = 200, 40
    x, A, d, s, lms, us, obs = make_inputs(N, 9, T)
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, defer_m=16)
    g.set_state_lowrank(x, s, A, d)
    o = O.EKF_SLAM_UC()
    o.x, o.P, o.s = x.copy(), lowrank_to_dense(A, d), list(s)
    for t in range(T):
        g.predict(us[t]); o.predict(us[t])
        rows = obs[t][:1 + (t * 5) % 9]
        Rd = np.column_stack([rows[:, 0] * g.Rc[0], rows[:, 1] * g.Rc[1]])
        if t % 4 == 3:
            for r, rd in zip(rows, Rd):
                g.update(r[:2], rd, int(r[2]))
        else:
            g.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
        for r, rd in zip(rows, Rd):
            o._update(r, np.diag(rd), int(r[2]))
        if t % 10 == 9:
            assert _rel(g.x, o.x) <= TOL and _rel(g.P, o.P) <= TOL, t
    assert _rel(g.x, o.x) <= TOL and _rel(g.P, o.P) <= TOL
    g.close()
