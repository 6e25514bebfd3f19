"""Monte-Carlo filter batch (ekfb_*, one CUDA block per filter, whole predict+measure on the device) vs the
dense-literal oracle run filter by filter, and vs the single-filter CUDA path (bit-identical: same primitives)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300))


def _worlds(nf, cls, N, T, k):
    from ekf_slam_b200.synth import make_world
    mode = 'prefix' if cls == 'EKF_SLAM' else 'nearest'
    return [make_world(N=N, T=T, k=k, mode=mode, seed=1000 + 17 * f) for f in range(nf)]


@pytest.mark.parametrize('storage', ['sym', 'full'])
@pytest.mark.parametrize('cls,defer_m,assoc', [('EKF_SLAM_UC', 8, 'signature'), ('EKF_SLAM_UC', 1, 'signature'),
                                                 ('EKF_SLAM', 8, 'signature'), ('EKF_SLAM_UC', 3, 'position')])
def test_batch_matches_oracle_per_filter(storage, cls, defer_m, assoc):
    import ekf_slam_b200 as E
    from oracle import dense_literal as O
    nf, N, T, k = 10, 14, 90, 5
    ws = _worlds(nf, cls, N, T, k)
    max_obs = max(len(o) for w in ws for o in w.observed)
    b = E.FilterBatch(cls, nf, capacity_landmarks=N, max_obs=max_obs, max_table=N, storage=storage, defer_m=defer_m,
                      assoc_mode=assoc)
    os_ = [getattr(O, cls)() for _ in range(nf)]
    if assoc == 'position':
        for o in os_:
            o.correspondence = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC', position_cost_in_score=True)
    acts = [[] for _ in range(nf)]
    for t in range(T):
        u = np.array([w.u[t] for w in ws])
        a = b.step(u, [w.observed[t] for w in ws], [list(w.table[:w.n_table[t]]) for w in ws], want_actions=True)
        for f, (w, o) in enumerate(zip(ws, os_)):
            o.predict(w.u[t]); o.measure(None, w.u[t], w.stub(t))
            acts[f] += [('update', int(v)) if v > 0 else ('append', int(-v)) for v in a[f, :len(w.observed[t])]]
    worst = [0.0, 0.0]
    for f, o in enumerate(os_):
        assert acts[f] == o.chosen, 'filter %d: actions differ' % f
        v = b[f]
        assert v.N == (o.x.size - 3) // 2 and np.array_equal(v.s, np.asarray(o.s, dtype=np.float64))
        ex, eP = _rel(v.x, o.x), _rel(v.P, o.P)
        worst = [max(worst[0], ex), max(worst[1], eP)]
        assert ex <= TOL and eP <= TOL, (f, ex, eP)
    print('batch %s/%s defer=%d %s: %d filters x %d steps, worst rel err x %.2e P %.2e' %
          (cls, storage, defer_m, assoc, nf, T, worst[0], worst[1]))
    b.close()


@pytest.mark.parametrize('impl', ['split-warp', 'split-warp-tc32', 'split-warp-tc128', 'split-tma', 'split-ldg', 'split-run', 'fused'])
@pytest.mark.parametrize('storage', ['sym', 'full'])
def test_batch_equals_single_filter_path_bitwise(storage, impl, monkeypatch):
    """All step implementations: split front + warp-per-column-group flush (default), the TMA-staged and the tiled
    register-path flush (EKF_MC_FLUSH=tma|ldg), the run-based front (EKF_MC_FRONT=run) and the one-kernel step
    (EKF_MC_FUSED=1); the warp flush also with 32 and 128 columns of a filter per CTA (EKF_MC_TC)."""
    import ekf_slam_b200 as E
    monkeypatch.setenv('EKF_MC_TC', {'split-warp-tc32': '32', 'split-warp-tc128': '128'}.get(impl, '64'))
    monkeypatch.setenv('EKF_MC_FUSED', '1' if impl == 'fused' else '0')
    monkeypatch.setenv('EKF_MC_FLUSH', {'split-ldg': 'ldg', 'split-tma': 'tma'}.get(impl, 'warp'))
    monkeypatch.setenv('EKF_MC_FRONT', 'run' if impl == 'split-run' else 'step')
    nf, N, T, k = 6, 12, 60, 4
    ws = _worlds(nf, 'EKF_SLAM_UC', N, T, k)
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=N, max_obs=k, max_table=N, storage=storage, defer_m=4)
    singles = [E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, defer_m=4) for _ in range(nf)]
    for t in range(T):
        b.step(np.array([w.u[t] for w in ws]), [w.observed[t] for w in ws], [list(w.table[:w.n_table[t]]) for w in ws])
        for w, g in zip(ws, singles):
            g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t))
    for f, g in enumerate(singles):
        assert np.array_equal(b[f].x, g.x), _rel(b[f].x, g.x)
        assert np.array_equal(b[f].P, g.P), _rel(b[f].P, g.P)
        g.close()
    b.close()


def test_batch_shared_table_ragged_counts_and_errors():
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=10, T=40, k=4, mode='nearest', seed=5)
    nf = 5
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=10, max_obs=4, max_table=10, storage='sym', defer_m=8)
    os_ = [O.EKF_SLAM_UC() for _ in range(nf)]
    for t in range(w.T):
        # filter f sees only the first (4 - f % 3) rows; filter 4 sees none on odd steps (empty observation list)
        obs = [w.observed[t][:max(0, 4 - f % 3)] if not (f == 4 and t % 2) else np.zeros((0, 3)) for f in range(nf)]
        b.step(np.tile(w.u[t], (nf, 1)), obs, list(w.table))          # one table object list -> shared by all filters
        for f, o in enumerate(os_):
            o.predict(w.u[t])
            from tests.golden_util import Stub
            o.measure(None, w.u[t], Stub(obs[f], w.table))
    for f, o in enumerate(os_):
        assert _rel(b[f].x, o.x) <= TOL and _rel(b[f].P, o.P) <= TOL
    b.close()
    # capacity: 3 landmarks offered to filters that can hold 2
    b = E.FilterBatch('EKF_SLAM_UC', 2, capacity_landmarks=2, max_obs=3, max_table=3)
    from types import SimpleNamespace
    tab = [SimpleNamespace(index=i + 1, loc=np.array([1.0 + i, 2.0])) for i in range(3)]
    obs = [np.array([[2.0, 30.0, 1], [3.0, 40.0, 2], [4.0, 50.0, 3]])] * 2
    b.step(np.zeros((2, 2)), obs, tab)
    with pytest.raises(E.EkfError) as e:
        b.sync()
    assert e.value.code == -2
    b.close()
    # unknown landmark index in the table
    b = E.FilterBatch('EKF_SLAM', 1, capacity_landmarks=4, max_obs=2, max_table=2)
    b.step(np.zeros((1, 2)), [np.array([[2.0, 30.0, 1], [3.0, 40.0, 7]])], [SimpleNamespace(index=1, loc=np.array([1.0, 2.0]))])
    with pytest.raises(E.EkfError) as e:
        b.sync()
    assert e.value.code == -1
    b.close()


@pytest.mark.parametrize('early', ['1', '0'])
@pytest.mark.parametrize('fused', ['0', '1'])
def test_poses_readback_early_and_late(early, fused, monkeypatch):
    """ekfb_get_poses: after a split step the poses come back on the side stream behind the front kernel (while the
    flush still runs); they must equal x(1:3) read filter by filter after a full sync, in every mode, also right after
    set_state, and a device-side error of the step must surface through the same call."""
    import ekf_slam_b200 as E
    from types import SimpleNamespace
    monkeypatch.setenv('EKF_MC_EARLY_POSES', early)
    monkeypatch.setenv('EKF_MC_FUSED', fused)
    nf, N, T, k = 7, 12, 25, 4
    ws = _worlds(nf, 'EKF_SLAM_UC', N, T, k)
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=N, max_obs=k, max_table=N, storage='sym', defer_m=4)
    assert np.array_equal(b.poses(), np.zeros((nf, 3)))
    for t in range(T):
        b.step(np.array([w.u[t] for w in ws]), [w.observed[t] for w in ws], [list(w.table[:w.n_table[t]]) for w in ws])
        p = b.poses()                                    # possibly before the flush of this step has finished
        if t % 6 == 0:
            b.sync()
            assert np.array_equal(p, np.array([b[f].x[:3] for f in range(nf)])), t
    p_last = b.poses()
    b.sync()
    assert np.array_equal(p_last, np.array([b[f].x[:3] for f in range(nf)]))
    # set_state invalidates the early path: the poses are those just written
    x0 = b[0].x.copy(); x0[:3] = [1.5, -2.5, 33.0]
    b.set_state(3, x0, b[0].P, b[0].s)
    assert np.array_equal(b.poses()[3], [1.5, -2.5, 33.0])
    b.close()
    # a capacity error raised by the front kernel is reported by poses() as well, and cleared
    b = E.FilterBatch('EKF_SLAM_UC', 2, capacity_landmarks=2, max_obs=3, max_table=3)
    tab = [SimpleNamespace(index=i + 1, loc=np.array([1.0 + i, 2.0])) for i in range(3)]
    obs = [np.array([[2.0, 30.0, 1], [3.0, 40.0, 2], [4.0, 50.0, 3]])] * 2
    b.step(np.zeros((2, 2)), obs, tab)
    with pytest.raises(E.EkfError) as e:
        b.poses()
    assert e.value.code == -2
    b.sync()                                             # error word was cleared
    b.close()


@pytest.mark.parametrize('storage', ['sym', 'full'])
def test_config5_full_size_4096_filters_x_200_landmarks(storage):
    """BASELINE config 5 at full size: 4096 independent EKF_SLAM_UC filters x 200 landmarks (n = 403), the populated
    synthetic map of the bench, 3 steps of predict + 5 associated updates per filter.  Sampled filters (first, last and
    three in between) against the dense-literal oracle stepping the same inputs: actions identical, x and P <= 1e-9
    relative.  Size-independent properties over ALL filters: the last filter is fed filter 0's inputs and must end
    bit-identical to it (no cross-filter interference anywhere in the batch), every filter keeps its 200 landmarks,
    every pose is finite, and poses() agrees with x(1:3) read filter by filter."""
    import ekf_slam_b200 as E
    from bench import make_mc_inputs
    from oracle import dense_literal as O
    from tests.golden_util import Stub
    from types import SimpleNamespace
    nf, N, k, T = 4096, 200, 5, 3
    x, P, s, lms, U, OBS, loc, idx = make_mc_inputs(N, k, nf, T)
    U[:, nf - 1] = U[:, 0]
    OBS[:, nf - 1] = OBS[:, 0]
    counts = np.full(nf, k, dtype=np.int32)
    ntab = np.array([N], dtype=np.int32)
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=N, max_obs=k, max_table=N, storage=storage, defer_m=8)
    b.set_state_all(x, P, s)
    acts = []
    for t in range(T):
        a = b.step(np.ascontiguousarray(U[t]), packed=(np.ascontiguousarray(OBS[t]), counts, loc, idx, ntab, True),
                   want_actions=True)
        acts.append(a.copy())
    poses = b.poses()
    b.sync()
    assert np.all(b.n_landmarks() == N) and np.all(np.isfinite(poses))
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    worst = [0.0, 0.0]
    for f in (0, 1, 1234, 2777, nf - 2):
        o = O.EKF_SLAM_UC()
        o.x, o.P, o.s = x.copy(), P.copy(), list(s)
        for t in range(T):
            o.predict(U[t, f]); o.measure(None, U[t, f], Stub(OBS[t, f].T.copy(), table))
        got = [('update', int(v)) for t in range(T) for v in acts[t][f, :k]]
        assert got == o.chosen, 'filter %d: actions differ' % f
        v = b[f]
        ex, eP = _rel(v.x, o.x), _rel(v.P, o.P)
        worst = [max(worst[0], ex), max(worst[1], eP)]
        assert ex <= TOL and eP <= TOL, (f, ex, eP)
        assert np.array_equal(poses[f], v.x[:3])
    first, last = b[0], b[nf - 1]
    assert np.array_equal(first.x, last.x) and np.array_equal(first.P, last.P)
    print('4096 x 200 %s: sampled filters vs oracle, worst rel err x %.2e P %.2e' % (storage, worst[0], worst[1]))
    b.close()


def test_config5_full_size_200_steps_sampled_filters():
    """BASELINE config 5 at full size over a long horizon: 4096 x EKF_SLAM_UC x 200 landmarks (SYM storage, the bench's
    configuration), 200 steps of predict + 5 associated updates per filter = 4.1 M updates; 8 sampled filters against
    oracle/structured.c stepping the same inputs (two of them also against the dense-literal oracle for the first 20
    steps): actions identical at every observation, x and P <= 1e-9 relative at steps 20, 50, 100, 150, 200.  The last
    filter is fed filter 0's inputs and must stay bit-identical to it throughout."""
    import ekf_slam_b200 as E
    from bench import make_mc_inputs
    from oracle import dense_literal as O
    from oracle import structured as S
    from tests.golden_util import Stub
    from types import SimpleNamespace
    nf, N, k, T = 4096, 200, 5, 200
    x, P, s, lms, U, OBS, loc, idx = make_mc_inputs(N, k, nf, T)
    U[:, nf - 1] = U[:, 0]
    OBS[:, nf - 1] = OBS[:, 0]
    counts = np.full(nf, k, dtype=np.int32)
    ntab = np.array([N], dtype=np.int32)
    b = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=N, max_obs=k, max_table=N, storage='sym', defer_m=8)
    b.set_state_all(x, P, s)
    table = [SimpleNamespace(index=i + 1, loc=lms[i]) for i in range(N)]
    sample = (0, 1, 511, 1234, 2048, 2777, 3333, nf - 2)
    os_, dense = {}, {}
    for f in sample:
        os_[f] = S.EKF_SLAM_UC(capacity_landmarks=N)
        os_[f].set_state(x, P, s)
    for f in (1, 2777):
        dense[f] = O.EKF_SLAM_UC()
        dense[f].x, dense[f].P, dense[f].s = x.copy(), P.copy(), list(s)
    worst = [0.0, 0.0]
    for t in range(T):
        a = b.step(np.ascontiguousarray(U[t]), packed=(np.ascontiguousarray(OBS[t]), counts, loc, idx, ntab, True),
                   want_actions=True)
        for f, o in os_.items():
            n0 = len(o.chosen)
            stub = Stub(OBS[t, f].T.copy(), table)
            o.predict(U[t, f]); o.measure(None, U[t, f], stub)
            assert [('update', int(v)) for v in a[f, :k]] == o.chosen[n0:], (t, f)
            if t < 20 and f in dense:
                dense[f].predict(U[t, f]); dense[f].measure(None, U[t, f], stub)
        if t + 1 in (20, 50, 100, 150, 200):
            for f, o in os_.items():
                v = b[f]
                ex, eP = _rel(v.x, o.x), _rel(v.P, o.P)
                worst = [max(worst[0], ex), max(worst[1], eP)]
                assert ex <= TOL and eP <= TOL, (t, f, ex, eP)
                if t + 1 == 20 and f in dense:
                    assert dense[f].chosen == o.chosen
                    assert _rel(v.x, dense[f].x) <= TOL and _rel(v.P, dense[f].P) <= TOL, (f, 'dense')
            first, last = b[0], b[nf - 1]
            assert np.array_equal(first.x, last.x) and np.array_equal(first.P, last.P)
    assert np.all(b.n_landmarks() == N) and np.all(np.isfinite(b.poses()))
    print('4096 x 200 x %d steps: 8 sampled filters vs oracle, worst rel err x %.2e P %.2e' % (T, worst[0], worst[1]))
    b.close()
