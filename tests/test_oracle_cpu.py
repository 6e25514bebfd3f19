"""CPU tests of the oracle itself (no GPU): golden-vector regression, MATLAB helper known answers, the reference's
quirks, and an independent 50-digit mpmath restatement of the update/predict/append algebra
that pins the oracle's floating-point path to the exact value of the source expressions."""
import os
import sys

import numpy as np
import pytest

from oracle import dense_literal as O
from oracle import matlab_compat as M
from tests import golden_util as G


# ---------------------------------------------------------------- golden vectors
@pytest.mark.parametrize('name', list(G.CASES))
def test_oracle_reproduces_golden(name):
    g = G.load(name)
    o = getattr(O, G.CASES[name])()
    seen = []

    def cp(i, f):
        xg = g['x_checkpoints'][i]
        live = ~np.isnan(xg)
        assert live.sum() == f.x.size
        seen.append(np.array_equal(xg[live], f.x))

    acts = G.replay(o, g, cp)
    assert np.array_equal(acts, g['actions'])
    assert all(seen) and len(seen) == len(g['checkpoint_steps'])
    assert np.array_equal(o.x, g['x_final']) and np.array_equal(o.P, g['P_final'])
    assert np.array_equal(np.asarray(o.s, dtype=np.float64), g['s_final'])


def test_golden_association_cases():
    g = G.load('association_N12')
    for q in range(len(g['z'])):
        z, R = g['z'][q], np.diag(g['Rdiag'][q])
        a = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC').estimateCorrespondence(z, R, g['x'], g['P'], g['s'])
        assert [int(a[0]), a[1]] == list(g['modeA'][q])
        zB = z.copy(); zB[2] = 1.0
        b = O.Correspondence(1e-11, 9.21, 'EKF_SLAM_UC', position_cost_in_score=True).estimateCorrespondence(
            zB, R, g['x'], g['P'], np.ones(len(g['s'])))
        assert [int(b[0]), b[1]] == list(g['modeB'][q])
    new = g['modeA'][:, 0] == 1
    assert new.sum() >= 8 and set(g['modeA'][new, 1]) == {13}     # unseen signature (incl. 5, which no landmark carries): default index N+1 (:40)
    assert 5 not in set(g['modeA'][:, 1])          # landmark 5 shares signature 3: the first minimum (3) wins (:81)


# ---------------------------------------------------------------- MATLAB helper semantics
def test_degree_trig_known_answers():
    for a, s, c in [(0, 0.0, 1.0), (90, 1.0, 0.0), (180, 0.0, -1.0), (270, -1.0, 0.0), (360, 0.0, 1.0), (-90, -1.0, 0.0),
                    (450, 1.0, 0.0)]:
        assert M.sind(a) == s and M.cosd(a) == c, a          # exact at multiples of 90 like MATLAB's sind/cosd
    assert abs(M.sind(30) - 0.5) < 1e-16 and abs(M.cosd(60) - 0.5) < 1.2e-16
    assert M.atan2d(1.0, 1.0) == 45.0 and M.atan2d(0.0, -1.0) == 180.0 and M.atan2d(-1.0, 0.0) == -90.0


def test_wrapTo360_and_mod_semantics():
    assert M.wrapTo360(360.0) == 360.0 and M.wrapTo360(0.0) == 0.0 and M.wrapTo360(-10.0) == 350.0
    assert M.wrapTo360(720.0) == 360.0 and M.wrapTo360(370.5) == 10.5 and M.wrapTo360(-360.0) == 0.0
    assert M.mod(-1.0, 360.0) == 359.0 and M.mod(725.0, 360.0) == 5.0   # sign follows the divisor


def test_inv2_is_an_inverse():
    A = np.array([[4.0, 1.0], [2.0, 3.0]])
    assert np.allclose(M.inv2(A) @ A, np.eye(2), atol=1e-15)


# ---------------------------------------------------------------- quirks the oracle must keep
def _stub(rows, table):
    from types import SimpleNamespace
    return G.Stub(np.array(rows, dtype=np.float64).reshape(-1, 3),
                  [SimpleNamespace(index=i, loc=np.array(l, dtype=np.float64)) for i, l in table])


def test_constructor_defaults():
    a, b = O.EKF_SLAM(), O.EKF_SLAM_UC()
    assert np.array_equal(a.x, [0, 0, 0]) and np.array_equal(a.P, 0.1 * np.eye(3))            # EKF_SLAM.m:28-31
    assert a.C == 0.2 and list(a.Rc) == [0.01, 5] and list(b.Rc) == [0.1, 5]                   # :12-13 / _UC:13
    assert b.correspondence.s_cost == 1e-11 and b.correspondence.s_thresh == 1e9               # _UC:16


def test_predict_uses_premotion_heading_and_leaves_landmarks():
    o = O.EKF_SLAM()
    o.measure(None, [1.0, 10.0], _stub([[2.0, 30.0, 1]], [(1, (2.0, 1.0))]))     # first sighting -> append (:110-111)
    o.measure(None, [1.0, 10.0], _stub([[2.0, 30.0, 1], [3.0, 60.0, 2]], [(1, (2.0, 1.0)), (2, (1.0, 3.0))]))
    assert o.chosen == [('append', 1), ('update', 1), ('append', 2)]
    P0, x0 = o.P.copy(), o.x.copy()
    u = [0.5, 20.0]
    o.predict(u)
    assert np.array_equal(o.P[3:, 3:], P0[3:, 3:]) and np.array_equal(o.x[3:], x0[3:])       # landmark block untouched
    th = x0[2]
    a, b = -u[0] * M.sind(th), u[0] * M.cosd(th)                                  # F(1,3), F(2,3) at the OLD heading
    F = np.eye(7); F[0, 2] = a; F[1, 2] = b
    W = np.array([u[0] * M.cosd(th), u[0] * M.sind(th), u[1]])
    Q = np.zeros((7, 7)); Q[:3, :3] = np.outer(W * 0.2, W)
    assert np.allclose(o.P, F @ P0 @ F.T + Q, rtol=1e-15, atol=1e-18)
    assert o.x[2] == M.wrapTo360(th + u[1]) and o.x[0] == x0[0] + u[0] * M.cosd(th + u[1])


def test_known_correspondence_updates_slot_ii_not_z3():
    """EKF_SLAM.m:123 idx = ii: the row number picks the slot, z(3) only decides append (F8)."""
    o = O.EKF_SLAM()
    tab = [(1, (2.0, 1.0)), (2, (1.0, 3.0))]
    o.measure(None, [1.0, 5.0], _stub([[2.0, 30.0, 1]], tab))
    o.measure(None, [1.0, 5.0], _stub([[9.9, 99.0, 7], [3.0, 70.0, 2]], tab[:1] + [(7, (5.0, 5.0))] + tab[1:]))
    assert o.chosen[1] == ('append', 7)            # z(3)=7 > N=1 -> append with signature 7
    o.measure(None, [1.0, 5.0], _stub([[3.0, 70.0, 2]], tab))
    assert o.chosen[-1] == ('update', 1)           # observation of landmark "2" in row 1 updates slot 1


def test_first_landmark_branch_uses_signature_one():
    o = O.EKF_SLAM_UC()
    o.measure(None, [0.3, 1.0], _stub([[4.0, 100.0, 9]], [(9, (1.5, -2.5))]))
    assert o.chosen == [('append', 1)] and list(o.s) == [1] and np.array_equal(o.x[3:], [1.5, -2.5])   # :112-113, F15


def test_association_scores_signature_only_as_shipped():
    """Correspondence.m:75: the Mahalanobis cost is computed and discarded; an absurd range still associates (F9)."""
    c = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC')
    x = np.array([0, 0, 0, 5.0, 0.0, 0.0, 5.0])
    P = 0.1 * np.eye(7)
    s = np.array([1.0, 2.0])
    assert c.estimateCorrespondence(np.array([500.0, 10.0, 2.0]), np.eye(2), x, P, s) == (False, 2)
    assert c.estimateCorrespondence(np.array([5.0, 0.0, 2.1]), np.eye(2), x, P, s) == (True, 3)     # (0.1)^2/1e-11 = 1e9 is NOT <= 1e9? it is; 0.1**2 rounds above -> gate fails
    assert O.Correspondence(1, 1, 'bogus').method == 'ML'                                            # :19-23


def test_innovation_is_not_wrapped():
    """EKF_SLAM.m:144: z - z_k across the 0/360 seam is ~-360 deg, and the oracle keeps it (F10)."""
    o = O.EKF_SLAM()
    o.measure(None, [0.0, 0.0], _stub([[5.0, 1.0, 1]], [(1, (5.0, 0.05))]))
    x0, P0 = o.x.copy(), o.P.copy()
    o.measure(None, [0.0, 0.0], _stub([[5.0, 359.9, 1]], [(1, (5.0, 0.05))]))    # predicted bearing ~0.57 deg
    d = x0[3:5] - x0[:2]
    q = d @ d
    zk = np.array([np.sqrt(q), M.wrapTo360(M.atan2d(d[1], d[0]) - x0[2])])
    H = (1 / q) * np.array([[-np.sqrt(q) * d[0], -np.sqrt(q) * d[1], 0, np.sqrt(q) * d[0], np.sqrt(q) * d[1]],
                            [d[1], -d[0], -q, -d[1], d[0]]])
    K = P0 @ H.T @ np.linalg.inv(H @ P0 @ H.T + np.diag([5.0 * 0.01, 359.9 * 5]))
    nu = np.array([5.0, 359.9]) - zk
    assert nu[1] > 359.0                                            # not wrapped to -0.67 deg
    assert np.allclose(o.x, x0 + K @ nu, rtol=1e-12, atol=1e-15)
    nu_wrapped = nu - np.array([0.0, 360.0])
    assert not np.allclose(o.x, x0 + K @ nu_wrapped, rtol=1e-3)


# ---------------------------------------------------------------- independent exact-arithmetic restatement
def test_oracle_matches_50_digit_restatement():
    """One predict + append + update sequence redone in mpmath (50 digits) straight from the reference expressions
    (EKF_SLAM.m:42-47, :84-97, :125-145); the float64 oracle must agree to rounding (<= 1e-12 relative)."""
    mp = pytest.importorskip('mpmath')
    mp.mp.dps = 50
    deg = mp.pi / 180
    sind = lambda a: mp.sin(mp.mpf(a) * deg)
    cosd = lambda a: mp.cos(mp.mpf(a) * deg)

    o = O.EKF_SLAM()
    x = mp.matrix([0, 0, 0]).T
    P = mp.eye(3) * mp.mpf('0.1000000000000000055511151231257827')     # the double nearest 0.1, as the oracle holds it
    seq = [('predict', [0.7, 12.0]), ('append', [0.7, 12.0], [0.04, 150.0], [3.0, 1.0]),
           ('predict', [0.5, -7.0]), ('append', [0.5, -7.0], [0.02, 300.0], [-1.0, 4.0]),
           ('update', [3.1, 14.0], [0.031, 70.0], 1), ('update', [4.2, 110.0], [0.042, 550.0], 2)]
    for op in seq:
        n = x.cols
        if op[0] == 'predict':
            u = [mp.mpf(v) for v in op[1]]
            th = x[0, 2]
            W = mp.matrix([u[0] * cosd(th), u[0] * sind(th), u[1]])
            Q = mp.zeros(n, n)
            Q[0:3, 0:3] = W * mp.mpf('0.2000000000000000111022302462515654') * W.T
            F = mp.eye(n)
            F[0, 2] = -u[0] * sind(th)
            F[1, 2] = u[0] * cosd(th)
            x[0, 0] += u[0] * cosd(th + u[1]); x[0, 1] += u[0] * sind(th + u[1]); x[0, 2] = th + u[1]
            P = F * P * F.T + Q
            o.predict(op[1])
        elif op[0] == 'append':
            u = [mp.mpf(v) for v in op[1]]
            R = mp.diag([mp.mpf(v) for v in op[2]])
            th = x[0, 2]
            jxr = mp.matrix([[1, 0, -u[0] * sind(th)], [0, 1, u[0] * cosd(th)]])
            jz = mp.matrix([[cosd(u[1]), -u[0] * sind(u[1])], [sind(u[1]), u[0] * cosd(u[1])]])
            Pn = mp.zeros(n + 2, n + 2)
            Pn[0:n, 0:n] = P
            Pn[n:n + 2, n:n + 2] = jxr * P[0:3, 0:3] * jxr.T + jz * R * jz.T
            Pn[0:3, n:n + 2] = P[0:3, 0:3] * jxr.T
            Pn[n:n + 2, 0:3] = Pn[0:3, n:n + 2].T
            for a in range(3, n, 2):
                Pn[n:n + 2, a:a + 2] = jxr * P[a:a + 2, 0:3].T
                Pn[a:a + 2, n:n + 2] = Pn[n:n + 2, a:a + 2].T
            P = Pn
            xn = mp.zeros(1, n + 2)
            xn[0, 0:n] = x
            xn[0, n] = mp.mpf(op[3][0]); xn[0, n + 1] = mp.mpf(op[3][1])
            x = xn
            o.append(op[1], np.diag(op[2]), op[3], (n - 3) // 2 + 1)
        else:
            z = [mp.mpf(v) for v in op[1]]
            R = mp.diag([mp.mpf(v) for v in op[2]])
            a = 3 + 2 * (op[3] - 1)
            d = mp.matrix([x[0, a] - x[0, 0], x[0, a + 1] - x[0, 1]])
            q = d[0] ** 2 + d[1] ** 2
            brg = mp.atan2(d[1], d[0]) / deg - x[0, 2]
            brg = brg - 360 * mp.floor(brg / 360)
            zk = mp.matrix([mp.sqrt(q), brg])
            Fk = mp.zeros(5, n)
            for i in range(3):
                Fk[i, i] = 1
            Fk[3, a] = 1; Fk[4, a + 1] = 1
            sq = mp.sqrt(q)
            H = (1 / q) * mp.matrix([[-sq * d[0], -sq * d[1], 0, sq * d[0], sq * d[1]],
                                     [d[1], -d[0], -q, -d[1], d[0]]]) * Fk
            S = H * P * H.T + R
            K = P * H.T * S ** -1
            x = x + (K * (mp.matrix(z) - zk)).T
            P = (mp.eye(n) - K * H) * P
            o._update(np.array(op[1] + [op[3]], dtype=np.float64), np.diag(op[2]), op[3])
    xe = np.array([float(v) for v in x], dtype=np.float64)
    Pe = np.array([[float(P[i, j]) for j in range(P.cols)] for i in range(P.rows)])
    assert G.rel(o.x, xe) <= 1e-12 and G.rel(o.P, Pe) <= 1e-12, (G.rel(o.x, xe), G.rel(o.P, Pe))


def test_ellipse_params_known_answers():
    """EKF_SLAM.m:180-229: a diagonal block draws an axis-aligned ellipse, extents 2*sqrt(2.2788*lambda)*scale."""
    x = np.array([1.0, 2.0, 30.0, 5.0, 6.0])
    P = np.diag([0.04, 0.01, 0.1, 0.09, 0.25])
    e = O.ellipse_params(x, P)
    assert e.shape == (2, 5) and np.array_equal(e[:, :2], [[1.0, 2.0], [5.0, 6.0]])
    assert np.allclose(e[0, 2:4], [0.25 * 2 * np.sqrt(2.2788 * 0.04), 0.25 * 2 * np.sqrt(2.2788 * 0.01)]) and e[0, 4] == 0.0
    assert np.allclose(e[1, 2:4], [0.5 * 2 * np.sqrt(2.2788 * 0.25), 0.5 * 2 * np.sqrt(2.2788 * 0.09)])
    assert abs(abs(e[1, 4]) - np.pi / 2) < 1e-15                       # larger variance along y: atan(1/0)
    P2 = P.copy(); P2[3, 4] = P2[4, 3] = 0.05
    a = O.ellipse_params(x, P2)[1, 4]
    lam, vec = np.linalg.eigh(P2[3:5, 3:5])
    assert abs(np.tan(a) - vec[1, 1] / vec[0, 1]) < 1e-12


# ---------------------------------------------------------------- structured C oracle (oracle/structured.c)
@pytest.mark.parametrize('name', list(G.CASES))
def test_structured_oracle_matches_golden(name):
    """The O(n^2) C oracle against the committed fixtures (made by the dense oracle): identical actions and
    signatures, x at every checkpoint and final x, P within 1e-12 (different grouping of the same products)."""
    from oracle import structured as S
    g = G.load(name)
    o = getattr(S, G.CASES[name])(capacity_landmarks=24)
    errs = []

    def cp(i, f):
        xg = g['x_checkpoints'][i]
        errs.append(G.rel(f.x, xg[~np.isnan(xg)]))

    acts = G.replay(o, g, cp)
    assert np.array_equal(acts, g['actions'])
    assert len(errs) == len(g['checkpoint_steps']) and max(errs) <= 1e-12, errs
    assert G.rel(o.x, g['x_final']) <= 1e-12 and G.rel(o.P, g['P_final']) <= 1e-12
    assert np.array_equal(o.s, g['s_final'])


def test_structured_oracle_association_cases():
    from oracle import structured as S
    g = G.load('association_N12')
    N = len(g['s'])
    for q in range(len(g['z'])):
        a = S.EKF_SLAM_UC(capacity_landmarks=N)
        a.set_state(g['x'], g['P'], g['s'])
        new, idx = a.associate(g['z'][q], g['Rdiag'][q])
        assert [int(new), idx] == list(g['modeA'][q])
        b = S.EKF_SLAM_UC(capacity_landmarks=N, assoc_mode=1)
        b.set_state(g['x'], g['P'], np.ones(N))
        b.s_thresh = 9.21
        zB = g['z'][q].copy(); zB[2] = 1.0
        new, idx = b.associate(zB, g['Rdiag'][q])
        assert [int(new), idx] == list(g['modeB'][q])


def test_structured_helpers_equal_numpy_restatement():
    """The C helpers and oracle/matlab_compat.py state the same MathWorks semantics: equal to the last bit (atan2d: to
    one ulp, glibc vs NumPy) on a grid that includes the exact cases (multiples of 90, +-360, 0)."""
    from oracle import structured as S
    L = S.lib()
    xs = np.concatenate([np.arange(-720.0, 721.0, 15.0), np.random.default_rng(3).uniform(-1e4, 1e4, 500)])
    for v in xs:
        assert L.so_sind(v) == M.sind(v) and L.so_cosd(v) == M.cosd(v), v
        assert L.so_wrapTo360(v) == M.wrapTo360(v), v
    for y, x in np.random.default_rng(4).normal(size=(200, 2)):
        assert abs(L.so_atan2d(y, x) - M.atan2d(y, x)) <= 2 * np.spacing(abs(M.atan2d(y, x)))   # libm vs NumPy atan2: last ulp of the radian value
    assert L.so_wrapTo360(360.0) == 360.0 and L.so_wrapTo360(0.0) == 0.0 and L.so_wrapTo360(-90.0) == 270.0


def test_structured_oracle_200_landmarks_vs_dense():
    """A longer map (N = 200 fills through append, 120 steps): the two oracles stay within 1e-12."""
    from oracle import structured as S
    from ekf_slam_b200.synth import make_world
    w = make_world(N=60, T=120, k=5, mode='nearest')
    a, b = S.EKF_SLAM_UC(capacity_landmarks=64), O.EKF_SLAM_UC()
    for t in range(w.T):
        a.predict(w.u[t]); a.measure(None, w.u[t], w.stub(t))
        b.predict(w.u[t]); b.measure(None, w.u[t], w.stub(t))
    assert a.chosen == b.chosen and a.N == (len(b.x) - 3) // 2
    assert G.rel(a.x, b.x) <= 1e-12 and G.rel(a.P, b.P) <= 1e-12


def test_structured_oracle_errors():
    from oracle import structured as S
    a = S.EKF_SLAM(capacity_landmarks=1)
    with pytest.raises(IndexError):
        a.update([1.0, 2.0], [0.1, 0.1], 1)                 # no landmark yet
    a.append([1.0, 0.0], [0.1, 0.1], [3.0, 4.0], 1)
    with pytest.raises(MemoryError):
        a.append([1.0, 0.0], [0.1, 0.1], [5.0, 6.0], 2)     # capacity


# ---------------------------------------------------------------- matrix-free oracle (oracle/lowrank_lazy.py)
def test_matrix_free_oracle_matches_structured_on_a_materialised_map():
    """The matrix-free oracle used for the 80 000-landmark rows, against oracle/structured.c on the same populated
    synthetic map small enough to materialise (N = 300): 3 steps of predict + 8 associated updates."""
    from oracle import structured as S
    from oracle.lowrank_lazy import LazyLowRankFilter
    from ekf_slam_b200.synth import lowrank_to_dense
    from bench import make_inputs
    N, k, T = 300, 8, 3
    x, A, d, s, lms, us, obs = make_inputs(N, k, T)
    a = S.EKF_SLAM_UC(capacity_landmarks=N)
    a.set_state(x, lowrank_to_dense(A, d), s)
    b = LazyLowRankFilter(x, A, d, s)
    for t in range(T):
        a.predict(us[t]); b.predict(us[t])
        acts = b.measure_rows(obs[t])
        for row, (_, idx) in zip(obs[t], acts):
            a.update(row[:2], [row[0] * a.Rc[0], row[1] * a.Rc[1]], idx)
        assert [i for _, i in acts] == [int(r[2]) for r in obs[t]]          # signature association finds the true slot
    assert G.rel(b.x, a.x) <= 1e-12
    P = a.P
    rows = [0, 1, 2, 3, 77, 300, 2 * N + 2]
    assert G.rel(np.vstack([b.row(i) for i in rows]), P[rows, :]) <= 1e-11
    assert G.rel(np.column_stack([b.col(j) for j in rows]), P[:, rows]) <= 1e-11
    assert G.rel(b.block(rows, 100, 140), P[np.ix_(rows, range(100, 140))]) <= 1e-11


def test_cross_step_deferral_identity_on_the_oracle():
    """The algebra behind defer_m = 9..16 (DESIGN.md "Cross-step deferral"), checked on the dense oracle's own matrices:
    with P = P0 - sum_q K_q B_q pending, predict gives F P F' + Q = (F P0 F' + Q) - sum_q (F K_q)(B_q F'), and F only
    changes rows 0, 1 of K_q and columns 0, 1 of B_q (EKF_SLAM.m:47, :62-64)."""
    from oracle import dense_literal as O
    from oracle.matlab_compat import sind, cosd
    rng = np.random.default_rng(3)
    n, m = 23, 5
    A = rng.standard_normal((n, n))
    P0 = A @ A.T / n + np.eye(n)
    Ks = [rng.standard_normal((n, 2)) * 0.1 for _ in range(m)]
    Bs = [rng.standard_normal((2, n)) * 0.1 for _ in range(m)]
    P = P0 - sum(K @ B for K, B in zip(Ks, Bs))
    o = O.EKF_SLAM()
    o.x = np.concatenate([[0.3, -0.2, 37.0], rng.standard_normal(n - 3)])
    o.P = P.copy()
    o.s = list(range(1, (n - 3) // 2 + 1))
    u = np.array([0.4, 2.0])
    th = o.x[2]
    o.predict(u)                                                   # the reference's dense F P F' + Q
    a, b = -1 * u[0] * sind(th), u[0] * cosd(th)
    F = np.eye(n); F[0, 2] = a; F[1, 2] = b
    W = np.array([[u[0] * cosd(th)], [u[0] * sind(th)], [u[1]]])
    Q = np.zeros((n, n)); Q[:3, :3] = (W * 0.2) @ W.T
    Kt, Bt = [], []
    for K, B in zip(Ks, Bs):
        K2, B2 = K.copy(), B.copy()
        K2[0] += a * K[2]; K2[1] += b * K[2]                       # K_q <- F K_q
        B2[:, 0] += a * B[:, 2]; B2[:, 1] += b * B[:, 2]           # B_q <- B_q F'
        assert np.allclose(K2, F @ K, rtol=0, atol=1e-15) and np.allclose(B2, B @ F.T, rtol=0, atol=1e-15)
        Kt.append(K2); Bt.append(B2)
    P_deferred = (F @ P0 @ F.T + Q) - sum(K @ B for K, B in zip(Kt, Bt))
    assert np.linalg.norm(P_deferred - o.P) / np.linalg.norm(o.P) < 1e-14


# ---------------------------------------------------------------- round 2: shared atan2, f(x,u)
def test_shared_atan2_accuracy_and_conventions():
    """ekf_math_shared.h (the atan2 both the CUDA library and oracle/structured.c evaluate, so that mode-B association
    costs agree bit for bit): <= 1.5 ulp against a 50-digit reference across the reduction breakpoints and all four
    quadrants, C99 conventions at zeros / infinities / NaN."""
    import math
    import mpmath as mp
    from oracle import structured as S
    L = S.lib()
    mp.mp.dps = 50
    rng = np.random.default_rng(12)
    cases = [tuple(rng.normal(size=2) * 10.0 ** rng.integers(-3, 4, 2)) for _ in range(3000)]
    for a in np.linspace(0.01, 3.0, 700):
        cases += [(a, 1.0), (-a, -1.0), (a, -1.0), (-a, 1.0)]
    worst = 0.0
    for y, x in cases:
        got = L.so_atan2d(float(y), float(x))
        ref = mp.atan2(mp.mpf(float(y)), mp.mpf(float(x))) * 180 / mp.pi
        worst = max(worst, float(abs(mp.mpf(got) - ref)) / math.ulp(float(ref)))
    # 1.3 ulp in radians; the degree conversion (x 180/pi, itself a rounded constant) may move the result up a binade
    # (ulp halves relative to the value) and rounds once more: <= 3.5 ulp of the result in degrees
    assert worst <= 3.5, worst
    d = 180.0 / math.pi
    for y, x in [(0.0, 1.0), (-0.0, 1.0), (0.0, -1.0), (-0.0, -1.0), (1.0, 0.0), (-1.0, 0.0), (1.0, -0.0),
                 (math.inf, 1.0), (1.0, math.inf), (1.0, -math.inf), (-1.0, -math.inf), (math.inf, math.inf),
                 (math.inf, -math.inf), (-math.inf, -math.inf), (1e-310, 1.0), (1e300, 1e-300), (1e-300, -1e300)]:
        g, r = L.so_atan2d(y, x), math.atan2(y, x) * d
        assert g == r and math.copysign(1.0, g) == math.copysign(1.0, r), (y, x, g, r)
    assert math.isnan(L.so_atan2d(math.nan, 1.0))


def test_f_motion_model_matches_the_oracle():
    """[x_new,F] = f(h,x,u), EKF_SLAM.m:56-65: the host mirror's public method (pure, needs no device) against the
    dense-literal oracle's, incl. multiples of 90 degrees where sind/cosd are exact, and the Jacobian taken at the
    heading BEFORE the motion (:63-64 use x(3), not x(3)+u(2))."""
    from ekf_slam_b200.filters import _FilterBase
    o = O.EKF_SLAM()
    rng = np.random.default_rng(3)
    for th in [0.0, 90.0, 180.0, 270.0, 360.0, 45.0, 359.5, -30.0, 725.0] + list(rng.uniform(-400, 400, 8)):
        N = int(rng.integers(0, 5))
        x = np.concatenate([[rng.normal(), rng.normal(), th], rng.normal(size=2 * N)])
        u = np.array([rng.uniform(0.0, 2.0), rng.uniform(-20, 20)])
        xg, Fg = _FilterBase.f(None, x, u)
        xo, Fo = o.f(x.copy(), u)
        assert np.array_equal(xg, np.asarray(xo).reshape(-1)) and np.array_equal(Fg, Fo), th
        assert xg[2] == th + u[1]                                      # f does not wrap (predict does, :50)
        assert np.array_equal(xg[3:], x[3:]) and Fg.shape == (3 + 2 * N, 3 + 2 * N)
    x = np.array([0.0, 0.0, 90.0])
    _, F = _FilterBase.f(None, x, [2.0, 45.0])
    assert F[0, 2] == -2.0 and F[1, 2] == 0.0                          # sind(90) = 1, cosd(90) = 0 exactly


def test_reference_witness_roundtrip_and_pin_if_present(tmp_path):
    """matlab/run_reference_fixture.m runs the UNMODIFIED reference on tests/golden/ref_inputs and writes CSVs;
    tests/golden/compare_reference_fixture.py turns them into the pin of the oracle.  (a) The exported inputs are the
    inputs of the reference-text run (refrun_*.npz: a first frame the reference can run) bit for bit; (b) the comparer accepts the oracle's own outputs written in the script's format (so a
    format slip cannot fake a mismatch); (c) if somebody has run the script under MATLAB/Octave and dropped the outputs
    into tests/golden/ref_outputs, they must match the oracle to 1e-9 -- that is the reference pin."""
    import os
    from tests.golden import compare_reference_fixture as C
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
    for name in G.CASES:
        g = dict(np.load(os.path.join(here, 'refrun_%s.npz' % name)))       # the in-domain worlds (make_reference_run.py)
        assert int(g['obs_counts'][0]) == 1 and int(g['n_table'][0]) == 1   # a first frame the reference can run
        d = os.path.join(here, 'ref_inputs', name)
        assert open(os.path.join(d, 'class.txt')).read().strip() == G.CASES[name]
        for key in ('u', 'obs_rows', 'table_loc', 'obs_counts', 'n_table', 'table_index', 'checkpoint_steps'):
            a = np.genfromtxt(os.path.join(d, key + '.csv'), delimiter=',').reshape(np.asarray(g[key]).shape)
            assert np.array_equal(a, g[key]), (name, key)
        out = tmp_path / name
        out.mkdir()
        wr = lambda fn, M: np.savetxt(str(out / fn), np.atleast_2d(M), delimiter=',', fmt='%.17g')
        wr('x_final.csv', g['x_final']); wr('P_final.csv', g['P_final']); wr('s_final.csv', g['s_final'])
        wr('x_checkpoints.csv', g['x_checkpoints'])
        wr('n_landmarks.csv', g['n_landmarks'])
        r = C.compare_case(name, str(tmp_path))
        assert r['pinned'] and r['x_final'] == 0.0 and r['P_final'] == 0.0, r
    real = os.path.join(here, 'ref_outputs')
    if os.path.isdir(real):
        for name in G.CASES:
            if os.path.isdir(os.path.join(real, name)):
                r = C.compare_case(name, real)
                assert r['pinned'], (name, r)


# ---------------------------------------------------------------- round 2: the continuous-flush schedule, modelled in NumPy
def test_continuous_flush_schedule_on_the_dense_oracle():
    """The bookkeeping behind the continuous flush, modelled in NumPy and checked against the
    reference's dense algebra (oracle/dense_literal.py) over a loop of predicts and updates:

      * the stored matrix `base` is only ever replaced by the output of a pass, base' = base - sum_{q in set} K_q B_q, computed
        from the terms AS THEY WERE WHEN THE SET WAS CLOSED and finishing an arbitrary number of steps later;
      * predict never touches `base`: it works on the robot strip (rows 0..2) and on the window of pending terms
        (K_q <- F K_q, B_q <- B_q F');
      * the current covariance every linearisation sees is  (base with rows / columns 0..2 taken from the strip)
        - sum over the whole window (sets in flight + the set that is filling up);
      * when a pass retires, the strip drops that set's terms (in their CURRENT, transformed form) and the pass's output
        becomes `base`; rows 0..2 of that output are never read.

    Only the upper triangle (i <= j) of anything is ever used, as in SYM storage."""
    from oracle import dense_literal as O
    from oracle.matlab_compat import sind, cosd, atan2d, wrapTo360, inv2
    rng = np.random.default_rng(11)
    N, T, k, cap = 9, 40, 3, 6                      # sets of <= 6 terms (the library: 16), 3 observations per step
    n = 3 + 2 * N
    A = rng.standard_normal((n, n))
    P0 = A @ A.T / n * 0.05 + np.diag(np.concatenate([[0.1, 0.1, 0.05], np.full(2 * N, 0.3)]))
    lms = rng.uniform(-20, 20, (N, 2))
    x0 = np.concatenate([[0.5, -0.3, 30.0], (lms + rng.normal(0, 0.2, lms.shape)).ravel()])
    o = O.EKF_SLAM()
    o.x, o.P, o.s = x0.copy(), P0.copy(), list(range(1, N + 1))
    C, Rc = o.C, o.Rc

    def upper(M):                                   # what SYM storage keeps
        return np.triu(M) + np.triu(M, 1).T

    # ---- the model's state
    x = x0.copy()
    base = P0.copy()
    strip = P0[:3, :].copy()                        # strip[r, j] = base(r, j), r < 3
    window = []                                     # [K (n x 2), B (2 x n)] of every pending term, in order, current frame
    passes = []                                     # in flight: (number of terms, output matrix computed at close time, steps left)
    n_closed = 0                                    # terms of the window that belong to passes in flight (a prefix)

    def current_P():
        M = base.copy()
        M[:3, :] = strip
        M = upper(M)                                # column j < 3 below row 2 comes from the strip by symmetry
        for K, B in window:
            M = M - upper(np.triu(K @ B))           # element (i, j), i <= j: K_q(i,:) B_q(:,j)
        return M

    def retire():
        nonlocal base, strip, window, n_closed
        m, out, _ = passes.pop(0)
        for K, B in window[:m]:
            strip = strip - (K @ B)[:3, :]          # the set's terms as they are NOW (every predict since has transformed them)
        base = out                                  # rows 0..2 of `out` are stale on purpose: never read
        window = window[m:]
        n_closed -= m

    def close():
        nonlocal n_closed
        fill = window[n_closed:]
        if not fill:
            return
        src = passes[-1][1] if passes else base     # a queued pass starts from the output of the one before it
        out = src.copy()
        for K, B in fill:
            out = out - K @ B                       # frozen: the terms as they are at close time
        out[:3, :] = np.nan                         # prove that nobody reads rows 0..2 of a pass's output ...
        out[:, :3] = np.nan                         # ... nor columns 0..2 (their live part is rows 0..2)
        passes.append((len(fill), out, int(rng.integers(1, 4))))
        n_closed += len(fill)

    for t in range(T):
        u = np.array([0.3 + 0.1 * rng.random(), rng.normal(0, 3.0)])
        # ---------------- predict: reference (dense) and model (strip + window)
        th = x[2]
        o.predict(u)
        a, b = -1 * u[0] * sind(th), u[0] * cosd(th)
        W = np.array([u[0] * cosd(th), u[0] * sind(th), u[1]])
        M3 = upper(strip[:, :3].copy())
        strip[0, 3:] += a * strip[2, 3:]; strip[1, 3:] += b * strip[2, 3:]
        F3 = np.eye(3); F3[0, 2] = a; F3[1, 2] = b
        strip[:, :3] = F3 @ M3 @ F3.T + np.outer(W * C, W)
        for K, B in window:
            K[0] += a * K[2]; K[1] += b * K[2]
            B[:, 0] += a * B[:, 2]; B[:, 1] += b * B[:, 2]
        x = x.copy()
        x[0] += u[0] * cosd(th + u[1]); x[1] += u[0] * sind(th + u[1]); x[2] = wrapTo360(th + u[1])
        # passes finish whenever they like
        for i in range(len(passes)):
            passes[i] = (passes[i][0], passes[i][1], passes[i][2] - 1)
        # ---------------- k observations of known landmarks, applied one after another
        ids = rng.choice(N, k, replace=False) + 1
        if len(passes) == 2:                        # a queued pass writes into the buffer the chain reads: retire first
            retire()
        if len(window) - n_closed + k > cap:        # the filling set cannot take this batch: close it
            close()
            if len(passes) == 2:
                retire()
        for idx in ids:
            tr = lms[idx - 1]
            d = tr - x[:2]
            z = np.array([np.hypot(*d) + rng.normal(0, 0.05), wrapTo360(atan2d(d[1], d[0]) - x[2]) + rng.normal(0, 0.5), idx])
            R = np.diag([z[0] * Rc[0], z[1] * Rc[1]])
            o._update(z, R, int(idx))
            # model: the same formulas on the reconstructed covariance, at 5 columns
            P = current_P()
            c = [0, 1, 2, 3 + 2 * (idx - 1), 4 + 2 * (idx - 1)]
            dk = x[c[3]:c[3] + 2] - x[:2]
            q = float(dk @ dk); sq = np.sqrt(q)
            zk = np.array([sq, wrapTo360(atan2d(dk[1], dk[0]) - x[2])])
            H = (1.0 / q) * np.array([[-sq * dk[0], -sq * dk[1], 0.0, sq * dk[0], sq * dk[1]],
                                      [dk[1], -dk[0], -q, -dk[1], dk[0]]])
            PHt = P[:, c] @ H.T                    # n x 2
            S = H @ P[np.ix_(c, c)] @ H.T + R
            K = PHt @ inv2(S)
            x = x + K @ (z[:2] - zk)
            window.append([K, PHt.T.copy()])        # B = (P H')' in SYM storage
        if len(window) - n_closed + k > cap and len(passes) < 2:
            close()                                 # eager: the pass of a full set is enqueued at once
        while passes and passes[0][2] <= 0 and len(passes) == 2:
            retire()
        if t % 7 == 6:                              # now and then somebody needs P whole: settle
            while passes:
                retire()
            for K, B in window:
                base = base - K @ B
                strip = strip - (K @ B)[:3, :]
            window, n_closed = [], 0
            base[:3, :] = strip
            base = upper(base)
            assert np.linalg.norm(base - o.P) / np.linalg.norm(o.P) < 1e-10, t
    while passes:
        retire()
    P_end = current_P()
    assert not np.isnan(P_end).any()
    assert np.linalg.norm(x - o.x) / np.linalg.norm(o.x) < 1e-10
    assert np.linalg.norm(P_end - o.P) / np.linalg.norm(o.P) < 1e-10


@pytest.mark.parametrize('cap', [16, 13, 11, 9])
def test_continuous_flush_queue_bookkeeping_bounds(cap):
    """A Python mirror of the host bookkeeping of the continuous flush (ekf_capi.cu: shadow_close / shadow_retire /
    shadow_settle and the arrival logic of fused_chunk), run over long random sequences of batch sizes and settles:
    the window of pending terms never exceeds EKF_MAX_WINDOW = 32 slabs (the kernels stage it in shared-memory arrays of
    that size), never reaches past the EKF_WIN_SLOTS = 256 slabs of its allocation, at most two passes are outstanding,
    and a closed set has 1..cap terms.  (The arithmetic of the schedule is the subject of the test above.)"""
    import random
    WIN, MAXW, CHUNK = 256, 32, 8
    for seed in range(4):
        rnd = random.Random(1000 * cap + seed)
        st = dict(m_pend=0, mA=0, mQ=0, nfl=0, kw0=0)

        def retire():
            if st['nfl'] == 0:
                return
            st['kw0'] += st['mA']; st['m_pend'] -= st['mA']; st['mA'] = st['mQ']; st['mQ'] = 0; st['nfl'] -= 1

        def close():
            if st['m_pend'] - st['mA'] - st['mQ'] <= 0:
                return
            assert st['nfl'] < 2 and st['mQ'] == 0
            if st['kw0'] + 2 * MAXW > WIN:                      # compaction: passes retired first, window back to the start
                while st['nfl'] > 0:
                    retire()
                st['kw0'] = 0
            m = st['m_pend'] - st['mA'] - st['mQ']
            assert 1 <= m <= cap
            if st['nfl'] == 0:
                st['mA'] = m
            else:
                st['mQ'] = m
            st['nfl'] += 1

        for _ in range(20000):
            if rnd.random() < 0.02:                             # a getter / append: settle
                while st['nfl'] > 0:
                    retire()
                st['m_pend'] = 0; st['kw0'] = 0
                continue
            k = rnd.choice([1, 1, 2, 3, 5, 8, 8, 8])
            if st['mQ'] > 0:
                retire()
            if st['m_pend'] - st['mA'] + k > cap:
                close()
                if st['mQ'] > 0:
                    retire()
            m0 = st['m_pend']
            assert m0 < MAXW and m0 + k <= MAXW                 # staged pending terms, and the slots the batch writes
            assert st['kw0'] + m0 + k <= WIN
            assert st['nfl'] <= 2
            st['m_pend'] += k
            if st['m_pend'] - st['mA'] - st['mQ'] + CHUNK > cap:
                close()


# ---------------------------------------------------------------- round 2: the reference's own source text, executed
REFRUN_CASES = ['known_prefix_N20', 'known_quirk_N20', 'uc_nearest_N20', 'uc_nearest_N8_seed7']


@pytest.mark.parametrize('name', REFRUN_CASES)
def test_oracle_matches_the_reference_text_run(name):
    """tests/golden/refrun_*.npz hold what the UNMODIFIED EKF_SLAM.m / EKF_SLAM_UC.m / Correspondence.m computed on these
    inputs when executed statement by statement by oracle/mini_matlab.py (generator: tests/golden/make_reference_run.py).
    The hand restatement oracle/dense_literal.py must agree: landmark count after every step and s exactly, x at every
    checkpoint and the final x and P to 1e-12 (both sides evaluate the same restated MathWorks built-ins; what is pinned
    is the reading of the reference's statements)."""
    from oracle import dense_literal as O
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_%s.npz' % name)))
    o = getattr(O, GU.CASES[name])()
    cps = {int(s): i for i, s in enumerate(g['checkpoint_steps'])}
    for t, u, stub in GU.steps(g):
        o.predict(u)
        o.measure(None, u, stub)
        assert (o.x.size - 3) // 2 == int(g['n_landmarks'][t]), (name, t)
        if (t + 1) in cps:
            want = g['x_checkpoints'][cps[t + 1]]
            want = want[~np.isnan(want)]
            assert want.size == o.x.size and GU.rel(o.x, want) <= 1e-12, (name, t)
    assert np.array_equal(np.asarray(o.s, dtype=np.float64).ravel(), g['s_final'].ravel())
    assert GU.rel(o.x, g['x_final']) <= 1e-12
    assert GU.rel(o.P, g['P_final']) <= 1e-12


@pytest.mark.parametrize('name', REFRUN_CASES)
def test_structured_oracle_matches_the_reference_text_run(name):
    """oracle/structured.c -- the O(n^2) form that checks the 10k-landmark runs, where the dense restatement cannot go --
    against what the reference's own text computed (refrun_*.npz): landmark counts and s exactly, x at every checkpoint,
    final x and P to 1e-11 (a different grouping of the same products than the reference's dense algebra)."""
    from oracle import structured as S
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_%s.npz' % name)))
    o = getattr(S, GU.CASES[name])(capacity_landmarks=24)
    cps = {int(s): i for i, s in enumerate(g['checkpoint_steps'])}
    for t, u, stub in GU.steps(g):
        o.predict(u)
        o.measure(None, u, stub)
        assert (o.x.size - 3) // 2 == int(g['n_landmarks'][t]), (name, t)
        if (t + 1) in cps:
            want = g['x_checkpoints'][cps[t + 1]]
            assert GU.rel(o.x, want[~np.isnan(want)]) <= 1e-11, (name, t)
    assert np.array_equal(np.asarray(o.s, dtype=np.float64).ravel(), g['s_final'].ravel())
    assert GU.rel(o.x, g['x_final']) <= 1e-11 and GU.rel(o.P, g['P_final']) <= 1e-11


@pytest.mark.skipif(not os.path.isdir('/root/reference'), reason='the reference checkout is not on this machine')
@pytest.mark.parametrize('seed', range(8))
def test_oracle_matches_the_reference_text_on_random_worlds(seed):
    """Beyond the four committed worlds: seeded random worlds (5 - 10 landmarks, 2 - 4 observations per frame, 40 steps,
    some frames empty -- the isempty(observed_LL) path, EKF_SLAM.m:105 -- known-correspondence class on 'prefix' worlds,
    unknown-correspondence class on 'nearest' worlds), the reference's text executed live against oracle/dense_literal.py:
    landmark count after every step and s exactly, x and P <= 1e-12 at every 10th step and at the end."""
    from ekf_slam_b200 import synth
    from oracle.mini_matlab import ReferenceFilter
    from tests import golden_util as GU
    rng = np.random.default_rng(1000 + seed)
    cls = 'EKF_SLAM' if seed % 2 == 0 else 'EKF_SLAM_UC'
    w = synth.make_world(N=int(rng.integers(5, 11)), T=40, k=int(rng.integers(2, 5)), seed=77 + seed,
                         mode='prefix' if cls == 'EKF_SLAM' else 'nearest')
    empty = set(int(t) for t in rng.choice(np.arange(3, 40), size=4, replace=False))
    r, o = ReferenceFilter('/root/reference', cls), getattr(O, cls)()
    for t in range(w.T):
        rows = w.observed[t]
        ntab = w.n_table[t]
        if t == 0:
            rows, ntab = rows[:1], 1                 # a first frame the reference can run (one new landmark)
        if t in empty:
            rows = np.zeros((0, 3))
        stub = GU.Stub(rows, w.table[:ntab])
        r.predict(w.u[t]); r.measure(None, w.u[t], stub)
        o.predict(w.u[t]); o.measure(None, w.u[t], stub)
        assert r.x.size == o.x.size, (seed, t)
        if t % 10 == 9 or t == w.T - 1:
            assert GU.rel(o.x, r.x) <= 1e-12 and GU.rel(o.P, r.P) <= 1e-12, (seed, t)
    assert np.array_equal(np.asarray(o.s, dtype=np.float64).ravel(), r.s) and o.x.size >= 3 + 2 * 2


def test_oracle_association_matches_the_reference_text_run():
    """Correspondence.estimateCorrespondence as shipped (Correspondence.m:28-88, signature score :75), executed from its
    own source on a fixed state with a duplicated signature (first minimum wins, :81) and unseen signatures (new landmark)."""
    from oracle import dense_literal as O
    from tests import golden_util as GU
    a = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_association_N12.npz')))
    c = O.Correspondence(1e-11, 1e9, 'EKF_SLAM_UC')
    got = []
    for z, Rd in zip(a['z'], a['Rdiag']):
        new, idx = c.estimateCorrespondence(z, np.diag(Rd), a['x'], a['P'], a['s'])
        got.append([int(bool(new)), int(idx)])
    assert np.array_equal(np.array(got, dtype=np.int32), a['mode_a'])
    assert 0 < int(a['mode_a'][:, 0].sum()) < len(got)
    # mode B: the reference's own commented-out score (Correspondence.m:74) enabled in memory, everything else as written
    # (make_reference_run.py): position cost + signature cost, all signatures equal, gate 9.21
    cb = O.Correspondence(1e-11, 9.21, 'EKF_SLAM_UC', position_cost_in_score=True)
    gotb = []
    for z, Rd in zip(a['z'], a['Rdiag']):
        zB = z.copy(); zB[2] = 1.0
        new, idx = cb.estimateCorrespondence(zB, np.diag(Rd), a['x'], a['P'], np.ones(len(a['s'])))
        gotb.append([int(bool(new)), int(idx)])
    assert np.array_equal(np.array(gotb, dtype=np.int32), a['mode_b'])
    assert np.array_equal(a['mode_b'], GU.load('association_N12')['modeB'])       # what the GPU association test is held to
    assert len(set(a['mode_b'][:, 1])) >= 10                                        # the Mahalanobis term really decides


@pytest.mark.parametrize('name', REFRUN_CASES)
def test_oracle_ellipses_match_the_reference_plot_text(name):
    """EKF_SLAM.plot / EKF_SLAM_UC.plot (EKF_SLAM.m:154-234) executed from source on the final state of each reference
    run, figure calls recorded instead of drawn: the polylines it hands to plot(xr, yr) are the ellipses
    oracle.ellipse_params describes (centre, semi-axes along the drawn x / y directions, rotation), robot first, to 1e-12
    of the map's extent.  (eig, cos, sin, atan are MathWorks built-ins, served by NumPy on both sides.)"""
    from oracle import dense_literal as O
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_%s.npz' % name)))
    e = O.ellipse_params(g['x_final'], g['P_final'])
    c = g['ellipse_curves']
    assert c.shape == (1 + (g['x_final'].size - 3) // 2, 2, 158)
    t = (-np.pi + 0.01 * np.arange(629))[::4]
    scale = max(1.0, float(np.abs(g['x_final']).max()))
    for k, (cx, cy, ax, ay, arc) in enumerate(e):
        X = cx + np.cos(arc) * ax * np.cos(t) - np.sin(arc) * ay * np.sin(t)
        Y = cy + np.sin(arc) * ax * np.cos(t) + np.cos(arc) * ay * np.sin(t)
        assert max(np.abs(X - c[k, 0]).max(), np.abs(Y - c[k, 1]).max()) <= 1e-12 * scale, (name, k)


@pytest.mark.skipif(not os.path.isdir('/root/reference'), reason='the reference checkout is not on this machine')
def test_witness_stub_drives_the_reference_text():
    """matlab/RefFixtureStub.m -- the landmark_list stand-in matlab/run_reference_fixture.m hands to the unmodified
    reference under MATLAB / Octave -- executed from its own text (struct array built with lm(k).f = v, wrapped in
    struct('landmark', lm)) and read by the reference's own expressions (EKF_SLAM_UC.m:113,123): the run reproduces
    refrun_uc_nearest_N8_seed7.npz bit for bit, so the witness's duck typing is what the reference expects."""
    from oracle.mini_matlab import ReferenceFilter, to_matlab
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_uc_nearest_N8_seed7.npz')))
    r = ReferenceFilter('/root/reference', 'EKF_SLAM_UC')
    r.it.load_file(os.path.join(os.path.dirname(GU.GOLDEN), os.pardir, 'matlab', 'RefFixtureStub.m'))
    off = np.concatenate([[0], np.cumsum(g['obs_counts'])])
    for t in range(len(g['obs_counts'])):
        nt = int(g['n_table'][t])
        stub = r.it.construct('RefFixtureStub', [to_matlab(g['obs_rows'][off[t]:off[t + 1]].reshape(-1, 3)),
                                                 to_matlab(g['table_index'][:nt].astype(np.float64).reshape(-1, 1)),
                                                 to_matlab(g['table_loc'][:nt].reshape(-1, 2))])
        r.predict(g['u'][t])
        r._method('measure', np.zeros((0, 0)), g['u'][t], stub)
    assert np.array_equal(r.x, g['x_final']) and np.array_equal(r.P, g['P_final']) and np.array_equal(r.s, g['s_final'].ravel())


@pytest.mark.skipif(not os.path.isdir('/root/reference'), reason='the reference checkout is not on this machine')
def test_witness_script_runs_end_to_end_on_the_reference_checkout(tmp_path):
    """matlab/run_reference_fixture.m executed as written (addpath of the reference checkout -- its own EKF_SLAM.m,
    EKF_SLAM_UC.m, Correspondence.m and RANSAC.m are found on the path and loaded unmodified --, the assert that EKF_SLAM
    resolves into the checkout, dir / fopen / sscanf / fprintf('%.17g') CSV handling, RefFixtureStub) on
    tests/golden/ref_inputs; tests/golden/compare_reference_fixture.py then reads what it wrote and finds the refrun
    vectors bit for bit: the CSV round trip loses nothing and the script is ready for whoever has MATLAB or Octave.
    (Run by oracle/mini_matlab.py, this pins nothing new -- the outputs are NOT committed as tests/golden/ref_outputs.)"""
    from oracle.mini_matlab import Interpreter
    from tests.golden import compare_reference_fixture as C
    from tests import golden_util as GU
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    it = Interpreter()
    it.load_file(os.path.join(root, 'matlab', 'run_reference_fixture.m'))
    it.call('run_reference_fixture', ['/root/reference', os.path.join(GU.GOLDEN, 'ref_inputs'), str(tmp_path)], 0, {})
    assert len(it.printed) == len(GU.CASES) and all('landmarks ->' in line for line in it.printed)
    assert any(f.startswith('/root/reference/') and f.endswith('RANSAC.m') for f in it._loaded)      # the reference's own constructor ran
    for name in GU.CASES:
        r = C.compare_case(name, str(tmp_path))
        assert r['pinned'] and r['x_final'] == 0.0 and r['P_final'] == 0.0 and r['x_checkpoints'] == 0.0, (name, r)


def test_oracle_matches_the_reference_free_functions():
    """append.m:1-27 and example.m:24-45 (predict, f) -- the free-function forms SURVEY 8(a) cites beside the class methods --
    executed from their own source (tests/golden/refrun_functions.npz): the oracle's append / predict / f give the same
    x, P and F bit for bit on states of 0..5 landmarks; an index inside the map leaves x and P untouched (append.m:4)."""
    from oracle import dense_literal as O
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_functions.npz')))
    for c in range(int(g['n_cases'])):
        x, P, u, R, pos = (g['%s%d' % (k, c)] for k in ('x', 'P', 'u', 'R', 'pos'))
        o = O.EKF_SLAM(); o.x = x.copy(); o.P = P.copy(); o.s = [float(i) for i in range((x.size - 3) // 2)]
        o.append(u, R, pos, 99.0)
        assert np.array_equal(o.x, g['append_x%d' % c]) and np.array_equal(o.P, g['append_P%d' % c]), c
        assert g['append_x%d' % c].size == x.size + 2
        assert np.array_equal(g['keep_x%d' % c], x) and np.array_equal(g['keep_P%d' % c], P), c
        o = O.EKF_SLAM(); o.x = x.copy(); o.P = P.copy()
        o.predict(u)
        assert np.array_equal(o.x, g['predict_x%d' % c]) and np.array_equal(o.P, g['predict_P%d' % c]), c
        assert 0.0 <= o.x[2] < 360.0 or o.x[2] == 360.0
        o = O.EKF_SLAM(); o.x = x.copy(); o.P = P.copy()
        xn, F = o.f(x, u)
        assert np.array_equal(np.asarray(xn).ravel(), g['f_x%d' % c]) and np.array_equal(F, g['f_F%d' % c]), c


def test_mini_matlab_grammar_known_answers():
    """The interpreter that executes the reference text (oracle/mini_matlab.py) against MATLAB's documented behaviour on
    the constructs the three files use: white space inside [ ], transpose versus string quote, precedence of unary minus
    and ^, an exponent with its own sign, ranges, orientation of vector indexing, growth on assignment, comma-separated
    lists from struct arrays, several outputs, switch on strings, handle versus value classes."""
    from types import SimpleNamespace
    from oracle.mini_matlab import Interpreter, MiniMatlabError, to_matlab
    src = """
classdef T < handle
    properties
        v = [1 2 3];
        n
    end
    methods
        function h = T()
            h.n = 0;
        end
        function bump(h)
            h.n = h.n + 1;
        end
        function [a,b] = two(h, x)
            a = x';
            b = [x(1) -x(2)];
        end
        function r = m1(h)
            r = [1 -2];
        end
        function r = m2(h)
            r = [1 - 2];
        end
        function r = m3(h)
            r = [1 -2 + 3, 4];
        end
        function r = m4(h)
            r = -2^2;
        end
        function r = m5(h)
            r = 2^-1;
        end
        function r = m6(h)
            r = [1 2; 3 4]^-1;
        end
        function r = m7(h)
            a = [1 2 3]';
            r = a(2:3)';
        end
        function r = m8(h)
            r = zeros(2,2); r(3,4) = 7; r(1,1) = -1;
        end
        function r = m9(h)
            r = []; r = [r; 5]; r = [r; 6];
        end
        function r = m10(h, lst)
            r = [lst.index];
        end
        function r = m11(h, lst)
            r = lst(find([lst.index]==2)).loc;
        end
        function r = m12(h, s)
            switch(s)
                case 'abc'
                    r = 1;
                otherwise
                    r = 2;
            end
        end
        function r = m13(h)
            x = [10 20 30 40];
            x(1:3) = [1; 2; 3];
            r = x;
        end
        function r = m14(h)
            r = 0;
            for ii = 1:3
                if(ii < 2 || ~isempty([]))
                    r = r + 100;
                elseif ii == 2
                    r = r + 10;
                else
                    r = r + 1;
                end
            end
        end
        function r = m15(h)
            A = eye(2); r = 2*A*[1;1] + (A(1,:))'*3;
        end
        function r = m16(h)
            r = size(zeros(3,5),1) + length(zeros(3,5))*10 + size(eye(size(zeros(4,4))),2)*100;
        end
        function r = m17(h)
            t = -1:0.5:1;
            r = [t(1) t(5) length(t)];
        end
        function r = m18(h)
            hold on;
            [V,D] = eig([2 0; 0 1]);
            hold off
            r = [D(1,1) D(2,2) abs(V(2,1)) abs(V(1,2))];
        end
        function r = m19(h)
            r = 0;
            for ii=4:2:9
                r = r*10 + ii;
            end
        end
    end
end
"""
    held = []
    it = Interpreter(stubs={'hold': lambda s: held.append(s)})
    it.load_source(src)
    t = it.construct('T', [])

    def call(name, *args, nargout=1):
        return it.call_function(t.cls['methods'][name], [t] + [to_matlab(a) for a in args], nargout)

    assert np.array_equal(call('m1')[0], [[1, -2]])                    # "1 -2": two elements
    assert np.array_equal(call('m2')[0], [[-1]])                       # "1 - 2": a subtraction
    assert np.array_equal(call('m3')[0], [[1, 1, 4]])                  # "1 -2 + 3, 4" -> [1, (-2+3), 4]
    assert call('m4')[0][0, 0] == -4.0                                 # ^ binds tighter than unary minus
    assert call('m5')[0][0, 0] == 0.5
    assert np.allclose(call('m6')[0], np.linalg.inv([[1, 2], [3, 4]]), rtol=0, atol=1e-15)
    assert call('m7')[0].shape == (1, 2) and np.array_equal(call('m7')[0], [[2, 3]])      # column indexed -> column, then '
    r8 = call('m8')[0]
    assert r8.shape == (3, 4) and r8[2, 3] == 7 and r8[0, 0] == -1 and r8.sum() == 6      # grown with zeros
    assert np.array_equal(call('m9')[0], [[5], [6]])
    lst = [SimpleNamespace(index=3, loc=np.array([1.0, 2.0])), SimpleNamespace(index=2, loc=np.array([5.0, 6.0]))]
    assert np.array_equal(call('m10', lst)[0], [[3, 2]])               # [lst.index]: comma-separated list
    assert np.array_equal(call('m11', lst)[0], [[5, 6]])
    assert call('m12', 'abc')[0][0, 0] == 1 and call('m12', 'zzz')[0][0, 0] == 2
    assert np.array_equal(call('m13')[0], [[1, 2, 3, 40]])             # column into a row slice: element count decides
    assert call('m14')[0][0, 0] == 111
    assert np.array_equal(call('m15')[0], [[5], [2]])
    assert call('m16')[0][0, 0] == 3 + 50 + 400
    assert np.array_equal(call('m17')[0], [[-1, 1, 5]])                # a:step:b
    assert np.array_equal(call('m19')[0], [[468]])                     # 4:2:9 stops at 8
    assert np.array_equal(call('m18')[0], [[1, 2, 1, 1]])              # eig of a symmetric matrix: ascending, D diagonal
    assert held == ['on', 'off']                                       # command syntax: `hold on` is hold('on')
    a, b = call('two', np.array([4.0, 9.0]), nargout=2)
    assert a.shape == (2, 1) and np.array_equal(b, [[4, -9]])
    call('bump', nargout=0); call('bump', nargout=0)
    assert t.props['n'][0, 0] == 2                                     # handle class: the method changed the object
    with pytest.raises(MiniMatlabError):                               # too many input arguments, as MATLAB
        it.call_function(t.cls['methods']['m1'], [t, to_matlab(1.0)], 1)


@pytest.mark.skipif(not os.path.isdir('/root/reference'), reason='the reference checkout is not on this machine')
def test_reference_run_fixture_is_reproducible():
    """Where the reference is present: executing its source again reproduces the committed refrun vectors bit for bit
    (shortest case), and a first frame with several non-zero table entries stops as MATLAB would."""
    from oracle.mini_matlab import ReferenceFilter, MiniMatlabError
    from tests import golden_util as GU
    g = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_known_quirk_N20.npz')))
    r = ReferenceFilter('/root/reference', 'EKF_SLAM')
    for t, u, stub in GU.steps(g):
        r.predict(u)
        r.measure(None, u, stub)
    assert np.array_equal(r.x, g['x_final']) and np.array_equal(r.P, g['P_final']) and np.array_equal(r.s, g['s_final'].ravel())
    assert np.array_equal(np.array(r.n_landmarks), g['n_landmarks'])
    assert np.array_equal(r.plot_curves()[:, :, ::4], g['ellipse_curves'])
    sys.path.insert(0, GU.GOLDEN)
    try:
        import make_reference_run as MR
        f = MR.functions_case('/root/reference')
    finally:
        sys.path.remove(GU.GOLDEN)
    gf = dict(np.load(os.path.join(GU.GOLDEN, 'refrun_functions.npz')))
    assert sorted(f) == sorted(gf) and all(np.array_equal(f[k], gf[k]) for k in gf)
    g0 = GU.load('known_quirk_N20')                  # the original world: 5 new landmarks in the first frame
    r = ReferenceFilter('/root/reference', 'EKF_SLAM')
    t, u, stub = next(GU.steps(g0))
    r.predict(u)
    with pytest.raises(MiniMatlabError, match='too many input arguments'):
        r.measure(None, u, stub)
