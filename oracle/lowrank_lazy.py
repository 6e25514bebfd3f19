"""TEST INFRASTRUCTURE ONLY -- matrix-free oracle for configurations whose P cannot be held on the host.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.

The throughput configurations start from the synthetic SPD map of SURVEY.md 8(d), P0 = A A'/r + diag(d) with A n x r.
Every operation of the filter then changes P by a few outer products:
  * predict, EKF_SLAM.m:47 with F = I + g e3' (g = [F(1,3); F(2,3); 0; ...], :63-64):
        F P F' + Q = P + g (e3' P) + (P e3) g' + (e3' P e3) g g' + (C W) W'                         (:44, :47)
  * update, EKF_SLAM.m:145:  (I - K H) P = P - K (H P) = P - K(:,1) B(1,:) - K(:,2) B(2,:)
so after a few steps P = P0 + sum_t u_t v_t' with explicit n-vectors, and any column, row or entry of the CURRENT P
costs O(n (r + T)).  The gain K (:143), the innovation covariance (:141) and x (:144) need only five columns / rows
of P, hence the whole step is O(n) per observation and a row of the 160 003 x 160 003 covariance of the 80 000-landmark
configuration (205 GB dense) can be compared with what the GPUs hold.
The algebra is the reference's, evaluated in a different order from every other implementation here (tolerance
tests only).  Appends are not supported (the large configurations never append).

parity unpinned (see dense_literal.py): pinned by agreement with oracle/structured.c on a materialised P
(tests/test_oracle_cpu.py).
"""
import numpy as np

from .matlab_compat import sind, cosd, atan2d, wrapTo360


class LazyLowRankFilter:
    def __init__(self, x, A, d, s, C=0.2, Rc=(0.1, 5.0)):
        self.x = np.array(x, dtype=np.float64)
        self.A = np.asarray(A, dtype=np.float64)
        self.d = np.asarray(d, dtype=np.float64)
        self.s = np.asarray(s, dtype=np.float64)
        self.r = self.A.shape[1]
        self.n = self.x.size
        self.C, self.Rc = C, list(Rc)                 # EKF_SLAM_UC.m:12-13 defaults
        self.U, self.V = [], []                       # P = P0 + sum_t U[t] V[t]'

    # ---- current P, matrix-free
    def col(self, c):
        v = self.A @ self.A[c] / self.r
        v[c] += self.d[c]
        for u, w in zip(self.U, self.V):
            v += u * w[c]
        return v

    def row(self, i):
        v = self.A @ self.A[i] / self.r
        v[i] += self.d[i]
        for u, w in zip(self.U, self.V):
            v += u[i] * w
        return v

    def block(self, rows, c0, c1):
        """P(rows, c0:c1) of the current covariance."""
        rows = np.asarray(rows)
        out = self.A[rows] @ self.A[c0:c1].T / self.r
        for a, i in enumerate(rows):
            if c0 <= i < c1:
                out[a, i - c0] += self.d[i]
        for u, w in zip(self.U, self.V):
            out += np.outer(u[rows], w[c0:c1])
        return out

    # ---- EKF_SLAM.m:40-51
    def predict(self, u):
        th = self.x[2]
        W = np.zeros(self.n)
        W[0:3] = [u[0] * cosd(th), u[0] * sind(th), u[1]]                     # :42
        g = np.zeros(self.n)
        g[0], g[1] = -1 * u[0] * sind(th), u[0] * cosd(th)                    # :63-64
        self.x[0] = self.x[0] + u[0] * cosd(th + u[1])                        # :58
        self.x[1] = self.x[1] + u[0] * sind(th + u[1])                        # :59
        self.x[2] = wrapTo360(th + u[1])                                      # :60, :50
        p2c, p2r = self.col(2), self.row(2)
        p22 = p2c[2]
        self.U += [g.copy(), p2c, g * p22, W * self.C]                        # g (e3'P) + (P e3) g' + p33 g g' + (C W) W'
        self.V += [p2r, g.copy(), g.copy(), W]

    # ---- EKF_SLAM.m:124-145
    def update(self, z, Rdiag, idx1):
        x = self.x
        al = 3 + 2 * (idx1 - 1)
        c = [0, 1, 2, al, al + 1]
        d0, d1 = x[al] - x[0], x[al + 1] - x[1]                               # :126
        q = d0 * d0 + d1 * d1                                                 # :127
        sq = np.sqrt(q)
        zk = np.array([sq, wrapTo360(atan2d(d1, d0) - x[2])])                 # :130
        H = (1.0 / q) * np.array([[-sq * d0, -sq * d1, 0.0, sq * d0, sq * d1],
                                  [d1, -d0, -q, -d1, d0]])                    # :137-138 (non-zero columns)
        Pc = np.column_stack([self.col(j) for j in c])                        # P(:, c)
        Pr = np.vstack([self.row(j) for j in c])                              # P(c, :)
        S = (H @ Pc[c, :]) @ H.T + np.diag(Rdiag)                             # :141
        K = (Pc @ H.T) @ np.linalg.inv(S)                                     # :143
        self.x = x + K @ (np.asarray(z[:2], dtype=np.float64) - zk)           # :144 (no wrap)
        B = H @ Pr                                                            # H P, :145
        self.U += [-K[:, 0].copy(), -K[:, 1].copy()]
        self.V += [B[0].copy(), B[1].copy()]

    # ---- EKF_SLAM_UC.m:102-152 restricted to updates (association by signature, Correspondence.m:71-81)
    def measure_rows(self, observed_LL, s_cost=1e-11, s_thresh=1e9):
        actions = []
        for row in np.asarray(observed_LL, dtype=np.float64).reshape(-1, 3):
            ll = (row[2] - self.s) * (1.0 / s_cost) * (row[2] - self.s)       # :71, :75
            ok = ll <= s_thresh                                               # :78
            if not ok.any():
                raise NotImplementedError('append in the matrix-free oracle')
            idx = int(np.argmin(np.where(ok, ll, np.inf))) + 1                # :81 (first minimum)
            self.update(row[:2], [row[0] * self.Rc[0], row[1] * self.Rc[1]], idx)
            actions.append(('update', idx))
        return actions
