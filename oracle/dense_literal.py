"""TEST INFRASTRUCTURE ONLY -- dense, line-by-line CPU oracle for the EKF-SLAM hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product (ekf_slam_b200) never does.

It restates the reference's MATLAB with dense matrices and MATLAB's left-to-right
evaluation order (dense F, Q, H_k, eye(n) -- O(n^3) exactly like the reference),
keeping every quirk.  Each function cites the
reference file:line it follows.

Parity status: the reference has no tests / golden vectors and MATLAB / Octave are
absent from this image.  Since round 2 this restatement is pinned to the reference's
SOURCE TEXT mechanically: oracle/mini_matlab.py executes the unmodified EKF_SLAM.m,
EKF_SLAM_UC.m and Correspondence.m on the golden worlds, tests/golden/refrun_*.npz
hold what they computed, and tests/test_oracle_cpu.py requires this module to agree
(landmark counts and s exactly, x and P to 1e-12).  What stays unpinned are the
MathWorks built-ins (matlab_compat.py), which both sides share: "parity unpinned" in
that sense, until matlab/run_reference_fixture.m is run under MATLAB or Octave.
"""
import numpy as np

from .matlab_compat import sind, cosd, atan2d, wrapTo360, inv2


class Correspondence:
    """Correspondence.m:1-92 (value class)."""

    def __init__(self, cost, thresh, method, position_cost_in_score=False):
        # Correspondence.m:12-25
        self.method = method
        self.s_cost = cost
        self.s_thresh = thresh
        if method != 'EKF_SLAM_UC':
            self.method = 'ML'          # :19-23 (warning + fallback)
        # mode A (shipped, :75) scores the signature cost only; mode B is the
        # commented-out line :74 (position_cost + signiture_cost).
        self.position_cost_in_score = position_cost_in_score

    def estimateCorrespondence(self, z, R, x, P, s):
        """Correspondence.m:28-88.  Returns (newLL, index) with a 1-based index."""
        newLL = True                                                # :33
        numOfLandmarks = (len(x) - 3) // 2                          # :39
        index = numOfLandmarks + 1                                  # :40
        min_log_likelihood = np.inf                                 # :43
        n = numOfLandmarks * 2 + 3
        for kk in range(1, numOfLandmarks + 1):                     # :49
            mu_k = np.array([x[(kk - 1) * 2 + 3], x[(kk - 1) * 2 + 4]])   # :50
            delta_k = mu_k - x[0:2]                                 # :52
            q_k = float(delta_k @ delta_k)                          # :53
            z_k = np.array([np.sqrt(q_k),
                            wrapTo360(atan2d(delta_k[1], delta_k[0]) - x[2])])  # :56
            F_k = np.zeros((5, n))                                  # :59
            F_k[0:3, 0:3] = np.eye(3)
            F_k[3:5, 3 + (kk - 1) * 2: 5 + (kk - 1) * 2] = np.eye(2)
            sq = np.sqrt(q_k)
            H_k = (1.0 / q_k) * np.array(
                [[-sq * delta_k[0], -sq * delta_k[1], 0.0, sq * delta_k[0], sq * delta_k[1]],
                 [delta_k[1], -delta_k[0], -q_k, -delta_k[1], delta_k[0]]]) @ F_k   # :62-63
            phi_k = H_k @ P @ H_k.T + R                             # :66
            nu = z[0:2] - z_k
            position_cost = float(nu @ inv2(phi_k) @ nu)            # :69
            d = z[2] - s[kk - 1]
            signiture_cost = d * (1.0 / self.s_cost) * d            # :71
            if self.position_cost_in_score:
                ll = position_cost + signiture_cost                 # :74 (commented out upstream)
            else:
                ll = signiture_cost                                 # :75
            if ll <= self.s_thresh:                                 # :78
                if ll < min_log_likelihood:                         # :81 strict <, first min wins
                    newLL = False
                    min_log_likelihood = ll
                    index = kk
        return newLL, index


class EKF_SLAM:
    """EKF_SLAM.m:1-151 (known correspondence; plot omitted -- graphics)."""

    def __init__(self):
        # EKF_SLAM.m:12-16, 26-34
        self.C = 0.2
        self.Rc = [0.01, 5.0]
        self.s_cost = 0.00000000001
        self.s_thresh = 1000000000
        self.x = np.array([0.0, 0.0, 0.0])          # row vector :28
        self.P = np.eye(3) * 0.1                    # :30
        self.Q = None
        self.s = np.zeros(0)
        self.observed = None
        self.chosen = []                            # oracle-side log: (action, idx) per observation

    # ---- prediction ---------------------------------------------------
    def predict(self, u):
        """EKF_SLAM.m:40-51 == EKF_SLAM_UC.m:42-53."""
        th = self.x[2]
        W = np.array([[u[0] * cosd(th)], [u[0] * sind(th)], [u[1]]])   # :42
        self.Q = np.zeros(self.P.shape)                                # :43
        self.Q[0:3, 0:3] = (W * self.C) @ W.T                          # :44
        self.x, F = self.f(self.x, u)                                  # :46
        self.P = (F @ self.P) @ F.T + self.Q                           # :47
        self.x[2] = wrapTo360(self.x[2])                               # :50

    def f(self, x, u):
        """EKF_SLAM.m:56-65."""
        x_new = x.copy()
        x_new[0] = x[0] + u[0] * cosd(x[2] + u[1])                     # :58
        x_new[1] = x[1] + u[0] * sind(x[2] + u[1])                     # :59
        x_new[2] = x[2] + u[1]                                         # :60
        F = np.eye(len(x))                                             # :62
        F[0, 2] = -1 * u[0] * sind(x[2])                               # :63
        F[1, 2] = u[0] * cosd(x[2])                                    # :64
        return x_new, F

    # ---- augmentation -------------------------------------------------
    def append(self, u, R, landmarkPos, signature):
        """EKF_SLAM.m:67-98 (append.m:1-27 is the function-style twin)."""
        self.s = np.append(self.s, float(signature))                   # :70
        numOfLandmarks = (len(self.x) - 3) // 2                        # :74
        n = self.P.shape[0]                                            # :76
        self.x = np.append(self.x, [landmarkPos[0], landmarkPos[1]])   # :79
        th = self.x[2]
        jxr = np.array([[1.0, 0.0, -u[0] * sind(th)],
                        [0.0, 1.0, u[0] * cosd(th)]])                  # :84-85
        jz = np.array([[cosd(u[1]), -u[0] * sind(u[1])],
                       [sind(u[1]), u[0] * cosd(u[1])]])               # :87-88
        P = np.zeros((n + 2, n + 2))                                   # MATLAB grows with zeros
        P[0:n, 0:n] = self.P
        P[n:n + 2, n:n + 2] = (jxr @ P[0:3, 0:3]) @ jxr.T + (jz @ R) @ jz.T   # :91
        P[0:3, n:n + 2] = P[0:3, 0:3] @ jxr.T                          # :92
        P[n:n + 2, 0:3] = P[0:3, n:n + 2].T                            # :93
        for idx in range(1, numOfLandmarks + 1):                       # :94
            a = (idx - 1) * 2 + 3
            P[n:n + 2, a:a + 2] = jxr @ P[a:a + 2, 0:3].T              # :95
            P[a:a + 2, n:n + 2] = P[n:n + 2, a:a + 2].T                # :96
        self.P = P

    # ---- measurement --------------------------------------------------
    def _loc_by_index(self, landmark_list, want):
        """find([landmark.index]==want) then .loc  (EKF_SLAM.m:120); first hit wins."""
        for lm in landmark_list.landmarkObj.landmark:
            if lm.index == want:
                return lm.loc
        raise IndexError('no landmark with index %r' % (want,))

    def _loc_first_nonzero(self, landmark_list):
        """find([landmark.index]) with no comparison (EKF_SLAM.m:111)."""
        for lm in landmark_list.landmarkObj.landmark:
            if lm.index != 0:
                return lm.loc
        raise IndexError('no indexed landmark')

    def _update(self, z, R, idx):
        """EKF_SLAM.m:124-145 == EKF_SLAM_UC.m:125-146 (dense, as written)."""
        x = self.x
        n = len(x)
        mu_k = np.array([x[3 + (idx - 1) * 2], x[3 + (idx - 1) * 2 + 1]])     # :125
        delta_k = mu_k - x[0:2]                                               # :126
        q_k = float(delta_k @ delta_k)                                        # :127
        z_k = np.array([np.sqrt(q_k),
                        wrapTo360(atan2d(delta_k[1], delta_k[0]) - x[2])])    # :130
        F_k = np.zeros((5, n))                                                # :134
        F_k[0:3, 0:3] = np.eye(3)
        F_k[3:5, 3 + (idx - 1) * 2: 5 + (idx - 1) * 2] = np.eye(2)
        sq = np.sqrt(q_k)
        H_k = (1.0 / q_k) * np.array(
            [[-sq * delta_k[0], -sq * delta_k[1], 0.0, sq * delta_k[0], sq * delta_k[1]],
             [delta_k[1], -delta_k[0], -q_k, -delta_k[1], delta_k[0]]]) @ F_k  # :137-138
        phi_k = (H_k @ self.P) @ H_k.T + R                                    # :141
        K = (self.P @ H_k.T) @ inv2(phi_k)                                    # :143
        self.x = x + (K @ (z[0:2] - z_k))                                     # :144 (no wrap, F10)
        self.P = (np.eye(n) - K @ H_k) @ self.P                               # :145

    def measure(self, laserData, u, landmark_list):
        """EKF_SLAM.m:100-151."""
        observed_LL = landmark_list.getLandmark(laserData, self.x)            # :102
        self.observed = observed_LL                                           # :103
        if observed_LL is None or len(observed_LL) == 0:                      # :105
            return
        observed_LL = np.asarray(observed_LL, dtype=np.float64).reshape(-1, 3)
        for ii in range(1, observed_LL.shape[0] + 1):                         # :107
            R = np.zeros((2, 2))                                              # :108
            R[0, 0] = observed_LL[ii - 1, 0] * self.Rc[0]
            R[1, 1] = observed_LL[ii - 1, 1] * self.Rc[1]
            if len(self.x) < 4:                                               # :110
                self.append(u, R, self._loc_first_nonzero(landmark_list), 1)  # :111
                self.chosen.append(('append', 1))
            else:
                z = observed_LL[ii - 1, :]                                    # :116
                numOfLandmarks = (len(self.x) - 3) // 2                       # :117
                if z[2] > numOfLandmarks:                                     # :118
                    self.append(u, R, self._loc_by_index(landmark_list, z[2]), z[2])   # :120
                    self.chosen.append(('append', int(z[2])))
                else:
                    idx = ii                                                  # :123 (F8 quirk)
                    self._update(z, R, idx)
                    self.chosen.append(('update', idx))


class EKF_SLAM_UC(EKF_SLAM):
    """EKF_SLAM_UC.m:1-152 (unknown correspondence)."""

    def __init__(self):
        super().__init__()
        self.Rc = [0.1, 5.0]                                                  # EKF_SLAM_UC.m:13
        self.correspondence = Correspondence(0.00000000001, 1000000000, 'EKF_SLAM_UC')  # :16

    def measure(self, laserData, u, landmark_list):
        """EKF_SLAM_UC.m:102-152."""
        observed_LL = landmark_list.getLandmark(laserData, self.x)            # :104
        self.observed = observed_LL                                           # :105
        if observed_LL is None or len(observed_LL) == 0:                      # :107
            return
        observed_LL = np.asarray(observed_LL, dtype=np.float64).reshape(-1, 3)
        for ii in range(1, observed_LL.shape[0] + 1):                         # :109
            R = np.zeros((2, 2))                                              # :110
            R[0, 0] = observed_LL[ii - 1, 0] * self.Rc[0]
            R[1, 1] = observed_LL[ii - 1, 1] * self.Rc[1]
            if len(self.x) < 4:                                               # :112
                self.append(u, R, self._loc_first_nonzero(landmark_list), 1)  # :113
                self.chosen.append(('append', 1))
            else:
                z = observed_LL[ii - 1, :]                                    # :118
                new_LM, idx = self.correspondence.estimateCorrespondence(
                    z, R, self.x, self.P, self.s)                             # :119
                if new_LM:                                                    # :121
                    self.append(u, R, self._loc_by_index(landmark_list, idx), idx)   # :123
                    self.chosen.append(('append', idx))
                else:
                    self._update(z, R, idx)                                   # :125-146
                    self.chosen.append(('update', idx))


def ellipse_params(x, P):
    """The ellipses EKF_SLAM.plot draws, EKF_SLAM.m:180-229, restated literally (eig of each 2x2 diagonal block taken
    with both off-diagonal entries as the reference does, the if/else on eigval(1,1) > eigval(2,2), atan of the
    eigenvector ratio, scale .25 for the robot and .50 for landmarks) but returning numbers instead of plotting:
    rows [cx, cy, semi-axis along the drawn x direction, semi-axis along y, arc_rad].  Test infrastructure only."""
    x = np.asarray(x, dtype=np.float64)
    out = []
    starts = [0] + list(range(3, x.size, 2))                                  # :181 robot, :203 ii = 4:2:end
    for b, i in enumerate(starts):
        Sigma = np.array([[P[i, i], P[i, i + 1]], [P[i + 1, i], P[i + 1, i + 1]]])   # :181 / :204
        mu = (x[i], x[i + 1])                                                 # :182 / :205
        if Sigma[0, 1] == Sigma[1, 0]:
            eigval, eigvec = np.linalg.eigh(Sigma)                            # MATLAB eig, symmetric input: ascending
        else:
            eigval, eigvec = np.linalg.eig(Sigma)                             # :183 / :207
        chi_square = 2.2788                                                   # :184 / :208
        major = 2 * np.sqrt(chi_square * eigval[0])                           # :185
        minor = 2 * np.sqrt(chi_square * eigval[1])                           # :186
        with np.errstate(divide='ignore'):
            if eigval[0] > eigval[1]:                                         # :188
                arc = np.arctan(np.divide(eigvec[1, 0], eigvec[0, 0]))        # :189
                ax, ay = major, minor                                         # :190-191
            else:
                arc = np.arctan(np.divide(eigvec[1, 1], eigvec[0, 1]))        # :193
                ax, ay = minor, major                                         # :194-195
        scale = 0.25 if b == 0 else 0.50                                      # :198 / :221
        out.append([mu[0], mu[1], ax * scale, ay * scale, arc])
    return np.array(out)
