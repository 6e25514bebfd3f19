"""Seeded synthetic range-bearing worlds + the duck-typed ``landmark_list`` stub.

The reference's filter receives its observations through
``landmark_list.getLandmark(laserData, x)`` -> ``observed_LL`` rows
``[range, bearing_deg, index]`` and reads ``landmark_list.landmarkObj.landmark(k).{index,loc}``
(EKF_SLAM.m:100-103,111,120; Landmark.m:25-33; row format RANSAC.m:275-283).
RANSAC/ROS are out of scope, so this module supplies that seam with synthetic
data.  Host-side NumPy only; no filter arithmetic here.
"""
import math
from types import SimpleNamespace

import numpy as np

DEFAULT_SEED = 20261104


def _wrap360(a):
    r = math.fmod(a, 360.0)
    if r < 0.0:
        r += 360.0
    if r == 0.0 and a > 0.0:
        r = 360.0
    return r


class LandmarkStub:
    """Stand-in for the reference's ``Landmark`` object (Landmark.m:1-37).

    ``getLandmark(laserdata, x)`` returns the pre-generated ``observed_LL`` (k x 3);
    ``landmarkObj.landmark`` is a list of entries with ``.index`` and ``.loc``
    (fields of RANSAC.m:238-241 that the filter reads).
    """

    def __init__(self, observed_LL, table, method='SYNTH'):
        self.method = method
        self._observed = observed_LL
        self.landmarkObj = SimpleNamespace(landmark=table)

    def getLandmark(self, laserdata, x):
        return self._observed

    def tables(self):
        """(index int32[T], loc float64[T,2]) of the landmark table.  The synthetic table never changes after it was
        built, so the flattened arrays are cached per table list; the filter classes use this optional method when a
        landmark_list offers it and otherwise flatten landmarkObj.landmark on every measure() (the reference's RANSAC
        database mutates entries in place)."""
        lst = self.landmarkObj.landmark
        hit = _TABLES.get(id(lst))
        if hit is None or hit[0] is not lst or hit[1] != len(lst):
            idx = np.fromiter((int(e.index) for e in lst), dtype=np.int32, count=len(lst))
            loc = np.asfortranarray(np.array([e.loc for e in lst], dtype=np.float64).reshape(-1, 2))
            if len(_TABLES) >= 8:
                _TABLES.clear()
            hit = (lst, len(lst), idx, loc)
            _TABLES[id(lst)] = hit
        return hit[2], hit[3]


_TABLES = {}


class World:
    """A pre-generated trajectory: per step ``u`` (2,), ``observed_LL`` (k_t x 3)."""

    def __init__(self, N, L, landmarks, table, u, observed, n_table, true_pose):
        self.N = N
        self.L = L
        self.landmarks = landmarks      # (N,2) true positions
        self.table = table              # entries in assigned-index order
        self.u = u                      # (T,2)
        self.observed = observed        # list of (k_t,3) arrays
        self.n_table = n_table          # table length visible at step t
        self.true_pose = true_pose      # (T,3)

    @property
    def T(self):
        return len(self.observed)

    def stub(self, t):
        return LandmarkStub(self.observed[t], self.table[:self.n_table[t]])

    def table_arrays(self, t=None):
        """(index_table int32[n], loc_table float64[n,2]) for the C-ABI batch call."""
        tab = self.table if t is None else self.table[:self.n_table[t]]
        idx = np.array([e.index for e in tab], dtype=np.int32)
        loc = np.array([e.loc for e in tab], dtype=np.float64).reshape(-1, 2)
        return idx, loc


def make_world(N=20, T=1000, k=5, seed=DEFAULT_SEED, mode='nearest', seam_deg=20.0,
               sensor_range=None):
    """Build a world per SURVEY.md 8(d).

    mode 'nearest': each step observes the k landmarks nearest the true pose, drops
        those whose relative bearing is within ``seam_deg`` of the 0/360 seam (the
        reference does not wrap the innovation, EKF_SLAM.m:144), rows sorted by
        assigned index.  Indices are assigned in order of first sighting, which is
        how RANSAC numbers landmarks (RANSAC.m:234-334) and what EKF_SLAM_UC's
        signature test (Correspondence.m:71) needs.
    mode 'prefix': rows for landmarks 1..m in index order, m = longest prefix with
        no landmark in the seam zone, at most one landmark beyond those already
        emitted (the known-correspondence filter updates slot ``ii`` = row number,
        EKF_SLAM.m:123, so row ii must carry landmark ii).
    """
    rng = np.random.default_rng(seed)
    L = 10.0 * math.sqrt(N / 20.0)
    Rr = 0.35 * L
    dth = 1.0
    dD = 2.0 * math.pi * Rr / 360.0
    # robot starts at the filter's initial state [0,0,0] (EKF_SLAM.m:28); circle centre (0,Rr)
    lms = np.column_stack([rng.uniform(-L / 2, L / 2, N), Rr + rng.uniform(-L / 2, L / 2, N)])
    loc_noise = rng.normal(0.0, 0.05, (N, 2))
    assigned = np.zeros(N, dtype=np.int64)       # 0 = not yet indexed
    table = []
    pose = np.zeros(3)
    us = np.zeros((T, 2))
    poses = np.zeros((T, 3))
    observed = []
    n_table = []
    n_emitted = 0                                 # prefix mode: landmarks emitted so far
    for t in range(T):
        # true motion follows the reference's model f (EKF_SLAM.m:57-60)
        th = pose[2] + dth
        pose = np.array([pose[0] + dD * math.cos(math.radians(th)),
                         pose[1] + dD * math.sin(math.radians(th)), th])
        poses[t] = pose
        us[t] = [dD + rng.normal(0.0, 0.01 * dD), dth + rng.normal(0.0, 0.05)]
        d = lms - pose[0:2]
        rng_true = np.hypot(d[:, 0], d[:, 1])
        if mode == 'nearest':
            kk = min(k, N)
            cand = np.argpartition(rng_true, kk - 1)[:kk] if kk < N else np.arange(N)
            cand = cand[np.argsort(rng_true[cand], kind='stable')]
        elif mode == 'prefix':
            # ids in a fixed global order; at most one new landmark per step
            cand = np.arange(min(n_emitted + 1, N))
        else:
            raise ValueError(mode)
        rows = []
        for j in cand:
            r = rng_true[j] + rng.normal(0.0, 0.01)
            b = _wrap360(math.degrees(math.atan2(d[j, 1], d[j, 0])) - pose[2] + rng.normal(0.0, 0.1))
            in_seam = (b < seam_deg) or (b > 360.0 - seam_deg)
            if sensor_range is not None and rng_true[j] > sensor_range:
                in_seam = True
            if in_seam:
                if mode == 'prefix':
                    break                       # truncate at the first seam landmark
                continue
            if assigned[j] == 0:
                assigned[j] = len(table) + 1
                table.append(SimpleNamespace(index=int(assigned[j]), loc=lms[j] + loc_noise[j],
                                             true_id=int(j)))
            rows.append([r, b, float(assigned[j])])
        if mode == 'prefix':
            n_emitted = max(n_emitted, len(rows))
        rows.sort(key=lambda q: q[2])
        observed.append(np.array(rows, dtype=np.float64).reshape(-1, 3))
        n_table.append(len(table))
    return World(N, L, lms, table, us, observed, n_table, poses)


def synthetic_spd_state(N, seed=DEFAULT_SEED, rank=8):
    """Fully-populated map for the throughput configs (SURVEY.md 8d "Map initialisation").

    P = A A^T / r + diag(d) with A n x r ~ N(0,1), d ~ U(0.01,0.1), robot block 0.1 I;
    x = [pose, landmark truth + noise]; s = 1..N.  Returns (x, A, d, landmarks): P is
    NOT materialised here (3.2 GB at 10k) -- callers build it on the device
    (ekf_set_state_lowrank) or on the host from (A, d) for the CPU baseline.
    """
    rng = np.random.default_rng(seed)
    n = 3 + 2 * N
    L = 10.0 * math.sqrt(N / 20.0)
    Rr = 0.35 * L
    lms = np.column_stack([rng.uniform(-L / 2, L / 2, N), Rr + rng.uniform(-L / 2, L / 2, N)])
    A = rng.standard_normal((n, rank)) * 0.05
    d = rng.uniform(0.01, 0.1, n)
    d[0:3] = 0.1
    x = np.zeros(n)
    x[3:] = (lms + rng.normal(0.0, 0.05, (N, 2))).reshape(-1)
    s = np.arange(1, N + 1, dtype=np.float64)
    return x, A, d, s, lms


def lowrank_to_dense(A, d):
    """Host materialisation of P = A A^T / r + diag(d) (CPU baseline / tests only)."""
    r = A.shape[1]
    P = (A @ A.T) / r
    P[np.diag_indices_from(P)] += d
    return P


def observe_from_truth(lms, pose, ids, rng, seam_deg=20.0):
    """Range-bearing rows [r, b, index] for landmark ids (0-based) seen from ``pose``;
    rows inside the seam zone are dropped.  Used by bench.py for the populated-map runs."""
    rows = []
    for j in ids:
        dx, dy = lms[j, 0] - pose[0], lms[j, 1] - pose[1]
        r = math.hypot(dx, dy) + rng.normal(0.0, 0.01)
        b = _wrap360(math.degrees(math.atan2(dy, dx)) - pose[2] + rng.normal(0.0, 0.1))
        if b < seam_deg or b > 360.0 - seam_deg:
            continue
        rows.append([r, b, float(j + 1)])
    return np.array(rows, dtype=np.float64).reshape(-1, 3)
