"""TEST INFRASTRUCTURE ONLY -- ctypes face of oracle/structured.c (the O(n^2)-per-update C oracle).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
Same class surface as oracle/dense_literal.py (EKF_SLAM, EKF_SLAM_UC: predict / append / measure, x, P, s, chosen),
so a test can drive both oracles and the GPU classes with the same loop.  State lives in caller-visible NumPy
arrays at a fixed capacity (column-major P), which is what makes the 10 000-landmark configuration affordable.

parity unpinned (see structured.c).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_D = C.POINTER(C.c_double)
_I = C.POINTER(C.c_int32)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, 'libekf_oracle.so')
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, 'structured.c')):
            subprocess.check_call(['make', '-C', _HERE])
        L = C.CDLL(path)
        L.so_predict.argtypes = [_D, _D, C.c_int64, C.c_int32, _D, C.c_double]
        L.so_predict.restype = None
        L.so_update.argtypes = [_D, _D, C.c_int64, C.c_int32, _D, _D, C.c_int32, _D]
        L.so_update.restype = C.c_int32
        L.so_append.argtypes = [_D, _D, C.c_int64, _D, C.c_int32, _D, _D, _D, C.c_double]
        L.so_append.restype = None
        L.so_associate.argtypes = [_D, _D, C.c_int64, _D, C.c_int32, _D, _D, C.c_double, C.c_double, C.c_int32, _I]
        L.so_associate.restype = C.c_int32
        L.so_measure.argtypes = [_D, _D, C.c_int64, _D, _I, C.c_int32, C.c_int32, C.c_int32, _D, C.c_int32, _D, _D, _I,
                                 C.c_int32, _D, C.c_double, C.c_double, _I, _D]
        L.so_measure.restype = C.c_int32
        for f in ('so_sind', 'so_cosd', 'so_wrapTo360'):
            getattr(L, f).argtypes = [C.c_double]
            getattr(L, f).restype = C.c_double
        L.so_set_threads.argtypes = [C.c_int32]
        L.so_set_threads.restype = None
        L.so_set_threads(min(16, os.cpu_count() or 1))
        L.so_atan2d.argtypes = [C.c_double, C.c_double]
        L.so_atan2d.restype = C.c_double
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(_D)


class EKF_SLAM:
    """EKF_SLAM.m:1-151 in structured form (variant 0: known correspondence, idx = ii quirk kept)."""
    VARIANT = 0

    def __init__(self, capacity_landmarks=64, assoc_mode=0):
        self.cap = int(capacity_landmarks)
        self.n_cap = 3 + 2 * self.cap
        self.C = 0.2                                   # EKF_SLAM.m:12
        self.Rc = [0.01, 5.0]                          # :13
        self.s_cost = 0.00000000001                    # :15
        self.s_thresh = 1000000000                     # :16
        self.assoc_mode = assoc_mode
        self._x = np.zeros(self.n_cap)
        self._P = np.zeros((self.n_cap, self.n_cap), order='F')
        self._P[0, 0] = self._P[1, 1] = self._P[2, 2] = 0.1       # :30
        self._s = np.zeros(self.cap)
        self.N = 0
        self.chosen = []
        self.observed = None
        self._work = np.zeros(4 * self.n_cap)

    # state views (copies of the live part)
    @property
    def n(self):
        return 3 + 2 * self.N

    @property
    def x(self):
        return self._x[:self.n].copy()

    @property
    def P(self):
        return np.array(self._P[:self.n, :self.n])

    @property
    def s(self):
        return self._s[:self.N].copy()

    def set_state(self, x, P, s):
        N = len(s)
        assert N <= self.cap and len(x) == 3 + 2 * N
        self.N = N
        self._x[:len(x)] = x
        self._P[:len(x), :len(x)] = P
        self._s[:N] = s

    def predict(self, u):
        u = np.ascontiguousarray(u[:2], dtype=np.float64)
        lib().so_predict(_p(self._x), _p(self._P), self.n_cap, self.N, _p(u), self.C)

    def update(self, z, Rdiag, idx1):
        z = np.ascontiguousarray(z, dtype=np.float64)
        R = np.ascontiguousarray(Rdiag, dtype=np.float64)
        rc = lib().so_update(_p(self._x), _p(self._P), self.n_cap, self.N, _p(z), _p(R), int(idx1), _p(self._work))
        if rc != 0:
            raise IndexError('slot %d of %d' % (idx1, self.N))

    def append(self, u, R, landmarkPos, signature):
        if self.N >= self.cap:
            raise MemoryError('capacity')
        u = np.ascontiguousarray(u[:2], dtype=np.float64)
        Rd = np.ascontiguousarray([R[0][0], R[1][1]] if np.ndim(R) == 2 else R, dtype=np.float64)
        pos = np.ascontiguousarray(landmarkPos[:2], dtype=np.float64)
        lib().so_append(_p(self._x), _p(self._P), self.n_cap, _p(self._s), self.N, _p(u), _p(Rd), _p(pos), float(signature))
        self.N += 1

    def associate(self, z, Rdiag):
        z = np.ascontiguousarray(z, dtype=np.float64)
        R = np.ascontiguousarray(Rdiag, dtype=np.float64)
        new = C.c_int32(0)
        idx = lib().so_associate(_p(self._x), _p(self._P), self.n_cap, _p(self._s), self.N, _p(z), _p(R),
                                 self.s_cost, self.s_thresh, self.assoc_mode, C.byref(new))
        return bool(new.value), int(idx)

    def measure(self, laserData, u, landmark_list):
        observed_LL = landmark_list.getLandmark(laserData, self.x)            # EKF_SLAM.m:102
        self.observed = observed_LL
        if observed_LL is None or len(observed_LL) == 0:                      # :105
            return
        obs = np.asfortranarray(np.asarray(observed_LL, dtype=np.float64).reshape(-1, 3))
        k = obs.shape[0]
        lm = landmark_list.landmarkObj.landmark
        ntab = len(lm)
        loc = np.asfortranarray(np.array([l.loc[:2] for l in lm], dtype=np.float64).reshape(ntab, 2))
        tidx = np.ascontiguousarray([int(l.index) for l in lm], dtype=np.int32)
        u = np.ascontiguousarray(u[:2], dtype=np.float64)
        Rc = np.ascontiguousarray(self.Rc, dtype=np.float64)
        actions = np.zeros(k, dtype=np.int32)
        N = C.c_int32(self.N)
        rc = lib().so_measure(_p(self._x), _p(self._P), self.n_cap, _p(self._s), C.byref(N), self.cap, self.VARIANT,
                              self.assoc_mode, _p(obs), k, _p(u), _p(loc), tidx.ctypes.data_as(_I), ntab, _p(Rc),
                              self.s_cost, self.s_thresh, actions.ctypes.data_as(_I), _p(self._work))
        self.N = N.value
        for a in actions:
            if a > 0:
                self.chosen.append(('update', int(a)))
            elif a < 0:
                self.chosen.append(('append', int(-a)))
        if rc != 0:
            raise {-1: IndexError, -2: IndexError, -3: MemoryError}[rc]('so_measure status %d' % rc)


class EKF_SLAM_UC(EKF_SLAM):
    """EKF_SLAM_UC.m:1-152 in structured form (variant 1: association through Correspondence.m:28-88)."""
    VARIANT = 1

    def __init__(self, capacity_landmarks=64, assoc_mode=0):
        super().__init__(capacity_landmarks, assoc_mode)
        self.Rc = [0.1, 5.0]                           # EKF_SLAM_UC.m:13
