"""Host-side mirror of the reference's class interface over the C ABI.

Same names, argument order and public properties as the MATLAB classes:
``EKF_SLAM`` (EKF_SLAM.m:1-151), ``EKF_SLAM_UC`` (EKF_SLAM_UC.m:1-152),
``Correspondence`` (Correspondence.m:1-92), ``Landmark`` (Landmark.m:1-37).
All filter arithmetic runs in libekfslam_b200.so (CUDA, sm_100a); this file only
marshals arguments.  ``plot`` is a no-op: no graphics stack (unavailable on this box).
"""
import ctypes as ct

import numpy as np

from . import _lib as L


def _f64(a, n=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))
    if n is not None and a.size < n:
        raise ValueError('expected at least %d elements, got %d' % (n, a.size))
    return a


def _rdiag(R):
    R = np.asarray(R, dtype=np.float64)
    if R.shape == (2, 2):
        return np.array([R[0, 0], R[1, 1]])      # the reference only ever builds diagonal R (EKF_SLAM.m:108)
    return _f64(R, 2)[:2].copy()


class _FilterBase:
    _VARIANT = L.VARIANT_KNOWN

    def __init__(self, capacity_landmarks=1024, storage='full', device=0, stream=0,
                 assoc_mode='signature', defer_m=0, rank=0, world=1, nccl_id=None, devices=None):
        self._lib = L.load()
        p = L.EkfParams()
        L.check(None, self._lib.ekf_params_default(ct.byref(p), self._VARIANT))
        p.capacity_landmarks = int(capacity_landmarks)
        p.storage = L.STORAGE_SYM if storage in ('sym', L.STORAGE_SYM) else L.STORAGE_FULL
        p.assoc_mode = L.ASSOC_POSITION if assoc_mode in ('position', L.ASSOC_POSITION) else L.ASSOC_SIGNATURE
        p.device = int(device)
        p.defer_m = int(defer_m)
        p.stream = int(stream)
        self._p = p
        self._h = L._H()
        if devices is not None and len(devices) > 1:       # one process, several GPUs: column slabs behind ONE handle
            dv = np.ascontiguousarray(devices, dtype=np.int32)
            rc = self._lib.ekf_create_multi(ct.byref(p), int(dv.size), L.iptr(dv), ct.byref(self._h))
        elif world > 1:
            idbuf = (ct.c_uint8 * 128).from_buffer_copy(bytes(nccl_id))
            rc = self._lib.ekf_create_sharded(ct.byref(p), int(rank), int(world), idbuf, ct.byref(self._h))
        else:
            rc = self._lib.ekf_create(ct.byref(p), ct.byref(self._h))
        self.rank, self.world = int(rank), int(world)
        if rc != L.EKF_OK:
            raise L.EkfError(rc, 'ekf_create failed (CUDA device required; there is no CPU fallback)')
        self._C = p.C
        self._Rc = [p.Rc[0], p.Rc[1]]
        self._s_cost = p.s_cost
        self._s_thresh = p.s_thresh
        self.landmark_list = None          # EKF_SLAM.m:18
        self.observed = None               # EKF_SLAM.m:19
        self._chosen = []                  # per observation: ('update'|'append', idx) -- parity log
        self._pending_act = None           # actions of the last measure(), filled by the library at its next call
        self._last_predict = None

    # ---- lifetime ------------------------------------------------------
    def close(self):
        if getattr(self, '_h', None):
            self._pending_act = None
            self._lib.ekf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        L.check(self._h, rc)

    def _collect_actions(self):
        """measure() does not wait for the device's decisions (ekf_measure_batch_async); fetch them when somebody looks."""
        act, self._pending_act = self._pending_act, None
        if act is not None:
            self._ck(self._lib.ekf_measure_wait(self._h))
            for a in act:
                self._chosen.append(('update', int(a)) if a > 0 else ('append', int(-a)))

    @property
    def chosen(self):
        self._collect_actions()
        return self._chosen

    # ---- public properties of the reference classes (EKF_SLAM.m:5-22) ----
    @property
    def N(self):
        n = ct.c_int32()
        self._ck(self._lib.ekf_n_landmarks(self._h, ct.byref(n)))
        return n.value

    @property
    def x(self):
        n = 3 + 2 * self.N
        out = np.empty(n)
        self._ck(self._lib.ekf_get_x(self._h, L.dptr(out), n))
        return out                         # row vector 1 x n (EKF_SLAM.m:28)

    @x.setter
    def x(self, v):
        """h.x = v (public read/write upstream, EKF_SLAM.m:6): same length as the current state; P and s untouched."""
        v = _f64(v)
        self._ck(self._lib.ekf_set_x(self._h, L.dptr(v), v.size))

    @property
    def P(self):
        n = 3 + 2 * self.N
        out = np.empty((n, n), order='F')
        self._ck(self._lib.ekf_get_P(self._h, L.dptr(out), n))
        return out

    @P.setter
    def P(self, M):
        """h.P = M (EKF_SLAM.m:7): n x n at the current n; x and s untouched."""
        M = np.asfortranarray(np.asarray(M, dtype=np.float64))
        n = 3 + 2 * self.N
        if M.shape != (n, n):
            raise ValueError('P must be %d x %d (use set_state to change the number of landmarks)' % (n, n))
        self._ck(self._lib.ekf_set_P(self._h, L.dptr(M), n))

    @property
    def s(self):
        N = self.N
        out = np.empty(N)
        self._ck(self._lib.ekf_get_s(self._h, L.dptr(out), N))
        return out

    @s.setter
    def s(self, v):
        """h.s = v (EKF_SLAM.m:9): N signatures; x and P untouched."""
        v = _f64(v) if np.size(v) else np.zeros(0)
        self._ck(self._lib.ekf_set_s(self._h, L.dptr(v) if v.size else None, v.size))

    def P_block(self, r0, c0, nr, nc):
        out = np.empty((nr, nc), order='F')
        self._ck(self._lib.ekf_get_P_block(self._h, r0, c0, nr, nc, L.dptr(out), nr))
        return out

    def set_state(self, x, P, s):
        x = _f64(x)
        N = (x.size - 3) // 2
        P = np.asfortranarray(np.asarray(P, dtype=np.float64))
        s = _f64(s) if N > 0 else np.zeros(1)
        self._ck(self._lib.ekf_set_state(self._h, L.dptr(x), L.dptr(P), P.shape[0], L.dptr(s), N))

    def set_state_lowrank(self, x, s, A, d):
        x = _f64(x)
        N = (x.size - 3) // 2
        A = np.asfortranarray(np.asarray(A, dtype=np.float64))
        self._ck(self._lib.ekf_set_state_lowrank(self._h, L.dptr(x), L.dptr(_f64(s)), N, L.dptr(A),
                                                 A.shape[1], L.dptr(_f64(d))))

    @property
    def Q(self):
        """Process-noise matrix of the last predict (EKF_SLAM.m:43-44): zeros(n) with a 3x3 corner."""
        if self._last_predict is None:
            return None
        u, th = self._last_predict
        if th is None:
            raise RuntimeError('Q needs track_Q=True on predict (heading before the motion step)')
        n = 3 + 2 * self.N
        W = np.array([[u[0] * _cosd(th)], [u[0] * _sind(th)], [u[1]]])
        Q = np.zeros((n, n))
        Q[0:3, 0:3] = (W * self._C) @ W.T
        return Q

    def _push_noise(self):
        rc2 = (ct.c_double * 2)(*self._Rc)
        self._ck(self._lib.ekf_set_noise(self._h, self._C, rc2, self._s_cost, self._s_thresh))

    C = property(lambda self: self._C, lambda self, v: (setattr(self, '_C', float(v)), self._push_noise())[0])
    Rc = property(lambda self: list(self._Rc),
                  lambda self, v: (setattr(self, '_Rc', [float(v[0]), float(v[1])]), self._push_noise())[0])

    # ---- methods -------------------------------------------------------
    def predict(self, u, track_Q=False):
        """predict(h,u) -- EKF_SLAM.m:40-51."""
        u = _f64(u, 2)
        th = float(self.x[2]) if track_Q else None
        self._ck(self._lib.ekf_predict(self._h, L.dptr(u)))
        self._last_predict = (u.copy(), th)

    def f(self, x, u):
        """[x_new,F] = f(h,x,u) -- EKF_SLAM.m:56-65.  Pure host helper (the device never forms dense F)."""
        x = np.asarray(x, dtype=np.float64).reshape(-1)
        x_new = x.copy()
        x_new[0] = x[0] + u[0] * _cosd(x[2] + u[1])
        x_new[1] = x[1] + u[0] * _sind(x[2] + u[1])
        x_new[2] = x[2] + u[1]
        F = np.eye(x.size)
        F[0, 2] = -1 * u[0] * _sind(x[2])
        F[1, 2] = u[0] * _cosd(x[2])
        return x_new, F

    def append(self, u, R, landmarkPos, signature):
        """append(h,u,R,landmarkPos,signature) -- EKF_SLAM.m:67-98."""
        self._ck(self._lib.ekf_append(self._h, L.dptr(_f64(u, 2)), L.dptr(_rdiag(R)),
                                      L.dptr(_f64(landmarkPos, 2)), float(signature)))

    def update(self, z, R, idx):
        """Update branch of measure for one observation -- EKF_SLAM.m:124-145."""
        self._ck(self._lib.ekf_update(self._h, L.dptr(_f64(z, 2)), L.dptr(_rdiag(R)), int(idx)))

    def update_batch(self, z, Rdiag, idx):
        """k updates with explicit slots (z, Rdiag: k x 2; idx: k, 1-based), then a flush."""
        z = np.asfortranarray(np.asarray(z, dtype=np.float64).reshape(-1, 2))
        Rd = np.asfortranarray(np.asarray(Rdiag, dtype=np.float64).reshape(-1, 2))
        ix = np.ascontiguousarray(np.asarray(idx, dtype=np.int32))
        self._ck(self._lib.ekf_update_batch(self._h, z.shape[0], L.dptr(z), L.dptr(Rd), L.iptr(ix)))

    def shard_info(self):
        """(rank, world, c_lo, c_hi): this rank stores columns [c_lo, c_hi) of P."""
        v = [ct.c_int32() for _ in range(4)]
        self._ck(self._lib.ekf_shard_info(self._h, *[ct.byref(q) for q in v]))
        return tuple(q.value for q in v)

    def P_slab(self):
        """Active part of this rank's column slab, n x (columns owned and < n), plus its first column."""
        _, _, c_lo, c_hi = self.shard_info()
        n = 3 + 2 * self.N
        hi = min(c_hi, n)
        if hi <= c_lo:
            return np.zeros((n, 0)), c_lo
        return self.P_block(0, c_lo, n, hi - c_lo), c_lo

    def _tables(self, landmark_list):
        # Rebuilt on every measure(): the reference's landmark DB mutates entries in place without changing the list or
        # its length (RANSAC.m:262 promotes index 0 to max+1, :268 rewrites .loc), so identity/length is no cache key.
        # A caller whose table is immutable can hand over a `tables()` method returning (index int32[T], loc float64[T,2]).
        fast = getattr(landmark_list, 'tables', None)
        if fast is not None:
            idx, loc = fast()
            return np.ascontiguousarray(idx, dtype=np.int32), np.asfortranarray(loc, dtype=np.float64).reshape(-1, 2, order='F')
        lst = landmark_list.landmarkObj.landmark
        idx = np.fromiter((int(e.index) for e in lst), dtype=np.int32, count=len(lst))
        loc = np.asfortranarray(np.array([e.loc for e in lst], dtype=np.float64).reshape(-1, 2))
        return idx, loc

    def measure(self, laserData, u, landmark_list, pass_x=True):
        """measure(h,laserData,u,landmark_list) -- EKF_SLAM.m:100-151 / EKF_SLAM_UC.m:102-152."""
        observed_LL = landmark_list.getLandmark(laserData, self.x if pass_x else None)   # :102
        self.observed = observed_LL                                                      # :103
        if observed_LL is None or len(observed_LL) == 0:                                 # :105
            return
        obs = np.asfortranarray(np.asarray(observed_LL, dtype=np.float64).reshape(-1, 3))
        k = obs.shape[0]
        idx, loc = self._tables(landmark_list)
        act = np.zeros(k, dtype=np.int32)
        n_tab = int(idx.size)
        if n_tab == 0:
            idx = np.zeros(1, dtype=np.int32)
            loc = np.zeros((1, 2), order='F')
        self._collect_actions()            # the previous call's (the library writes them at its next call: read first)
        rc = self._lib.ekf_measure_batch_async(self._h, L.dptr(obs), k, L.dptr(_f64(u, 2)), L.dptr(loc),
                                               L.iptr(idx), n_tab, L.iptr(act))
        self._pending_act = act            # keeps the buffer alive until the actions are in
        self._ck(rc)

    def plot(self, landmark_list=None):
        """EKF_SLAM.m:154-234 draws MATLAB figures; unavailable on this box (no graphics stack).  The numbers it
        would draw are available from covariance_ellipses()."""
        return None

    def covariance_ellipses(self):
        """The ellipses plot() draws (EKF_SLAM.m:180-229) as an (N+1) x 5 array [cx, cy, semi-axis along the major
        eigenvector, semi-axis across, arc_rad]; row 0 = robot (scale .25), row b = landmark b (scale .50)."""
        nb = self.N + 1
        out = np.empty((nb, 5), order='F')
        self._ck(self._lib.ekf_get_ellipses(self._h, L.dptr(out), nb))
        return out

    def flush(self):
        self._ck(self._lib.ekf_flush(self._h))

    def flush_info(self):
        """(continuous flush in use at the current size, K3 passes outstanding on its stream)."""
        a, b = ct.c_int32(), ct.c_int32()
        self._ck(self._lib.ekf_flush_info(self._h, ct.byref(a), ct.byref(b)))
        return bool(a.value), b.value

    def pend_cap(self):
        """Rank-2 terms that may be pending before a pass over P is forced (what defer_m became on this handle)."""
        b, c = ct.c_int32(), ct.c_int32()
        self._ck(self._lib.ekf_defer_info(self._h, ct.byref(b), ct.byref(c)))
        return c.value

    def sync(self):
        self._ck(self._lib.ekf_sync(self._h))

    # ---- measurement hooks (bench.py) -----------------------------------
    def profile_enable(self, on=True):
        self._ck(self._lib.ekf_profile_enable(self._h, 1 if on else 0))

    def profile_read(self):
        ms = ct.c_double()
        nd = ct.c_int64()
        nl = ct.c_int64()
        self._ck(self._lib.ekf_profile_read(self._h, ct.byref(ms), ct.byref(nd), ct.byref(nl)))
        return ms.value, nd.value, nl.value

    def profile_exchange(self):
        """(total ms, count, p2p flag) of the sharded column exchanges since profiling was enabled."""
        ms, n, p = ct.c_double(), ct.c_int64(), ct.c_int32()
        self._ck(self._lib.ekf_profile_exchange(self._h, ct.byref(ms), ct.byref(n), ct.byref(p)))
        return ms.value, n.value, bool(p.value)

    def downdate_bytes(self):
        b = ct.c_double()
        self._ck(self._lib.ekf_downdate_bytes(self._h, ct.byref(b)))
        return b.value


def _reduce90(x):
    x = np.fmod(float(x), 360.0)
    n = np.rint(x / 90.0)
    return int(n) & 3, (x - 90.0 * n) * (np.pi / 180.0)


def _sind(x):
    m, a = _reduce90(x)
    return float((np.sin(a), np.cos(a), -np.sin(a), -np.cos(a))[m])


def _cosd(x):
    m, a = _reduce90(x)
    return float((np.cos(a), -np.sin(a), -np.cos(a), np.sin(a))[m])


class EKF_SLAM(_FilterBase):
    """Known-correspondence filter, EKF_SLAM.m (Rc=[.01,5], idx=ii quirk kept)."""
    _VARIANT = L.VARIANT_KNOWN

    s_cost = property(lambda self: self._s_cost,
                      lambda self, v: (setattr(self, '_s_cost', float(v)), self._push_noise())[0])
    s_thresh = property(lambda self: self._s_thresh,
                        lambda self, v: (setattr(self, '_s_thresh', float(v)), self._push_noise())[0])


class Correspondence:
    """Correspondence.m:1-92.  Value class: estimateCorrespondence(z,R,x,P,s) takes the state explicitly.

    When bound to a filter (``EKF_SLAM_UC.correspondence``) and called with x=P=s=None it scores against the
    device-resident state without any copy; with explicit arrays it uploads them into a scratch handle.
    """

    def __init__(self, cost, thresh, method, _owner=None, assoc_mode='signature', storage='full', device=0):
        self.method = method if method == 'EKF_SLAM_UC' else 'ML'     # :14-24 (fallback with a warning upstream)
        self.s_cost = cost
        self.s_thresh = thresh
        self._owner = _owner
        self._assoc_mode = assoc_mode
        self._storage = storage
        self._device = device

    def estimateCorrespondence(self, z, R, x=None, P=None, s=None):
        z = _f64(z, 3)
        Rd = _rdiag(R)
        new = ct.c_int32()
        idx = ct.c_int32()
        if x is None and self._owner is not None:
            flt = self._owner
        else:
            x = _f64(x)
            N = (x.size - 3) // 2
            flt = EKF_SLAM_UC(capacity_landmarks=max(N, 1), storage=self._storage, device=self._device,
                              assoc_mode=self._assoc_mode)
            flt.correspondence.s_cost = self.s_cost
            flt.correspondence.s_thresh = self.s_thresh
            flt._sync_corr()
            flt.set_state(x, P, s)
        flt._ck(flt._lib.ekf_associate(flt._h, L.dptr(z), L.dptr(Rd), ct.byref(new), ct.byref(idx)))
        if flt is not self._owner:
            flt.close()
        return bool(new.value), int(idx.value)


class EKF_SLAM_UC(_FilterBase):
    """Unknown-correspondence filter, EKF_SLAM_UC.m (Rc=[.1,5], owns a Correspondence, :16)."""
    _VARIANT = L.VARIANT_UC

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.correspondence = Correspondence(0.00000000001, 1000000000, 'EKF_SLAM_UC', _owner=self)

    def _sync_corr(self):
        self._s_cost = float(self.correspondence.s_cost)
        self._s_thresh = float(self.correspondence.s_thresh)
        self._push_noise()

    def measure(self, laserData, u, landmark_list, pass_x=True):
        if (self.correspondence.s_cost, self.correspondence.s_thresh) != (self._s_cost, self._s_thresh):
            self._sync_corr()
        super().measure(laserData, u, landmark_list, pass_x)


class Landmark:
    """Landmark.m:1-37 interface: ``.method``, ``.landmarkObj``, ``getLandmark(laserdata,x)``.

    The reference's only extractor (RANSAC over ROS LaserScan) is out of scope; 'SYNTH' wraps a
    synthetic sensor object exposing the same two members (see synth.LandmarkStub)."""

    def __init__(self, method, landmarkObj=None):
        self.method = method
        self.landmarkObj = landmarkObj

    def getLandmark(self, laserdata, x):
        return self.landmarkObj.getLandmark(laserdata, x)
