"""Monte-Carlo batch of independent filters over the ekfb_* C ABI (BASELINE config 5).

``FilterBatch('EKF_SLAM_UC', n_filters, ...)`` stands for n_filters reference objects driven in lock step:
``step(u, observed, tables)`` == ``for f: slam[f].predict(u[f]); slam[f].measure(laserData, u[f], LM[f])``
(SLAM.m:110,116), executed by one CUDA block per filter with no host round trip.  ``batch[f]`` is a read-only view
with the reference's property names (x, P, s).  Across GPUs each rank builds its own FilterBatch over its share of
the filters (``shard_filters``): there is no communication.
"""
import ctypes as ct

import numpy as np

from . import _lib as L


def shard_filters(n_filters, rank, world):
    """Global filter ids owned by `rank` (filter f lives on rank f mod world)."""
    return list(range(rank, n_filters, world))


class _View:
    def __init__(self, batch, f):
        self._b, self._f = batch, f

    @property
    def N(self):
        return int(self._b.n_landmarks()[self._f])

    @property
    def x(self):
        n = 3 + 2 * self.N
        out = np.empty(n)
        self._b._ck(self._b._lib.ekfb_get_x(self._b._h, self._f, L.dptr(out), n))
        return out

    @property
    def s(self):
        N = self.N
        out = np.empty(max(N, 1))
        self._b._ck(self._b._lib.ekfb_get_s(self._b._h, self._f, L.dptr(out), N))
        return out[:N]

    @property
    def P(self):
        n = 3 + 2 * self.N
        out = np.empty((n, n), order='F')
        self._b._ck(self._b._lib.ekfb_get_P(self._b._h, self._f, L.dptr(out), n))
        return out


class FilterBatch:
    def __init__(self, cls='EKF_SLAM_UC', n_filters=1, capacity_landmarks=200, max_obs=8, max_table=None, storage='sym',
                 assoc_mode='signature', defer_m=8, device=0, stream=0):
        self._lib = L.load()
        variant = L.VARIANT_UC if cls in ('EKF_SLAM_UC', L.VARIANT_UC) else L.VARIANT_KNOWN
        p = L.EkfParams()
        L.check(None, self._lib.ekf_params_default(ct.byref(p), variant))
        p.capacity_landmarks = int(capacity_landmarks)
        p.storage = L.STORAGE_SYM if storage in ('sym', L.STORAGE_SYM) else L.STORAGE_FULL
        p.assoc_mode = L.ASSOC_POSITION if assoc_mode in ('position', L.ASSOC_POSITION) else L.ASSOC_SIGNATURE
        p.device, p.defer_m, p.stream = int(device), int(defer_m), int(stream)
        self.params = p
        self.n_filters, self.max_obs = int(n_filters), int(max_obs)
        self.max_table = int(max_table if max_table is not None else capacity_landmarks)
        self._h = L._H()
        rc = self._lib.ekfb_create(ct.byref(p), self.n_filters, self.max_obs, self.max_table, ct.byref(self._h))
        if rc != L.EKF_OK:
            raise L.EkfError(rc, 'ekfb_create failed (CUDA device required; there is no CPU fallback)')

    def _ck(self, rc):
        if rc != L.EKF_OK:
            raise L.EkfError(rc, (self._lib.ekfb_last_error(self._h) or b'').decode('utf-8', 'replace'))

    def close(self):
        if getattr(self, '_h', None):
            self._lib.ekfb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self):
        return self.n_filters

    def __getitem__(self, f):
        if not 0 <= f < self.n_filters:
            raise IndexError(f)
        return _View(self, f)

    # ---- packing helpers -------------------------------------------------
    def pack_observations(self, observed):
        """list (per filter) of k_f x 3 arrays -> (obs [nf,3,max_obs] float64, counts [nf] int32)."""
        obs = np.zeros((self.n_filters, 3, self.max_obs))
        counts = np.zeros(self.n_filters, dtype=np.int32)
        for f, o in enumerate(observed):
            o = np.zeros((0, 3)) if o is None else np.asarray(o, dtype=np.float64).reshape(-1, 3)
            if o.shape[0] > self.max_obs:
                raise ValueError('filter %d: %d observations > max_obs=%d' % (f, o.shape[0], self.max_obs))
            counts[f] = o.shape[0]
            obs[f, :, :o.shape[0]] = o.T
        return obs, counts

    def pack_tables(self, tables):
        """list (one per filter, or a single shared list) of landmark entries with .index/.loc ->
        (loc [nt,2,max_table], index [nt,max_table] int32, n_table [nt] int32, shared flag)."""
        shared = len(tables) > 0 and not isinstance(tables[0], (list, tuple))
        tabs = [tables] if shared else tables
        nt = len(tabs)
        loc = np.zeros((nt, 2, self.max_table))
        idx = np.zeros((nt, self.max_table), dtype=np.int32)
        ntab = np.zeros(nt, dtype=np.int32)
        for t, tab in enumerate(tabs):
            if len(tab) > self.max_table:
                raise ValueError('table %d: %d entries > max_table=%d' % (t, len(tab), self.max_table))
            ntab[t] = len(tab)
            for i, e in enumerate(tab):
                idx[t, i] = int(e.index)
                loc[t, :, i] = e.loc
        return loc, idx, ntab, shared

    # ---- the step ----------------------------------------------------------
    def step(self, u, observed=None, tables=None, packed=None, predict=True, measure=True, want_actions=False):
        """One SLAM iteration of every filter.  u: [nf,2].  Either `observed` + `tables` (Python objects, packed here)
        or `packed` = (obs, counts, loc, idx, ntab, shared) from the pack_* helpers."""
        u = np.ascontiguousarray(np.asarray(u, dtype=np.float64).reshape(self.n_filters, 2))
        if measure:
            if packed is None:
                packed = self.pack_observations(observed) + self.pack_tables(tables)
            obs, counts, loc, idx, ntab, shared = packed
        else:
            obs = counts = loc = idx = ntab = None
            shared = True
        act = np.zeros((self.n_filters, self.max_obs), dtype=np.int32) if want_actions else None
        vp = lambda a: None if a is None else a.ctypes.data_as(ct.c_void_p)
        self._ck(self._lib.ekfb_step(self._h, 0, vp(u), vp(obs), vp(counts), vp(loc), vp(idx), vp(ntab),
                                     1 if shared else 0, 1 if predict else 0, 1 if measure else 0,
                                     None if act is None else L.iptr(act)))
        self._keep = (u, packed)                      # host buffers must outlive the asynchronous copies
        return act

    def step_device(self, u_ptr, obs_ptr, counts_ptr, loc_ptr, idx_ptr, ntab_ptr, shared, predict=True, measure=True):
        """Same, inputs already resident in HBM (raw device pointers, layouts as pack_* produce)."""
        self._ck(self._lib.ekfb_step(self._h, 1, u_ptr, obs_ptr, counts_ptr, loc_ptr, idx_ptr, ntab_ptr,
                                     1 if shared else 0, 1 if predict else 0, 1 if measure else 0, None))

    def sync(self):
        self._ck(self._lib.ekfb_sync(self._h))

    def n_landmarks(self):
        out = np.zeros(self.n_filters, dtype=np.int32)
        self._ck(self._lib.ekfb_n_landmarks(self._h, L.iptr(out)))
        return out

    def set_state(self, f, x, P, s):
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1))
        N = (x.size - 3) // 2
        P = np.asfortranarray(np.asarray(P, dtype=np.float64))
        s = np.ascontiguousarray(np.asarray(s, dtype=np.float64).reshape(-1)) if N > 0 else np.zeros(1)
        self._ck(self._lib.ekfb_set_state(self._h, int(f), L.dptr(x), L.dptr(P), P.shape[0], L.dptr(s), N))

    def set_state_all(self, x, P, s):
        """Every filter starts from the same (x, P, s): one upload, replicated on the device."""
        self.set_state(-1, x, P, s)

    def poses(self):
        """x(1:3) of every filter, [nf,3].  A small kernel packs the strided triples, one copy brings them to the host.
        After a split step the call returns as soon as the front kernel of that step is done (the poses are final
        then), while its flush still streams the covariances; device-side errors of the step are raised here too."""
        out = np.empty((self.n_filters, 3))
        self._ck(self._lib.ekfb_get_poses(self._h, L.dptr(out)))
        return out

    def downdate_bytes(self):
        b = ct.c_double()
        self._ck(self._lib.ekfb_downdate_bytes(self._h, ct.byref(b)))
        return b.value

    def profile(self, enable=None):
        """Returns (total ms of the step kernels, steps, total ms of the flush kernel alone) since profiling was
        enabled; enable=True/False (re)arms it."""
        ms, n, fl = ct.c_double(), ct.c_int64(), ct.c_double()
        self._ck(self._lib.ekfb_profile(self._h, -1 if enable is None else int(bool(enable)), ct.byref(ms), ct.byref(n),
                                        ct.byref(fl)))
        return ms.value, n.value, fl.value
