"""ekf_associate_static: the association kernel (K4) with its scratch in a file-scope __device__ array, for callers
that may not allocate.  No counterpart in the reference; checked against
the existing entry point ekf_associate through the C ABI.  Collected last on purpose (pytest -x)."""
import ctypes as ct
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_LANDMARKS = 1048576            # EKF_ASSOC_STATIC_MAX_LANDMARKS (include/ekfslam_b200.h)


def _static(lib, t, N, z, Rd, s_cost, s_thresh, mode, storage, res, stream=0):
    err = ct.create_string_buffer(256)
    zz = np.ascontiguousarray(z, dtype=np.float64)
    rr = np.ascontiguousarray(Rd, dtype=np.float64)
    from ekf_slam_b200 import _lib as L
    rc = lib.ekf_associate_static(t['x'].data_ptr(), t['P'].data_ptr() if t['P'] is not None else None, t['ld'],
                                  t['s'].data_ptr(), N, L.dptr(zz), L.dptr(rr), s_cost, s_thresh, mode, storage,
                                  res.data_ptr(), stream, err, 256)
    return rc, err.value.decode()


def _read(res):
    raw = res.cpu().numpy().tobytes()
    new, idx, cost = struct.unpack('<iid', raw)
    return bool(new), idx, cost


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('mode', ['signature', 'position'])
def test_static_scratch_association_matches_handle_entry_point(lib, mode, storage):
    import torch
    import ekf_slam_b200 as E
    from ekf_slam_b200 import _lib as L
    from ekf_slam_b200.synth import synthetic_spd_state, lowrank_to_dense, observe_from_truth
    torch.cuda.set_device(0)
    N = 700                                                           # 3 blocks of 256: the cross-block fold runs
    x, A, d, s, lms = synthetic_spd_state(N, seed=11)
    P = lowrank_to_dense(A, d)
    s_cost, s_thresh = 1e-11, 1e9
    if mode == 'position':
        s = np.ones(N)
        s_thresh = 9.21
        x[3 + 2 * 600: 5 + 2 * 600] = x[3 + 2 * 7: 5 + 2 * 7]        # duplicated landmark in another block: first-minimum tie rule
        P[:, 3 + 2 * 600: 5 + 2 * 600] = P[:, 3 + 2 * 7: 5 + 2 * 7]
        P[3 + 2 * 600: 5 + 2 * 600, :] = P[3 + 2 * 7: 5 + 2 * 7, :]
        P[3 + 2 * 600: 5 + 2 * 600, 3 + 2 * 600: 5 + 2 * 600] = P[3 + 2 * 7: 5 + 2 * 7, 3 + 2 * 7: 5 + 2 * 7]
    g = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, assoc_mode=mode)
    g.correspondence.s_cost, g.correspondence.s_thresh = s_cost, s_thresh
    g._sync_corr()
    g.set_state(x, P, s)
    n = 3 + 2 * N
    # the caller's own device buffers (torch is the test's allocator; the entry point allocates nothing)
    t = {'x': torch.from_numpy(x.copy()).cuda(), 's': torch.from_numpy(np.asarray(s, dtype=np.float64).copy()).cuda(),
         'P': torch.from_numpy(np.asfortranarray(P).T.copy()).cuda(), 'ld': n}     # column-major n x n
    res = torch.zeros(2, dtype=torch.float64, device='cuda')
    mode_i = L.ASSOC_POSITION if mode == 'position' else L.ASSOC_SIGNATURE
    stor_i = L.STORAGE_SYM if storage == 'sym' else L.STORAGE_FULL
    rng = np.random.default_rng(3)
    ids = np.concatenate([[7, 600], rng.choice(N, 40, replace=False)])
    rows = observe_from_truth(lms, x[:3], ids, rng, seam_deg=0.0)
    n_new = 0
    free0 = None
    for q, r in enumerate(rows):
        z = r.copy()
        if mode == 'position':
            z[2] = 1.0
        elif q % 5 == 4:
            z[2] = N + 5.0                                            # unseen signature -> newLL
        Rd = np.array([z[0] * 0.1, z[1] * 5.0])
        new, idx = ct.c_int32(), ct.c_int32()
        g._ck(lib.ekf_associate(g._h, L.dptr(np.ascontiguousarray(z)), L.dptr(Rd), ct.byref(new), ct.byref(idx)))
        want = (bool(new.value), int(idx.value))
        got = []
        for rep in range(2):                                          # two calls in a row share the scratch array
            res.fill_(-1.0)
            rc, msg = _static(lib, t, N, z, Rd, s_cost, s_thresh, mode_i, stor_i, res)
            assert rc == 0, msg
            got.append(_read(res))
        assert got[0][:2] == want and got[1][:2] == want, (q, z, got, want)      # index work: bit-exact
        assert got[0][2] == got[1][2] or (np.isnan(got[0][2]) and np.isnan(got[1][2]))
        assert np.isinf(got[0][2]) == want[0]
        n_new += int(want[0])
        if q == 1:
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0                      # nothing was allocated on the device by the calls
    assert mode == 'position' or 0 < n_new < len(rows)
    g.close()


def test_static_scratch_on_a_side_stream_and_mode_a_without_P(lib):
    import torch
    from ekf_slam_b200 import _lib as L
    torch.cuda.set_device(0)
    N = 5000
    rng = np.random.default_rng(0)
    s = np.arange(1, N + 1, dtype=np.float64)
    t = {'x': torch.zeros(3 + 2 * N, dtype=torch.float64, device='cuda'), 's': torch.from_numpy(s).cuda(), 'P': None, 'ld': 0}
    res = torch.zeros(2, dtype=torch.float64, device='cuda')
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    for sig in (1.0, 4097.0, float(N), N + 1.0):
        for rep in range(2):
            rc, msg = _static(lib, t, N, [1.0, 10.0, sig], [0.1, 50.0], 1e-11, 1e9, L.ASSOC_SIGNATURE, L.STORAGE_SYM, res,
                              stream=st.cuda_stream)
            assert rc == 0, msg
            st.synchronize()
            new, idx, cost = _read(res)
            assert (new, idx) == ((False, int(sig)) if sig <= N else (True, N + 1)), (sig, new, idx, cost)


def test_static_scratch_rejects_what_it_cannot_hold(lib):
    import torch
    from ekf_slam_b200 import _lib as L
    torch.cuda.set_device(0)
    t = {'x': torch.zeros(16, dtype=torch.float64, device='cuda'), 's': torch.zeros(16, dtype=torch.float64, device='cuda'),
         'P': None, 'ld': 0}
    res = torch.zeros(2, dtype=torch.float64, device='cuda')
    rc, msg = _static(lib, t, MAX_LANDMARKS + 1, [1, 2, 3], [1, 1], 1e-11, 1e9, L.ASSOC_SIGNATURE, L.STORAGE_SYM, res)
    assert rc == L.EKF_EINVAL and 'EKF_ASSOC_STATIC_MAX_LANDMARKS' in msg and str(MAX_LANDMARKS) in msg
    rc, msg = _static(lib, t, 0, [1, 2, 3], [1, 1], 1e-11, 1e9, L.ASSOC_SIGNATURE, L.STORAGE_SYM, res)
    assert rc == L.EKF_EINVAL
    rc, msg = _static(lib, t, 4, [1, 2, 3], [1, 1], 1e-11, 1e9, L.ASSOC_POSITION, L.STORAGE_SYM, res)      # mode B without P
    assert rc == L.EKF_EINVAL
    host = torch.zeros(16, dtype=torch.float64)                                                             # a host pointer is refused, not dereferenced
    rc, msg = _static(lib, {'x': host, 's': t['s'], 'P': None, 'ld': 0}, 4, [1, 2, 3], [1, 1], 1e-11, 1e9,
                      L.ASSOC_SIGNATURE, L.STORAGE_SYM, res)
    assert rc == L.EKF_EINVAL and 'd_x' in msg
