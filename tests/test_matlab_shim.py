"""The MEX gateway (matlab/ekfslam_mex.c) compiled against tests/mex_stub/mex.h and driven the way the MATLAB
classes in matlab/ drive it: ekfslam_mex('verb', handle, args...).  No MATLAB/Octave exists in this image, so the
stand-in mxArray runtime (tests/mex_stub/mex_stub.c) plays the interpreter."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB_DIR = os.path.join(ROOT, 'tests', 'mex_stub')


class Mex:
    def __init__(self):
        import ekf_slam_b200
        ekf_slam_b200.load_library()
        subprocess.check_call(['make', '-s', '-C', STUB_DIR])
        self.so = C.CDLL(os.path.join(STUB_DIR, 'ekfslam_mex_stub.so'))
        so = self.so
        so.stub_double.restype = C.c_void_p
        so.stub_double.argtypes = [C.c_size_t, C.c_size_t, C.POINTER(C.c_double)]
        so.stub_uint64.restype = C.c_void_p
        so.stub_uint64.argtypes = [C.c_uint64]
        so.stub_string.restype = C.c_void_p
        so.stub_string.argtypes = [C.c_char_p]
        so.stub_struct.restype = C.c_void_p
        so.stub_set_field.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        so.stub_free.argtypes = [C.c_void_p]
        so.stub_class.argtypes = [C.c_void_p]
        so.stub_error_id.restype = C.c_char_p
        so.stub_error_msg.restype = C.c_char_p
        so.stub_mex.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
        for f in ('mxGetM', 'mxGetN'):
            getattr(so, f).restype = C.c_size_t
            getattr(so, f).argtypes = [C.c_void_p]
        so.mxGetData.restype = C.c_void_p
        so.mxGetData.argtypes = [C.c_void_p]
        so.mxGetScalar.restype = C.c_double
        so.mxGetScalar.argtypes = [C.c_void_p]

    def _to_mx(self, a):
        so = self.so
        if isinstance(a, str):
            return so.stub_string(a.encode())
        if isinstance(a, Handle):
            return so.stub_uint64(a.v)
        if isinstance(a, dict):
            s = so.stub_struct()
            for k, v in a.items():
                so.stub_set_field(s, k.encode(), self._to_mx(float(v)))
            return s
        arr = np.asfortranarray(np.atleast_2d(np.asarray(a, dtype=np.float64)))
        return so.stub_double(arr.shape[0], arr.shape[1], arr.ctypes.data_as(C.POINTER(C.c_double)))

    def _from_mx(self, p):
        so = self.so
        cls = so.stub_class(p)
        m, n = so.mxGetM(p), so.mxGetN(p)
        if cls == 13:
            return Handle(C.cast(so.mxGetData(p), C.POINTER(C.c_uint64))[0])
        if cls == 3:
            return bool(so.mxGetScalar(p))
        if cls == 4:
            return C.string_at(so.mxGetData(p), m * n).decode()
        buf = C.cast(so.mxGetData(p), C.POINTER(C.c_double))
        return np.ctypeslib.as_array(buf, shape=(m * n,)).copy().reshape((m, n), order='F') if m * n else np.zeros((m, n))

    def __call__(self, *args, nlhs=1):
        """ekfslam_mex(args...) -> list of nlhs outputs; raises MexError like mexErrMsgIdAndTxt would."""
        prhs = (C.c_void_p * len(args))(*[self._to_mx(a) for a in args])
        plhs = (C.c_void_p * max(nlhs, 1))()
        rc = self.so.stub_mex(nlhs, plhs, len(args), prhs)
        for p in prhs:
            self.so.stub_free(p)
        if rc:
            raise MexError(self.so.stub_error_id().decode(), self.so.stub_error_msg().decode())
        out = [self._from_mx(p) for p in plhs[:nlhs] if p]
        for p in plhs[:nlhs]:
            if p:
                self.so.stub_free(p)
        return out


class Handle:
    def __init__(self, v):
        self.v = int(v)


class MexError(RuntimeError):
    def __init__(self, ident, msg):
        super().__init__('%s: %s' % (ident, msg))
        self.ident = ident


@pytest.fixture(scope='module')
def mex():
    return Mex()


def test_gateway_compiles_and_reports_version(mex):
    assert mex('version')[0].startswith('ekfslam_b200')


def test_gateway_rejects_unknown_verb_and_bad_handle(mex):
    with pytest.raises(MexError) as e:
        mex('predict', 3.0, [1.0, 1.0])
    assert e.value.ident == 'ekfslam:EINVAL'
    with pytest.raises(MexError):
        mex(1.0)


def test_create_without_device_raises_ecuda(mex):
    import torch
    if torch.cuda.is_available():
        pytest.skip('CUDA device present')
    with pytest.raises(MexError) as e:
        mex('create', 0.0, {'capacity_landmarks': 4})
    assert e.value.ident == 'ekfslam:ECUDA'


def test_matlab_sources_cover_reference_interface():
    """The classdefs keep the reference's names: classes, methods and public properties."""
    src = {f: open(os.path.join(ROOT, 'matlab', f)).read() for f in os.listdir(os.path.join(ROOT, 'matlab')) if f.endswith('.m')}
    for cls in ('EKF_SLAM', 'EKF_SLAM_UC', 'Correspondence', 'Landmark'):
        assert ('classdef %s' % cls) in src[cls + '.m']
    base = src['EKFSLAM_B200_Base.m']
    for meth in ('function predict(h, u)', 'function [x_new, F] = f(', 'function append(h, u, R, landmarkPos, signature)',
                 'function measure(h, laserData, u, landmark_list)', 'function plot('):
        assert meth in base, meth
    for prop in ('x;', 'P;', 'Q;', 's;', 'landmark_list;', 'observed;', 'C;', 'Rc;'):
        assert prop in base, prop
    assert 'estimateCorrespondence(h, z, R, x, P, s)' in src['Correspondence.m']
    assert 'getLandmark(h, laserdata, x)' in src['Landmark.m']
    verbs_used = set()
    import re
    for s in src.values():
        verbs_used |= set(re.findall(r"ekfslam_mex\('(\w+)'", s))
    gateway = open(os.path.join(ROOT, 'matlab', 'ekfslam_mex.c')).read()
    for v in verbs_used:
        assert ('"%s"' % v) in gateway, 'verb %s used by a classdef but not implemented by the gateway' % v


@pytest.mark.gpu
@pytest.mark.parametrize('variant', [0, 1])
def test_gateway_full_loop_matches_oracle(mex, variant):
    """predict/measure through the MEX verbs == the oracle (<= 1e-9) and == the ctypes class path bit for bit."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    from oracle import dense_literal as O
    w = make_world(N=12, T=60, k=4, mode='prefix' if variant == 0 else 'nearest')
    h = mex('create', float(variant), {'capacity_landmarks': 16, 'storage': 1, 'defer_m': 4})[0]
    g = (E.EKF_SLAM if variant == 0 else E.EKF_SLAM_UC)(capacity_landmarks=16, storage='sym', defer_m=4)
    o = (O.EKF_SLAM if variant == 0 else O.EKF_SLAM_UC)()
    acts = []
    for t in range(w.T):
        stub = w.stub(t)
        mex('predict', h, w.u[t], nlhs=0)
        obs = stub.getLandmark(None, None)
        tab = stub.landmarkObj.landmark
        if len(obs):
            a = mex('measure', h, obs, w.u[t], np.array([e.loc for e in tab]).reshape(-1, 2),
                    np.array([[float(e.index)] for e in tab]).reshape(-1, 1))[0]
            acts += [('update', int(v)) if v > 0 else ('append', int(-v)) for v in a.ravel()]
        g.predict(w.u[t]); g.measure(None, w.u[t], stub)
        o.predict(w.u[t]); o.measure(None, w.u[t], stub)
    assert acts == o.chosen == g.chosen
    x = mex('get_x', h)[0].ravel()
    P = mex('get_P', h)[0]
    s = mex('get_s', h)[0].ravel()
    assert int(mex("n_landmarks", h)[0][0, 0]) == (o.x.size - 3) // 2
    assert np.array_equal(x, g.x) and np.array_equal(P, g.P) and np.array_equal(s, g.s)
    assert np.linalg.norm(x - o.x) / np.linalg.norm(o.x) <= 1e-9
    assert np.linalg.norm(P - o.P) / np.linalg.norm(o.P) <= 1e-9
    blk = mex('get_P_block', h, 1.0, 1.0, 3.0, 3.0)[0]
    assert np.array_equal(blk, P[:3, :3])
    newLL, idx = mex('associate', h, [1.0, 50.0, 2.0], np.diag([0.1, 250.0]), nlhs=2)
    assert newLL is False and idx[0, 0] == 2.0
    with pytest.raises(MexError) as e:
        mex('update', h, [1.0, 2.0], [0.1, 0.1], 99.0, nlhs=0)
    assert e.value.ident == 'ekfslam:EINVAL'
    mex('destroy', h, nlhs=0)
    g.close()


@pytest.mark.gpu
def test_gateway_batch_verbs_match_filter_batch(mex):
    """The Monte-Carlo batch through the MEX verbs ('batch_create' / 'batch_step' / 'batch_get_*' / 'batch_poses' /
    'batch_set_state'), the way matlab/EKFSLAM_B200_Batch.m drives them, == ekf_slam_b200.FilterBatch bit for bit
    (same library underneath), incl. the update_batch / flush_info / defer_info verbs of a single filter."""
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world
    nf, max_obs, cap = 6, 4, 12
    w = make_world(N=cap, T=40, k=max_obs, mode='nearest', seed=9)
    hb = mex('batch_create', 1.0, float(nf), float(max_obs), float(cap), {'capacity_landmarks': cap, 'storage': 1, 'defer_m': 8})[0]
    fb = E.FilterBatch('EKF_SLAM_UC', nf, capacity_landmarks=cap, max_obs=max_obs, max_table=cap, storage='sym', defer_m=8)
    scale = 1.0 + 0.01 * np.arange(nf)
    for t in range(w.T):
        stub = w.stub(t)
        obs = np.asarray(stub.getLandmark(None, None), dtype=np.float64).reshape(-1, 3)
        tab = stub.landmarkObj.landmark
        U = np.outer(scale, w.u[t])                                   # nf x 2
        packed = np.zeros((max_obs, 3, nf), order='F')
        packed[:obs.shape[0], :, :] = obs[:, :, None]
        acts = mex('batch_step', hb, U.T, packed.reshape(max_obs, 3 * nf, order='F'), np.full((nf, 1), float(obs.shape[0])),
                   np.array([e.loc for e in tab]).reshape(-1, 2), np.array([[float(e.index)] for e in tab]).reshape(-1, 1), 1.0, 1.0)[0]
        want = fb.step(U, [obs] * nf, tab, want_actions=True)
        assert np.array_equal(acts.T.astype(np.int32), want), t
    N = mex('batch_n_landmarks', hb)[0].ravel().astype(int)
    assert np.array_equal(N, fb.n_landmarks())
    assert np.array_equal(mex('batch_poses', hb)[0].T, fb.poses())
    for f in range(nf):
        assert np.array_equal(mex('batch_get_x', hb, float(f + 1))[0].ravel(), fb[f].x)
        assert np.array_equal(mex('batch_get_P', hb, float(f + 1))[0], fb[f].P)
        assert np.array_equal(mex('batch_get_s', hb, float(f + 1))[0].ravel(), fb[f].s)
    x0, P0, s0 = fb[0].x, fb[0].P, fb[0].s
    mex('batch_set_state', hb, 0.0, x0, P0, s0.reshape(-1, 1), nlhs=0)        # every filter <- filter 1's state
    mex('batch_sync', hb, nlhs=0)
    assert np.array_equal(mex('batch_get_P', hb, float(nf))[0], P0)
    with pytest.raises(MexError) as e:
        mex('batch_get_x', hb, float(nf + 1))
    assert e.value.ident == 'ekfslam:EINVAL'
    mex('batch_destroy', hb, nlhs=0)
    fb.close()
    # single-filter verbs added in round 2
    h = mex('create', 0.0, {'capacity_landmarks': cap, 'storage': 1, 'defer_m': 16})[0]
    g = E.EKF_SLAM(capacity_landmarks=cap, storage='sym', defer_m=16)
    mex('set_state', h, x0, P0, s0.reshape(-1, 1), nlhs=0)
    g.set_state(x0, P0, s0)
    k = min(3, int(N[0]))
    z = np.column_stack([np.linspace(1.0, 2.0, k), np.linspace(10.0, 40.0, k)])
    Rd = np.column_stack([z[:, 0] * 0.01, z[:, 1] * 5.0])
    mex('update_batch', h, z, Rd, np.arange(1.0, k + 1).reshape(-1, 1), nlhs=0)
    g.update_batch(z, Rd, np.arange(1, k + 1))
    assert np.array_equal(mex('get_x', h)[0].ravel(), g.x)
    bsz, capm = mex('defer_info', h, nlhs=2)
    assert (bsz[0, 0], capm[0, 0]) == (8.0, 16.0)
    cont, outst = mex('flush_info', h, nlhs=2)
    assert cont is False and outst[0, 0] == 0.0                   # 12 landmarks: far below the continuous-flush threshold
    mex('destroy', h, nlhs=0)
    g.close()
