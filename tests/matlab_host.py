"""TEST INFRASTRUCTURE.  The MATLAB host classes of this repo (matlab/EKFSLAM_B200_Base.m, EKF_SLAM.m, EKF_SLAM_UC.m,
Correspondence.m, Landmark.m) executed from their own text by oracle/mini_matlab.py -- no MATLAB or Octave exists in this
image or on the GPU box -- with `ekfslam_mex` bound to a gateway of the caller's choice:

  * OracleGateway (CPU): every verb of matlab/ekfslam_mex.c answered by oracle/dense_literal.py, with the same argument
    checks the C gateway makes (shapes, element counts).  It exercises the classdef text -- constructor chaining, the
    Dependent x / P / s / C / Rc accessors, the flattening of the landmark table in measure(), the value-class
    Correspondence and its bind() -- not the filter arithmetic.
  * MexGateway: the real gateway, matlab/ekfslam_mex.c compiled against tests/mex_stub/mex.h (tests/test_matlab_shim.Mex),
    i.e. classdef text -> gateway C -> C ABI -> CUDA.
"""
import os
from types import SimpleNamespace

import numpy as np

from oracle import dense_literal as O
from oracle.mini_matlab import Interpreter, MiniMatlabError, MStruct, PyObj, to_matlab

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MATLAB_DIR = os.path.join(ROOT, 'matlab')
CLASS_FILES = ('EKFSLAM_B200_Base.m', 'EKF_SLAM.m', 'EKF_SLAM_UC.m', 'Correspondence.m', 'Landmark.m')


class GatewayError(MiniMatlabError):
    """What mexErrMsgIdAndTxt raises inside MATLAB."""


class _Stub:
    def __init__(self, rows, table):
        self._rows = rows
        self.landmarkObj = SimpleNamespace(landmark=table)

    def getLandmark(self, laserdata, x):
        return self._rows


class OracleGateway:
    """ekfslam_mex('verb', handle, args...) answered by the CPU oracle; arguments arrive as the interpreter holds them
    (2-D float64 arrays, 1x1 uint64 handles, str, MStruct), so the shape checks are those of the C gateway."""
    wants_nargout = True

    def __init__(self):
        self.filters = {}
        self.opts = {}
        self.calls = []                 # verb log
        self._next = 0x1000

    def __call__(self, nargout, verb, *a):
        self.calls.append(verb)
        fn = getattr(self, 'v_' + verb, None)
        if fn is None:
            raise GatewayError("ekfslam:EINVAL: unknown verb '%s'" % verb)
        if verb != 'create':
            if not a or not (isinstance(a[0], np.ndarray) and a[0].dtype == np.uint64):
                raise GatewayError('ekfslam:EINVAL: handle must be a uint64 scalar')
            h = int(a[0].ravel()[0])
            if h not in self.filters:
                raise GatewayError('ekfslam:EINVAL: null or stale handle')
            a = (self.filters[h],) + tuple(a[1:])
        out = fn(nargout, *a)
        return [] if out is None else list(out)

    @staticmethod
    def _vec(v, n, what):
        if not isinstance(v, np.ndarray) or v.dtype != np.float64 or v.size < n:
            raise GatewayError('ekfslam:EINVAL: %s must be a real double array with >= %d elements' % (what, n))
        return v.ravel(order='F')[:n].copy() if n else np.zeros(0)

    @staticmethod
    def _R(v):
        if v.size == 2:
            return np.diag(v.ravel())
        if v.shape != (2, 2):
            raise GatewayError('ekfslam:EINVAL: R must be 2 x 2 or its diagonal')
        return np.diag(np.diag(v))

    def v_create(self, nargout, variant, opts=None):
        o = (O.EKF_SLAM_UC if int(variant.ravel()[0]) == 1 else O.EKF_SLAM)()
        if opts is not None and not isinstance(opts, MStruct):
            raise GatewayError('ekfslam:EINVAL: opts must be a struct')
        self._next += 8
        self.filters[self._next] = o
        self.opts[self._next] = {} if opts is None else {k: v for k, v in opts.f.items()}
        o.capacity = int(self.opts[self._next].get('capacity_landmarks', np.array([[1024.0]])).ravel()[0])
        return [np.array([[self._next]], dtype=np.uint64)]

    def v_destroy(self, nargout, o):
        for k in [k for k, v in self.filters.items() if v is o]:
            del self.filters[k]

    def v_predict(self, nargout, o, u):
        o.predict(self._vec(u, 2, 'u'))

    def v_append(self, nargout, o, u, R, pos, sig):
        if (o.x.size - 3) // 2 >= o.capacity:
            raise GatewayError('ekfslam:ECAPACITY: landmark capacity exhausted')
        o.append(self._vec(u, 2, 'u'), self._R(R), self._vec(pos, 2, 'landmarkPos'), float(sig.ravel()[0]))

    def v_associate(self, nargout, o, z, R):
        new, idx = o.correspondence.estimateCorrespondence(self._vec(z, 3, 'z'), self._R(R), o.x, o.P, o.s)
        return [bool(new), float(idx)][:max(nargout, 1)]

    def v_measure(self, nargout, o, obs, u, loc, idx):
        k = obs.shape[0]
        if k and obs.shape[1] != 3:
            raise GatewayError('ekfslam:EINVAL: observed_LL must be k x 3')
        T = idx.size
        idx = self._vec(idx, T, 'index table')
        loc = self._vec(loc, 2 * T, 'loc table').reshape((T, 2), order='F')        # T x 2 column-major, as the C ABI reads it
        table = [SimpleNamespace(index=int(i), loc=l.copy()) for i, l in zip(idx, loc)]
        before = len(o.chosen)
        o.measure(None, self._vec(u, 2, 'u'), _Stub(obs.copy(), table))
        act = [(1 if a == 'update' else -1) * int(i) for a, i in o.chosen[before:]]
        return [np.array(act, dtype=np.float64).reshape(-1, 1)]

    def v_n_landmarks(self, nargout, o):
        return [float((o.x.size - 3) // 2)]

    def v_get_x(self, nargout, o):
        return [o.x.reshape(1, -1).copy()]

    def v_get_s(self, nargout, o):
        return [np.asarray(o.s, dtype=np.float64).reshape(-1, 1).copy()]

    def v_get_P(self, nargout, o):
        return [o.P.copy()]

    def v_get_P_block(self, nargout, o, r1, c1, nr, nc):
        r1, c1, nr, nc = (int(v.ravel()[0]) for v in (r1, c1, nr, nc))
        if r1 < 1 or c1 < 1 or r1 - 1 + nr > o.P.shape[0] or c1 - 1 + nc > o.P.shape[1]:
            raise GatewayError('ekfslam:EINVAL: block out of range')
        return [o.P[r1 - 1:r1 - 1 + nr, c1 - 1:c1 - 1 + nc].copy()]

    def v_get_ellipses(self, nargout, o):
        return [O.ellipse_params(o.x, o.P)]

    def v_set_state(self, nargout, o, x, P, s):
        n = x.size
        if n < 3 or (n - 3) % 2 or P.shape != (n, n):
            raise GatewayError('ekfslam:EINVAL: set_state: x must have 3+2N elements and P must be n x n')
        N = (n - 3) // 2
        o.x, o.P, o.s = self._vec(x, n, 'x'), P.copy(), self._vec(s, N, 's')

    def v_set_x(self, nargout, o, x):
        if x.size != o.x.size:
            raise GatewayError('ekfslam:EINVAL: set_x: size must match the current state')
        o.x = self._vec(x, x.size, 'x')

    def v_set_s(self, nargout, o, s):
        if s.size != (o.x.size - 3) // 2:
            raise GatewayError('ekfslam:EINVAL: set_s: size must match the current state')
        o.s = self._vec(s, s.size, 's')

    def v_set_P(self, nargout, o, P):
        if P.shape != (o.x.size, o.x.size):
            raise GatewayError('ekfslam:EINVAL: set_P: P must be n x n')
        o.P = P.copy()

    def v_set_noise(self, nargout, o, C, Rc, s_cost, s_thresh):
        o.C = float(C.ravel()[0])
        o.Rc = list(self._vec(Rc, 2, 'Rc'))
        o.s_cost, o.s_thresh = float(s_cost.ravel()[0]), float(s_thresh.ravel()[0])
        if hasattr(o, 'correspondence'):
            o.correspondence.s_cost, o.correspondence.s_thresh = o.s_cost, o.s_thresh

    def v_flush(self, nargout, o):
        pass

    v_sync = v_flush


class OracleBatchGateway(OracleGateway):
    """The batch_* verbs (matlab/ekfslam_mex.c: n_filters objects of one reference class stepped in lock step) answered by
    n_filters oracle objects, behind the gateway's own size checks."""

    def __call__(self, nargout, verb, *a):
        if not verb.startswith('batch_'):
            return super().__call__(nargout, verb, *a)
        self.calls.append(verb)
        if verb == 'batch_create':
            if len(a) < 4:
                raise GatewayError('ekfslam:EINVAL: batch_create needs variant, n_filters, max_obs, max_table')
            variant, nf, mo, mt = (int(v.ravel()[0]) for v in a[:4])
            cap = 1024 if len(a) < 5 else int(a[4].f.get('capacity_landmarks', np.array([[1024.0]])).ravel()[0])
            fl = [(O.EKF_SLAM_UC if variant == 1 else O.EKF_SLAM)() for _ in range(nf)]
            for o in fl:
                o.capacity = cap
            self._next += 8
            self.filters[self._next] = SimpleNamespace(f=fl, nf=nf, mo=mo, mt=mt)
            return [np.array([[self._next]], dtype=np.uint64)]
        if not a or not (isinstance(a[0], np.ndarray) and a[0].dtype == np.uint64 and a[0].size == 1):
            raise GatewayError('ekfslam:EINVAL: batch handle must be a uint64 scalar')
        b = self.filters.get(int(a[0].ravel()[0]))
        if b is None:
            raise GatewayError('ekfslam:EINVAL: batch handle is null or destroyed')
        a = a[1:]
        if verb == 'batch_destroy':
            del self.filters[[k for k, v in self.filters.items() if v is b][0]]
            return []
        if verb == 'batch_sync':
            return []
        if verb == 'batch_step':
            if len(a) < 7:
                raise GatewayError('ekfslam:EINVAL: batch_step needs u, observed_LL, counts, loc, index, do_predict, do_measure')
            U, obs, counts, loc, idx = a[:5]
            do_p, do_m = (float(v.ravel()[0]) != 0.0 for v in a[5:7])
            T = idx.size
            if T > b.mt:
                raise GatewayError('ekfslam:EINVAL: landmark table has %d entries, max_table is %d' % (T, b.mt))
            counts = self._vec(counts, b.nf, 'counts').astype(int)
            U = self._vec(U, 2 * b.nf, 'u').reshape((2, b.nf), order='F')
            obs = self._vec(obs, 3 * b.mo * b.nf, 'observed_LL').reshape((b.mo, 3, b.nf), order='F')
            idx = self._vec(idx, T, 'index table')
            loc = self._vec(loc, 2 * T, 'loc table').reshape((T, 2), order='F')
            table = [SimpleNamespace(index=int(i), loc=l.copy()) for i, l in zip(idx, loc)]
            act = np.zeros((b.mo, b.nf))
            for f, o in enumerate(b.f):
                if do_p:
                    o.predict(U[:, f])
                if do_m:
                    before = len(o.chosen)
                    o.measure(None, U[:, f], _Stub(obs[:counts[f], :, f].copy(), table))
                    got = [(1 if w == 'update' else -1) * int(i) for w, i in o.chosen[before:]]
                    act[:len(got), f] = got
            return [act]
        if verb == 'batch_n_landmarks':
            return [np.array([[(o.x.size - 3) // 2] for o in b.f], dtype=np.float64)]
        if verb == 'batch_poses':
            return [np.array([o.x[:3] for o in b.f]).T.copy()]
        if verb in ('batch_get_x', 'batch_get_s', 'batch_get_P'):
            if not a:
                raise GatewayError("ekfslam:EINVAL: verb '%s' needs a filter index" % verb)
            f = int(a[0].ravel()[0])
            if f < 1 or f > b.nf:
                raise GatewayError('ekfslam:EINVAL: filter index %d outside 1..%d' % (f, b.nf))
            o = b.f[f - 1]
            return [{'x': o.x.reshape(1, -1), 's': np.asarray(o.s, dtype=np.float64).reshape(-1, 1), 'P': o.P}[verb[10]].copy()]
        if verb == 'batch_set_state':
            if len(a) < 4:
                raise GatewayError('ekfslam:EINVAL: batch_set_state needs f, x, P, s')
            f = int(a[0].ravel()[0])
            if f < 0 or f > b.nf:
                raise GatewayError('ekfslam:EINVAL: filter index %d outside 1..%d' % (f, b.nf))
            for o in (b.f if f == 0 else [b.f[f - 1]]):
                self.v_set_state(0, o, *a[1:4])
            return []
        raise GatewayError("ekfslam:EINVAL: unknown verb '%s'" % verb)


class MexGateway:
    """The compiled gateway (tests/test_matlab_shim.Mex): interpreter values <-> the stand-in mxArray runtime."""
    wants_nargout = True

    def __init__(self, mex):
        from tests.test_matlab_shim import Handle, MexError
        self.mex, self.Handle, self.MexError = mex, Handle, MexError
        self.calls = []

    def __call__(self, nargout, verb, *a):
        self.calls.append(verb)
        args = []
        for v in a:
            if isinstance(v, np.ndarray) and v.dtype == np.uint64:
                args.append(self.Handle(int(v.ravel()[0])))
            elif isinstance(v, MStruct):
                args.append({k: float(e.ravel()[0]) for k, e in v.f.items()})
            elif isinstance(v, np.ndarray):
                args.append(v.astype(np.float64))
            else:
                args.append(v)
        try:
            out = self.mex(verb, *args, nlhs=max(nargout, 0))
        except self.MexError as e:
            raise GatewayError(str(e))
        return [np.array([[o.v]], dtype=np.uint64) if isinstance(o, self.Handle) else o for o in out]


class MatlabHost:
    """The repo's MATLAB classes, loaded from matlab/*.m, bound to `gateway` as their ekfslam_mex."""

    def __init__(self, gateway):
        from oracle import matlab_compat as MC
        self.gateway = gateway
        self.it = Interpreter(stubs={'ekfslam_mex': gateway})
        for f in CLASS_FILES:
            self.it.load_file(os.path.join(MATLAB_DIR, f))

    def load(self, *files):
        for f in files:
            self.it.load_file(os.path.join(MATLAB_DIR, f))
        return self

    def with_rng(self, seed=0):
        """rng / rand / randn served by NumPy (not MATLAB's streams: runs are reproducible here, not comparable with MATLAB's)
        and fprintf recorded in self.printed as (format, args...)."""
        rs = np.random.RandomState(seed)

        def dims(a):
            d = [int(v.ravel()[0]) for v in a]
            return (1, 1) if not d else ((d[0], d[0]) if len(d) == 1 else tuple(d))
        self.printed = []
        self.it.builtins.update({'rng': lambda sd, kind=None: rs.seed(int(sd.ravel()[0]) % (2 ** 32)),
                                 'rand': lambda *a: rs.random_sample(dims(a)),
                                 'randn': lambda *a: rs.standard_normal(dims(a)),
                                 'fprintf': lambda *a: self.printed.append(a)})
        return self

    def new(self, cls, *args):
        return self.it.construct(cls, [to_matlab(a) for a in args])

    def call(self, obj, method, *args, nargout=1):
        f = self.it.find_method(obj.cls, method)
        if f is None:
            raise MiniMatlabError('no method %r' % method)
        return self.it.call_function(f, [obj] + [to_matlab(a) for a in args], nargout)

    def get(self, obj, prop):
        return self.it.eval(('field', ('id', 'o'), prop), {'o': obj})

    def set(self, obj, prop, value):
        self.it.assign(('field', ('id', 'o'), prop), to_matlab(value), {'o': obj})

    def run(self, src, **env):
        """Execute a few statements of MATLAB text (the driver loop of a test) with `env` as the workspace."""
        from oracle.mini_matlab import Parser, tokenize
        ws = {k: to_matlab(v) for k, v in env.items()}
        p = Parser(tokenize(src))
        stmts = []
        p.skip_nl()
        while p.peek().kind != 'EOF':
            stmts.append(p.parse_statement())
            p.skip_nl()
        self.it.exec_block(stmts, ws)
        return ws
