#!/usr/bin/env python
"""Reference-run fixtures: the UNMODIFIED source text of the reference's filter classes, executed on golden inputs.

    python tests/golden/make_reference_run.py [/path/to/SamShue/EKF_SLAM]        (default /root/reference)

oracle/mini_matlab.py (test infrastructure) interprets EKF_SLAM.m, EKF_SLAM_UC.m and Correspondence.m where they lie in
the reference checkout -- nothing is copied -- and drives `slam.predict(u); slam.measure(laserData, u, LM)` (SLAM.m:110,116)
over the worlds of the golden vectors.  What comes out (x, P, s, the landmark count after every step, checkpoints of x,
and the association decisions of Correspondence.estimateCorrespondence on a fixed state) is committed as
tests/golden/refrun_*.npz together with the inputs, because the reference does not travel to the GPU box.
tests/test_oracle_cpu.py then holds oracle/dense_literal.py to these vectors, tests/test_parity_gpu.py the CUDA path.

Input domain.  The worlds are those of make_golden.py with ONE change: the first frame is cut down to one observation and
one table entry.  The reference's first-landmark branch (EKF_SLAM.m:110-111, EKF_SLAM_UC.m:112-113) passes
`landmark(find([landmark.index])).loc` to append; with more than one entry carrying a non-zero index that expression is
a comma-separated list and MATLAB stops with "too many input arguments" (the interpreter does the same), so a frame of
several new landmarks at the very first measure() is outside what the reference can run.  (oracle/dense_literal.py takes
the first such entry there -- an interpretation, flagged F9 in SURVEY.md, which these fixtures therefore do not cover.)

What this pins: the oracle's reading of the reference's statements (control flow, precedence, indexing, growth, quirks).
What it does not: MathWorks' built-ins (sind, cosd, atan2d, wrapTo360, inv, BLAS), served by oracle/matlab_compat.py on both sides.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.mini_matlab import ReferenceFilter, Interpreter, to_matlab     # noqa: E402
from tests import golden_util as GU                                       # noqa: E402

INPUT_KEYS = ('u', 'obs_rows', 'obs_counts', 'table_index', 'table_loc', 'n_table')


def in_domain(g):
    """First frame: one observation, one table entry (see the module docstring)."""
    g = {k: g[k].copy() for k in INPUT_KEYS}
    c0 = int(g['obs_counts'][0])
    if c0 > 1:
        g['obs_rows'] = np.concatenate([g['obs_rows'][:1], g['obs_rows'][c0:]])
        g['obs_counts'][0] = 1
    g['n_table'][0] = 1
    return g


def run_case(ref_dir, cls, g, every=25):
    r = ReferenceFilter(ref_dir, cls)
    T = len(g['obs_counts'])
    nmax = 3 + 2 * int(g['table_index'].size)
    xs, cps = [], []
    for t, u, stub in GU.steps(g):
        r.predict(u)
        r.measure(None, u, stub)
        if (t + 1) % every == 0 or t == T - 1:
            cps.append(t + 1)
            xs.append(np.pad(r.x, (0, nmax - r.x.size), constant_values=np.nan))
    out = dict(g)
    out.update(x_final=r.x, P_final=r.P, s_final=r.s, n_landmarks=np.array(r.n_landmarks, dtype=np.int32),
               x_checkpoints=np.array(xs), checkpoint_steps=np.array(cps, dtype=np.int32))
    # the class's own plot() on the final state (EKF_SLAM.m:154-234), figure calls recorded instead of drawn: the
    # covariance-ellipse polylines, robot first; every 4th of the 629 points of t = -pi:0.01:pi is kept
    out['ellipse_curves'] = r.plot_curves()[:, :, ::4]
    return out


def association_case(ref_dir):
    """Correspondence.estimateCorrespondence as shipped (signature score, Correspondence.m:75) on the fixed state of
    tests/golden/association_N12.npz."""
    a = dict(np.load(os.path.join(HERE, 'association_N12.npz')))
    it = Interpreter()
    it.load_file(os.path.join(ref_dir, 'Correspondence.m'))
    c = it.construct('Correspondence', [to_matlab(1e-11), to_matlab(1e9), 'EKF_SLAM_UC'])
    f = c.cls['methods']['estimateCorrespondence']
    out = []
    for z, Rd in zip(a['z'], a['Rdiag']):
        new, idx = it.call_function(f, [c, to_matlab(z), to_matlab(np.diag(Rd)), to_matlab(a['x']), to_matlab(a['P']),
                                        to_matlab(a['s'].reshape(-1, 1))], 2)
        out.append([int(bool(new.ravel()[0])), int(idx.ravel()[0])])
    # Mode B: the reference's own alternative score, Correspondence.m:74 `%log_likelihood(kk)=position_cost+signiture_cost;`,
    # which ships commented out.  The text is taken from the checkout and changed IN MEMORY by exactly two characters
    # -- the `%` in front of line 74 goes, line 75 gets one -- and everything else (position_cost at :69 with the dense
    # H_k*P*H_k', the gate, the strict-< running minimum) is executed as written.  All signatures equal so that the
    # Mahalanobis term decides; s_thresh = 9.21 as in association_N12.npz.
    lines = open(os.path.join(ref_dir, 'Correspondence.m')).read().split('\n')
    i74 = next(i for i, l in enumerate(lines) if l.strip() == '%log_likelihood(kk)=position_cost+signiture_cost;')
    assert lines[i74 + 1].strip() == 'log_likelihood(kk)=signiture_cost;'
    lines[i74] = lines[i74].replace('%', '', 1)
    lines[i74 + 1] = lines[i74 + 1].replace('log_likelihood', '%log_likelihood', 1)
    itb = Interpreter()
    itb.load_source('\n'.join(lines))
    cb = itb.construct('Correspondence', [to_matlab(1e-11), to_matlab(9.21), 'EKF_SLAM_UC'])
    fb = cb.cls['methods']['estimateCorrespondence']
    outb = []
    sB = np.ones((len(a['s']), 1))
    for z, Rd in zip(a['z'], a['Rdiag']):
        zB = z.copy(); zB[2] = 1.0
        new, idx = itb.call_function(fb, [cb, to_matlab(zB), to_matlab(np.diag(Rd)), to_matlab(a['x']), to_matlab(a['P']),
                                          to_matlab(sB)], 2)
        outb.append([int(bool(new.ravel()[0])), int(idx.ravel()[0])])
    return dict(x=a['x'], P=a['P'], s=a['s'], z=a['z'], Rdiag=a['Rdiag'], mode_a=np.array(out, dtype=np.int32),
                mode_b=np.array(outb, dtype=np.int32))


def functions_case(ref_dir, seed=3, cases=6):
    """The free-function forms of the path the reference also ships: append.m:1-27 `[x,P]=append(x,P,u,idx,R,pos)`
    and the local functions of example.m:24-45, `[x,P]=predict(x,u,P,C)` and `[x_new,F]=f(x,u)`, executed from their own
    source on seeded states of 0..5 landmarks (append: once with idx beyond the map -- the landmark is added -- and once
    with idx inside it -- nothing changes)."""
    it = Interpreter()
    it.load_file(os.path.join(ref_dir, 'append.m'))
    it.load_file(os.path.join(ref_dir, 'example.m'))
    rng = np.random.default_rng(seed)
    out = {'n_cases': np.array(cases, dtype=np.int32)}
    for c in range(cases):
        nl = c
        n = 3 + 2 * nl
        A = rng.standard_normal((n, n))
        P = 0.01 * (A @ A.T) + 0.1 * np.eye(n)
        x = np.concatenate([[rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(0, 720)], rng.uniform(-8, 8, 2 * nl)])
        u = np.array([rng.uniform(0.1, 1.5), rng.uniform(-40, 40)])
        R = np.diag([0.01 * rng.uniform(0.5, 3), 5.0])
        pos = rng.uniform(-8, 8, 2)
        M = [to_matlab(v) for v in (x, P, u)]
        xa, Pa = it.call('append', M + [to_matlab(float(nl + 1)), to_matlab(R), to_matlab(pos)], 2, {})
        xk, Pk = it.call('append', M + [to_matlab(float(max(nl, 0))), to_matlab(R), to_matlab(pos)], 2, {})
        xp, Pp = it.call('predict', [to_matlab(x), to_matlab(u), to_matlab(P), to_matlab(0.2)], 2, {})
        xf, F = it.call('f', [to_matlab(x), to_matlab(u)], 2, {})
        out.update({'x%d' % c: x, 'P%d' % c: P, 'u%d' % c: u, 'R%d' % c: R, 'pos%d' % c: pos,
                    'append_x%d' % c: xa.ravel(), 'append_P%d' % c: Pa, 'keep_x%d' % c: xk.ravel(), 'keep_P%d' % c: Pk,
                    'predict_x%d' % c: xp.ravel(), 'predict_P%d' % c: Pp, 'f_x%d' % c: xf.ravel(), 'f_F%d' % c: F})
    return out


def main():
    ref_dir = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
    np.savez_compressed(os.path.join(HERE, 'refrun_functions.npz'), **functions_case(ref_dir))
    print('functions: append.m, example.m predict / f on 6 seeded states')
    for name, cls in GU.CASES.items():
        out = run_case(ref_dir, cls, in_domain(GU.load(name)))
        np.savez_compressed(os.path.join(HERE, 'refrun_%s.npz' % name), **out)
        print('%-22s %-12s %3d steps, %2d landmarks, |x| %.6g' % (name, cls, len(out['obs_counts']), (out['x_final'].size - 3) // 2,
                                                                  np.linalg.norm(out['x_final'])))
    np.savez_compressed(os.path.join(HERE, 'refrun_association_N12.npz'), **association_case(ref_dir))
    print('association_N12: %d decisions' % 40)


if __name__ == '__main__':
    main()
