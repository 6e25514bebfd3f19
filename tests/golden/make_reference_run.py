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
    return dict(x=a['x'], P=a['P'], s=a['s'], z=a['z'], Rdiag=a['Rdiag'], mode_a=np.array(out, dtype=np.int32))


def main():
    ref_dir = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
    for name, cls in GU.CASES.items():
        out = run_case(ref_dir, cls, in_domain(GU.load(name)))
        np.savez_compressed(os.path.join(HERE, 'refrun_%s.npz' % name), **out)
        print('%-22s %-12s %3d steps, %2d landmarks, |x| %.6g' % (name, cls, len(out['obs_counts']), (out['x_final'].size - 3) // 2,
                                                                  np.linalg.norm(out['x_final'])))
    np.savez_compressed(os.path.join(HERE, 'refrun_association_N12.npz'), **association_case(ref_dir))
    print('association_N12: %d decisions' % 40)


if __name__ == '__main__':
    main()
