#!/usr/bin/env python
"""Turn the outputs of matlab/run_reference_fixture.m into the missing pin of the oracle:
    python tests/golden/compare_reference_fixture.py [tests/golden/ref_outputs]
For every case directory found: x at every checkpoint, final x / P / s and the landmark count after every step, as the
UNMODIFIED reference produced them under MATLAB / Octave, against tests/golden/refrun_<case>.npz -- the same source text
executed on the same inputs by oracle/mini_matlab.py with restated built-ins, which oracle/dense_literal.py reproduces to
1e-12 (tests/test_oracle_cpu.py).  Agreement therefore pins what is still open: MathWorks' built-ins.
Bar: 1e-9 relative (north star), landmark counts and signatures identical."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden_util import CASES, rel       # noqa: E402


def _csv(path):
    return np.atleast_2d(np.genfromtxt(path, delimiter=',', dtype=np.float64))


def compare_case(name, out_dir, tol=1e-9):
    g = dict(np.load(os.path.join(HERE, 'refrun_%s.npz' % name)))
    d = os.path.join(out_dir, name)
    res = {}
    x = _csv(os.path.join(d, 'x_final.csv')).reshape(-1)
    P = _csv(os.path.join(d, 'P_final.csv'))
    s = _csv(os.path.join(d, 's_final.csv')).reshape(-1) if g['s_final'].size else np.zeros(0)
    xs = _csv(os.path.join(d, 'x_checkpoints.csv'))
    nl = _csv(os.path.join(d, 'n_landmarks.csv')).reshape(-1).astype(int)
    res['x_final'] = rel(x, g['x_final'])
    res['P_final'] = rel(P, g['P_final'])
    res['s_identical'] = bool(np.array_equal(s, g['s_final']))
    cp = 0.0
    for i in range(len(g['checkpoint_steps'])):
        want = g['x_checkpoints'][i]; want = want[~np.isnan(want)]
        got = xs[i][~np.isnan(xs[i])]
        cp = max(cp, rel(got, want) if got.size == want.size else np.inf)
    res['x_checkpoints'] = cp
    res['landmark_counts_identical'] = bool(np.array_equal(nl, g['n_landmarks']))
    res['pinned'] = bool(res['x_final'] <= tol and res['P_final'] <= tol and cp <= tol and res['s_identical'] and res['landmark_counts_identical'])
    return res


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, 'ref_outputs')
    ok, n = True, 0
    for name in CASES:
        if not os.path.isdir(os.path.join(out_dir, name)):
            continue
        r = compare_case(name, out_dir)
        n += 1
        ok &= r['pinned']
        print(name, r)
    print('no reference outputs found under %s' % out_dir if n == 0 else ('ORACLE PINNED TO THE REFERENCE' if ok else 'MISMATCH'))
    return 0 if ok else 1


if __name__ == '__main__':
    sys.exit(main())
