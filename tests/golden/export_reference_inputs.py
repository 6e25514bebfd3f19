#!/usr/bin/env python
"""Export the inputs of the reference-run fixtures as plain CSV for matlab/run_reference_fixture.m (the reference-side witness):
    python tests/golden/export_reference_inputs.py          -> tests/golden/ref_inputs/<case>/*.csv
17 significant digits, so MATLAB/Octave read back the very doubles the interpreter run, the oracle and the GPU were fed.

The worlds are those of tests/golden/refrun_<case>.npz (make_reference_run.py), i.e. the golden worlds with the first
frame cut down to one observation and one table entry: with several new landmarks in the very first measure() the
reference stops with "too many input arguments" (EKF_SLAM.m:110-111 hands a comma-separated list to append), so the
original worlds are outside what the unmodified reference can run -- under MATLAB just as under oracle/mini_matlab.py."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden_util import CASES       # noqa: E402


def main():
    for name, cls in CASES.items():
        g = dict(np.load(os.path.join(HERE, 'refrun_%s.npz' % name)))
        d = os.path.join(HERE, 'ref_inputs', name)
        os.makedirs(d, exist_ok=True)
        open(os.path.join(d, 'class.txt'), 'w').write(cls + '\n')
        for key in ('u', 'obs_rows', 'obs_counts', 'table_index', 'table_loc', 'n_table', 'checkpoint_steps'):
            a = np.asarray(g[key], dtype=np.float64)
            np.savetxt(os.path.join(d, key + '.csv'), a.reshape(a.shape[0], -1), delimiter=',', fmt='%.17g')
        print('wrote', d)


if __name__ == '__main__':
    main()
