#!/usr/bin/env python
"""Export the inputs of the golden cases as plain CSV for matlab/run_reference_fixture.m (the reference-side witness):
    python tests/golden/export_reference_inputs.py          -> tests/golden/ref_inputs/<case>/*.csv
17 significant digits, so MATLAB/Octave read back the very doubles the oracle and the GPU were fed."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.golden_util import CASES, load       # noqa: E402


def main():
    for name, cls in CASES.items():
        g = load(name)
        d = os.path.join(HERE, 'ref_inputs', name)
        os.makedirs(d, exist_ok=True)
        open(os.path.join(d, 'class.txt'), 'w').write(cls + '\n')
        for key in ('u', 'obs_rows', 'obs_counts', 'table_index', 'table_loc', 'n_table', 'checkpoint_steps'):
            a = np.asarray(g[key], dtype=np.float64)
            np.savetxt(os.path.join(d, key + '.csv'), a.reshape(a.shape[0], -1), delimiter=',', fmt='%.17g')
        print('wrote', d)


if __name__ == '__main__':
    main()
