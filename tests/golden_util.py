"""Replay helpers for the golden vectors in tests/golden/ (inputs are stored as flat arrays)."""
import os
from types import SimpleNamespace

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
CASES = {'known_prefix_N20': 'EKF_SLAM', 'known_quirk_N20': 'EKF_SLAM', 'uc_nearest_N20': 'EKF_SLAM_UC',
         'uc_nearest_N8_seed7': 'EKF_SLAM_UC'}


def load(name):
    return dict(np.load(os.path.join(GOLDEN, name + '.npz')))


class Stub:
    """landmark_list duck type (Landmark.m:25-33): getLandmark + landmarkObj.landmark(i).{index,loc}."""

    def __init__(self, rows, table):
        self._rows = rows
        self.landmarkObj = SimpleNamespace(landmark=table)

    def getLandmark(self, laserdata, x):
        return self._rows


def steps(g):
    """Yield (t, u_t, stub_t) for every step of a golden world."""
    table = [SimpleNamespace(index=int(i), loc=l.copy()) for i, l in zip(g['table_index'], g['table_loc'])]
    off = np.concatenate([[0], np.cumsum(g['obs_counts'])])
    for t in range(len(g['obs_counts'])):
        yield t, g['u'][t], Stub(g['obs_rows'][off[t]:off[t + 1]], table[:int(g['n_table'][t])])


def replay(filt, g, on_checkpoint=None):
    """Drive `filt` (oracle or GPU class: predict/measure/chosen) through the golden world `g`."""
    cps = {int(s): i for i, s in enumerate(g['checkpoint_steps'])}
    for t, u, stub in steps(g):
        filt.predict(u)
        filt.measure(None, u, stub)
        if on_checkpoint is not None and (t + 1) in cps:
            on_checkpoint(cps[t + 1], filt)
    return np.array([(1 if a == 'update' else -1) * i for a, i in filt.chosen], dtype=np.int32)


def rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300))
