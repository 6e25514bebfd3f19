"""TEST INFRASTRUCTURE ONLY -- CPU oracle helpers, never imported by the product path.

Restatement of the MATLAB runtime / toolbox functions the reference's hot path
calls.  Their source is NOT in /root/reference (MathWorks core + Mapping
Toolbox, version unpinned by the reference), so the published semantics are
restated here; call sites in the reference are cited per function.

parity unpinned (these built-ins; the reading of the reference's own statements is pinned by oracle/mini_matlab.py):
the reference ships no golden vectors and MATLAB/Octave are
absent from this image, so these helpers are pinned only by MathWorks'
documented behaviour (exact values at multiples of 90 deg, wrapTo360(360)=360).
"""
import numpy as np

_DEG = np.pi / 180.0


def _reduce90(x):
    """Quadrant reduction used by MATLAB's sind/cosd: x = 90*n + r, |r| <= 45."""
    x = np.fmod(float(x), 360.0)          # exact
    n = np.rint(x / 90.0)
    r = x - 90.0 * n                      # exact (Sterbenz-range subtraction)
    return int(n) & 3, r


def sind(x):
    """sind -- EKF_SLAM.m:42,59,63,84,87-88 call sites. Exact 0/+-1 at k*90."""
    m, r = _reduce90(x)
    if m == 0:
        return float(np.sin(r * _DEG))
    if m == 1:
        return float(np.cos(r * _DEG))
    if m == 2:
        return float(-np.sin(r * _DEG))
    return float(-np.cos(r * _DEG))


def cosd(x):
    """cosd -- EKF_SLAM.m:42,58,64,85,87-88 call sites."""
    m, r = _reduce90(x)
    if m == 0:
        return float(np.cos(r * _DEG))
    if m == 1:
        return float(-np.sin(r * _DEG))
    if m == 2:
        return float(-np.cos(r * _DEG))
    return float(np.sin(r * _DEG))


def atan2d(y, x):
    """atan2d(y,x) = atan2(y,x)*180/pi -- EKF_SLAM.m:130, Correspondence.m:56."""
    return float(np.arctan2(y, x) * (180.0 / np.pi))


def mod(a, m):
    """MATLAB mod: result carries the sign of the divisor."""
    r = np.fmod(a, m)
    if r != 0.0 and (r < 0.0) != (m < 0.0):
        r = r + m
    return float(r)


def wrapTo360(a):
    """Mapping Toolbox wrapTo360 -- EKF_SLAM.m:50,130; Correspondence.m:56.

    positive multiples of 360 map to 360, zero stays 0, negatives wrap up.
    """
    r = mod(a, 360.0)
    if r == 0.0 and a > 0.0:
        r = 360.0
    return r


def inv2(A):
    """phi_k^-1 -- EKF_SLAM.m:143, Correspondence.m:69 (MATLAB mpower(-1) == inv, LAPACK LU)."""
    return np.linalg.inv(A)
