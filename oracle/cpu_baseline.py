"""TEST/BENCH INFRASTRUCTURE ONLY -- timed CPU baseline of the reference's dense algorithm.

Used by bench.py's ``cpu_baseline`` leg and ``--impl reference`` (the only places besides tests/ and
smoke() allowed to execute oracle/).  It times the reference's own dense expressions
(EKF_SLAM.m:47 ``F*P*F'+Q`` and EKF_SLAM.m:141-145 ``K=P*H'*phi^-1``, ``(eye(n)-K*H_k)*P``) with
NumPy/OpenBLAS on all host threads.  At 10k landmarks one dense update is 1.6e13 flop, so the O(n^3)
products are evaluated on a ROW SAMPLE of the output (the first m rows of the left factor, exactly the
reference's arithmetic for those rows) and scaled by n/m; the O(n^2) parts are evaluated in full.
"""
import time

import numpy as np

from .matlab_compat import sind, cosd, atan2d, wrapTo360, inv2


def _H_dense(x, idx):
    """H_k as written, EKF_SLAM.m:125-138 (dense 2 x n)."""
    n = len(x)
    mu_k = np.array([x[3 + (idx - 1) * 2], x[4 + (idx - 1) * 2]])
    delta_k = mu_k - x[0:2]
    q_k = float(delta_k @ delta_k)
    sq = np.sqrt(q_k)
    z_k = np.array([sq, wrapTo360(atan2d(delta_k[1], delta_k[0]) - x[2])])
    H = np.zeros((2, n))
    a = 3 + (idx - 1) * 2
    small = (1.0 / q_k) * np.array([[-sq * delta_k[0], -sq * delta_k[1], 0.0, sq * delta_k[0], sq * delta_k[1]],
                                    [delta_k[1], -delta_k[0], -q_k, -delta_k[1], delta_k[0]]])
    H[:, 0:3] = small[:, 0:3]
    H[:, a:a + 2] = small[:, 3:5]
    return H, z_k


def pick_rows(n, budget_s, gemm_flops_per_s):
    """Rows m of the left factor so that 2*m*n*n flop fits the budget."""
    m = int(budget_s * gemm_flops_per_s / (2.0 * n * n))
    return max(8, min(n, m))


def gemm_rate(n=1536, reps=2):
    a = np.random.default_rng(0).standard_normal((n, n))
    b = a.T.copy()
    a @ b
    t0 = time.perf_counter()
    for _ in range(reps):
        a @ b
    return reps * 2.0 * n ** 3 / (time.perf_counter() - t0)


def sampled_update_seconds(x, P, z, Rdiag, idx, m):
    """Seconds for ONE dense reference update at this n, big product on m sampled rows scaled by n/m.
    Does not modify x/P (the sample cannot produce the full new P)."""
    n = len(x)
    t0 = time.perf_counter()
    H, z_k = _H_dense(x, idx)
    phi = (H @ P) @ H.T + np.diag(Rdiag)                 # EKF_SLAM.m:141
    K = (P @ H.T) @ inv2(phi)                            # :143
    _ = x + (K @ (np.asarray(z[:2]) - z_k))              # :144
    t_small = time.perf_counter() - t0
    t0 = time.perf_counter()
    L = -(K[:m, :] @ H)                                  # rows of eye(n) - K*H_k, :145
    L[np.arange(m), np.arange(m)] += 1.0
    _ = L @ P                                            # m x n x n of the n x n x n product
    t_big = time.perf_counter() - t0
    return t_small + t_big * (n / float(m)), t_small, t_big


def sampled_predict_seconds(x, P, u, C, m):
    """Seconds for ONE dense reference predict (EKF_SLAM.m:42-47), two n^3 products on m rows each."""
    n = len(x)
    t0 = time.perf_counter()
    Ft = np.zeros((n, n))                                # dense F' (allocated like eye(n), :62), O(n^2): unscaled
    Ft[np.diag_indices(n)] = 1.0
    Ft[2, 0] = -1 * u[0] * sind(x[2])
    Ft[2, 1] = u[0] * cosd(x[2])
    t_alloc = time.perf_counter() - t0
    t0 = time.perf_counter()
    F = np.zeros((m, n))
    F[np.arange(m), np.arange(m)] = 1.0
    F[0, 2] = Ft[2, 0]
    F[1, 2] = Ft[2, 1]
    FP = F @ P                                           # m rows of F*P
    _ = FP @ Ft                                          # m rows of (F*P)*F'
    t = time.perf_counter() - t0
    return t_alloc + t * (n / float(m))
