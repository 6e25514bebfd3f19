// ekf_kernels.cu -- hand-written sm_100a kernels of the EKF-SLAM hot path.
//   K1 k_predict_P / k_predict_x  : EKF_SLAM.m:40-65   (F P F' + Q without touching the landmark block)
//   K2 k_linearize / k_gain       : EKF_SLAM.m:124-144 (5-column gather, S, S^-1, K, x)
//   K3 k_downdate                 : EKF_SLAM.m:145     (P -= K * (H P), streaming, HBM-bound)
//   K4 k_associate                : Correspondence.m:28-88 (gating + first-argmin)
//   K5 k_append                   : EKF_SLAM.m:67-98
#include "ekf_kernels.h"
#include "ekf_device.cuh"
#include <math_constants.h>
#include <limits.h>
#include <cstdlib>
#include <cstdio>
#include <mutex>
#include <string>
#include "../../include/ekfslam_b200.h"

namespace ekf {

// ------------------------------------------------------------------ K1 predict
// F = I + a e1 e3' + b e2 e3' (EKF_SLAM.m:62-64): (F P) adds a,b * row 3 to rows 1,2; (.) F' adds a,b * col 3
// to cols 1,2.  Thread t >= 3 owns column t of the row pass and row t of the column pass; thread 0 of block 0
// owns the 3x3 corner (both passes, then + Q = (W C) W', EKF_SLAM.m:42-47).  x is NOT written here so that
// every block sees the pre-motion heading (EKF_SLAM.m:42,63-64 use theta before f()).
__global__ void k_predict_P(double* __restrict__ P, int64_t ld, const double* __restrict__ x, int n,
                            double u1, double u2, double C, int sym,
                            double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds, int m) {
    const double th = x[2];
    const double st = sind(th), ct = cosd(th);
    const double a = -1.0 * u1 * st;      // F(1,3), EKF_SLAM.m:63
    const double b = u1 * ct;             // F(2,3), EKF_SLAM.m:64
    if (m > 0 && blockIdx.x == gridDim.x - 1) {      // extra block: the pending slabs take K_q <- F K_q, B_q <- B_q F' (k_predict_slabs)
        const int t = threadIdx.x;
        if (t < 2 * m) {
            double2* S = (t & 1 ? Ball : Kall) + (int64_t)(t >> 1) * lds;
            const double2 v0 = S[0], v1 = S[1], v2 = S[2];
            S[0] = make_double2(v0.x + a * v2.x, v0.y + a * v2.y);
            S[1] = make_double2(v1.x + b * v2.x, v1.y + b * v2.y);
        }
        return;
    }
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) {
        double M[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) M[i][j] = ldP(P, ld, i, j, sym);
        for (int j = 0; j < 3; ++j) {     // F*P
            M[0][j] = M[0][j] + a * M[2][j];
            M[1][j] = M[1][j] + b * M[2][j];
        }
        for (int i = 0; i < 3; ++i) {     // (F*P)*F'
            M[i][0] = M[i][0] + M[i][2] * a;
            M[i][1] = M[i][1] + M[i][2] * b;
        }
        double W[3] = {u1 * ct, u1 * st, u2};      // EKF_SLAM.m:42
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) P[i + j * ld] = M[i][j] + (W[i] * C) * W[j];   // :44,47
    } else if (t >= 3 && t < n) {
        double* col = P + (int64_t)t * ld;         // rows 0..2 of column t: one 32 B sector
        double p2 = col[2];
        col[0] = col[0] + a * p2;
        col[1] = col[1] + b * p2;
        if (!sym) {                                // row t, columns 0..2 (dead triangle in SYM storage)
            double q2 = P[t + 2 * ld];
            P[t] = P[t] + q2 * a;
            P[t + ld] = P[t + ld] + q2 * b;
        }
    }
}

// `prev` (may be null): receives the pre-motion pose.  The decoupled predict (ekf_capi.cu, async_x) runs this kernel
// FIRST on the side stream and hands k_predict_P `prev` as its x, so that it still sees the heading before f().
// `pose_host` (may be null): mapped pinned host memory that receives the new pose as well, so that a caller who reads x
// right after predict (the reference's loop does: getLandmark(laserData, h.x), EKF_SLAM.m:102) needs no copy, only the event.
__global__ void k_predict_x(double* __restrict__ x, double u1, double u2, double* __restrict__ prev, double* __restrict__ pose_host) {   // EKF_SLAM.m:57-60, :50
    double th = x[2];
    if (prev) { prev[0] = x[0]; prev[1] = x[1]; prev[2] = th; }
    const double x0 = x[0] + u1 * cosd(th + u2);
    const double x1 = x[1] + u1 * sind(th + u2);
    const double x2 = wrapTo360(th + u2);
    x[0] = x0; x[1] = x1; x[2] = x2;
    if (pose_host) { pose_host[0] = x0; pose_host[1] = x1; pose_host[2] = x2; __threadfence_system(); }
}

// predict while m rank-2 terms are pending (cross-step deferral): the covariance is P0 - sum_q K_q B_q, so
// F (.) F' + Q applies to P0 through k_predict_P and to the terms as K_q <- F K_q, B_q <- B_q F'.  F = I + a e1 e3' +
// b e2 e3' touches rows 0, 1 of K_q and columns 0, 1 of B_q only (slab rows 0, 1; both components).  xpre = pose
// BEFORE the motion step (EKF_SLAM.m:63-64).  Thread = (q, K or B).
__global__ void k_predict_slabs(double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds, int m,
                                const double* __restrict__ xpre, double u1) {
    const int t = threadIdx.x;
    if (t >= 2 * m) return;
    const double th = xpre[2];
    const double a = -1.0 * u1 * sind(th), b = u1 * cosd(th);
    double2* S = (t & 1 ? Ball : Kall) + (int64_t)(t >> 1) * lds;
    const double2 v0 = S[0], v1 = S[1], v2 = S[2];
    S[0] = make_double2(v0.x + a * v2.x, v0.y + a * v2.y);
    S[1] = make_double2(v1.x + b * v2.x, v1.y + b * v2.y);
}

void launch_predict_slabs(cudaStream_t st, double2* Kall, double2* Ball, int64_t lds, int m, const double* xpre, double u1) {
    if (m > 0) k_predict_slabs<<<1, 32, 0, st>>>(Kall, Ball, lds, m, xpre, u1);
}

// ------------------------------------------------------------------ K2 linearize + gain
// Both kernels read the CURRENT covariance = P0 - (m pending rank-2 terms) through ldP_def, so they serve the
// immediate path (m = 0) and the deferred path alike, and write K, B of this observation into slab slot m.
// One warp: lanes 0..24 fetch the 5x5 block P(c,c); lane 0 builds z_k, H_k, phi_k, phi_k^-1.
__global__ void k_linearize(const double* __restrict__ P, int64_t ld, const double* __restrict__ x,
                            int a, double z0, double z1, double R0, double R1, int sym,
                            const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int m,
                            Lin* __restrict__ lin) {
    __shared__ double Pcc[5][5];
    int lane = threadIdx.x;
    int c[5] = {0, 1, 2, a, a + 1};
    if (lane < 25) {
        int i = lane / 5, j = lane % 5;
        Pcc[i][j] = ldP_def(P, ld, c[i], c[j], sym, Kall, Ball, lds, m);
    }
    __syncwarp();
    if (lane == 0) {
        double xr[3] = {x[0], x[1], x[2]};
        double zk[2], Hs[2][5], S[2][2], Si[2][2];
        measurement_model(xr, x[a], x[a + 1], zk, Hs);
        double Rd[2] = {R0, R1};
        double Pl[5][5];
        for (int i = 0; i < 5; ++i)
            for (int j = 0; j < 5; ++j) Pl[i][j] = Pcc[i][j];
        innovation_cov(Hs, Pl, Rd, S, Si);
        int bad = 0;
        for (int r = 0; r < 2; ++r) {
            for (int k = 0; k < 5; ++k) lin->Hs[r][k] = Hs[r][k];
            for (int t = 0; t < 2; ++t) {
                lin->S[r][t] = S[r][t];
                lin->Sinv[r][t] = Si[r][t];
                if (!isfinite(S[r][t]) || !isfinite(Si[r][t])) bad = 1;
            }
        }
        lin->nu[0] = z0 - zk[0];
        lin->nu[1] = z1 - zk[1];
        for (int k = 0; k < 5; ++k) lin->c[k] = c[k];
        lin->nonfinite = bad;
        if (bad) report_nonfinite(lin->nf);
    }
}

// Row i of P H' (5-column gather), K = (P H') phi^-1, x += K nu, and column i of B = H P.
__global__ void k_gain(const double* __restrict__ P, int64_t ld, double* __restrict__ x, int n, int a, int sym,
                       const Lin* __restrict__ lin, double2* __restrict__ Kall, double2* __restrict__ Ball,
                       int64_t lds, int m) {
    __shared__ Lin L;
    if (threadIdx.x < sizeof(Lin) / sizeof(double)) ((double*)&L)[threadIdx.x] = ((const double*)lin)[threadIdx.x];
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c[5] = {0, 1, 2, a, a + 1};
    double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        double pc = ldP_def(P, ld, i, c[k], sym, Kall, Ball, lds, m);   // P(i, c_k)
        ph_acc(ph0, ph1, pc, L.Hs[0][k], L.Hs[1][k]);
        if (!sym) {
            double pr = ldP_def(P, ld, c[k], i, 0, Kall, Ball, lds, m); // P(c_k, i): 2 sectors per column i
            b_acc(b0, b1, pr, L.Hs[0][k], L.Hs[1][k]);
        }
    }
    if (sym) { b0 = ph0; b1 = ph1; }                           // B = (P H')' when P is symmetric
    double k0, k1, dx;
    gain_from_ph(ph0, ph1, L.Sinv, L.nu, k0, k1, dx);          // EKF_SLAM.m:143
    x[i] = __dadd_rn(x[i], dx);                                // EKF_SLAM.m:144
    Kall[m * lds + i] = make_double2(k0, k1);                  // slot m is not read by ldP_def (q < m)
    Ball[m * lds + i] = make_double2(b0, b1);
}

// ------------------------------------------------------------------ K2b fused batch linearise + gain
// ekf_update_batch on one GPU.  The k observations of a batch are applied strictly in order (each linearised at
// the x and P left by the previous one, EKF_SLAM.m:107), but everything observation q+1 needs from observation q
// lives in the slot set S = {0,1,2} + {a_p, a_p+1 : p < k}: x at S and the |S| x |S| block of P.  k_batch_lin
// (one warp) walks the k linearisations on that block, keeping it current with the same rank2_sub chain the
// deferred gathers use; k_batch_gain (one thread per state row) then rebuilds row i of every K_q / B_q from the
// |S| entries P0(i, S) it loads once.  Same operands, same order, same primitives as k_linearize + k_gain run k
// times: the results are bit-identical (tests/test_parity_gpu.py::test_fused_batch_is_bit_identical).
// Sharded filters pass Pc/ldc: the exchanged current columns of the slot set (Pc[slot][i] = P(i, col(slot)), see
// ekf_shard.cu) replace the gathers from P, and [c_lo, c_hi) names the columns this rank stores.
// View of the 5 x 5 block P(c,c) of observation q inside the slot block G (row stride EKF_BATCH_SLOTS + 1).
struct GSel {
    const double* g; const int* sl;
    struct Row { const double* g; const int* sl; int a; __device__ double operator[](int b) const { return g[sl[a] * (EKF_BATCH_SLOTS + 1) + sl[b]]; } };
    __device__ Row operator[](int a) const { return Row{g, sl, a}; }
};

// The batch description reaches the kernels either by value (slots decided on the host) or from device memory (`src.ob`,
// written by k_measure_decide: slots decided on the device, the host never sees them before it enqueues the update).
// Either way it is staged in shared memory once per CTA.  `src.skip` (device word, may be null): the decision kernel
// found a row the device path cannot serve; the batch then contributes ZERO terms (k_batch_gain fills its slots with
// zeros and leaves x alone), so every flush the host has already enqueued behind it stays valid.
__device__ __forceinline__ void stage_batch_obs(BatchObs& sh, const BatchObs& arg, const BatchObs* dev) {
    const int q = threadIdx.x;
    if (q < EKF_BATCH_MAX) {
        if (dev) { sh.z0[q] = dev->z0[q]; sh.z1[q] = dev->z1[q]; sh.R0[q] = dev->R0[q]; sh.R1[q] = dev->R1[q]; sh.a[q] = dev->a[q]; }
        else { sh.z0[q] = arg.z0[q]; sh.z1[q] = arg.z1[q]; sh.R0[q] = arg.R0[q]; sh.R1[q] = arg.R1[q]; sh.a[q] = arg.a[q]; }
    }
    if (q == 0) sh.k = dev ? dev->k : arg.k;
    __syncthreads();
}

// The linearisation chain of a batch on the (3+2k)^2 slot block, for a CTA of NT threads (shared by k_batch_lin, NT = 384: one
// G entry per thread, and by block 0 of k_batch_front, NT = 128: three).  Warp 0 walks the chain, warp 1 evaluates the bearing
// (atan2d, the longest single piece) beside it -- lanes of ONE warp would run the two branches one after the other.  Same
// pieces, same bits as the one-warp form.  Kp / Bp: shared-memory staging for the pending terms' slot entries.
// Raw slot block: exchanged / gathered copies Pc when given, else the stored matrix (rows / columns 0..2 from the strip S
// when given).
template <int NT>
__device__ __forceinline__ void batch_lin_body(const double* __restrict__ P, int64_t ld, const double* __restrict__ x, int sym,
                                               const BatchObs& ob, BatchLin* __restrict__ out, const double* __restrict__ Pc, int64_t ldc,
                                               const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int m0,
                                               const double* __restrict__ S,
                                               double2 (*Kp)[EKF_BATCH_SLOTS], double2 (*Bp)[EKF_BATCH_SLOTS]) {
    __shared__ double G[EKF_BATCH_SLOTS][EKF_BATCH_SLOTS + 1];
    __shared__ double xs[EKF_BATCH_SLOTS];
    __shared__ double2 Ks[EKF_BATCH_SLOTS], Bs[EKF_BATCH_SLOTS];
    __shared__ double Hs[2][5], HPs[2][5], Ss[2][2], Sinv[2][2], nu[2];
    __shared__ int col[EKF_BATCH_SLOTS];
    __shared__ int bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int k = ob.k, ns = 3 + 2 * k;
    if (tid < ns) col[tid] = tid < 3 ? tid : ob.a[(tid - 3) >> 1] + ((tid - 3) & 1);
    if (tid == 0) bad = 0;
    __syncthreads();
    // m0 = 0: nothing is pending at batch start; m0 > 0 (cross-step deferral): current = stored - the pending terms.
    // The pending K_q(col_s), B_q(col_s) are staged in shared memory first -- one round trip to memory for all of them
    // instead of one per term and entry (beside a K3 pass that saturates HBM a dependent load costs microseconds);
    // every entry then subtracts the same terms in the same order as ldP_def.
    for (int e = tid; e < m0 * ns; e += NT) {
        const int q = e / ns, sl = e - q * ns;
        Kp[q][sl] = Kall[q * lds + col[sl]];
        Bp[q][sl] = Ball[q * lds + col[sl]];
    }
    for (int e = tid; e < ns * ns; e += NT) {                      // raw stored / exchanged element
        const int gi = e / ns, gj = e - gi * ns;
        double g0;
        if (Pc) g0 = Pc[(int64_t)gj * ldc + col[gi]];
        else {
            const int ci = col[gi], cj = col[gj];
            const int lo = (sym && ci > cj) ? cj : ci, hi = (sym && ci > cj) ? ci : cj;
            g0 = (S && lo < 3) ? S[4 * (int64_t)hi + lo] : P[(int64_t)lo + (int64_t)hi * ld];
        }
        G[gi][gj] = g0;
    }
    if (tid < ns) xs[tid] = x[col[tid]];
    __syncthreads();
    for (int e = tid; e < ns * ns; e += NT) {
        const int gi = e / ns, gj = e - gi * ns;
        const bool sw = sym && col[gi] > col[gj];                  // SYM: the live element is (min, max)
        const int si = sw ? gj : gi, sj = sw ? gi : gj;
        double g0 = G[gi][gj];
        for (int q = 0; q < m0; ++q) g0 = rank2_sub(g0, Kp[q][si], Bp[q][sj]);
        G[gi][gj] = g0;
    }
    __syncthreads();
    for (int q = 0; q < k; ++q) {
        const int sa = 3 + 2 * q;
        const int sl[5] = {0, 1, 2, sa, sa + 1};
        if (warp == 0) {   // the linearisation, entry-wise across the warp (same pieces, same bits as measurement_model / innovation_cov)
            const double xr[3] = {xs[0], xs[1], xs[2]};
            const MMPrep mp = mm_prep(xr, xs[sa], xs[sa + 1]);               // redundantly on every lane
            if (lane < 10) { const double hv = mm_H(mp, lane); Hs[lane / 5][lane % 5] = hv; out->Hs[q][lane / 5][lane % 5] = hv; }
            else if (lane == 10) { nu[0] = ob.z0[q] - mp.sq; out->nu[q][0] = nu[0]; }                   // EKF_SLAM.m:144, not wrapped (F10)
            __syncwarp();
            if (lane < 10) HPs[lane / 5][lane % 5] = ic_HP(Hs, GSel{&G[0][0], sl}, lane / 5, lane % 5);
            __syncwarp();
            const double Rd[2] = {ob.R0[q], ob.R1[q]};
            if (lane < 4) Ss[lane / 2][lane % 2] = ic_S(HPs, Hs, Rd, lane / 2, lane % 2);
            __syncwarp();
            if (lane == 0) {
                double Si[2][2];
                ic_inv(Ss, Si);
                for (int r = 0; r < 2; ++r)
                    for (int t = 0; t < 2; ++t) {
                        Sinv[r][t] = Si[r][t]; out->Sinv[q][r][t] = Si[r][t];
                        if (!isfinite(Ss[r][t]) || !isfinite(Si[r][t])) bad = 1;
                    }
            }
        } else if (tid == 32) {
            const double xr[3] = {xs[0], xs[1], xs[2]};
            const MMPrep mp = mm_prep(xr, xs[sa], xs[sa + 1]);
            nu[1] = ob.z1[q] - mm_bearing(mp, xr[2]);
            out->nu[q][1] = nu[1];
        }
        __syncthreads();
        if (tid < ns) {
            double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                ph_acc(ph0, ph1, G[tid][sl[c]], Hs[0][c], Hs[1][c]);          // P(col_i, c_k)
                if (!sym) b_acc(b0, b1, G[sl[c]][tid], Hs[0][c], Hs[1][c]);   // P(c_k, col_i)
            }
            if (sym) { b0 = ph0; b1 = ph1; }
            double k0, k1, dx;
            gain_from_ph(ph0, ph1, Sinv, nu, k0, k1, dx);
            xs[tid] = __dadd_rn(xs[tid], dx);
            Ks[tid] = make_double2(k0, k1);
            Bs[tid] = make_double2(b0, b1);
            out->Kslot[q][tid] = Ks[tid];
            out->Bslot[q][tid] = Bs[tid];
        }
        __syncthreads();
        for (int e = tid; e < ns * ns; e += NT) {                  // keep the block current: P <- P - K_q B_q
            const int gi = e / ns, gj = e - gi * ns;
            const bool gup = !sym || col[gi] <= col[gj];           // SYM: the live element is (min, max)
            G[gi][gj] = rank2_sub(G[gi][gj], gup ? Ks[gi] : Ks[gj], gup ? Bs[gj] : Bs[gi]);
        }
        __syncthreads();
    }
    if (tid == 0) { out->nonfinite = bad; if (bad) report_nonfinite(out->nf); }
}

// __launch_bounds__(384, 4): <= 42 registers, i.e. 16 K registers per CTA.  Under the lazy flush this kernel has to start
// beside a K3 grid that fills every SM (3 CTAs x 288 threads x 72 registers): the slot one retiring flush CTA frees is
// 20.7 K registers, and a CTA that does not fit it waits for the flush's tail instead of overlapping with it.
__global__ void __launch_bounds__(384, 4)
k_batch_lin(const double* __restrict__ P, int64_t ld, const double* __restrict__ x, int sym,
                            const BatchObs ob_arg, BatchLin* __restrict__ out, const double* __restrict__ Pc, int64_t ldc,
                            const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int m0,
                            ObSrc src) {
    __shared__ BatchObs ob;
    stage_batch_obs(ob, ob_arg, src.ob);
    if (src.skip && *src.skip) return;
    __shared__ double2 Kp[EKF_MAX_WINDOW][EKF_BATCH_SLOTS], Bp[EKF_MAX_WINDOW][EKF_BATCH_SLOTS];
    batch_lin_body<384>(P, ld, x, sym, ob, out, Pc, ldc, Kall, Ball, lds, m0, nullptr, Kp, Bp);
}

// Front of a batch in ONE launch (SYM storage): block 0 walks the linearisation chain on the slot block while blocks 1.. bring
// every state row's 3+2k slot entries up to date -- gathered from the stored matrix + robot strip (gather != 0: continuous
// flush) or taken from the exchanged copies (slabs), then corrected by the m0 pending terms exactly as k_batch_gain would
// (same terms, same order, same operands) -- and leave them in Pc.  k_batch_gain then runs with nothing pending (m0 = 0,
// output slots o0 = m0): the pending-term phase of the gains, which does not depend on the linearisation, has left the
// serial chain, and so have the gather kernel's launch and its round trip.  Bit-identical to lin + gain(m0).
__global__ void __launch_bounds__(128)
k_batch_front(const double* __restrict__ P, int64_t ld, const double* __restrict__ S, const double* __restrict__ x, int n,
              const BatchObs ob_arg, BatchLin* __restrict__ out, double* __restrict__ Pc, int64_t ldc, int gather,
              const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int m0, ObSrc src) {
    __shared__ BatchObs ob;
    stage_batch_obs(ob, ob_arg, src.ob);
    if (src.skip && *src.skip) return;
    __shared__ double2 Kc[EKF_MAX_WINDOW][EKF_BATCH_SLOTS], Bc[EKF_MAX_WINDOW][EKF_BATCH_SLOTS];   // pending K_q(col_s,:), B_q(:,col_s)
    if (blockIdx.x == 0) {
        batch_lin_body<128>(P, ld, x, 1, ob, out, gather ? nullptr : Pc, ldc, Kall, Ball, lds, m0, S, Kc, Bc);
        return;
    }
    const int k = ob.k, ns = 3 + 2 * k;
    for (int e = threadIdx.x; e < m0 * ns; e += blockDim.x) {
        const int q = e / ns, sl = e - q * ns;
        const int c = sl < 3 ? sl : ob.a[(sl - 3) >> 1] + ((sl - 3) & 1);
        Kc[q][sl] = Kall[q * lds + c];
        Bc[q][sl] = Ball[q * lds + c];
    }
    __syncthreads();
    const int i = (blockIdx.x - 1) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double pc[EKF_BATCH_SLOTS];
    bool own[EKF_BATCH_SLOTS];                                     // element (i, col_s) is stored as (i, col_s), i <= col_s
#pragma unroll
    for (int s = 0; s < EKF_BATCH_SLOTS; ++s) {
        pc[s] = 0.0; own[s] = true;
        if (s < ns) {
            const int c = s < 3 ? s : ob.a[(s - 3) >> 1] + ((s - 3) & 1);
            own[s] = i <= c;
            if (gather) {
                const int lo = i < c ? i : c, hi = i < c ? c : i;
                pc[s] = (S && lo < 3) ? S[4 * (int64_t)hi + lo] : P[(int64_t)lo + (int64_t)hi * ld];
            } else pc[s] = Pc[(int64_t)s * ldc + i];
        }
    }
    for (int q0 = 0; q0 < m0; q0 += 4) {                           // as in k_batch_gain: four terms' operands in flight
        double2 kq[4], bq[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (q0 + u < m0) { kq[u] = Kall[(q0 + u) * lds + i]; bq[u] = Ball[(q0 + u) * lds + i]; }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (q0 + u < m0) {
                const int q = q0 + u;
#pragma unroll
                for (int s = 0; s < EKF_BATCH_SLOTS; ++s)
                    if (s < ns) pc[s] = own[s] ? rank2_sub(pc[s], kq[u], Bc[q][s]) : rank2_sub(pc[s], Kc[q][s], bq[u]);
            }
        }
    }
#pragma unroll
    for (int s = 0; s < EKF_BATCH_SLOTS; ++s)
        if (s < ns) Pc[(int64_t)s * ldc + i] = pc[s];
}

__global__ void __launch_bounds__(128)
k_batch_gain(const double* __restrict__ P, int64_t ld, double* __restrict__ x, int n, int sym, const BatchObs ob_arg,
             const BatchLin* __restrict__ lin, double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds,
             const double* __restrict__ Pc, int64_t ldc, int c_lo, int c_hi, int m0,
             double2* __restrict__ Kout, double2* __restrict__ Bout, int o0, ObSrc src, double* __restrict__ xhost) {
    // xhost (may be null): mapped pinned host copy of x, written beside x so that a caller who reads x after the batch waits
    // for this kernel's event and needs no device-to-host copy behind it
    __shared__ BatchObs ob;
    stage_batch_obs(ob, ob_arg, src.ob);
    if (src.skip && *src.skip) {           // the batch contributes nothing: zero terms in its slots, x untouched
        const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
        if (i0 < n) {
            for (int q = 0; q < ob.k; ++q) { Kout[(o0 + q) * lds + i0] = make_double2(0.0, 0.0); Bout[(o0 + q) * lds + i0] = make_double2(0.0, 0.0); }
            if (xhost) xhost[i0] = x[i0];
        }
        return;
    }
    // Kall / Ball: the m0 pending terms (read only).  The k new terms go to slots o0 .. o0+k-1 of Kout / Bout: the same
    // slabs with o0 = m0 normally; the OTHER pair of slabs with o0 = 0 when the pending terms are being flushed by a
    // concurrent K3 launch (lazy flush, ekf_capi.cu).
    __shared__ BatchLin L;
    __shared__ double2 Kc[EKF_MAX_WINDOW][EKF_BATCH_SLOTS], Bc[EKF_MAX_WINDOW][EKF_BATCH_SLOTS];   // pending K_q(col_s,:), B_q(:,col_s)
    for (int e = threadIdx.x; e < (int)(sizeof(BatchLin) / sizeof(double)); e += blockDim.x)
        ((double*)&L)[e] = ((const double*)lin)[e];
    const int k = ob.k, ns = 3 + 2 * k;
    const bool pend = m0 > 0;                                      // cross-step deferral: current = stored (or exchanged) P0 - pending terms
    if (pend)
        for (int e = threadIdx.x; e < m0 * ns; e += blockDim.x) {
            const int q = e / ns, sl = e - q * ns;
            const int c = sl < 3 ? sl : ob.a[(sl - 3) >> 1] + ((sl - 3) & 1);
            Kc[q][sl] = Kall[q * lds + c];
            Bc[q][sl] = Ball[q * lds + c];
        }
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool mine = i >= c_lo && i < c_hi;                       // column i of P is stored here (always, on one GPU)
    double pc[EKF_BATCH_SLOTS], pr[EKF_BATCH_SLOTS];               // current P(i, col_s) and (FULL) P(col_s, i)
    bool own[EKF_BATCH_SLOTS];                                     // SYM: element (i, col_s) is stored as (i, col_s), i <= col_s
#pragma unroll
    for (int s = 0; s < EKF_BATCH_SLOTS; ++s) {
        pc[s] = 0.0; pr[s] = 0.0; own[s] = true;
        if (s < ns) {
            const int c = s < 3 ? s : ob.a[(s - 3) >> 1] + ((s - 3) & 1);
            own[s] = !sym || i <= c;
            pc[s] = Pc ? Pc[(int64_t)s * ldc + i] : ldP(P, ld, i, c, sym);
            if (!sym && mine) pr[s] = P[(int64_t)c + (int64_t)(i - c_lo) * ld];    // row c of the (local) column i
        }
    }
    if (pend) {
        // same terms, same order as ldP_def element by element; this row's K_q(i,:), B_q(:,i) are loaded once per term,
        // four terms ahead (beside a K3 pass that saturates HBM every dependent load costs microseconds)
        for (int q0 = 0; q0 < m0; q0 += 4) {
            double2 kq[4], bq[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (q0 + u < m0) { kq[u] = Kall[(q0 + u) * lds + i]; bq[u] = Ball[(q0 + u) * lds + i]; }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (q0 + u < m0) {
                    const int q = q0 + u;
                    const double2 ki = kq[u], bi = bq[u];
#pragma unroll
                    for (int s = 0; s < EKF_BATCH_SLOTS; ++s) {
                        if (s < ns) {
                            pc[s] = own[s] ? rank2_sub(pc[s], ki, Bc[q][s]) : rank2_sub(pc[s], Kc[q][s], bi);
                            if (!sym) pr[s] = rank2_sub(pr[s], Kc[q][s], bi);
                        }
                    }
                }
            }
        }
    }
    double xi = x[i];
#pragma unroll
    for (int q = 0; q < EKF_BATCH_MAX; ++q) {
        if (q < k) {
            double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                const int s = c < 3 ? c : 3 + 2 * q + (c - 3);
                ph_acc(ph0, ph1, pc[s], L.Hs[q][0][c], L.Hs[q][1][c]);
                if (!sym) b_acc(b0, b1, pr[s], L.Hs[q][0][c], L.Hs[q][1][c]);
            }
            if (sym) { b0 = ph0; b1 = ph1; }
            double k0, k1, dx;
            gain_from_ph(ph0, ph1, L.Sinv[q], L.nu[q], k0, k1, dx);
            xi = __dadd_rn(xi, dx);
            const double2 Ki = make_double2(k0, k1), Bi = make_double2(b0, b1);
            Kout[(o0 + q) * lds + i] = Ki;
            if (sym || mine) Bout[(o0 + q) * lds + i] = Bi;        // FULL slabs: B_q(:, i) only where column i lives
#pragma unroll
            for (int s = 0; s < EKF_BATCH_SLOTS; ++s) {
                if (s < ns) {
                    pc[s] = own[s] ? rank2_sub(pc[s], Ki, L.Bslot[q][s]) : rank2_sub(pc[s], L.Kslot[q][s], Bi);
                    if (!sym && mine) pr[s] = rank2_sub(pr[s], L.Kslot[q][s], Bi);
                }
            }
        }
    }
    x[i] = xi;
    if (xhost) xhost[i] = xi;
}

// Raw copies of the slot columns of a batch, Pc[slot][i] = stored P(i, col(slot)) (SYM: the live element (min, max)):
// what the sharded exchange delivers, gathered locally.  Taken BEFORE a flush is launched, so that the batch can be
// linearised from them (plus the pending terms) while that flush rewrites P.
__global__ void k_gather_cols(const double* __restrict__ P, int64_t ld, int n, int sym, const BatchObs ob_arg,
                              double* __restrict__ Pc, int64_t ldc, ObSrc src) {
    __shared__ BatchObs ob;
    stage_batch_obs(ob, ob_arg, src.ob);
    const int slot = blockIdx.y;
    if (slot >= 3 + 2 * ob.k) return;
    const int c = slot < 3 ? slot : ob.a[(slot - 3) >> 1] + ((slot - 3) & 1);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) Pc[(int64_t)slot * ldc + i] = ldP(P, ld, i, c, sym);
}

void launch_gather_cols(cudaStream_t st, const double* P, int64_t ld, int n, int sym, const BatchObs& ob, double* Pc, int64_t ldc,
                        ObSrc src) {
    dim3 grid((n + 255) / 256, 3 + 2 * ob.k);      // ob.k on the host is the chunk size in both modes
    k_gather_cols<<<grid, 256, 0, st>>>(P, ld, n, sym, ob, Pc, ldc, src);
}

// ------------------------------------------------------------------ robot strip (continuous flush, ekf_capi.cu "shadow")
// While a K3 pass streams the whole of P on its own stream, nothing else may write P -- but predict rewrites rows 0, 1
// of every column (F P F' + Q).  In that mode the authoritative copy of rows 0..2 of the stored matrix lives in a strip,
// S[4 j + r] = P0(r, j) (one 32-byte sector per column; the 3 x 3 corner is stored in full): predict works on the strip,
// the gather of a batch's slot columns takes every element with a row or column index < 3 from it, and when a pass
// retires the strip catches up with it (k_strip_fix: S <- S - sum_q K_q B_q over the retired terms, which the
// predicts since have transformed together with the strip).  k_strip_store writes it back when P must be whole.
// All strip kernels take a column range: the whole matrix on one GPU (c_lo = 0, ncols = n), the local columns [c_lo, c_lo + ncols)
// of a slab otherwise (P = local slab, S[4 t + r] = P0(r, c_lo + t)); the 3 x 3 corner lives on the rank with c_lo = 0.
__global__ void k_strip_load(const double* __restrict__ P, int64_t ld, int c_lo, int ncols, double* __restrict__ S) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncols) return;
    const int j = c_lo + t;
#pragma unroll
    for (int r = 0; r < 3; ++r) S[4 * (int64_t)t + r] = r <= j ? P[(int64_t)r + (int64_t)t * ld] : P[(int64_t)j + (int64_t)r * ld];   // r > j only in the corner (c_lo = 0)
    S[4 * (int64_t)t + 3] = 0.0;
}

__global__ void k_strip_store(double* __restrict__ P, int64_t ld, int c_lo, int ncols, const double* __restrict__ S) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncols) return;
    const int j = c_lo + t;
#pragma unroll
    for (int r = 0; r < 3; ++r)
        if (r <= j) P[(int64_t)r + (int64_t)t * ld] = S[4 * (int64_t)t + r];
}

// k_predict_P / k_predict_shard on the strip (same expressions), plus the pending window's K_q <- F K_q, B_q <- B_q F' in an
// extra block (every rank transforms its own replica of the slabs).
__global__ void k_predict_strip(double* __restrict__ S, const double* __restrict__ x, int c_lo, int ncols, double u1, double u2, double C,
                                double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds, int m) {
    const double th = x[2];
    const double st = sind(th), ct = cosd(th);
    const double a = -1.0 * u1 * st;      // F(1,3), EKF_SLAM.m:63
    const double b = u1 * ct;             // F(2,3), EKF_SLAM.m:64
    if (m > 0 && blockIdx.x == gridDim.x - 1) {
        const int t = threadIdx.x;
        if (t < 2 * m) {
            double2* Sl = (t & 1 ? Ball : Kall) + (int64_t)(t >> 1) * lds;
            const double2 v0 = Sl[0], v1 = Sl[1], v2 = Sl[2];
            Sl[0] = make_double2(v0.x + a * v2.x, v0.y + a * v2.y);
            Sl[1] = make_double2(v1.x + b * v2.x, v1.y + b * v2.y);
        }
        return;
    }
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = c_lo + t;
    if (t == 0 && c_lo == 0) {
        double M[3][3];
        for (int i = 0; i < 3; ++i)
            for (int jj = 0; jj < 3; ++jj) M[i][jj] = S[4 * (i > jj ? i : jj) + (i > jj ? jj : i)];      // the live (min, max) entry, like ldP
        for (int jj = 0; jj < 3; ++jj) {     // F*P
            M[0][jj] = M[0][jj] + a * M[2][jj];
            M[1][jj] = M[1][jj] + b * M[2][jj];
        }
        for (int i = 0; i < 3; ++i) {     // (F*P)*F'
            M[i][0] = M[i][0] + M[i][2] * a;
            M[i][1] = M[i][1] + M[i][2] * b;
        }
        double W[3] = {u1 * ct, u1 * st, u2};      // EKF_SLAM.m:42
        for (int i = 0; i < 3; ++i)
            for (int jj = 0; jj < 3; ++jj) S[4 * jj + i] = M[i][jj] + (W[i] * C) * W[jj];   // :44,47
    } else if (j >= 3 && t < ncols) {
        double* col = S + 4 * (int64_t)t;
        const double p2 = col[2];
        col[0] = col[0] + a * p2;
        col[1] = col[1] + b * p2;
    }
}

// The m retired terms leave the strip: S(r, j) <- S(r, j) - sum_q K_q(min) B_q(max), in ldP_def's order (K, B: all rows, global index).
__global__ void k_strip_fix(double* __restrict__ S, int c_lo, int ncols, const double2* __restrict__ Kall, const double2* __restrict__ Ball,
                            int64_t lds, int m) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = t < ncols;
    const int j = c_lo + t;
    double v[3];
    int lo[3], hi[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        lo[r] = r < j ? r : j; hi[r] = r < j ? j : r;
        v[r] = in ? S[4 * (int64_t)(hi[r] - c_lo) + lo[r]] : 0.0;     // hi != j only in the corner, where c_lo = 0
    }
    __syncthreads();                      // corner: every thread of block 0 has read before any of them writes a mirror entry
    if (!in) return;
    for (int q0 = 0; q0 < m; q0 += 4) {                    // four terms' operands in flight at a time
        double2 kq[4][3], bq[4][3];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (q0 + u < m) {
#pragma unroll
                for (int r = 0; r < 3; ++r) { kq[u][r] = Kall[(q0 + u) * lds + lo[r]]; bq[u][r] = Ball[(q0 + u) * lds + hi[r]]; }
            }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (q0 + u < m) {
#pragma unroll
                for (int r = 0; r < 3; ++r) v[r] = rank2_sub(v[r], kq[u][r], bq[u][r]);
            }
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) S[4 * (int64_t)t + r] = v[r];
}

// k_gather_cols with the strip in charge of rows / columns 0..2.  One thread = one row of ALL slot columns (<= 19 independent
// loads): the grid is n / 256 CTAs instead of 19 n / 256 -- beside a K3 pass that fills every SM a CTA can only start when a
// pass CTA retires, so the number of CTAs, not the work, sets this kernel's latency.
__global__ void k_gather_cols_strip(const double* __restrict__ P, int64_t ld, int n, const BatchObs ob_arg,
                                    double* __restrict__ Pc, int64_t ldc, ObSrc src, const double* __restrict__ S) {
    __shared__ BatchObs ob;
    stage_batch_obs(ob, ob_arg, src.ob);
    const int ns = 3 + 2 * ob.k;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v[EKF_BATCH_SLOTS];
#pragma unroll
    for (int slot = 0; slot < EKF_BATCH_SLOTS; ++slot) {
        if (slot < ns) {
            const int c = slot < 3 ? slot : ob.a[(slot - 3) >> 1] + ((slot - 3) & 1);
            const int lo = i < c ? i : c, hi = i < c ? c : i;
            v[slot] = lo < 3 ? S[4 * (int64_t)hi + lo] : P[(int64_t)lo + (int64_t)hi * ld];
        }
    }
#pragma unroll
    for (int slot = 0; slot < EKF_BATCH_SLOTS; ++slot)
        if (slot < ns) Pc[(int64_t)slot * ldc + i] = v[slot];
}

// Window bookkeeping: terms [mA, mA + mB) move to the front of the window (the mA retired ones drop out) and, when Kf is
// given, are also frozen into (Kf, Bf) for the K3 pass that is about to start (predict keeps transforming the window).
__global__ void k_rotate_slabs(double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds, int rows, int mA, int mB,
                               double2* __restrict__ Kf, double2* __restrict__ Bf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
#pragma unroll 4
    for (int q = 0; q < mB; ++q) {        // ascending q: reads stay ahead of writes
        const double2 kv = Kall[(int64_t)(mA + q) * lds + i], bv = Ball[(int64_t)(mA + q) * lds + i];
        if (mA > 0) { Kall[(int64_t)q * lds + i] = kv; Ball[(int64_t)q * lds + i] = bv; }
        if (Kf) { Kf[(int64_t)q * lds + i] = kv; Bf[(int64_t)q * lds + i] = bv; }
    }
}

void launch_strip_load(cudaStream_t st, const double* P, int64_t ld, int c_lo, int ncols, double* S) {
    if (ncols > 0) k_strip_load<<<(ncols + 255) / 256, 256, 0, st>>>(P, ld, c_lo, ncols, S);
}
void launch_strip_store(cudaStream_t st, double* P, int64_t ld, int c_lo, int ncols, const double* S) {
    if (ncols > 0) k_strip_store<<<(ncols + 255) / 256, 256, 0, st>>>(P, ld, c_lo, ncols, S);
}
void launch_predict_strip(cudaStream_t st, double* S, const double* xprev, int c_lo, int ncols, double u1, double u2, double C,
                          double2* Kall, double2* Ball, int64_t lds, int m) {
    const int nb = (ncols + 255) / 256 + (m > 0 ? 1 : 0);
    if (nb > 0) k_predict_strip<<<nb, 256, 0, st>>>(S, xprev, c_lo, ncols, u1, u2, C, Kall, Ball, lds, m);
}
void launch_strip_fix(cudaStream_t st, double* S, int c_lo, int ncols, const double2* Kall, const double2* Ball, int64_t lds, int m) {
    if (m > 0 && ncols > 0) k_strip_fix<<<(ncols + 255) / 256, 256, 0, st>>>(S, c_lo, ncols, Kall, Ball, lds, m);
}
void launch_gather_cols_strip(cudaStream_t st, const double* P, int64_t ld, int n, const BatchObs& ob, double* Pc, int64_t ldc,
                              ObSrc src, const double* S) {
    k_gather_cols_strip<<<(n + 255) / 256, 256, 0, st>>>(P, ld, n, ob, Pc, ldc, src, S);
}
void launch_rotate_slabs(cudaStream_t st, double2* Kall, double2* Ball, int64_t lds, int rows, int mA, int mB, double2* Kf,
                         double2* Bf) {
    if (mB > 0 && (mA > 0 || Kf)) k_rotate_slabs<<<(rows + 255) / 256, 256, 0, st>>>(Kall, Ball, lds, rows, mA, mB, Kf, Bf);
}

// ------------------------------------------------------------------ K5 append
// New landmark at state rows n, n+1 (0-based), n = 3 + 2N.  EKF_SLAM.m:84-97.
__global__ void k_append(double* __restrict__ P, int64_t ld, double* __restrict__ x, double* __restrict__ s,
                         int N, double u1, double u2, double R0, double R1, double px, double py, double sig, int sym) {
    const int n = 3 + 2 * N;
    const double th = x[2];
    const double j02 = -u1 * sind(th), j12 = u1 * cosd(th);    // jxr = [1 0 j02; 0 1 j12], :84-85
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < N) {                                               // old landmark t: blocks F and G, :94-97
        int a = 3 + 2 * t;
        for (int m = 0; m < 2; ++m) {
            double p0 = ldP(P, ld, a + m, 0, sym), p1 = ldP(P, ld, a + m, 1, sym), p2 = ldP(P, ld, a + m, 2, sym);
            double f0 = dot3(1.0, p0, 0.0, p1, j02, p2);        // jxr * P(lm,1:3)'
            double f1 = dot3(0.0, p0, 1.0, p1, j12, p2);
            P[(int64_t)(a + m) + (int64_t)n * ld] = f0;        // G (upper part)
            P[(int64_t)(a + m) + (int64_t)(n + 1) * ld] = f1;
            P[(int64_t)n + (int64_t)(a + m) * ld] = f0;        // F (lower part; don't-care in SYM)
            P[(int64_t)(n + 1) + (int64_t)(a + m) * ld] = f1;
        }
    } else if (t == N) {                                       // robot block, new diagonal block, x, s
        double Prr[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Prr[i][j] = ldP(P, ld, i, j, sym);
        const double jxr[2][3] = {{1.0, 0.0, j02}, {0.0, 1.0, j12}};
        const double cu = cosd(u2), su = sind(u2);
        const double jz[2][2] = {{cu, -u1 * su}, {su, u1 * cu}};     // :87-88
        double JP[2][3];
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j) JP[r][j] = dot3(jxr[r][0], Prr[0][j], jxr[r][1], Prr[1][j], jxr[r][2], Prr[2][j]);
        double JR[2][2] = {{__dmul_rn(jz[0][0], R0), __dmul_rn(jz[0][1], R1)}, {__dmul_rn(jz[1][0], R0), __dmul_rn(jz[1][1], R1)}};
        for (int r = 0; r < 2; ++r)
            for (int c = 0; c < 2; ++c) {
                double v = __dadd_rn(dot3(JP[r][0], jxr[c][0], JP[r][1], jxr[c][1], JP[r][2], jxr[c][2]),
                           __dadd_rn(__dmul_rn(JR[r][0], jz[c][0]), __dmul_rn(JR[r][1], jz[c][1])));                  // :91
                P[(int64_t)(n + r) + (int64_t)(n + c) * ld] = v;
            }
        for (int i = 0; i < 3; ++i)
            for (int c = 0; c < 2; ++c) {
                double v = dot3(Prr[i][0], jxr[c][0], Prr[i][1], jxr[c][1], Prr[i][2], jxr[c][2]);   // :92
                P[(int64_t)i + (int64_t)(n + c) * ld] = v;
                P[(int64_t)(n + c) + (int64_t)i * ld] = v;                                          // :93
            }
        x[n] = px; x[n + 1] = py;                              // :79
        s[N] = sig;                                            // :70
    }
}

// ------------------------------------------------------------------ K4 association
struct Cand { double cost; int idx; };
__device__ __forceinline__ bool cand_less(double c1, int i1, double c2, int i2) {
    return (c1 < c2) || (c1 == c2 && i1 < i2);     // strict '<' scan == lexicographic (cost, index) min
}

// Slab form (sharded filters, mode B): the grid covers the `lm_cnt` landmarks starting at `lm_lo` whose columns this
// rank stores (P = local slab, column c at P + (c - c_lo) * ld); everything that lives in the robot columns comes from
// the exchanged copies Pc[j][i] = P(i, j), j < 3; the block minima fold into `cand_out` (this rank's candidate) instead
// of the final result.  lm_lo = 0, lm_cnt = N, c_lo = 0, Pc = cand_out = null is the single-GPU kernel, bit for bit.
__global__ void k_associate(const double* __restrict__ P, int64_t ld, const double* __restrict__ x,
                            const double* __restrict__ s, int N, double z0, double z1, double z2,
                            double R0, double R1, double s_cost, double s_thresh, int mode, int sym,
                            Cand* __restrict__ partial, unsigned int* __restrict__ ticket,
                            AssocResult* __restrict__ out, int lm_lo, int lm_cnt, int c_lo,
                            const double* __restrict__ Pc, int64_t ldc, Cand* __restrict__ cand_out) {
    __shared__ double Prr[3][3];
    __shared__ double wc[32];
    __shared__ int wi[32];
    __shared__ bool is_last;
    if (mode == 1 && threadIdx.x < 9)
        Prr[threadIdx.x / 3][threadIdx.x % 3] = Pc ? Pc[(int64_t)(threadIdx.x % 3) * ldc + threadIdx.x / 3]
                                                   : ldP(P, ld, threadIdx.x / 3, threadIdx.x % 3, sym);
    __syncthreads();
    const int kl = blockIdx.x * blockDim.x + threadIdx.x;
    int kk = lm_lo + kl;                                       // 0-based landmark
    double best = CUDART_INF;
    int bidx = INT_MAX;
    P -= (int64_t)c_lo * ld;                                   // global column index -> local slab
    if (kl < lm_cnt) {
        double d = __dsub_rn(z2, s[kk]);
        double sig = __dmul_rn(__dmul_rn(d, 1.0 / s_cost), d); // Correspondence.m:71
        double ll = sig;                                       // :75
        if (mode == 1) {                                       // :74 (commented out upstream)
            int a = 3 + 2 * kk;
            double xr[3] = {x[0], x[1], x[2]};
            double zk[2], Hs[2][5], S[2][2], Si[2][2], Pcc[5][5];
            measurement_model(xr, x[a], x[a + 1], zk, Hs);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) Pcc[i][j] = Prr[i][j];
            for (int i = 0; i < 3; ++i)
                for (int m = 0; m < 2; ++m) {
                    Pcc[i][3 + m] = P[(int64_t)i + (int64_t)(a + m) * ld];
                    Pcc[3 + m][i] = sym ? Pcc[i][3 + m] : (Pc ? Pc[(int64_t)i * ldc + (a + m)] : P[(int64_t)(a + m) + (int64_t)i * ld]);
                }
            Pcc[3][3] = P[(int64_t)a + (int64_t)a * ld];
            Pcc[3][4] = P[(int64_t)a + (int64_t)(a + 1) * ld];
            Pcc[4][3] = sym ? Pcc[3][4] : P[(int64_t)(a + 1) + (int64_t)a * ld];
            Pcc[4][4] = P[(int64_t)(a + 1) + (int64_t)(a + 1) * ld];
            double Rd[2] = {R0, R1};
            innovation_cov(Hs, Pcc, Rd, S, Si);
            double n0 = __dsub_rn(z0, zk[0]), n1 = __dsub_rn(z1, zk[1]);
            double t0 = __dadd_rn(__dmul_rn(n0, Si[0][0]), __dmul_rn(n1, Si[1][0]));   // (nu' * phi^-1) * nu, :69
            double t1 = __dadd_rn(__dmul_rn(n0, Si[0][1]), __dmul_rn(n1, Si[1][1]));
            double pos = __dadd_rn(__dmul_rn(t0, n0), __dmul_rn(t1, n1));
            ll = __dadd_rn(pos, sig);
        }
        if (ll <= s_thresh && ll < CUDART_INF) { best = ll; bidx = kk + 1; }   // :78,:81 (NaN fails the gate; Inf never beats the initial Inf)
    }
    // warp-level first-argmin
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double oc = __shfl_down_sync(0xffffffffu, best, o);
        int oi = __shfl_down_sync(0xffffffffu, bidx, o);
        if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
    }
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { wc[w] = best; wi[w] = bidx; }
    __syncthreads();
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        best = (lane < nw) ? wc[lane] : CUDART_INF;
        bidx = (lane < nw) ? wi[lane] : INT_MAX;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double oc = __shfl_down_sync(0xffffffffu, best, o);
            int oi = __shfl_down_sync(0xffffffffu, bidx, o);
            if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
        }
        if (lane == 0) {
            partial[blockIdx.x].cost = best;
            partial[blockIdx.x].idx = bidx;
            __threadfence();
            unsigned int tk = atomicAdd(ticket, 1u);
            is_last = (tk == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (is_last && w == 0) {                                   // last block folds the per-block minima
        __threadfence();
        best = CUDART_INF; bidx = INT_MAX;
        for (int b = lane; b < (int)gridDim.x; b += 32) {
            double oc = ((volatile Cand*)partial)[b].cost;
            int oi = ((volatile Cand*)partial)[b].idx;
            if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double oc = __shfl_down_sync(0xffffffffu, best, o);
            int oi = __shfl_down_sync(0xffffffffu, bidx, o);
            if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
        }
        if (lane == 0) {
            bool found = (bidx != INT_MAX);
            if (cand_out) { cand_out->cost = best; cand_out->idx = bidx; }     // slab: this rank's candidate
            else {
                out->newLL = found ? 0 : 1;                    // :33,82
                out->index1 = found ? bidx : N + 1;            // :40,84
                out->cost = best;
            }
            *ticket = 0;
        }
    }
}

// Candidate exchange of the sharded mode-B association: rank r stores its (cost, index) into slot r of every peer's
// candidate area (inside the flag allocation: EKF_XF_CAND, double-buffered by epoch parity) and publishes the epoch in
// the peers' association flag words (EKF_XF_ASSOC + r); k_cand_fold waits (bounded, like k_wait_flags) for all ranks
// and takes the lexicographic (cost, index) minimum == the strict-'<' first-minimum scan of Correspondence.m:81 over
// all landmarks.  Every rank computes the same result.  `gathered` (NCCL fallback): candidates already all-gathered.
__global__ void k_cand_none(Cand* c) { c->cost = CUDART_INF; c->idx = INT_MAX; }

__global__ void k_cand_send(const Cand* __restrict__ mine, PeerSet ps, int epoch) {
    const int g = threadIdx.x;
    if (g < ps.world) {
        Cand* dst = (Cand*)(ps.flags[g] + EKF_XF_CAND) + (epoch & 1) * EKF_MAX_PEERS + ps.rank;
        *dst = *mine;
        __threadfence_system();
        asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(ps.flags[g] + EKF_XF_ASSOC + ps.rank), "r"(epoch) : "memory");
    }
}

__global__ void k_cand_fold(const int* __restrict__ flags, const Cand* __restrict__ gathered, int world, int epoch,
                            long long timeout, int* __restrict__ err, int N, AssocResult* __restrict__ out) {
    const int g = threadIdx.x;
    double best = CUDART_INF;
    int bidx = INT_MAX;
    if (g < world) {
        if (gathered) { best = gathered[g].cost; bidx = gathered[g].idx; }
        else {
            const long long t0 = clock64();
            int v;
            do {
                asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(flags + EKF_XF_ASSOC + g) : "memory");
                if (v >= epoch) break;
                if (*(volatile int*)err) break;
                if (clock64() - t0 > timeout) { *(volatile int*)err = 1; __threadfence_system(); break; }
            } while (true);
            const volatile Cand* c = (const volatile Cand*)(flags + EKF_XF_CAND) + (epoch & 1) * EKF_MAX_PEERS + g;
            best = c->cost; bidx = c->idx;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double oc = __shfl_down_sync(0xffffffffu, best, o);
        int oi = __shfl_down_sync(0xffffffffu, bidx, o);
        if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
    }
    if (g == 0) {
        const bool found = (bidx != INT_MAX);
        out->newLL = found ? 0 : 1;                            // Correspondence.m:33,82
        out->index1 = found ? bidx : N + 1;                    // :40,84
        out->cost = best;
    }
}

// Mode-A association for a whole measure() batch: the shipped score (Correspondence.m:75) depends only on z(3) and
// the signature vector s, not on x or P, so the k observations of one measure() call can be gated in ONE launch
// (block b = observation b) and read back with ONE copy; the host re-runs it only after an append changed s.
__global__ void k_associate_sig_batch(const double* __restrict__ s, int N, AssocBatchZ zb, double s_cost,
                                      double s_thresh, AssocResult* __restrict__ out) {
    __shared__ double wc[32];
    __shared__ int wi[32];
    const double z2 = zb.z2[blockIdx.x];
    const double inv = 1.0 / s_cost;
    double best = CUDART_INF;
    int bidx = INT_MAX;
    for (int kk = threadIdx.x; kk < N; kk += blockDim.x) {         // ascending kk per thread: strict '<' keeps the first
        double d = __dsub_rn(z2, s[kk]);
        double ll = __dmul_rn(__dmul_rn(d, inv), d);               // Correspondence.m:71,75
        if (ll <= s_thresh && ll < best) { best = ll; bidx = kk + 1; }   // :78,:81
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double oc = __shfl_down_sync(0xffffffffu, best, o);
        int oi = __shfl_down_sync(0xffffffffu, bidx, o);
        if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { wc[w] = best; wi[w] = bidx; }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = (lane < nw) ? wc[lane] : CUDART_INF;
        bidx = (lane < nw) ? wi[lane] : INT_MAX;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double oc = __shfl_down_sync(0xffffffffu, best, o);
            int oi = __shfl_down_sync(0xffffffffu, bidx, o);
            if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
        }
        if (lane == 0) {
            const bool found = (bidx != INT_MAX);
            out[blockIdx.x].newLL = found ? 0 : 1;                 // :33,82
            out[blockIdx.x].index1 = found ? bidx : N + 1;         // :40,84
            out[blockIdx.x].cost = best;
        }
    }
}


// ------------------------------------------------------------------ measure() decisions on the device
// EKF_SLAM_UC.measure, mode A (EKF_SLAM_UC.m:107-124 with Correspondence.m:28-88 as shipped): block b associates row b
// of observed_LL (the k_associate_sig_batch scan), the last block to finish turns the k results into the batch
// descriptions the update kernels read from device memory -- chunk c = rows 8c .. 8c+7: z, R = z .* Rc (:108), slot
// a = 3 + 2 (idx - 1) -- and writes the actions (+idx) into mapped host memory.  The host enqueues the update kernels
// behind this one WITHOUT waiting for it.  It only takes this path when every row's signature is already in the map
// (it keeps the set of signatures it appended), so no row can need an append; should one disagree (newLL), `skip` is
// raised, the update kernels contribute nothing, and status[0] = 1 tells the host at its next call.
__global__ void k_measure_decide(const double* __restrict__ s, int N, const double* __restrict__ obs, int k, double Rc0,
                                 double Rc1, double s_cost, double s_thresh, BatchObs* __restrict__ chunks,
                                 int* __restrict__ skip, int* __restrict__ status, int* __restrict__ scratch,
                                 unsigned int* __restrict__ ticket) {
    __shared__ double wc[32];
    __shared__ int wi[32];
    __shared__ bool is_last;
    const int b = blockIdx.x;
    const double z2 = obs[b + 2 * k];
    const double inv = 1.0 / s_cost;
    double best = CUDART_INF;
    int bidx = INT_MAX;
    for (int kk = threadIdx.x; kk < N; kk += blockDim.x) {         // ascending kk per thread: strict '<' keeps the first
        double d = __dsub_rn(z2, s[kk]);
        double ll = __dmul_rn(__dmul_rn(d, inv), d);               // Correspondence.m:71,75
        if (ll <= s_thresh && ll < best) { best = ll; bidx = kk + 1; }   // :78,:81
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double oc = __shfl_down_sync(0xffffffffu, best, o);
        int oi = __shfl_down_sync(0xffffffffu, bidx, o);
        if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { wc[w] = best; wi[w] = bidx; }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = (lane < nw) ? wc[lane] : CUDART_INF;
        bidx = (lane < nw) ? wi[lane] : INT_MAX;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double oc = __shfl_down_sync(0xffffffffu, best, o);
            int oi = __shfl_down_sync(0xffffffffu, bidx, o);
            if (cand_less(oc, oi, best, bidx)) { best = oc; bidx = oi; }
        }
        if (lane == 0) {
            scratch[b] = bidx;                                     // INT_MAX: newLL (EKF_SLAM_UC.m:121 would append)
            __threadfence();
            is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        }
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    const int t = threadIdx.x;
    int bad = 0;
    if (t < k) {
        const int idx = ((volatile int*)scratch)[t];
        const int c = t / EKF_BATCH_MAX, q = t - c * EKF_BATCH_MAX;
        const double z0 = obs[t], z1 = obs[t + k];
        bad = (idx == INT_MAX || idx < 1 || idx > N) ? 1 : 0;
        chunks[c].z0[q] = z0;
        chunks[c].z1[q] = z1;
        chunks[c].R0[q] = __dmul_rn(z0, Rc0);                     // EKF_SLAM_UC.m:110, R = diag(z(1) Rc(1), z(2) Rc(2))
        chunks[c].R1[q] = __dmul_rn(z1, Rc1);
        chunks[c].a[q] = bad ? 3 : 3 + 2 * (idx - 1);
        status[1 + t] = bad ? 0 : idx;                             // action: +idx = update of slot idx
        if (q == 0) chunks[c].k = min(EKF_BATCH_MAX, k - c * EKF_BATCH_MAX);
    }
    const int any_bad = __syncthreads_or(bad);
    if (t == 0) {
        *skip = any_bad;
        status[0] = any_bad;
        *ticket = 0;
        __threadfence_system();
    }
}

void launch_measure_decide(cudaStream_t st, const double* s, int N, const double* obs, int k, double Rc0, double Rc1,
                           double s_cost, double s_thresh, BatchObs* ob_chunks, int* skip, int* status_host, int* scratch,
                           unsigned int* ticket) {
    // one block per observation scans all N signatures: 1024 threads on a large map (10 loop trips per thread at 10k instead of 40)
    k_measure_decide<<<k, N >= 4096 ? 1024 : 256, 0, st>>>(s, N, obs, k, Rc0, Rc1, s_cost, s_thresh, ob_chunks, skip, status_host, scratch, ticket);
}

// ------------------------------------------------------------------ state initialisation helpers
__global__ void k_init_lowrank(double* __restrict__ P, int64_t ld, int n, int c_lo, const double* __restrict__ A, int r,
                               const double* __restrict__ d) {
    int i = blockIdx.y * blockDim.x + threadIdx.x;   // columns on grid.x: a slab can hold > 65535 of them (80k landmarks)
    int j = c_lo + blockIdx.x;
    if (i >= n) return;
    double acc = 0.0;
    for (int q = 0; q < r; ++q) acc += A[i + (int64_t)q * n] * A[j + (int64_t)q * n];
    acc = acc / (double)r;
    if (i == j) acc += d[i];
    P[(int64_t)i + (int64_t)(j - c_lo) * ld] = acc;
}

// SYM storage read-back: copy the live upper triangle onto the dead lower one (32x32 smem transpose).
__global__ void k_mirror_upper(double* __restrict__ P, int64_t ld, int n) {
    __shared__ double t[32][33];
    int bi = blockIdx.x, bj = blockIdx.y;
    if (bi > bj) return;
    int i = bi * 32 + threadIdx.x;
    for (int jj = threadIdx.y; jj < 32; jj += blockDim.y) {
        int j = bj * 32 + jj;
        t[jj][threadIdx.x] = (i < n && j < n) ? P[(int64_t)i + (int64_t)j * ld] : 0.0;
    }
    __syncthreads();
    int i2 = bj * 32 + threadIdx.x;                 // destination row (old column)
    for (int jj = threadIdx.y; jj < 32; jj += blockDim.y) {
        int j2 = bi * 32 + jj;                      // destination column (old row)
        if (i2 < n && j2 < n && i2 > j2) P[(int64_t)i2 + (int64_t)j2 * ld] = t[threadIdx.x][jj];
    }
}

// ------------------------------------------------------------------ launchers
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ------------------------------------------------------------------ K6 covariance ellipses (headless plot data)
// EKF_SLAM.m:180-229 draws, for the robot and for every landmark, the ellipse of the 2x2 diagonal block of P:
// eig of the block, extents 2*sqrt(chi_square*lambda) with chi_square = 2.2788, rotated by arc = atan(v_y / v_x) of
// the eigenvector of the LARGER eigenvalue (both branches of the reference's if/else reduce to that), scaled by .25
// (robot, :198) or .50 (landmarks, :221) and centred on the estimate.  There is no graphics stack here, so the kernel
// returns the numbers instead of drawing: out(b, :) = [cx, cy, semi-axis along v_max, semi-axis across, arc (rad)].
__global__ void k_ellipses(const double* __restrict__ P, int64_t ld, const double* __restrict__ x, int N,
                           double* __restrict__ out, int64_t ldo) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;           // block 0 = robot, b >= 1 = landmark b
    if (b > N) return;
    const int i = b == 0 ? 0 : 3 + 2 * (b - 1);
    const double a = P[(int64_t)i + (int64_t)i * ld];
    const double c = P[(int64_t)i + (int64_t)(i + 1) * ld];        // upper element: live in both storages
    const double d = P[(int64_t)(i + 1) + (int64_t)(i + 1) * ld];
    const double h = 0.5 * (a - d);
    const double rad = sqrt(h * h + c * c);
    const double mid = 0.5 * (a + d);
    const double lmax = mid + rad, lmin = mid - rad;
    double vx, vy;                                                 // eigenvector of lmax
    if (c == 0.0) { vx = a > d ? 1.0 : 0.0; vy = a > d ? 0.0 : 1.0; }     // a == d: eig returns I, column 2 is used
    else if (h >= 0.0) { vx = lmax - d; vy = c; }                  // (h + rad, c): no cancellation for h >= 0
    else { vx = c; vy = lmax - a; }                                // (c, rad - h)
    const double chi = 2.2788;
    const double scale = b == 0 ? 0.25 : 0.50;
    out[b] = x[i];
    out[b + ldo] = x[i + 1];
    out[b + 2 * ldo] = scale * (2.0 * sqrt(chi * lmax));
    out[b + 3 * ldo] = scale * (2.0 * sqrt(chi * lmin));
    out[b + 4 * ldo] = atan(vy / vx);
}

void launch_ellipses(cudaStream_t st, const double* P, int64_t ld, const double* x, int N, double* out, int64_t ldo) {
    k_ellipses<<<cdiv(N + 1, 128), 128, 0, st>>>(P, ld, x, N, out, ldo);
}


void launch_predict(cudaStream_t st, double* P, int64_t ld, double* x, int n, double u1, double u2, double C, int sym) {
    k_predict_P<<<cdiv(n, 256), 256, 0, st>>>(P, ld, x, n, u1, u2, C, sym, nullptr, nullptr, 0, 0);
    k_predict_x<<<1, 1, 0, st>>>(x, u1, u2, nullptr, nullptr);
}

void launch_predict_x(cudaStream_t st, double* x, double u1, double u2) { k_predict_x<<<1, 1, 0, st>>>(x, u1, u2, nullptr, nullptr); }

// The two halves of launch_predict for the decoupled path: the state half saves the pre-motion pose into xprev, the
// covariance half reads its heading from there (same arithmetic, bit-identical results).
void launch_predict_x_save(cudaStream_t st, double* x, double u1, double u2, double* xprev, double* pose_host) {
    k_predict_x<<<1, 1, 0, st>>>(x, u1, u2, xprev, pose_host);
}
void launch_predict_P(cudaStream_t st, double* P, int64_t ld, const double* xprev, int n, double u1, double u2, double C, int sym,
                      double2* Kall, double2* Ball, int64_t lds, int m) {
    k_predict_P<<<cdiv(n, 256) + (m > 0 ? 1 : 0), 256, 0, st>>>(P, ld, xprev, n, u1, u2, C, sym, Kall, Ball, lds, m);
}

void launch_linearize_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int a, double z0, double z1,
                           double R0, double R1, int sym, Lin* lin, double2* Kall, double2* Ball, int64_t lds, int m) {
    k_linearize<<<1, 32, 0, st>>>(P, ld, x, a, z0, z1, R0, R1, sym, Kall, Ball, lds, m, lin);
    k_gain<<<cdiv(n, 256), 256, 0, st>>>(P, ld, x, n, a, sym, lin, Kall, Ball, lds, m);
}

void launch_batch_lin_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int sym, const BatchObs& ob,
                           BatchLin* lin, double2* Kall, double2* Ball, int64_t lds, const double* Pc, int64_t ldc,
                           int c_lo, int c_hi, int m0, double2* Kout, double2* Bout, int o0, ObSrc src, double* xhost) {
    if (!Pc) { c_lo = 0; c_hi = n; }
    if (!Kout) { Kout = Kall; Bout = Ball; o0 = m0; }
    k_batch_lin<<<1, 384, 0, st>>>(P, ld, x, sym, ob, lin, Pc, ldc, Kall, Ball, lds, m0, src);   // >= EKF_BATCH_SLOTS^2 = 361 threads
    static int gb = 0;                     // rows per CTA of the gain kernel (EKF_GAIN_BLOCK = 32 | 64 | 128, default 128)
    if (gb == 0) { const char* v = getenv("EKF_GAIN_BLOCK"); gb = v ? atoi(v) : 128; if (gb != 32 && gb != 64) gb = 128; }
    k_batch_gain<<<cdiv(n, gb), gb, 0, st>>>(P, ld, x, n, sym, ob, lin, Kall, Ball, lds, Pc, ldc, c_lo, c_hi, m0, Kout, Bout, o0, src, xhost);
}


// front (linearisation + slot entries brought up to date) and gains of a batch, SYM storage: see k_batch_front
void launch_batch_front_only(cudaStream_t st, const double* P, int64_t ld, const double* S, double* x, int n, const BatchObs& ob,
                             BatchLin* lin, double* Pc, int64_t ldc, int gather, double2* Kall, double2* Ball, int64_t lds, int m0,
                             ObSrc src) {
    const int rows = (gather || m0 > 0) ? cdiv(n, 128) : 0;        // nothing pending and nothing to gather: the chain alone
    k_batch_front<<<1 + rows, 128, 0, st>>>(P, ld, S, x, n, ob, lin, Pc, ldc, gather, Kall, Ball, lds, m0, src);
}
void launch_batch_gain_only(cudaStream_t st, const double* P, int64_t ld, double* x, int n, const BatchObs& ob, BatchLin* lin,
                            double* Pc, int64_t ldc, double2* Kall, double2* Ball, int64_t lds, int m0, ObSrc src, double* xhost) {
    static int gb = 0;
    if (gb == 0) { const char* v = getenv("EKF_GAIN_BLOCK"); gb = v ? atoi(v) : 128; if (gb != 32 && gb != 64) gb = 128; }
    k_batch_gain<<<cdiv(n, gb), gb, 0, st>>>(P, ld, x, n, 1, ob, lin, Kall, Ball, lds, Pc, ldc, 0, n, 0, Kall, Ball, m0, src, xhost);
}
void launch_batch_front_gain(cudaStream_t st, const double* P, int64_t ld, const double* S, double* x, int n, const BatchObs& ob,
                             BatchLin* lin, double* Pc, int64_t ldc, int gather, double2* Kall, double2* Ball, int64_t lds, int m0,
                             ObSrc src, double* xhost) {
    launch_batch_front_only(st, P, ld, S, x, n, ob, lin, Pc, ldc, gather, Kall, Ball, lds, m0, src);
    launch_batch_gain_only(st, P, ld, x, n, ob, lin, Pc, ldc, Kall, Ball, lds, m0, src, xhost);
}

void launch_append(cudaStream_t st, double* P, int64_t ld, double* x, double* s, int N, double u1, double u2,
                   double R0, double R1, double px, double py, double sig, int sym) {
    k_append<<<cdiv(N + 1, 128), 128, 0, st>>>(P, ld, x, s, N, u1, u2, R0, R1, px, py, sig, sym);
}

int assoc_blocks(int N) { return cdiv(N, 256); }

void launch_associate(cudaStream_t st, const double* P, int64_t ld, const double* x, const double* s, int N,
                      double z0, double z1, double z2, double R0, double R1, double s_cost, double s_thresh,
                      int mode, int sym, void* partial, unsigned int* ticket, AssocResult* out) {
    k_associate<<<assoc_blocks(N), 256, 0, st>>>(P, ld, x, s, N, z0, z1, z2, R0, R1, s_cost, s_thresh, mode, sym,
                                                 (Cand*)partial, ticket, out, 0, N, 0, nullptr, 0, nullptr);
}

// ---- association mode B on a slab: local candidate, then a 16-byte (cost, index) exchange over the peer flag words ----
void launch_associate_slab(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int lm_lo, int lm_cnt, const double* x,
                           const double* s, int N, double z0, double z1, double z2, double R0, double R1, double s_cost,
                           double s_thresh, int sym, const double* Pc, int64_t ldc, void* partial, unsigned int* ticket,
                           void* cand_out) {
    if (lm_cnt <= 0) { k_cand_none<<<1, 1, 0, st>>>((Cand*)cand_out); return; }
    k_associate<<<assoc_blocks(lm_cnt), 256, 0, st>>>(Ploc, ld, x, s, N, z0, z1, z2, R0, R1, s_cost, s_thresh, 1, sym,
                                                      (Cand*)partial, ticket, nullptr, lm_lo, lm_cnt, c_lo, Pc, ldc, (Cand*)cand_out);
}

void launch_cand_send(cudaStream_t st, const void* mine, const PeerSet& ps, int epoch) {
    k_cand_send<<<1, 32, 0, st>>>((const Cand*)mine, ps, epoch);
}
void launch_cand_fold(cudaStream_t st, const int* flags, int world, int epoch, long long timeout_cycles, int* err_word, int N,
                      AssocResult* out) {
    k_cand_fold<<<1, 32, 0, st>>>(flags, nullptr, world, epoch, timeout_cycles, err_word, N, out);
}
void launch_cand_fold_gathered(cudaStream_t st, const void* all, int world, int N, AssocResult* out) {
    k_cand_fold<<<1, 32, 0, st>>>(nullptr, (const Cand*)all, world, 0, 0, nullptr, N, out);
}

void launch_associate_sig_batch(cudaStream_t st, const double* s, int N, const AssocBatchZ& zb, int k, double s_cost,
                                double s_thresh, AssocResult* out) {
    k_associate_sig_batch<<<k, 256, 0, st>>>(s, N, zb, s_cost, s_thresh, out);
}

void launch_init_lowrank(cudaStream_t st, double* P, int64_t ld, int n, int c_lo, int ncols, const double* A, int r,
                         const double* d) {
    if (ncols <= 0) return;
    dim3 grid(ncols, cdiv(n, 256));
    k_init_lowrank<<<grid, 256, 0, st>>>(P, ld, n, c_lo, A, r, d);
}

void launch_mirror_upper(cudaStream_t st, double* P, int64_t ld, int n) {
    dim3 grid(cdiv(n, 32), cdiv(n, 32));
    k_mirror_upper<<<grid, dim3(32, 8), 0, st>>>(P, ld, n);
}

}  // namespace ekf

// ------------------------------------------------------------------ K4 without allocation (ekf_associate_static)
// No counterpart in the reference: added for callers that may not call
// cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).  K4 needs scratch memory -- one 16-byte
// (cost, index) candidate per block of 256 landmarks plus the ticket word of the last-block fold; the handle-based
// ekf_associate keeps that scratch in the handle's own allocations, this entry point keeps it in a file-scope device
// array of fixed size, so the call allocates nothing (no cudaMalloc / cudaFree, no library that allocates): the caller
// passes device pointers of the state and of the 16-byte result and a stream, the kernel is the same k_associate.
// The array exists once per GPU that loads the library (65 792 bytes, until exit) and is shared by every call: calls
// that overlap on one GPU (two streams, two threads) race on it -- serialise them.  The ticket word is cleared before
// every launch (a call that was cut short must not poison the next); the candidates need no clearing, every block of
// the grid writes its own before the fold reads them.
#define EKF_STATIC_ASSOC_MAX_BLOCKS 4096                      // x 256 landmarks = EKF_ASSOC_STATIC_MAX_LANDMARKS (header)
#define EKF_STATIC_ASSOC_TICKET_OFF (EKF_STATIC_ASSOC_MAX_BLOCKS * 16)
static __device__ __align__(256) unsigned char g_assoc_static_scratch[EKF_STATIC_ASSOC_TICKET_OFF + 256];

namespace {
// In host code `g_assoc_static_scratch` is only a stand-in for the symbol, not a device address: the address every
// launch passes to the kernel comes from cudaGetSymbolAddress, resolved once per device with that device current.
constexpr int kMaxDev = 64;
void* g_static_ptr[kMaxDev];
bool  g_static_ok[kMaxDev];
std::mutex g_static_mu;

void put_err(char* errbuf, int32_t errlen, const std::string& m) {
    if (errbuf && errlen > 0) { snprintf(errbuf, (size_t)errlen, "%s", m.c_str()); }
}
}  // namespace

extern "C" int32_t ekf_associate_static(const double* d_x, const double* d_P, int64_t ld, const double* d_s, int32_t N,
                                        const double z[3], const double Rdiag[2], double s_cost, double s_thresh,
                                        int32_t assoc_mode, int32_t storage, void* d_result, uint64_t stream,
                                        char* errbuf, int32_t errlen) {
    static_assert(sizeof(ekf::Cand) == 16 && sizeof(ekf::AssocResult) == 16, "scratch / result layout");
    if (!d_x || !d_s || !z || !Rdiag || !d_result) { put_err(errbuf, errlen, "ekf_associate_static: null argument"); return EKF_EINVAL; }
    if (assoc_mode != EKF_ASSOC_SIGNATURE && assoc_mode != EKF_ASSOC_POSITION) { put_err(errbuf, errlen, "ekf_associate_static: assoc_mode must be EKF_ASSOC_SIGNATURE or EKF_ASSOC_POSITION"); return EKF_EINVAL; }
    if (storage != EKF_STORAGE_FULL && storage != EKF_STORAGE_SYM) { put_err(errbuf, errlen, "ekf_associate_static: storage must be EKF_STORAGE_FULL or EKF_STORAGE_SYM"); return EKF_EINVAL; }
    if (N < 1) { put_err(errbuf, errlen, "ekf_associate_static: N must be >= 1 (an empty map has nothing to scan: newLL = 1, index = 1)"); return EKF_EINVAL; }
    if (N > EKF_ASSOC_STATIC_MAX_LANDMARKS) {
        put_err(errbuf, errlen, "ekf_associate_static: N = " + std::to_string(N) + " landmarks exceeds the static scratch (EKF_ASSOC_STATIC_MAX_LANDMARKS = " +
                                    std::to_string(EKF_ASSOC_STATIC_MAX_LANDMARKS) + "); use ekf_associate on a handle");
        return EKF_EINVAL;
    }
    if (assoc_mode == EKF_ASSOC_POSITION && (!d_P || ld < 3 + 2 * (int64_t)N)) { put_err(errbuf, errlen, "ekf_associate_static: mode B needs P with ld >= 3+2N"); return EKF_EINVAL; }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDev) { put_err(errbuf, errlen, std::string("ekf_associate_static: cudaGetDevice: ") + cudaGetErrorString(e)); return EKF_ECUDA; }
    {   // the caller's buffers must be device memory of the CURRENT device (a host pointer would fault or, where the GPU can
        // reach host memory, be read silently); cudaPointerGetAttributes allocates nothing
        const void* ptrs[4] = {d_x, d_s, d_result, assoc_mode == EKF_ASSOC_POSITION ? d_P : nullptr};
        const char* names[4] = {"d_x", "d_s", "d_result", "d_P"};
        for (int i = 0; i < 4; ++i) {
            if (!ptrs[i]) continue;
            cudaPointerAttributes at;
            e = cudaPointerGetAttributes(&at, ptrs[i]);
            if (e != cudaSuccess || (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) ||
                (at.type == cudaMemoryTypeDevice && at.device != dev)) {
                cudaGetLastError();
                put_err(errbuf, errlen, std::string("ekf_associate_static: ") + names[i] + " is not device memory of the current device " + std::to_string(dev));
                return EKF_EINVAL;
            }
        }
    }
    void* base = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ok[dev]) {                               // set-up, once per device, that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_assoc_static_scratch);
            if (e != cudaSuccess || !p) { put_err(errbuf, errlen, std::string("ekf_associate_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); return EKF_ECUDA; }
            g_static_ptr[dev] = p;
            g_static_ok[dev] = true;
        }
        base = g_static_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)(uintptr_t)stream;
    unsigned int* ticket = (unsigned int*)((unsigned char*)base + EKF_STATIC_ASSOC_TICKET_OFF);
    e = cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st);
    if (e != cudaSuccess) { put_err(errbuf, errlen, std::string("ekf_associate_static: cudaMemsetAsync: ") + cudaGetErrorString(e)); return EKF_ECUDA; }
    ekf::launch_associate(st, d_P, ld, d_x, d_s, N, z[0], z[1], z[2], Rdiag[0], Rdiag[1], s_cost, s_thresh, assoc_mode,
                          storage == EKF_STORAGE_SYM ? 1 : 0, base, ticket, (ekf::AssocResult*)d_result);
    e = cudaGetLastError();
    if (e != cudaSuccess) { put_err(errbuf, errlen, std::string("ekf_associate_static: launch: ") + cudaGetErrorString(e)); return EKF_ECUDA; }
    return EKF_OK;
}
