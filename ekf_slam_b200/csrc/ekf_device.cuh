// ekf_device.cuh -- device-side helpers shared by the EKF-SLAM kernels (sm_100a).
//
// Layout in HBM (DESIGN.md "Data layout"):
//   P   : column-major fp64, leading dimension ld (multiple of 16 doubles = 128 B, so every
//         column starts on a cache-line boundary); element (i,j) at P[i + j*ld]; allocated once
//         at capacity (the reference reallocates on every append, EKF_SLAM.m:91).
//   x   : ld doubles (n active), s : capacity doubles (N active).
//   Kd  : ld x double2  = rows of the Kalman gain K (n x 2), zero beyond n.
//   Bd  : ld x double2  = columns of B = H_k * P (2 x n).
//   EKF_STORAGE_SYM: only elements with i <= j are live; (i > j) is don't-care.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "ekf_math_shared.h"

namespace ekf {

// Host helper for the dynamic-shared-memory opt-in (cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device
// attribute): true when `smem` exceeds what was requested so far on the CURRENT device, and records it.
inline bool smem_grows(size_t (&seen)[64], size_t smem) {
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    if (smem <= seen[dev]) return false;
    seen[dev] = smem;
    return true;
}

struct Lin {            // written by k_linearize, consumed by k_gain (one observation)
    double Hs[2][5];    // the 5 non-zero columns of H_k (EKF_SLAM.m:137-138)
    double nu[2];       // innovation z - z_k (EKF_SLAM.m:144), NOT angle wrapped
    double Sinv[2][2];  // phi_k^-1 (EKF_SLAM.m:143)
    double S[2][2];     // phi_k (EKF_SLAM.m:141)
    int    c[5];        // 0-based state indices [0,1,2,a,a+1], a = 3+2(idx-1)
    int    nonfinite;   // 1 when phi_k or its inverse is not finite
    int*   nf;          // host-mapped sticky flag of the handle (set by report_nonfinite when nonfinite; read at sync points)
};

struct AssocResult {    // written by k_associate's last block
    int    newLL;       // Correspondence.m:33/82
    int    index1;      // 1-based, Correspondence.m:40/84
    double cost;        // min_log_likelihood (Inf when newLL)
};

// ---- MATLAB degree helpers (restated semantics; toolbox source is not in the reference) ----
__device__ __forceinline__ void reduce90(double x, int& m, double& r) {
    x = fmod(x, 360.0);
    double n = rint(x / 90.0);
    r = x - 90.0 * n;
    m = ((int)n) & 3;
}
__device__ __forceinline__ double sind(double x) {      // EKF_SLAM.m:42,59,63,84,87-88
    int m; double r; reduce90(x, m, r);
    const double d2r = 0.017453292519943295;
    double a = r * d2r;
    switch (m) { case 0: return sin(a); case 1: return cos(a); case 2: return -sin(a); default: return -cos(a); }
}
__device__ __forceinline__ double cosd(double x) {      // EKF_SLAM.m:42,58,64,85,87-88
    int m; double r; reduce90(x, m, r);
    const double d2r = 0.017453292519943295;
    double a = r * d2r;
    switch (m) { case 0: return cos(a); case 1: return -sin(a); case 2: return -cos(a); default: return sin(a); }
}
__device__ __forceinline__ double atan2d(double y, double x) {   // EKF_SLAM.m:130
    return __dmul_rn(ekm_atan2(y, x), 57.29577951308232);   // shared restatement: same bits as the structured C oracle
}
__device__ __forceinline__ double wrapTo360(double a) {          // EKF_SLAM.m:50,130; Correspondence.m:56
    double r = fmod(a, 360.0);
    if (r != 0.0 && r < 0.0) r = __dadd_rn(r, 360.0);
    if (r == 0.0 && a > 0.0) r = 360.0;
    return r;
}

// Element read honouring the storage mode (SYM keeps only i <= j live).
__device__ __forceinline__ double ldP(const double* __restrict__ P, int64_t ld, int i, int j, int sym) {
    if (sym && i > j) { int t = i; i = j; j = t; }
    return P[(int64_t)i + (int64_t)j * ld];
}

// One rank-2 term of the covariance downdate, v - K(i,:) * B(:,j), as two FMAs in a fixed order.  Every
// consumer (flush kernel, deferred gathers) goes through this so that the deferred path reproduces the
// immediate path bit for bit.
__device__ __forceinline__ double rank2_sub(double v, double2 k, double2 b) {
    return __fma_rn(-k.y, b.y, __fma_rn(-k.x, b.x, v));
}

// Deferred downdate: HBM holds P0 and m pending rank-2 terms (Kall/Ball slabs, slab stride lds);
// the current covariance element is P0(i,j) - sum_q K_q(i,:) B_q(:,j), subtracted in observation order.
__device__ __forceinline__ double ldP_def(const double* __restrict__ P, int64_t ld, int i, int j, int sym,
                                          const double2* __restrict__ Kall, const double2* __restrict__ Ball,
                                          int64_t lds, int m) {
    if (sym && i > j) { int t = i; i = j; j = t; }
    double v = P[(int64_t)i + (int64_t)j * ld];
    for (int q = 0; q < m; ++q) v = rank2_sub(v, Kall[q * lds + i], Ball[q * lds + j]);
    return v;
}

// Range-bearing measurement model about landmark slot a (0-based state index of its x coordinate).
// Follows EKF_SLAM.m:125-138 == Correspondence.m:50-63 with explicit (non-contracted) roundings so a
// plain C restatement compiled with -ffp-contract=off reproduces every bit (atan2 included: ekf_math_shared.h).
// It is written as entry-wise pieces (mm_* / ic_*) so that a warp can evaluate the entries on different lanes
// (k_batch_lin) and still produce exactly the bits of the one-thread form below.
struct MMPrep { double d0, d1, q, sq, iq; };
__device__ __forceinline__ MMPrep mm_prep(const double xr[3], double mx, double my) {
    MMPrep p;
    p.d0 = __dsub_rn(mx, xr[0]);
    p.d1 = __dsub_rn(my, xr[1]);
    p.q = __dadd_rn(__dmul_rn(p.d0, p.d0), __dmul_rn(p.d1, p.d1));
    p.sq = sqrt(p.q);
    p.iq = 1.0 / p.q;
    return p;
}
__device__ __forceinline__ double mm_bearing(const MMPrep& p, double theta) {          // z_k(2), EKF_SLAM.m:130
    return wrapTo360(__dsub_rn(atan2d(p.d1, p.d0), theta));
}
__device__ __forceinline__ double mm_H(const MMPrep& p, int e) {                        // entry e = 5 r + c of H_k's 2 x 5 block, :137-138
    switch (e) {
        case 0: return __dmul_rn(p.iq, __dmul_rn(-p.sq, p.d0));
        case 1: return __dmul_rn(p.iq, __dmul_rn(-p.sq, p.d1));
        case 2: return __dmul_rn(p.iq, 0.0);
        case 3: return __dmul_rn(p.iq, __dmul_rn(p.sq, p.d0));
        case 4: return __dmul_rn(p.iq, __dmul_rn(p.sq, p.d1));
        case 5: return __dmul_rn(p.iq, p.d1);
        case 6: return __dmul_rn(p.iq, -p.d0);
        case 7: return __dmul_rn(p.iq, -p.q);
        case 8: return __dmul_rn(p.iq, -p.d1);
        default: return __dmul_rn(p.iq, p.d0);
    }
}
__device__ __forceinline__ void measurement_model(const double xr[3], double mx, double my,
                                                  double zk[2], double Hs[2][5]) {
    const MMPrep p = mm_prep(xr, mx, my);
    zk[0] = p.sq;
    zk[1] = mm_bearing(p, xr[2]);
#pragma unroll
    for (int e = 0; e < 10; ++e) Hs[e / 5][e % 5] = mm_H(p, e);
}

// phi = (Hs * Pcc) * Hs' + diag(R)   (EKF_SLAM.m:141, left-to-right like MATLAB), 2x2 inverse.
template <typename HT, typename PT>
__device__ __forceinline__ double ic_HP(const HT& Hs, const PT& Pcc, int r, int b) {    // (Hs * Pcc)(r, b)
    double acc = __dmul_rn(Hs[r][0], Pcc[0][b]);
#pragma unroll
    for (int a = 1; a < 5; ++a) acc = __dadd_rn(acc, __dmul_rn(Hs[r][a], Pcc[a][b]));
    return acc;
}
template <typename HPT, typename HT>
__device__ __forceinline__ double ic_S(const HPT& HP, const HT& Hs, const double Rd[2], int r, int t) {
    double acc = __dmul_rn(HP[r][0], Hs[t][0]);
#pragma unroll
    for (int b = 1; b < 5; ++b) acc = __dadd_rn(acc, __dmul_rn(HP[r][b], Hs[t][b]));
    return (r == t) ? __dadd_rn(acc, Rd[r]) : acc;
}
template <typename ST>
__device__ __forceinline__ void ic_inv(const ST& S, double Si[2][2]) {
    double det = __dsub_rn(__dmul_rn(S[0][0], S[1][1]), __dmul_rn(S[0][1], S[1][0]));
    double id = 1.0 / det;
    Si[0][0] = __dmul_rn(S[1][1], id);
    Si[0][1] = __dmul_rn(-S[0][1], id);
    Si[1][0] = __dmul_rn(-S[1][0], id);
    Si[1][1] = __dmul_rn(S[0][0], id);
}
__device__ __forceinline__ void innovation_cov(const double Hs[2][5], const double Pcc[5][5],
                                               const double Rd[2], double S[2][2], double Si[2][2]) {
    double HP[2][5];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int b = 0; b < 5; ++b) HP[r][b] = ic_HP(Hs, Pcc, r, b);
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int t = 0; t < 2; ++t) S[r][t] = ic_S(HP, Hs, Rd, r, t);
    ic_inv(S, Si);
}

// Three-term dot product with one rounding per operation, left to right (MATLAB's scalar order; never contracted): the
// append arithmetic (EKF_SLAM.m:91-97) of the single-GPU, slab and Monte-Carlo kernels goes through it, so that the
// three stay bit-identical whatever each translation unit's compiler would have fused.
__device__ __forceinline__ double dot3(double a0, double b0, double a1, double b1, double a2, double b2) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a0, b0), __dmul_rn(a1, b1)), __dmul_rn(a2, b2));
}

// Gain arithmetic shared by every path (per-observation k_gain, fused k_batch_*, sharded k_gain_shard): explicit FMA
// forms in a fixed order, so that the three paths stay bit-identical to each other whatever the compiler would contract.
//   ph  += P(i, c_k) * Hs(:,k)                 row i of P H'  (EKF_SLAM.m:143, left factor)
//   b   += Hs(:,k) * P(c_k, i)                 column i of H P (EKF_SLAM.m:145, right factor; FULL storage)
//   K(i,:) = ph * phi^-1 ; dx = K(i,:) * nu    (EKF_SLAM.m:143-144)
__device__ __forceinline__ void ph_acc(double& ph0, double& ph1, double pc, double h0k, double h1k) {
    ph0 = __fma_rn(pc, h0k, ph0);
    ph1 = __fma_rn(pc, h1k, ph1);
}
__device__ __forceinline__ void b_acc(double& b0, double& b1, double pr, double h0k, double h1k) {
    b0 = __fma_rn(h0k, pr, b0);
    b1 = __fma_rn(h1k, pr, b1);
}
__device__ __forceinline__ void gain_from_ph(double ph0, double ph1, const double Sinv[2][2], const double nu[2],
                                             double& k0, double& k1, double& dx) {
    k0 = __fma_rn(ph1, Sinv[1][0], __dmul_rn(ph0, Sinv[0][0]));
    k1 = __fma_rn(ph1, Sinv[1][1], __dmul_rn(ph0, Sinv[0][1]));
    dx = __fma_rn(k1, nu[1], __dmul_rn(k0, nu[0]));
}

// Fused batch (ekf_update_batch on one GPU): k <= EKF_MAX_DEFER observations with known slots are linearised
// back to back by ONE warp (k_batch_lin) from the (3+2k)^2 sub-block of P, then ONE grid-wide kernel (k_batch_gain)
// produces all k gain slabs -- 2 launches per batch instead of 2 per observation.
#define EKF_BATCH_MAX 8
#define EKF_BATCH_SLOTS (3 + 2 * EKF_BATCH_MAX)
struct BatchObs { double z0[EKF_BATCH_MAX], z1[EKF_BATCH_MAX], R0[EKF_BATCH_MAX], R1[EKF_BATCH_MAX]; int a[EKF_BATCH_MAX]; int k; };
struct BatchLin {                       // written by k_batch_lin, read by k_batch_gain
    double  Hs[EKF_BATCH_MAX][2][5];
    double  Sinv[EKF_BATCH_MAX][2][2];
    double  nu[EKF_BATCH_MAX][2];
    double2 Kslot[EKF_BATCH_MAX][EKF_BATCH_SLOTS];   // K_q(col(slot), :)
    double2 Bslot[EKF_BATCH_MAX][EKF_BATCH_SLOTS];   // B_q(:, col(slot))
    int     nonfinite;
    int*    nf;                         // host-mapped sticky flag of the handle, like Lin::nf
};

// phi_k or its inverse is not finite (EKF_SLAM.m:137 divides by q = 0 when the robot stands on the landmark): raise the
// handle's sticky flag in mapped host memory; the next synchronising call returns EKF_ENONFINITE.  Costs nothing unless hit.
__device__ __forceinline__ void report_nonfinite(int* nf) {
    if (nf) { *(volatile int*)nf = 1; __threadfence_system(); }
}

// ---- streaming vector access for the downdate (LDG.E.128/256 with L1 no-allocate) ----
struct __align__(32) double4a { double x, y, z, w; };

__device__ __forceinline__ double2 ld_stream2(const double* p) {
    double2 v;
    asm volatile("ld.global.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream2(double* p, double2 v) {
    asm volatile("st.global.L1::no_allocate.v2.f64 [%0], {%1,%2};" :: "l"(p), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ double4a ld_stream4(const double* p) {
    double4a v;
    asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream4(double* p, double4a v) {
    asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}

}  // namespace ekf
