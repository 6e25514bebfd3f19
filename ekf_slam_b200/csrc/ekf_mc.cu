// ekf_mc.cu -- Monte-Carlo batch of independent EKF-SLAM filters (BASELINE config 5: 4096 x EKF_SLAM_UC x 200
// landmarks).  One CTA owns one filter and runs a whole SLAM iteration for it on the device:
//     predict(u)                                        EKF_SLAM.m:40-51
//     measure(): for every observation                  EKF_SLAM.m:100-151 / EKF_SLAM_UC.m:102-152
//         R = diag(range*Rc1, bearing*Rc2)              :108
//         first-landmark branch / z(3) > N append / idx = ii     (known correspondence, :110-123)
//         estimateCorrespondence -> append | update     (unknown correspondence, _UC:119-125; Correspondence.m:28-88)
//         update: linearise, S, K, x                    :124-144
//     P <- P - sum_q K_q B_q                            :145, deferred: ONE pass over this filter's P per measure()
// No host round trip happens inside a step: association results, append decisions and the per-filter landmark count
// N live on the device.  Filters never communicate; across GPUs filter f simply lives on rank f mod G.
// The pending gain slabs K_q, B_q (<= maxd observations) sit in shared memory, so the deferred gathers
// P0(i,c) - sum_q K_q(i) B_q(c) cost no extra HBM traffic and the flush streams P exactly once.
#include "ekf_kernels.h"
#include "ekf_device.cuh"
#include <math_constants.h>
#include <limits.h>
#include <cstdlib>

namespace ekf {


struct McShared {                                // small per-CTA scratch (the slabs follow in dynamic shared memory)
    double Pcc[5][5];
    double Hs[2][5], Sinv[2][2], nu[2];
    double HP[2][5], S[2][2];
    double2 Kc[EKF_MAX_DEFER][5], Bc[EKF_MAX_DEFER][5];   // pending gains at the 5 slot rows / columns of the current observation
    double wc[32];
    int wi[32];
    double best; int bidx;
    int N, m, action, fatal;
    int rr_ok;                                   // the shared-memory copy of the robot rows matches P0
    double pos[2];
};

__device__ __forceinline__ double mc_cur(const double* __restrict__ P, int64_t ldp, int i, int j, int sym,
                                         const double2* Ks, const double2* Bs, int64_t lds, int m) {
    if (sym && i > j) { int t = i; i = j; j = t; }
    double v = P[(int64_t)i + (int64_t)j * ldp];
    for (int q = 0; q < m; ++q) v = rank2_sub(v, Ks[q * lds + i], Bs[q * lds + j]);
    return v;
}

// P(i,j) -= sum_q K_q(i,:) B_q(:,j) for the whole filter; row pairs as double2, UNR column-steps in flight per thread.
__device__ __noinline__ void mc_flush(double* __restrict__ P, int64_t ldp, int n, int sym, const double2* Ks,
                                         const double2* Bs, int64_t lds, int m) {
    if (m == 0) return;
    const int nrp = (n + 1) >> 1;                              // row pairs (row n of an odd n is padding: K = 0 there)
    const int total = nrp * n;
    constexpr int UNR = 8;
    for (int e0 = threadIdx.x; e0 < total; e0 += UNR * blockDim.x) {
        double2 v[UNR];
        int jj[UNR], rr[UNR];
        bool ok[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const int e = e0 + u * blockDim.x;
            jj[u] = e / nrp;
            rr[u] = 2 * (e - jj[u] * nrp);
            ok[u] = e < total && !(sym && rr[u] > jj[u]);
            if (ok[u]) v[u] = *(const double2*)(P + (int64_t)rr[u] + (int64_t)jj[u] * ldp);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (ok[u]) {
                for (int q = 0; q < m; ++q) {
                    const double2 b = Bs[q * lds + jj[u]];
                    v[u].x = rank2_sub(v[u].x, Ks[q * lds + rr[u]], b);
                    v[u].y = rank2_sub(v[u].y, Ks[q * lds + rr[u] + 1], b);
                }
                *(double2*)(P + (int64_t)rr[u] + (int64_t)jj[u] * ldp) = v[u];
            }
        }
    }
}

// Block-wide lexicographic (cost, index) minimum == the strict-'<' first-minimum scan of Correspondence.m:81.
__device__ __forceinline__ void mc_argmin(double best, int bidx, McShared& sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double oc = __shfl_down_sync(0xffffffffu, best, o);
        int oi = __shfl_down_sync(0xffffffffu, bidx, o);
        if ((oc < best) || (oc == best && oi < bidx)) { best = oc; bidx = oi; }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sh.wc[w] = best; sh.wi[w] = bidx; }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = (lane < nw) ? sh.wc[lane] : CUDART_INF;
        bidx = (lane < nw) ? sh.wi[lane] : INT_MAX;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double oc = __shfl_down_sync(0xffffffffu, best, o);
            int oi = __shfl_down_sync(0xffffffffu, bidx, o);
            if ((oc < best) || (oc == best && oi < bidx)) { best = oc; bidx = oi; }
        }
        if (lane == 0) { sh.best = best; sh.bidx = bidx; }
    }
    __syncthreads();
}

// MODE_B = 1 compiles the Mahalanobis association (Correspondence.m:74) in.
// SPLIT = 0: everything in one kernel, gain slabs in shared memory, the CTA also streams its P (256 threads).
// SPLIT = 1: "front" half only -- the latency-bound sequential part (associate / linearise / gain per observation) with
//            the slabs in global memory (L2 resident), 128 threads and no big shared-memory footprint so that many
//            filters are resident per SM and hide each other's fp64 trig / dependent-chain latency; the pending count
//            goes to Mpend[f] and the bandwidth-bound pass over every P is left to k_mc_flush at full HBM rate.
template <int MODE_B, int SPLIT>
#ifndef MC_FRONT_MINB
#define MC_FRONT_MINB 4
#endif
__global__ void __launch_bounds__(SPLIT ? 128 : 256, SPLIT ? (MODE_B ? 2 : MC_FRONT_MINB) : (MODE_B ? 1 : 2))
k_mc_step(McArgs a) {
    extern __shared__ __align__(16) unsigned char mc_dyn[];
    __shared__ McShared sh;
    const int f = blockIdx.x, tid = threadIdx.x, T = blockDim.x;
    double* __restrict__ P = a.P + (int64_t)f * a.pstride;
    double* __restrict__ x = a.x + (int64_t)f * a.xstride;
    double* __restrict__ s = a.s + (int64_t)f * a.sstride;
    const int64_t ld = a.ldp, lds = a.ldp;
    const int sym = a.sym;
    double2* Ks = SPLIT ? a.Kg + (int64_t)f * a.maxd * lds : (double2*)mc_dyn;
    double2* Bs = SPLIT ? a.Bg + (int64_t)f * a.maxd * lds : Ks + (int64_t)a.maxd * lds;
    if (!SPLIT)
        for (int64_t e = tid; e < 2 * (int64_t)a.maxd * lds; e += T) Ks[e] = make_double2(0.0, 0.0);
    if (tid == 0) { sh.N = a.N[f]; sh.m = 0; sh.fatal = 0; sh.rr_ok = 0; }
    __syncthreads();
    const double u1 = a.u[2 * f], u2 = a.u[2 * f + 1];
    // Robot-row cache (split front): every update needs P(0:2, i) for all i -- in column-major storage one 32-byte sector
    // per column at stride ld, the same 400 scattered sectors for each of the step's observations.  P0 does not change
    // during a split step (the terms stay pending), so the three rows are read ONCE per step -- predict has them in its
    // hands anyway -- and served from shared memory afterwards: rr[k * ld + i] = stored P0(k, i).  Same values, same
    // arithmetic: bit-identical.  An append or an in-kernel flush rewrites P0 and invalidates the copy.
    double* rr = (SPLIT && a.robot_cache) ? (double*)mc_dyn : nullptr;
    auto rr_fill = [&](int n_) {
        for (int t = tid; t < n_; t += T) {
            const double* col = P + (int64_t)t * ld;
            rr[t] = col[0]; rr[ld + t] = col[1]; rr[2 * ld + t] = col[2];
        }
        __syncthreads();
        if (tid == 0) sh.rr_ok = 1;
        __syncthreads();
    };

    // ------------------------------------------------------------ predict (EKF_SLAM.m:40-51), nothing pending
    if (a.do_predict) {
        const int n = 3 + 2 * sh.N;
        const double th = x[2];
        const double st = sind(th), ct = cosd(th);
        const double fa = -1.0 * u1 * st, fb = u1 * ct;          // F(1,3), F(2,3): heading BEFORE the motion (:63-64)
        __syncthreads();
        if (tid == 0) {
            double M[3][3];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) M[i][j] = ldP(P, ld, i, j, sym);
            for (int j = 0; j < 3; ++j) { M[0][j] = M[0][j] + fa * M[2][j]; M[1][j] = M[1][j] + fb * M[2][j]; }
            for (int i = 0; i < 3; ++i) { M[i][0] = M[i][0] + M[i][2] * fa; M[i][1] = M[i][1] + M[i][2] * fb; }
            const double W[3] = {u1 * ct, u1 * st, u2};
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) P[i + j * ld] = M[i][j] + (W[i] * a.C) * W[j];
            x[0] = x[0] + u1 * cosd(th + u2);                    // :57-60
            x[1] = x[1] + u1 * sind(th + u2);
            x[2] = wrapTo360(th + u2);                           // :50
        }
        for (int t = 3 + tid; t < n; t += T) {
            double* col = P + (int64_t)t * ld;
            const double p2 = col[2];
            const double c0 = col[0] + fa * p2, c1 = col[1] + fb * p2;
            col[0] = c0;
            col[1] = c1;
            if (rr) { rr[t] = c0; rr[ld + t] = c1; rr[2 * ld + t] = p2; }
            if (!sym) {
                const double q2 = P[t + 2 * ld];
                P[t] = P[t] + q2 * fa;
                P[t + ld] = P[t + ld] + q2 * fb;
            }
        }
        __syncthreads();
        if (rr) {                                                // columns 0..2 (the corner thread 0 just wrote)
            if (tid < 3) { const double* col = P + (int64_t)tid * ld; rr[tid] = col[0]; rr[ld + tid] = col[1]; rr[2 * ld + tid] = col[2]; }
            if (tid == 0) sh.rr_ok = 1;
            __syncthreads();
        }
    }

    // ------------------------------------------------------------ measure (EKF_SLAM.m:100-151, EKF_SLAM_UC.m:102-152)
    const int cnt = a.do_measure ? a.counts[f] : 0;
    const double* ob = a.obs + (int64_t)f * 3 * a.max_obs;
    const double* loc = a.loc + (int64_t)f * a.loc_stride;
    const int* tix = a.idx + (int64_t)f * a.idx_stride;
    const int ntab = a.ntab[f * a.ntab_stride];
    for (int ii = 1; ii <= cnt; ++ii) {                          // :107
        const double z0 = ob[ii - 1], z1 = ob[ii - 1 + a.max_obs], z2 = ob[ii - 1 + 2 * a.max_obs];
        const double R0 = z0 * a.Rc0, R1 = z1 * a.Rc1;           // :108
        const int N = sh.N, n = 3 + 2 * N, m = sh.m;
        // ---- decide: append (with pos, signature) or update (slot idx)
        int do_append = 0, idx = 0;
        double want = 0.0, sig = 0.0;
        int lookup = 0;                                          // 1: first non-zero index, 2: index == want
        if (N == 0) { do_append = 1; sig = 1.0; lookup = 1; }    // :110-111 (F15)
        else if (a.variant == 0 /* EKF_VARIANT_KNOWN */) {
            if (z2 > (double)N) { do_append = 1; sig = z2; want = z2; lookup = 2; }      // :118-120
            else idx = ii;                                       // :123 (F8)
        } else {
            double best = CUDART_INF;
            int bidx = INT_MAX;
            for (int kk = tid; kk < N; kk += T) {                // Correspondence.m:49
                const double d = __dsub_rn(z2, s[kk]);
                double ll = __dmul_rn(__dmul_rn(d, 1.0 / a.s_cost), d);      // :71, :75
                if (MODE_B) {                                    // :74 (commented out upstream): + Mahalanobis cost
                    const int al = 3 + 2 * kk;
                    const int c[5] = {0, 1, 2, al, al + 1};
                    double xr[3] = {x[0], x[1], x[2]};
                    double zk[2], Hs[2][5], S[2][2], Si[2][2], Pcc[5][5];
                    measurement_model(xr, x[al], x[al + 1], zk, Hs);
                    for (int i = 0; i < 5; ++i)
                        for (int j = 0; j < 5; ++j) Pcc[i][j] = mc_cur(P, ld, c[i], c[j], sym, Ks, Bs, lds, m);
                    const double Rd[2] = {R0, R1};
                    innovation_cov(Hs, Pcc, Rd, S, Si);
                    const double n0 = __dsub_rn(z0, zk[0]), n1 = __dsub_rn(z1, zk[1]);
                    const double t0 = __dadd_rn(__dmul_rn(n0, Si[0][0]), __dmul_rn(n1, Si[1][0]));   // :69
                    const double t1 = __dadd_rn(__dmul_rn(n0, Si[0][1]), __dmul_rn(n1, Si[1][1]));
                    ll = __dadd_rn(__dadd_rn(__dmul_rn(t0, n0), __dmul_rn(t1, n1)), ll);
                }
                if (ll <= a.s_thresh && ll < best) { best = ll; bidx = kk + 1; }         // :78, :81
            }
            mc_argmin(best, bidx, sh);
            if (sh.bidx == INT_MAX) { do_append = 1; idx = N + 1; sig = (double)idx; want = (double)idx; lookup = 2; }   // :40, :121-123
            else idx = sh.bidx;
        }
        if (tid == 0) {
            int fatal = 0;
            if (do_append) {
                int hit = -1;
                for (int i = 0; i < ntab && hit < 0; ++i)
                    if (lookup == 1 ? (tix[i] != 0) : ((double)tix[i] == want)) hit = i;
                if (hit < 0) fatal = MC_ERR_TABLE;
                else { sh.pos[0] = loc[hit]; sh.pos[1] = loc[hit + a.max_table]; }
                if (N >= a.cap) fatal = MC_ERR_CAPACITY;
                sh.action = -(int)sig;
            } else {
                if (idx < 1 || idx > N) fatal = MC_ERR_INDEX;
                sh.action = idx;
            }
            sh.fatal = fatal;
        }
        __syncthreads();
        if (sh.fatal) break;
        if (a.actions && tid == 0) a.actions[(int64_t)f * a.max_obs + ii - 1] = sh.action;

        if (do_append) {
            // ---------------------------------------------------- append (EKF_SLAM.m:67-98), needs the whole P
            mc_flush(P, ld, n, sym, Ks, Bs, lds, m);
            __syncthreads();
            const double th = x[2];
            const double j02 = -u1 * sind(th), j12 = u1 * cosd(th);          // jxr, :84-85
            for (int t = tid; t <= N; t += T) {
                if (t < N) {
                    const int al = 3 + 2 * t;
                    for (int mm = 0; mm < 2; ++mm) {
                        const double p0 = ldP(P, ld, al + mm, 0, sym), p1 = ldP(P, ld, al + mm, 1, sym), p2 = ldP(P, ld, al + mm, 2, sym);
                        const double f0 = dot3(1.0, p0, 0.0, p1, j02, p2);    // jxr * P(lm,1:3)', :95
                        const double f1 = dot3(0.0, p0, 1.0, p1, j12, p2);
                        P[(int64_t)(al + mm) + (int64_t)n * ld] = f0;
                        P[(int64_t)(al + mm) + (int64_t)(n + 1) * ld] = f1;
                        P[(int64_t)n + (int64_t)(al + mm) * ld] = f0;
                        P[(int64_t)(n + 1) + (int64_t)(al + mm) * ld] = f1;
                    }
                } else {
                    double Prr[3][3];
                    for (int i = 0; i < 3; ++i)
                        for (int j = 0; j < 3; ++j) Prr[i][j] = ldP(P, ld, i, j, sym);
                    const double jxr[2][3] = {{1.0, 0.0, j02}, {0.0, 1.0, j12}};
                    const double cu = cosd(u2), su = sind(u2);
                    const double jz[2][2] = {{cu, -u1 * su}, {su, u1 * cu}};             // :87-88
                    double JP[2][3];
                    for (int r = 0; r < 2; ++r)
                        for (int j = 0; j < 3; ++j) JP[r][j] = dot3(jxr[r][0], Prr[0][j], jxr[r][1], Prr[1][j], jxr[r][2], Prr[2][j]);
                    const double JR[2][2] = {{__dmul_rn(jz[0][0], R0), __dmul_rn(jz[0][1], R1)}, {__dmul_rn(jz[1][0], R0), __dmul_rn(jz[1][1], R1)}};
                    for (int r = 0; r < 2; ++r)
                        for (int c = 0; c < 2; ++c)
                            P[(int64_t)(n + r) + (int64_t)(n + c) * ld] =
                                __dadd_rn(dot3(JP[r][0], jxr[c][0], JP[r][1], jxr[c][1], JP[r][2], jxr[c][2]),
                                __dadd_rn(__dmul_rn(JR[r][0], jz[c][0]), __dmul_rn(JR[r][1], jz[c][1])));             // :91
                    for (int i = 0; i < 3; ++i)
                        for (int c = 0; c < 2; ++c) {
                            const double v = dot3(Prr[i][0], jxr[c][0], Prr[i][1], jxr[c][1], Prr[i][2], jxr[c][2]);   // :92
                            P[(int64_t)i + (int64_t)(n + c) * ld] = v;
                            P[(int64_t)(n + c) + (int64_t)i * ld] = v;                   // :93
                        }
                    x[n] = sh.pos[0]; x[n + 1] = sh.pos[1];                              // :79
                    s[N] = sig;                                                          // :70
                }
            }
            __syncthreads();
            if (tid == 0) { sh.N = N + 1; sh.m = 0; sh.rr_ok = 0; }
            __syncthreads();
            continue;
        }

        // -------------------------------------------------------- update (EKF_SLAM.m:124-145)
        const int al = 3 + 2 * (idx - 1);
        const int c[5] = {0, 1, 2, al, al + 1};
        if (rr && !sh.rr_ok) rr_fill(n);                 // block-uniform: no predict in this call, or P0 was rewritten
        // pending gains of the 5 slot rows / columns, gathered once per observation: every later correction then
        // costs shared-memory reads instead of a chain of dependent L2 round trips (same terms, same order as mc_cur)
        if (tid < 10 * m) {
            const int q = tid / 10, r = tid - 10 * q;
            if (r < 5) sh.Kc[q][r] = Ks[q * lds + c[r]];
            else sh.Bc[q][r - 5] = Bs[q * lds + c[r - 5]];
        }
        __syncthreads();
        if (tid < 32) {
            // one warp, one entry per lane (the pieces of ekf_device.cuh give the bits of the one-thread form)
            if (tid < 25) {
                int i = tid / 5, j = tid - 5 * (tid / 5);
                if (sym && i > j) { const int t = i; i = j; j = t; }     // c[] is increasing: stored element is (c[i], c[j]), i <= j
                double v = P[(int64_t)c[i] + (int64_t)c[j] * ld];
                for (int q = 0; q < m; ++q) v = rank2_sub(v, sh.Kc[q][i], sh.Bc[q][j]);
                sh.Pcc[tid / 5][tid - 5 * (tid / 5)] = v;
            }
            const double xr[3] = {x[0], x[1], x[2]};
            const MMPrep mp = mm_prep(xr, x[al], x[al + 1]);             // every lane: cheaper than a broadcast
            if (tid < 10) sh.Hs[tid / 5][tid % 5] = mm_H(mp, tid);
            if (tid == 31) sh.nu[0] = z0 - mp.sq;                        // :144, not wrapped (F10)
            __syncwarp();
            if (tid < 10) sh.HP[tid / 5][tid % 5] = ic_HP(sh.Hs, sh.Pcc, tid / 5, tid % 5);
            __syncwarp();
            const double Rd[2] = {R0, R1};
            if (tid < 4) sh.S[tid >> 1][tid & 1] = ic_S(sh.HP, sh.Hs, Rd, tid >> 1, tid & 1);
            __syncwarp();
            if (tid == 0) {
                double Si[2][2];
                ic_inv(sh.S, Si);
                sh.Sinv[0][0] = Si[0][0]; sh.Sinv[0][1] = Si[0][1]; sh.Sinv[1][0] = Si[1][0]; sh.Sinv[1][1] = Si[1][1];
            }
        } else if (tid == 32) {
            // the bearing (atan2d, the longest single piece) on a second warp, beside the S chain of the first:
            // lanes of one warp would run the two branches one after the other
            const double xr[3] = {x[0], x[1], x[2]};
            const MMPrep mp = mm_prep(xr, x[al], x[al + 1]);
            sh.nu[1] = z1 - mm_bearing(mp, xr[2]);
        }
        __syncthreads();
        for (int i = tid; i < n; i += T) {
            // current P(i, c[k]) (and P(c[k], i) in FULL storage): all loads of the row are issued before any is used
            double pc[5], pr[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const bool cached = rr && k < 3;         // P(k, i): the robot rows, from shared memory
                if (sym) pc[k] = i <= c[k] ? P[(int64_t)i + (int64_t)c[k] * ld] : (cached ? rr[(int64_t)k * ld + i] : P[(int64_t)c[k] + (int64_t)i * ld]);
                else { pc[k] = P[(int64_t)i + (int64_t)c[k] * ld]; pr[k] = cached ? rr[(int64_t)k * ld + i] : P[(int64_t)c[k] + (int64_t)i * ld]; }
            }
            for (int q0 = 0; q0 < m; q0 += 4) {                  // pending terms four at a time: their slab loads overlap too
                double2 kq[4], bq[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (q0 + u < m) { kq[u] = Ks[(q0 + u) * lds + i]; bq[u] = Bs[(q0 + u) * lds + i]; }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (q0 + u < m) {
                        const int q = q0 + u;
#pragma unroll
                        for (int k = 0; k < 5; ++k) {
                            if (sym) pc[k] = i <= c[k] ? rank2_sub(pc[k], kq[u], sh.Bc[q][k]) : rank2_sub(pc[k], sh.Kc[q][k], bq[u]);
                            else { pc[k] = rank2_sub(pc[k], kq[u], sh.Bc[q][k]); pr[k] = rank2_sub(pr[k], sh.Kc[q][k], bq[u]); }
                        }
                    }
                }
            }
            double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                ph_acc(ph0, ph1, pc[k], sh.Hs[0][k], sh.Hs[1][k]);
                if (!sym) b_acc(b0, b1, pr[k], sh.Hs[0][k], sh.Hs[1][k]);
            }
            if (sym) { b0 = ph0; b1 = ph1; }
            double k0, k1, dx;
            gain_from_ph(ph0, ph1, sh.Sinv, sh.nu, k0, k1, dx);                          // :143
            x[i] = __dadd_rn(x[i], dx);                                                  // :144
            Ks[m * lds + i] = make_double2(k0, k1);
            Bs[m * lds + i] = make_double2(b0, b1);
        }
        if (tid == 0) {                                  // n is odd: row n pads the last row pair of the flush
            Ks[m * lds + n] = make_double2(0.0, 0.0);
            Bs[m * lds + n] = make_double2(0.0, 0.0);
        }
        __syncthreads();
        if (m + 1 >= a.maxd && !(SPLIT && ii == cnt)) {        // slabs full (split: the last batch is left to k_mc_flush)
            mc_flush(P, ld, n, sym, Ks, Bs, lds, m + 1);
            __syncthreads();
            if (tid == 0) { sh.m = 0; sh.rr_ok = 0; }
        } else if (tid == 0) sh.m = m + 1;
        __syncthreads();
    }
    if (!SPLIT) {   // P is whole again when measure() returns; on a fatal row the terms applied so far are kept
        const int n = 3 + 2 * sh.N;
        mc_flush(P, ld, n, sym, Ks, Bs, lds, sh.m);
    }
    if (tid == 0) {
        if (SPLIT) a.Mpend[f] = sh.m;                    // k_mc_flush applies them
        a.N[f] = sh.N;
        if (sh.fatal && atomicCAS(a.err, 0, sh.fatal) == 0) a.err[1] = f;
    }
}

// Bandwidth half of the split step: P_f(i,j) -= sum_{q<Mpend[f]} K_q(i,:) B_q(:,j) for every filter.
// One 64-thread CTA per (128-row x TC-column tile, filter): a thread owns a row pair (double2, 16 B) with its <= M gains
// in registers and keeps UNR column loads in flight; the tile's B columns sit in shared memory.  Small CTAs with few
// registers -> ~12 resident per SM -> ~100 KB of reads in flight per SM.  SYM drops tiles below the diagonal whole and
// masks the diagonal tiles element by element.
template <int M, int TC, int UNR>
__global__ void __launch_bounds__(64)
k_mc_flush(McArgs a, int row_tiles) {
    const int f = blockIdx.y, tid = threadIdx.x;
    const int rt = blockIdx.x % row_tiles, ct = blockIdx.x / row_tiles;
    const int r0 = rt * 128, c0 = ct * TC;
    const int n = 3 + 2 * a.N[f], m = min(a.Mpend[f], M);
    const int sym = a.sym;
    if (m == 0 || c0 >= n || r0 >= n) return;
    const int nc = min(TC, n - c0);
    if (sym && r0 > c0 + nc - 1) return;                 // tile strictly below the diagonal
    const int64_t ld = a.ldp, lds = a.ldp;
    double* __restrict__ P = a.P + (int64_t)f * a.pstride;
    const double2* __restrict__ Kg = a.Kg + (int64_t)f * a.maxd * lds;
    const double2* __restrict__ Bg = a.Bg + (int64_t)f * a.maxd * lds;
    __shared__ double2 Bsh[TC * M];
    for (int e = tid; e < nc * M; e += 64) {
        const int j = e / M, q = e - j * M;
        Bsh[e] = q < m ? Bg[q * lds + c0 + j] : make_double2(0.0, 0.0);
    }
    __syncthreads();
    const int r = r0 + 2 * tid;
    if (r >= n || (sym && r > c0 + nc - 1)) return;
    double2 k0[M], k1[M];
#pragma unroll
    for (int q = 0; q < M; ++q) {
        k0[q] = q < m ? Kg[q * lds + r] : make_double2(0.0, 0.0);
        k1[q] = q < m ? Kg[q * lds + r + 1] : make_double2(0.0, 0.0);     // row n (n is odd) is written as zero by the front
    }
    double* p = P + (int64_t)r + (int64_t)c0 * ld;
    const int j0 = sym ? max(0, r - c0) : 0;             // first live column of this row pair within the tile
    for (int j = j0; j < nc; j += UNR) {
        double2 v[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u)
            if (j + u < nc) v[u] = *(const double2*)(p + (int64_t)(j + u) * ld);
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (j + u < nc) {
#pragma unroll
                for (int q = 0; q < M; ++q) {
                    const double2 b = Bsh[(j + u) * M + q];
                    v[u].x = rank2_sub(v[u].x, k0[q], b);
                    v[u].y = rank2_sub(v[u].y, k1[q], b);
                }
                *(double2*)(p + (int64_t)(j + u) * ld) = v[u];
            }
        }
    }
}

// Third flush variant, "warp streams whole columns": a warp owns 4 adjacent columns of one filter and walks down their
// rows (lane = row pair, 64 rows per step), so every warp request is a long contiguous run of a column (the
// microbenchmark tools/tri_stream.cu reaches 6.0 TB/s on exactly this traffic with that mapping, against 4.1 TB/s
// for the tiled kernels above).  The gains cannot sit in registers then (a lane meets every row), so the K slab of
// the filter is staged in shared memory once per CTA ([row][q], 2M LDS.128 per row pair, amortised over the 4
// columns) and B_q of the 4 columns is read as shared-memory broadcasts.
template <int M, int TC>
__global__ void __launch_bounds__(256, 4)
k_mc_flush_w(McArgs a) {
    extern __shared__ __align__(16) unsigned char fw_dyn[];
    const int f = blockIdx.y, c0 = blockIdx.x * TC, tid = threadIdx.x;
    const int n = 3 + 2 * a.N[f], m = min(a.Mpend[f], M);
    if (m == 0 || c0 >= n) return;
    const int sym = a.sym;
    const int64_t ld = a.ldp, lds = a.ldp;
    double* __restrict__ P = a.P + (int64_t)f * a.pstride;
    const double2* __restrict__ Kg = a.Kg + (int64_t)f * a.maxd * lds;
    const double2* __restrict__ Bg = a.Bg + (int64_t)f * a.maxd * lds;
    const int nc = min(TC, n - c0);
    const int nrows = sym ? min(n + 1, ((c0 + nc + 1) & ~1)) : n + 1;     // rows this column range can touch (n is odd: row n pads)
    const int NRH = a.cap + 2;                                             // row pairs per (q, parity) plane
    double2* Ksh = (double2*)fw_dyn;                                       // [q][row parity][row / 2]: a lane's K_q(r), K_q(r+1) are two conflict-free LDS.128
    double2* Bsh = Ksh + (size_t)2 * NRH * M;                              // [TC][M]
    const int warp = tid >> 5, lane = tid & 31;
    // work list of the warp: (column group g, 64-row step rb), flattened and software-pipelined one item ahead so the
    // loads of the next item are in flight while this one is computed and stored
    auto grp_rend = [&](int g) { const int jmax = c0 + min(g + 3, nc - 1); return sym ? min(n, jmax + 1) : n; };
    auto load = [&](double2 (&v)[4], int g, int rb) {
        const int r = rb + 2 * lane;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int col = c0 + g + c;
            if (g + c < nc && (sym ? r <= col : r < n)) v[c] = ld_stream2(P + (int64_t)r + (int64_t)col * ld);
        }
    };
    int g = warp * 4, rb = 0;
    bool have = g < nc;
    double2 vn[4];
    if (have) load(vn, g, rb);                                             // first item's loads overlap the gain staging below
    for (int e = tid; e < nrows * M; e += 256) {
        const int q = e / nrows, i = e - q * nrows;
        Ksh[(q * 2 + (i & 1)) * NRH + (i >> 1)] = q < m ? Kg[q * lds + i] : make_double2(0.0, 0.0);   // row n was written as zero by the front
    }
    for (int e = tid; e < nc * M; e += 256) {
        const int j = e / M, q = e - j * M;
        Bsh[e] = q < m ? Bg[q * lds + c0 + j] : make_double2(0.0, 0.0);
    }
    __syncthreads();
    while (have) {
        double2 v[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] = vn[c];
        const int cg = g, r = rb + 2 * lane;
        rb += 64;
        if (rb >= grp_rend(g)) { g += 32; rb = 0; }
        have = g < nc;
        if (have) load(vn, g, rb);
        if (r <= min(NRH * 2 - 2, n)) {                                    // lanes past the matrix have nothing to do (and no K rows staged)
#pragma unroll
            for (int q = 0; q < M; ++q) {
                const double2 k0 = Ksh[(q * 2) * NRH + (r >> 1)], k1 = Ksh[(q * 2 + 1) * NRH + (r >> 1)];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const double2 b = Bsh[min(cg + c, nc - 1) * M + q];
                    v[c].x = rank2_sub(v[c].x, k0, b);
                    v[c].y = rank2_sub(v[c].y, k1, b);
                }
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int col = c0 + cg + c;
                if (cg + c < nc && (sym ? r <= col : r < n)) st_stream2(P + (int64_t)r + (int64_t)col * ld, v[c]);
            }
        }
    }
}

// Pose read-back: x(1:3) of every filter (stride xstride) packed into out[3 nf]; the device error word rides behind it
// so that one contiguous copy brings both to the host.
__global__ void k_mc_pack_poses(const double* __restrict__ x, int64_t xstride, int nf, double* __restrict__ out,
                                const int* __restrict__ err) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < 3 * nf) {
        const int f = e / 3, c = e - 3 * f;
        out[e] = x[(int64_t)f * xstride + c];
    }
    if (e == 0) {
        int* w = (int*)(out + 3 * (int64_t)nf);
        w[0] = err[0];
        w[1] = err[1];
    }
}

void launch_mc_pack_poses(cudaStream_t st, const double* x, int64_t xstride, int n_filters, double* out, const int* err) {
    k_mc_pack_poses<<<(3 * n_filters + 255) / 256, 256, 0, st>>>(x, xstride, n_filters, out, err);
}

template <int M, int TC>
static void launch_mc_flush_w_m(cudaStream_t st, const McArgs& a, int n_filters, int cap_n) {
    const size_t smem = ((size_t)(a.cap + 2) * 2 * M + (size_t)TC * M) * sizeof(double2);
    static size_t smem_set[64] = {0};                    // per device: the attribute belongs to the device's context
    if (smem_grows(smem_set, smem))
        // no max-shared carve-out here: the LDG stream needs the L1 as its landing buffer (tools/tri_stream.cu: 4.2 -> 5.5 TB/s)
        cudaFuncSetAttribute(k_mc_flush_w<M, TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((cap_n + TC - 1) / TC, n_filters);
    k_mc_flush_w<M, TC><<<grid, 256, smem, st>>>(a);
}

template <int TC>
static void launch_mc_flush_w_tc(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max) {
    switch (m_max) {
        case 1: launch_mc_flush_w_m<1, TC>(st, a, n_filters, cap_n); break;
        case 2: launch_mc_flush_w_m<2, TC>(st, a, n_filters, cap_n); break;
        case 3: launch_mc_flush_w_m<3, TC>(st, a, n_filters, cap_n); break;
        case 4: launch_mc_flush_w_m<4, TC>(st, a, n_filters, cap_n); break;
        case 5: launch_mc_flush_w_m<5, TC>(st, a, n_filters, cap_n); break;
        case 6: launch_mc_flush_w_m<6, TC>(st, a, n_filters, cap_n); break;
        case 7: launch_mc_flush_w_m<7, TC>(st, a, n_filters, cap_n); break;
        default: launch_mc_flush_w_m<8, TC>(st, a, n_filters, cap_n); break;
    }
}

// Columns of one filter per CTA (EKF_MC_TC = 32 | 64 | 128, default 64): wider CTAs stage the filter's K slab fewer
// times, narrower ones balance the triangle better; all give bit-identical results.
void launch_mc_flush_w(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max, int tc) {
    if (tc == 128) launch_mc_flush_w_tc<128>(st, a, n_filters, cap_n, m_max);
    else if (tc == 32) launch_mc_flush_w_tc<32>(st, a, n_filters, cap_n, m_max);
    else launch_mc_flush_w_tc<64>(st, a, n_filters, cap_n, m_max);
}

// ---------------------------------------------------------------------------------------------------------------
// Run-based front half (split step, association mode A).  k_mc_step<.,1> above pays 3-4 dependent memory round trips
// per observation (signature scan, slot block, gain rows); here nothing on the per-observation chain touches memory:
//   * the remaining observations are associated together, one warp each (the mode-A cost depends only on s);
//   * a RUN = the following consecutive updates (up to the free slab depth).  Its 3 + 2 len slot columns of the
//     stored P -- every column any of its observations reads -- are loaded once into shared memory;
//   * per observation: warp 0 linearises from that cache, every thread turns its rows of the cache into K, B, dx,
//     the cache columns later observations need are brought up to date with the same rank2_sub terms, in the same
//     order, as the lazy mc_cur() evaluation (bit-identical results);
//   * an append ends the run (flush of the pending terms by this CTA, EKF_SLAM.m:67-98), then association restarts.
// A run always starts with nothing pending: it ends by an append, by full slabs (inline flush) or with the step.
template <int SYM, int RPT>   // RPT: rows of the state per thread (n <= 256 RPT)
__global__ void __launch_bounds__(256, RPT <= 2 ? (SYM ? 3 : 2) : 1)
k_mc_front(McArgs a) {
    extern __shared__ __align__(16) unsigned char fr_dyn[];
    __shared__ McShared sh;
    __shared__ double2 Kc[EKF_MAX_SLOTS], Bc[EKF_MAX_SLOTS];       // gains of the current observation at the slot rows
    __shared__ int sc[EKF_MAX_SLOTS];                              // slot -> state index (column of P)
    const int f = blockIdx.x, tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
    const int64_t ld = a.ldp, lds = a.ldp;
    const int NS = 3 + 2 * a.maxd;
    double* xs = (double*)fr_dyn;                                  // state vector
    double* pcs = xs + ld;                                         // [NS][ld]: current P(i, sc[k])
    double* prs = pcs + (int64_t)NS * ld;                          // [NS][ld]: current P(sc[k], i), FULL storage only
    int* act = (int*)(SYM ? prs : prs + (int64_t)NS * ld);         // [max_obs]: slot (1-based) to update, 0 = append
    double* __restrict__ P = a.P + (int64_t)f * a.pstride;
    double* __restrict__ x = a.x + (int64_t)f * a.xstride;
    double* __restrict__ s = a.s + (int64_t)f * a.sstride;
    double2* Ks = a.Kg + (int64_t)f * a.maxd * lds;
    double2* Bs = a.Bg + (int64_t)f * a.maxd * lds;
    if (tid == 0) { sh.N = a.N[f]; sh.m = 0; sh.fatal = 0; }
    __syncthreads();
    {
        const int n0 = 3 + 2 * sh.N;
        for (int i = tid; i < n0; i += T) xs[i] = x[i];
    }
    const double u1 = a.u[2 * f], u2 = a.u[2 * f + 1];
    __syncthreads();

    // ------------------------------------------------------------ predict (EKF_SLAM.m:40-51), nothing pending
    if (a.do_predict) {
        const int n = 3 + 2 * sh.N;
        const double th = xs[2];
        const double st = sind(th), ct = cosd(th);
        const double fa = -1.0 * u1 * st, fb = u1 * ct;          // F(1,3), F(2,3): heading BEFORE the motion (:63-64)
        __syncthreads();
        if (tid == 0) {
            double M[3][3];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) M[i][j] = ldP(P, ld, i, j, SYM);
            for (int j = 0; j < 3; ++j) { M[0][j] = M[0][j] + fa * M[2][j]; M[1][j] = M[1][j] + fb * M[2][j]; }
            for (int i = 0; i < 3; ++i) { M[i][0] = M[i][0] + M[i][2] * fa; M[i][1] = M[i][1] + M[i][2] * fb; }
            const double W[3] = {u1 * ct, u1 * st, u2};
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) P[i + j * ld] = M[i][j] + (W[i] * a.C) * W[j];
            xs[0] = xs[0] + u1 * cosd(th + u2);                  // :57-60
            xs[1] = xs[1] + u1 * sind(th + u2);
            xs[2] = wrapTo360(th + u2);                          // :50
        }
        for (int t = 3 + tid; t < n; t += T) {
            double* col = P + (int64_t)t * ld;
            const double p2 = col[2];
            col[0] = col[0] + fa * p2;
            col[1] = col[1] + fb * p2;
            if (!SYM) {
                const double q2 = P[t + 2 * ld];
                P[t] = P[t] + q2 * fa;
                P[t + ld] = P[t + ld] + q2 * fb;
            }
        }
        __syncthreads();
    }
    // ------------------------------------------------------------ measure (EKF_SLAM.m:100-151, EKF_SLAM_UC.m:102-152)
    const int cnt = a.do_measure ? a.counts[f] : 0;
    const double* ob = a.obs + (int64_t)f * 3 * a.max_obs;
    const double* loc = a.loc + (int64_t)f * a.loc_stride;
    const int* tix = a.idx + (int64_t)f * a.idx_stride;
    const int ntab = a.ntab[f * a.ntab_stride];
    int ii = 1, mpend = 0;                                       // both uniform over the CTA
    bool need_assoc = true;
    while (ii <= cnt) {                                          // :107
        const int N = sh.N, n = 3 + 2 * N;
        if (need_assoc) {
            // ---- associate observations ii..cnt against the current map, one warp each
            for (int t = ii + warp; t <= cnt; t += NW) {
                const double z2 = ob[t - 1 + 2 * a.max_obs];
                int r;
                if (N == 0) r = 0;                                                   // :110-111 (F15)
                else if (a.variant == 0 /* EKF_VARIANT_KNOWN */) r = (z2 > (double)N) ? 0 : t;   // :118-123 (F8: idx = ii)
                else {
                    double best = CUDART_INF;
                    int bidx = INT_MAX;
                    for (int kk = lane; kk < N; kk += 32) {                          // Correspondence.m:49
                        const double d = __dsub_rn(z2, s[kk]);
                        const double ll = __dmul_rn(__dmul_rn(d, 1.0 / a.s_cost), d);            // :71, :75
                        if (ll <= a.s_thresh && ll < best) { best = ll; bidx = kk + 1; }         // :78, :81
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {                               // lexicographic (cost, index) minimum
                        const double oc = __shfl_xor_sync(0xffffffffu, best, o);
                        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
                        if ((oc < best) || (oc == best && oi < bidx)) { best = oc; bidx = oi; }
                    }
                    r = bidx == INT_MAX ? 0 : bidx;                                  // :40: no match -> new landmark
                }
                if (lane == 0) act[t - 1] = r;
            }
            need_assoc = false;
            __syncthreads();
        }
        // ---- the run: consecutive updates of valid slots, as many as the slabs hold (nothing is pending here)
        int len = 0;
        while (ii + len <= cnt && len < a.maxd) {
            const int r = act[ii + len - 1];
            if (r < 1 || r > N) break;
            ++len;
        }
        if (len > 0) {
            const int NSr = 3 + 2 * len;
            if (tid < NSr) sc[tid] = tid < 3 ? tid : 3 + 2 * (act[ii - 1 + ((tid - 3) >> 1)] - 1) + ((tid - 3) & 1);
            __syncthreads();
            constexpr int SB = SYM ? 8 : 4;
            for (int k0 = 0; k0 < NSr; k0 += SB) {               // slot columns of the stored P, SB RPT loads in flight per thread
                double v[SB][RPT], w[SYM ? 1 : SB][RPT];
#pragma unroll
                for (int q = 0; q < SB; ++q) {
                    const int c = sc[min(k0 + q, NSr - 1)];
#pragma unroll
                    for (int r = 0; r < RPT; ++r) {
                        const int i = tid + r * T;
                        if (k0 + q < NSr && i < n) {
                            if (SYM) v[q][r] = i <= c ? P[(int64_t)i + (int64_t)c * ld] : P[(int64_t)c + (int64_t)i * ld];
                            else { v[q][r] = P[(int64_t)i + (int64_t)c * ld]; w[SYM ? 0 : q][r] = P[(int64_t)c + (int64_t)i * ld]; }
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < SB; ++q) {
#pragma unroll
                    for (int r = 0; r < RPT; ++r) {
                        const int i = tid + r * T;
                        if (k0 + q < NSr && i < n) {
                            pcs[(k0 + q) * ld + i] = v[q][r];
                            if (!SYM) prs[(k0 + q) * ld + i] = w[SYM ? 0 : q][r];
                        }
                    }
                }
            }
            __syncthreads();
            for (int t = 0; t < len; ++t) {
                const int o = ii + t, m = t;                     // observation ordinal; its slab index
                const double z0 = ob[o - 1], z1 = ob[o - 1 + a.max_obs];
                const double R0 = z0 * a.Rc0, R1 = z1 * a.Rc1;   // :108
                const int ks[5] = {0, 1, 2, 3 + 2 * t, 4 + 2 * t};
                const int al = sc[3 + 2 * t];
                if (tid < 32) {                                  // linearise: one entry per lane, bits of the one-thread form
                    if (tid < 25) {
                        const int pi = tid / 5, pj = tid - 5 * pi;
                        const int si = pi < 3 ? pi : 2 * t + pi, sj = pj < 3 ? pj : 2 * t + pj;   // slots {0, 1, 2, 3 + 2t, 4 + 2t}
                        sh.Pcc[pi][pj] = pcs[sj * ld + sc[si]];
                    }
                    const double xr[3] = {xs[0], xs[1], xs[2]};
                    const MMPrep mp = mm_prep(xr, xs[al], xs[al + 1]);
                    if (tid < 10) sh.Hs[tid / 5][tid % 5] = mm_H(mp, tid);
                    if (tid == 31) {
                        sh.nu[0] = z0 - mp.sq;                   // :144, not wrapped (F10)
                        sh.nu[1] = z1 - mm_bearing(mp, xr[2]);
                    }
                    __syncwarp();
                    if (tid < 10) sh.HP[tid / 5][tid % 5] = ic_HP(sh.Hs, sh.Pcc, tid / 5, tid % 5);
                    __syncwarp();
                    const double Rd[2] = {R0, R1};
                    if (tid < 4) sh.S[tid >> 1][tid & 1] = ic_S(sh.HP, sh.Hs, Rd, tid >> 1, tid & 1);
                    __syncwarp();
                    if (tid == 0) {
                        double Si[2][2];
                        ic_inv(sh.S, Si);
                        sh.Sinv[0][0] = Si[0][0]; sh.Sinv[0][1] = Si[0][1]; sh.Sinv[1][0] = Si[1][0]; sh.Sinv[1][1] = Si[1][1];
                        if (a.actions) a.actions[(int64_t)f * a.max_obs + o - 1] = act[o - 1];
                    }
                }
                __syncthreads();
                double2 kown[RPT], bown[RPT];
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    const int i = tid + r * T;
                    if (i < n) {
                        double ph0 = 0.0, ph1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
                        for (int k = 0; k < 5; ++k) {
                            ph_acc(ph0, ph1, pcs[ks[k] * ld + i], sh.Hs[0][k], sh.Hs[1][k]);
                            if (!SYM) b_acc(b0, b1, prs[ks[k] * ld + i], sh.Hs[0][k], sh.Hs[1][k]);
                        }
                        if (SYM) { b0 = ph0; b1 = ph1; }
                        double k0, k1, dx;
                        gain_from_ph(ph0, ph1, sh.Sinv, sh.nu, k0, k1, dx);              // :143
                        xs[i] = __dadd_rn(xs[i], dx);                                    // :144
                        kown[r] = make_double2(k0, k1);
                        bown[r] = make_double2(b0, b1);
                        Ks[m * lds + i] = kown[r];
                        Bs[m * lds + i] = bown[r];
                        if (t + 1 < len)
                            for (int k = 0; k < NSr; ++k)
                                if (sc[k] == i) { Kc[k] = kown[r]; Bc[k] = bown[r]; }
                    }
                }
                if (tid == 0) {                                  // n is odd: row n pads the last row pair of the flush
                    Ks[m * lds + n] = make_double2(0.0, 0.0);
                    Bs[m * lds + n] = make_double2(0.0, 0.0);
                }
                __syncthreads();
                if (t + 1 < len) {                               // columns the later observations of the run read: same terms as mc_cur
#pragma unroll
                    for (int r = 0; r < RPT; ++r) {
                        const int i = tid + r * T;
                        if (i < n) {
#pragma unroll 4
                            for (int kk = 0; kk < NSr - 2 - 2 * t; ++kk) {
                                const int k = kk < 3 ? kk : kk + 2 + 2 * t;      // robot slots, then the slots of observations > t
                                const int c = sc[k];
                                if (SYM) pcs[k * ld + i] = i <= c ? rank2_sub(pcs[k * ld + i], kown[r], Bc[k]) : rank2_sub(pcs[k * ld + i], Kc[k], bown[r]);
                                else {
                                    pcs[k * ld + i] = rank2_sub(pcs[k * ld + i], kown[r], Bc[k]);
                                    prs[k * ld + i] = rank2_sub(prs[k * ld + i], Kc[k], bown[r]);
                                }
                            }
                        }
                    }
                    __syncthreads();
                }
            }
            ii += len;
            mpend = len;
            if (len >= a.maxd && ii <= cnt) {                    // slabs full and the step goes on: this CTA streams its P
                mc_flush(P, ld, n, SYM, Ks, Bs, lds, mpend);
                __syncthreads();
                mpend = 0;
            }
            if (ii > cnt) break;
            const int rn = act[ii - 1];
            if (rn >= 1 && rn <= N) continue;                    // the run was cut by the slab depth only
        }
        // ---- observation ii is an append (or names a slot that does not exist)
        const int r = act[ii - 1];
        const double z0 = ob[ii - 1], z1 = ob[ii - 1 + a.max_obs], z2 = ob[ii - 1 + 2 * a.max_obs];
        const double R0 = z0 * a.Rc0, R1 = z1 * a.Rc1;
        double want = 0.0, sig = 0.0;
        int lookup = 2;                                          // 1: first non-zero index, 2: index == want
        if (N == 0) { sig = 1.0; lookup = 1; }                   // :110-111 (F15)
        else if (a.variant == 0) { sig = z2; want = z2; }        // :118-120
        else { sig = (double)(N + 1); want = sig; }              // :121-123
        if (tid == 0) {
            int fatal = 0;
            if (r > 0) fatal = MC_ERR_INDEX;
            else {
                int hit = -1;
                for (int i = 0; i < ntab && hit < 0; ++i)
                    if (lookup == 1 ? (tix[i] != 0) : ((double)tix[i] == want)) hit = i;
                if (hit < 0) fatal = MC_ERR_TABLE;
                else { sh.pos[0] = loc[hit]; sh.pos[1] = loc[hit + a.max_table]; }
                if (N >= a.cap) fatal = MC_ERR_CAPACITY;
            }
            sh.fatal = fatal;
        }
        __syncthreads();
        if (sh.fatal) break;
        if (a.actions && tid == 0) a.actions[(int64_t)f * a.max_obs + ii - 1] = -(int)sig;
        // ---------------------------------------------------- append (EKF_SLAM.m:67-98), needs the whole P
        mc_flush(P, ld, n, SYM, Ks, Bs, lds, mpend);
        mpend = 0;
        __syncthreads();
        {
            const double th = xs[2];
            const double j02 = -u1 * sind(th), j12 = u1 * cosd(th);          // jxr, :84-85
            for (int t = tid; t <= N; t += T) {
                if (t < N) {
                    const int al = 3 + 2 * t;
                    for (int mm = 0; mm < 2; ++mm) {
                        const double p0 = ldP(P, ld, al + mm, 0, SYM), p1 = ldP(P, ld, al + mm, 1, SYM), p2 = ldP(P, ld, al + mm, 2, SYM);
                        const double f0 = dot3(1.0, p0, 0.0, p1, j02, p2);    // jxr * P(lm,1:3)', :95
                        const double f1 = dot3(0.0, p0, 1.0, p1, j12, p2);
                        P[(int64_t)(al + mm) + (int64_t)n * ld] = f0;
                        P[(int64_t)(al + mm) + (int64_t)(n + 1) * ld] = f1;
                        P[(int64_t)n + (int64_t)(al + mm) * ld] = f0;
                        P[(int64_t)(n + 1) + (int64_t)(al + mm) * ld] = f1;
                    }
                } else {
                    double Prr[3][3];
                    for (int i = 0; i < 3; ++i)
                        for (int j = 0; j < 3; ++j) Prr[i][j] = ldP(P, ld, i, j, SYM);
                    const double jxr[2][3] = {{1.0, 0.0, j02}, {0.0, 1.0, j12}};
                    const double cu = cosd(u2), su = sind(u2);
                    const double jz[2][2] = {{cu, -u1 * su}, {su, u1 * cu}};             // :87-88
                    double JP[2][3];
                    for (int rr = 0; rr < 2; ++rr)
                        for (int j = 0; j < 3; ++j) JP[rr][j] = dot3(jxr[rr][0], Prr[0][j], jxr[rr][1], Prr[1][j], jxr[rr][2], Prr[2][j]);
                    const double JR[2][2] = {{__dmul_rn(jz[0][0], R0), __dmul_rn(jz[0][1], R1)}, {__dmul_rn(jz[1][0], R0), __dmul_rn(jz[1][1], R1)}};
                    for (int rr = 0; rr < 2; ++rr)
                        for (int c = 0; c < 2; ++c)
                            P[(int64_t)(n + rr) + (int64_t)(n + c) * ld] =
                                __dadd_rn(dot3(JP[rr][0], jxr[c][0], JP[rr][1], jxr[c][1], JP[rr][2], jxr[c][2]),
                                __dadd_rn(__dmul_rn(JR[rr][0], jz[c][0]), __dmul_rn(JR[rr][1], jz[c][1])));   // :91
                    for (int i = 0; i < 3; ++i)
                        for (int c = 0; c < 2; ++c) {
                            const double v = dot3(Prr[i][0], jxr[c][0], Prr[i][1], jxr[c][1], Prr[i][2], jxr[c][2]);   // :92
                            P[(int64_t)i + (int64_t)(n + c) * ld] = v;
                            P[(int64_t)(n + c) + (int64_t)i * ld] = v;                   // :93
                        }
                    xs[n] = sh.pos[0]; xs[n + 1] = sh.pos[1];                            // :79
                    s[N] = sig;                                                          // :70
                }
            }
        }
        __syncthreads();
        if (tid == 0) sh.N = N + 1;
        __syncthreads();
        ++ii;
        need_assoc = true;
    }
    {
        const int n1 = 3 + 2 * sh.N;
        for (int i = tid; i < n1; i += T) x[i] = xs[i];
    }
    if (tid == 0) {
        a.Mpend[f] = mpend;                              // k_mc_flush applies them
        a.N[f] = sh.N;
        if (sh.fatal && atomicCAS(a.err, 0, sh.fatal) == 0) a.err[1] = f;
    }
}

void launch_mc_step(cudaStream_t st, const McArgs& a, int n_filters, int threads, size_t smem) {
    static size_t smem_set[2][64] = {{0}, {0}};
    const int mb = a.assoc_mode == 1 && a.variant != 0;
    if (smem_grows(smem_set[mb], smem)) {
        if (mb) cudaFuncSetAttribute(k_mc_step<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        else cudaFuncSetAttribute(k_mc_step<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    if (mb) k_mc_step<1, 0><<<n_filters, threads, smem, st>>>(a);
    else k_mc_step<0, 0><<<n_filters, threads, smem, st>>>(a);
}

template <int SYM, int RPT>
static void launch_mc_front_t(cudaStream_t st, const McArgs& a, int n_filters, size_t smem) {
    static size_t smem_set[64] = {0};
    if (smem_grows(smem_set, smem))
        cudaFuncSetAttribute(k_mc_front<SYM, RPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_mc_front<SYM, RPT><<<n_filters, 256, smem, st>>>(a);
}

// smem_limit: the device's opt-in shared memory per block; front_kind 0 = run-based kernel when it applies, 1 = k_mc_step<.,1>
void launch_mc_front(cudaStream_t st, const McArgs& a, int n_filters, size_t smem_limit, int front_kind) {
    const int mb = a.assoc_mode == 1 && a.variant != 0;
    const int n_cap = 3 + 2 * a.cap, NS = 3 + 2 * a.maxd;
    const size_t smem = sizeof(double) * (size_t)a.ldp * (1 + (a.sym ? 1 : 2) * NS) + sizeof(int) * (size_t)((a.max_obs + 3) & ~3);
    if (front_kind == 0 && !mb && n_cap <= 2048 && smem + 4096 <= smem_limit) {
        const int rpt = (n_cap + 255) / 256;
        if (a.sym) {
            if (rpt <= 2) launch_mc_front_t<1, 2>(st, a, n_filters, smem);
            else if (rpt <= 4) launch_mc_front_t<1, 4>(st, a, n_filters, smem);
            else launch_mc_front_t<1, 8>(st, a, n_filters, smem);
        } else {
            if (rpt <= 2) launch_mc_front_t<0, 2>(st, a, n_filters, smem);
            else if (rpt <= 4) launch_mc_front_t<0, 4>(st, a, n_filters, smem);
            else launch_mc_front_t<0, 8>(st, a, n_filters, smem);
        }
        return;
    }
    // robot-row cache: 3 ld doubles of dynamic shared memory per filter (9.9 KB at 200 landmarks; 4 CTAs/SM stay resident)
    static int rc_env = -1;
    if (rc_env < 0) { const char* v = getenv("EKF_MC_ROBOT_CACHE"); rc_env = v ? atoi(v) : 1; }
    McArgs b = a;
    const size_t rsm = 3 * sizeof(double) * (size_t)a.ldp;
    b.robot_cache = rc_env && rsm <= 40 * 1024;
    const size_t dyn = b.robot_cache ? rsm : 0;
    if (mb) k_mc_step<1, 1><<<n_filters, 128, dyn, st>>>(b);
    else k_mc_step<0, 1><<<n_filters, 128, dyn, st>>>(b);
}

void launch_mc_flush(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max) {
    constexpr int TC = 64;
    const int row_tiles = (cap_n + 127) / 128, col_tiles = (cap_n + TC - 1) / TC;
    dim3 grid(row_tiles * col_tiles, n_filters);
    // terms with q >= Mpend[f] are zero-filled, so any M >= the largest pending count is exact
    switch (m_max) {
        case 1: k_mc_flush<1, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 2: k_mc_flush<2, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 3: k_mc_flush<3, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 4: k_mc_flush<4, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 5: k_mc_flush<5, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 6: k_mc_flush<6, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        case 7: k_mc_flush<7, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
        default: k_mc_flush<8, TC, 8><<<grid, 64, 0, st>>>(a, row_tiles); break;
    }
}

}  // namespace ekf
