// ekf_shard.cu -- kernels of the column-slab sharded filter (one large P across G GPUs).
//
// Rank g stores columns [c_lo, c_hi) of the column-major P at full height (by symmetry of P this is the "row
// slab" of the survey, but contiguous in memory), EKF_STORAGE_FULL semantics.  x, s, the gains K_q (all n rows)
// and the linearisations are replicated: every rank computes them redundantly from identical inputs, so they
// stay bit-identical without any exchange.  The only data that crosses NVLink is, once per batch of <= 8
// observations, the 3 robot columns (from rank 0) and the 2 landmark columns of each observation (from their
// owner): Pc[slot][0..n).  Pending rank-2 terms are folded in on the fly:
//      cur(i, slot) = Pc[slot][i] - sum_{q<m} K_q(i,:) * Bc[q][slot]          (Bc[q][slot] = B_q(:, col(slot)))
// in the same order and with the same primitive (rank2_sub) as the single-GPU path.
#include "ekf_kernels.h"
#include "ekf_device.cuh"

namespace ekf {

// SYM storage: Pc[slot] is the assembled symmetric column c = col(slot); its element i is the live element
// (min(i,c), max(i,c)), whose pending correction is K_q(min) * B_q(max) with B_q = (P H')' known for all rows on
// every rank (Ball) -- exactly what ldP_def computes on one GPU, so sharded SYM == single-GPU SYM bit for bit.
__device__ __forceinline__ double cur_col(const double* __restrict__ Pc, int64_t ldc, int slot, int i,
                                          const double2* __restrict__ Kall, int64_t lds,
                                          const double2* __restrict__ Bc, int m,
                                          int sym = 0, int c = 0, const double2* __restrict__ Ball = nullptr) {
    double v = Pc[(int64_t)slot * ldc + i];
    if (sym) {
        const int lo = i <= c ? i : c, hi = i <= c ? c : i;
        for (int q = 0; q < m; ++q) v = rank2_sub(v, Kall[q * lds + lo], Ball[q * lds + hi]);
    } else {
        for (int q = 0; q < m; ++q) v = rank2_sub(v, Kall[q * lds + i], Bc[q * EKF_MAX_SLOTS + slot]);
    }
    return v;
}

// SYM slabs: every rank writes the part of each symmetric column it holds (zeros elsewhere); an all-reduce(sum)
// then assembles the columns exactly (each element has one non-zero contributor).
// S (continuous flush, may be null): strip of the local columns, S[4 t + r] = P0(r, c_lo + t) -- in charge of rows 0..2.
__global__ void k_pack_cols_sym(const double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, SlotSet ss, int n,
                                double* __restrict__ Pc, int64_t ldc, const double* __restrict__ S) {
    const int slot = blockIdx.y;
    const int c = ss.col[slot];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = 0.0;
    if (i <= c) { if (c >= c_lo && c < c_hi) v = (S && i < 3) ? S[4 * (int64_t)(c - c_lo) + i] : Ploc[(int64_t)i + (int64_t)(c - c_lo) * ld]; }
    else if (i >= c_lo && i < c_hi) v = (S && c < 3) ? S[4 * (int64_t)(i - c_lo) + c] : Ploc[(int64_t)c + (int64_t)(i - c_lo) * ld];
    Pc[(int64_t)slot * ldc + i] = v;
}

// Owner ranks copy their columns of the batch set into Pc (P0 is current: nothing is pending at batch start).
__global__ void k_pack_cols(const double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, SlotSet ss, int n,
                            double* __restrict__ Pc, int64_t ldc) {
    const int slot = blockIdx.y;
    const int c = ss.col[slot];
    if (c < c_lo || c >= c_hi) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) Pc[(int64_t)slot * ldc + i] = Ploc[(int64_t)i + (int64_t)(c - c_lo) * ld];
}

// ---- peer-memory exchange (default on one box): no collective library call on the data path ----
// Every element of an exchanged column has exactly one rank that stores it, so the "all-reduce" is really a scatter:
// the pack kernel writes each element it owns straight into the Pc buffer of EVERY rank over NVLink (plain stores into
// cudaIpc-mapped peer memory, coalesced along the column), then publishes the batch number in every peer's flag
// word (system-scope release).  Consumers wait on their local flag words (k_wait_flags).  Pc is double-buffered by
// batch parity: a rank can only start writing batch e+2 after it consumed batch e+1, which every peer signalled
// after it had finished reading batch e.
__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One thread = one row of all slot columns (grid n / 256 CTAs: see k_gather_cols_strip for why the CTA count matters).
__global__ void k_scatter_cols(const double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, SlotSet ss, int n,
                               PeerSet ps, int64_t buf_off, int sym, int epoch, unsigned int* __restrict__ ticket,
                               const double* __restrict__ S) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        double v[EKF_MAX_SLOTS];
        unsigned int have = 0u;
#pragma unroll
        for (int slot = 0; slot < EKF_MAX_SLOTS; ++slot) {
            if (slot < ss.n) {
                const int c = ss.col[slot];
                if (!sym || i <= c) { if (c >= c_lo && c < c_hi) { v[slot] = (S && i < 3) ? S[4 * (int64_t)(c - c_lo) + i] : Ploc[(int64_t)i + (int64_t)(c - c_lo) * ld]; have |= 1u << slot; } }
                else if (i >= c_lo && i < c_hi) { v[slot] = (S && c < 3) ? S[4 * (int64_t)(i - c_lo) + c] : Ploc[(int64_t)c + (int64_t)(i - c_lo) * ld]; have |= 1u << slot; }   // (c, i) lives in column i
            }
        }
#pragma unroll
        for (int slot = 0; slot < EKF_MAX_SLOTS; ++slot)
            if (have & (1u << slot))
                for (int g = 0; g < ps.world; ++g) ps.pc[g][buf_off + (int64_t)slot * ld + i] = v[slot];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                                // cumulative: covers the block's stores ordered before it by the barrier
        const unsigned int total = gridDim.x * gridDim.y;
        if (atomicAdd(ticket, 1u) == total - 1) {              // last block: every block's stores are fenced
            *ticket = 0;
            __threadfence_system();
            for (int g = 0; g < ps.world; ++g) st_release_sys(ps.flags[g] + ps.rank, epoch);
        }
    }
}

// Bounded spin: a peer that died (or never made the matching call) must not hang this rank's stream for ever.  After
// `timeout` clock cycles the waiter raises the handle's sticky exchange-error word in mapped host memory and gives up;
// the work queued behind it runs on incomplete columns, and every later synchronising call returns EKF_ENCCL.
__global__ void k_wait_flags(const int* __restrict__ flags, int world, int epoch, long long timeout, int* __restrict__ err) {
    if ((int)threadIdx.x < world) {
        const long long t0 = clock64();
        while (ld_acquire_sys(flags + threadIdx.x) < epoch) {
            if (*(volatile int*)err) break;                                   // already failed: do not wait again
            if (clock64() - t0 > timeout) { *(volatile int*)err = 1; __threadfence_system(); break; }
        }
    }
}

void launch_scatter_cols(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, const SlotSet& ss, int n,
                         const PeerSet& ps, int64_t buf_off, int sym, int epoch, unsigned int* ticket, const double* strip) {
    k_scatter_cols<<<(n + 255) / 256, 256, 0, st>>>(Ploc, ld, c_lo, c_hi, ss, n, ps, buf_off, sym, epoch, ticket, strip);
}

void launch_wait_flags(cudaStream_t st, const int* flags, int world, int epoch, long long timeout_cycles, int* err_word) {
    k_wait_flags<<<1, 32, 0, st>>>(flags, world, epoch, timeout_cycles, err_word);
}

// One warp, replicated on every rank: EKF_SLAM.m:125-143 from the broadcast columns, plus B_m(:, c) for every
// column c of the batch set (needed to correct the columns of later observations).
__global__ void k_lin_shard(const double* __restrict__ Pc, int64_t ldc, SlotSet ss, const double* __restrict__ x,
                            int a, int sa, double z0, double z1, double R0, double R1,
                            const double2* __restrict__ Kall, int64_t lds, double2* __restrict__ Bc, int m,
                            Lin* __restrict__ lin, int sym, const double2* __restrict__ Ball) {
    __shared__ double Pcc[5][5];
    __shared__ double Hsm[2][5];
    const int lane = threadIdx.x;
    const int c[5] = {0, 1, 2, a, a + 1};
    const int sl[5] = {0, 1, 2, sa, sa + 1};
    if (lane < 25) {
        int i = lane / 5, j = lane % 5;
        Pcc[i][j] = cur_col(Pc, ldc, sl[j], c[i], Kall, lds, Bc, m, sym, c[j], Ball);   // P(c_i, c_j)
    }
    __syncwarp();
    if (lane == 0) {
        double xr[3] = {x[0], x[1], x[2]};
        double zk[2], Hs[2][5], S[2][2], Si[2][2], Pl[5][5];
        measurement_model(xr, x[a], x[a + 1], zk, Hs);
        double Rd[2] = {R0, R1};
        for (int i = 0; i < 5; ++i)
            for (int j = 0; j < 5; ++j) Pl[i][j] = Pcc[i][j];
        innovation_cov(Hs, Pl, Rd, S, Si);
        int bad = 0;
        for (int r = 0; r < 2; ++r) {
            for (int k = 0; k < 5; ++k) { lin->Hs[r][k] = Hs[r][k]; Hsm[r][k] = Hs[r][k]; }
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
    __syncwarp();
    if (!sym && lane < ss.n) {                           // B_m(:, col(slot)) = H_m * cur(c, col(slot))
        double b0 = 0.0, b1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            double pr = cur_col(Pc, ldc, lane, c[k], Kall, lds, Bc, m);
            b_acc(b0, b1, pr, Hsm[0][k], Hsm[1][k]);
        }
        Bc[m * EKF_MAX_SLOTS + lane] = make_double2(b0, b1);
    }
}

// All n rows on every rank: K_m(i,:), x(i); owned columns only: B_m(:, i) from the local slab.
__global__ void k_gain_shard(const double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, double* __restrict__ x,
                             int n, int a, int sa, const Lin* __restrict__ lin, const double* __restrict__ Pc,
                             int64_t ldc, double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds,
                             const double2* __restrict__ Bc, int m, int sym) {
    __shared__ Lin L;
    if (threadIdx.x < sizeof(Lin) / sizeof(double)) ((double*)&L)[threadIdx.x] = ((const double*)lin)[threadIdx.x];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c[5] = {0, 1, 2, a, a + 1};
    const int sl[5] = {0, 1, 2, sa, sa + 1};
    double ph0 = 0.0, ph1 = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        double pc = cur_col(Pc, ldc, sl[k], i, Kall, lds, Bc, m, sym, c[k], Ball);   // P(i, c_k)
        ph_acc(ph0, ph1, pc, L.Hs[0][k], L.Hs[1][k]);
    }
    double k0, k1, dx;
    gain_from_ph(ph0, ph1, L.Sinv, L.nu, k0, k1, dx);
    x[i] = __dadd_rn(x[i], dx);
    Kall[m * lds + i] = make_double2(k0, k1);
    if (sym) {
        Ball[m * lds + i] = make_double2(ph0, ph1);      // B = (P H')', all rows on every rank
    } else if (i >= c_lo && i < c_hi) {
        const double* col = Ploc + (int64_t)(i - c_lo) * ld;
        double b0 = 0.0, b1 = 0.0;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            double pr = col[c[k]];                                            // P0(c_k, i), local
            for (int q = 0; q < m; ++q) pr = rank2_sub(pr, Kall[q * lds + c[k]], Ball[q * lds + i]);
            b_acc(b0, b1, pr, L.Hs[0][k], L.Hs[1][k]);
        }
        Ball[m * lds + i] = make_double2(b0, b1);
    }
}

// predict on a slab: the row pass is local to every owned column, the column pass and the corner live on rank 0.
__global__ void k_predict_shard(double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, const double* __restrict__ x,
                                int n, double u1, double u2, double C, int sym,
                                double2* __restrict__ Kall, double2* __restrict__ Ball, int64_t lds, int m) {
    const double th = x[2];
    const double st = sind(th), ct = cosd(th);
    const double a = -1.0 * u1 * st, b = u1 * ct;
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
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0 && c_lo == 0) {
        double M[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) M[i][j] = (sym && i > j) ? Ploc[j + i * ld] : Ploc[i + j * ld];
        for (int j = 0; j < 3; ++j) { M[0][j] = M[0][j] + a * M[2][j]; M[1][j] = M[1][j] + b * M[2][j]; }
        for (int i = 0; i < 3; ++i) { M[i][0] = M[i][0] + M[i][2] * a; M[i][1] = M[i][1] + M[i][2] * b; }
        double W[3] = {u1 * ct, u1 * st, u2};
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Ploc[i + j * ld] = M[i][j] + (W[i] * C) * W[j];
    } else if (t >= 3 && t < n) {
        if (t >= c_lo && t < c_hi) {
            double* col = Ploc + (int64_t)(t - c_lo) * ld;
            double p2 = col[2];
            col[0] = col[0] + a * p2;
            col[1] = col[1] + b * p2;
        }
        if (c_lo == 0 && !sym) {                         // rank 0 owns columns 0..2 (dead triangle in SYM)
            double q2 = Ploc[t + 2 * ld];
            Ploc[t] = Ploc[t] + q2 * a;
            Ploc[t + ld] = Ploc[t + ld] + q2 * b;
        }
    }
}

// append on a slab; Pc slots 0..2 hold the broadcast columns 0..2 (P(i,k) = Pc[k][i]).  EKF_SLAM.m:84-97.
__global__ void k_append_shard(double* __restrict__ Ploc, int64_t ld, int c_lo, int c_hi, double* __restrict__ x,
                               double* __restrict__ s, int N, double u1, double u2, double R0, double R1, double px,
                               double py, double sig, const double* __restrict__ Pc, int64_t ldc) {
    const int n = 3 + 2 * N;
    const double th = x[2];
    const double j02 = -u1 * sind(th), j12 = u1 * cosd(th);
    const bool own_new = (n >= c_lo && n < c_hi);
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < N) {
        const int a = 3 + 2 * t;
        for (int m = 0; m < 2; ++m) {
            double p0 = Pc[a + m], p1 = Pc[ldc + a + m], p2 = Pc[2 * ldc + a + m];
            double f0 = dot3(1.0, p0, 0.0, p1, j02, p2);
            double f1 = dot3(0.0, p0, 1.0, p1, j12, p2);
            if (own_new) {
                Ploc[(int64_t)(a + m) + (int64_t)(n - c_lo) * ld] = f0;
                Ploc[(int64_t)(a + m) + (int64_t)(n + 1 - c_lo) * ld] = f1;
            }
            if (a + m >= c_lo && a + m < c_hi) {
                Ploc[(int64_t)n + (int64_t)(a + m - c_lo) * ld] = f0;
                Ploc[(int64_t)(n + 1) + (int64_t)(a + m - c_lo) * ld] = f1;
            }
        }
    } else if (t == N) {
        double Prr[3][3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Prr[i][j] = Pc[(int64_t)j * ldc + i];
        const double jxr[2][3] = {{1.0, 0.0, j02}, {0.0, 1.0, j12}};
        const double cu = cosd(u2), su = sind(u2);
        const double jz[2][2] = {{cu, -u1 * su}, {su, u1 * cu}};
        double JP[2][3];
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j) JP[r][j] = dot3(jxr[r][0], Prr[0][j], jxr[r][1], Prr[1][j], jxr[r][2], Prr[2][j]);
        double JR[2][2] = {{__dmul_rn(jz[0][0], R0), __dmul_rn(jz[0][1], R1)}, {__dmul_rn(jz[1][0], R0), __dmul_rn(jz[1][1], R1)}};
        for (int r = 0; r < 2; ++r)
            for (int c = 0; c < 2; ++c) {
                double v = __dadd_rn(dot3(JP[r][0], jxr[c][0], JP[r][1], jxr[c][1], JP[r][2], jxr[c][2]),
                           __dadd_rn(__dmul_rn(JR[r][0], jz[c][0]), __dmul_rn(JR[r][1], jz[c][1])));
                if (own_new) Ploc[(int64_t)(n + r) + (int64_t)(n + c - c_lo) * ld] = v;
            }
        for (int i = 0; i < 3; ++i)
            for (int c = 0; c < 2; ++c) {
                double v = dot3(Prr[i][0], jxr[c][0], Prr[i][1], jxr[c][1], Prr[i][2], jxr[c][2]);
                if (own_new) Ploc[(int64_t)i + (int64_t)(n + c - c_lo) * ld] = v;
                if (c_lo == 0) Ploc[(int64_t)(n + c) + (int64_t)i * ld] = v;
            }
        x[n] = px; x[n + 1] = py;
        s[N] = sig;
    }
}

// ------------------------------------------------------------------ launchers
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

void launch_pack_cols(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, const SlotSet& ss, int n,
                      double* Pc, int64_t ldc, int sym, const double* strip) {
    dim3 grid(cdiv(n, 256), ss.n);
    if (sym) k_pack_cols_sym<<<grid, 256, 0, st>>>(Ploc, ld, c_lo, c_hi, ss, n, Pc, ldc, strip);
    else k_pack_cols<<<grid, 256, 0, st>>>(Ploc, ld, c_lo, c_hi, ss, n, Pc, ldc);
}

void launch_lin_gain_shard(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, int a,
                           int sa, double z0, double z1, double R0, double R1, const SlotSet& ss, const double* Pc,
                           int64_t ldc, Lin* lin, double2* Kall, double2* Ball, int64_t lds, double2* Bc, int m, int sym) {
    k_lin_shard<<<1, 32, 0, st>>>(Pc, ldc, ss, x, a, sa, z0, z1, R0, R1, Kall, lds, Bc, m, lin, sym, Ball);
    k_gain_shard<<<cdiv(n, 256), 256, 0, st>>>(Ploc, ld, c_lo, c_hi, x, n, a, sa, lin, Pc, ldc, Kall, Ball, lds, Bc, m, sym);
}

void launch_predict_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, double u1,
                          double u2, double C, int sym) {
    k_predict_shard<<<cdiv(n, 256), 256, 0, st>>>(Ploc, ld, c_lo, c_hi, x, n, u1, u2, C, sym, nullptr, nullptr, 0, 0);
    launch_predict_x(st, x, u1, u2);
}

// covariance half alone (decoupled predict): the heading comes from the pre-motion pose the state half saved
void launch_predict_shard_P(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, const double* xprev, int n,
                            double u1, double u2, double C, int sym, double2* Kall, double2* Ball, int64_t lds, int m) {
    k_predict_shard<<<cdiv(n, 256) + (m > 0 ? 1 : 0), 256, 0, st>>>(Ploc, ld, c_lo, c_hi, xprev, n, u1, u2, C, sym, Kall, Ball, lds, m);
}

void launch_append_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, double* s, int N,
                         double u1, double u2, double R0, double R1, double px, double py, double sig, const double* Pc,
                         int64_t ldc) {
    k_append_shard<<<cdiv(N + 1, 128), 128, 0, st>>>(Ploc, ld, c_lo, c_hi, x, s, N, u1, u2, R0, R1, px, py, sig, Pc, ldc);
}

}  // namespace ekf
