// ekf_flush.cu -- K3, the covariance downdate / rank-2M flush kernels (EKF_SLAM.m:145 `P=(eye(n)-K*H_k)*P`), sm_100a.
//   k_downdate        CUDA-core DFMA, LDG/STG streaming                       (variants 0-3)
//   k_downdate_mma    FP64 DMMA on LDG fragments                              (variants 4-7)
//   k_downdate_tma    cp.async.bulk staging in and out + FP64 DMMA            (variants 8-24; default for rank <= 16)
//   k_downdate_hyb    cp.async.bulk staging in, DMMA, direct 256-bit stores out (variants 30+; rank-32 flush)
//   k_mc_flush_tma    the tile pipeline over a batch of small matrices (Monte-Carlo filters)
#include "ekf_kernels.h"
#include "ekf_device.cuh"
#include <cstdlib>
#include <atomic>

namespace ekf {

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device (context): remember it per (kernel, device), so that a process
// that creates handles on several devices (ekf_params.device, ekf_create_multi) opts in on each of them.
template <auto KERNEL>
static void smem_optin(int bytes) {
    static std::atomic<unsigned long long> done{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_relaxed) & bit)) {
        cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        done.fetch_or(bit, std::memory_order_relaxed);
    }
}

// ------------------------------------------------------------------ K3 covariance downdate (rank-2M flush)
// P(i,j) -= sum_{q<M} K_q(i,0) B_q(0,j) + K_q(i,1) B_q(1,j), terms subtracted in observation order (rank2_sub),
// == M successive applications of P <- (I - K H) P (EKF_SLAM.m:145) in ONE pass over P: 16 B of HBM traffic per
// live element for M observations.  M = 1 is the immediate path.  Column-major P: a block owns TR = BLOCK*VEC
// contiguous rows and TC columns; each thread keeps its VEC rows of the M gains in registers (VEC*M double2),
// the TC columns of the M B-rows sit in shared memory (broadcast LDS.128), and the loop issues UNROLL
// independent 128/256-bit loads per thread before the first use.  SYM storage launches the same grid and drops
// tiles strictly below the diagonal, halving the traffic.  4M flop per 16 B: HBM-bound up to M ~ 20 on B200.
template <int VEC> struct VecIO;
template <> struct VecIO<2> {
    typedef double2 T;
    static __device__ __forceinline__ T ld(const double* p) { return ld_stream2(p); }
    static __device__ __forceinline__ void st(double* p, T v) { st_stream2(p, v); }
    static __device__ __forceinline__ void sub(T& v, const double2* k, int M, int q, double2 b) {
        v.x = rank2_sub(v.x, k[0 * M + q], b);
        v.y = rank2_sub(v.y, k[1 * M + q], b);
    }
};
template <> struct VecIO<4> {
    typedef double4a T;
    static __device__ __forceinline__ T ld(const double* p) { return ld_stream4(p); }
    static __device__ __forceinline__ void st(double* p, T v) { st_stream4(p, v); }
    static __device__ __forceinline__ void sub(T& v, const double2* k, int M, int q, double2 b) {
        v.x = rank2_sub(v.x, k[0 * M + q], b);
        v.y = rank2_sub(v.y, k[1 * M + q], b);
        v.z = rank2_sub(v.z, k[2 * M + q], b);
        v.w = rank2_sub(v.w, k[3 * M + q], b);
    }
};

template <int M, int VEC, int BLOCK, int TC, int UNROLL>
__global__ void __launch_bounds__(BLOCK)
k_downdate(double* __restrict__ P, int64_t ld, int n, int ncols, int sym,
           const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int col_off) {
    constexpr int TR = BLOCK * VEC;
    typedef VecIO<VEC> IO;
    const int r0 = blockIdx.x * TR;
    const int c0 = blockIdx.y * TC;
    if (sym && r0 > col_off + c0 + TC - 1) return;    // tile strictly below the diagonal: dead in SYM storage
    __shared__ double2 Bs[TC * M];                    // Bs[j*M + q] = B_q(:, c0+j)
    for (int t = threadIdx.x; t < TC * M; t += BLOCK) {
        int j = t / M, q = t - j * M;
        Bs[t] = (c0 + j < ncols) ? Ball[q * lds + c0 + j] : make_double2(0.0, 0.0);
    }
    __syncthreads();
    const int r = r0 + threadIdx.x * VEC;
    if (r >= n) return;                               // rows [n, ld) are padding; the K slabs are zero there
    double2 k[VEC * M];
#pragma unroll
    for (int v = 0; v < VEC; ++v)
#pragma unroll
        for (int q = 0; q < M; ++q) k[v * M + q] = Kall[q * lds + r + v];
    double* p = P + (int64_t)r + (int64_t)c0 * ld;
    const int nc = min(TC, ncols - c0);
    int j = 0;
    for (; j + UNROLL <= nc; j += UNROLL) {
        typename IO::T v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = IO::ld(p + (int64_t)(j + u) * ld);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
#pragma unroll
            for (int q = 0; q < M; ++q) IO::sub(v[u], k, M, q, Bs[(j + u) * M + q]);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) IO::st(p + (int64_t)(j + u) * ld, v[u]);
    }
    for (; j < nc; ++j) {
        typename IO::T v = IO::ld(p + (int64_t)j * ld);
#pragma unroll
        for (int q = 0; q < M; ++q) IO::sub(v, k, M, q, Bs[j * M + q]);
        IO::st(p + (int64_t)j * ld, v);
    }
}

// ------------------------------------------------------------------ K3b rank-2M flush on FP64 tensor cores
// Same update as k_downdate, contracted with mma.sync.m8n8k4.f64 (DMMA): for M >= 2 the flush is a SYRK-like
// n x 2M x n contraction and the operands fit in a few fragment registers instead of VEC*M gains per thread,
// so more warps (and more bytes) are in flight.  Mapping (per warp: 16 rows x 8 columns of P per group):
//   MMA row r = T/4   <-> P column j0 + r          A[r][k] = B_q(comp, j0+r)          (from shared memory)
//   MMA col c         <-> P row, permuted so that a thread's two C registers of tile 0 and of tile 1 are four
//                         consecutive rows i0+4t..i0+4t+3 (t = T%4): one LDG.E.256 / STG.E.256 per thread,
//                         four lanes = one full 128-byte line per column
//   k = 2*(q%2)+comp, k-step s = q/2               Bop[k][c] = -K_q(row(c), comp)    (registers, loaded once)
// Terms enter each accumulator in the order (q, x), (q, y), (q+1, x), ... like rank2_sub.
__device__ __forceinline__ void dmma8x8x4(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int KS, int TC, int UNROLL>      // KS = ceil(M/2) k-steps; 128 threads = 4 warps x 16 rows
__global__ void __launch_bounds__(128)
k_downdate_mma(double* __restrict__ P, int64_t ld, int n, int ncols, int sym,
               const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int M, int col_off) {
    constexpr int TR = 64;
    constexpr int SROW = 4 * KS + 4;                   // doubles per column in smem (+4 pad: 2-way max conflict)
    constexpr int NG = TC / 8;
    const int r0 = blockIdx.x * TR;
    const int c0 = blockIdx.y * TC;
    if (sym && r0 > col_off + c0 + TC - 1) return;
    __shared__ double Bs[TC * SROW];
    for (int e = threadIdx.x; e < TC * 2 * KS; e += 128) {
        int j = e / (2 * KS), q = e - j * (2 * KS);
        double2 b = (c0 + j < ncols && q < M) ? Ball[q * lds + c0 + j] : make_double2(0.0, 0.0);
        Bs[j * SROW + 2 * q] = b.x;
        Bs[j * SROW + 2 * q + 1] = b.y;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = lane >> 2, t = lane & 3;
    const int i0 = r0 + warp * 16;
    if (i0 >= n) return;
    double kf[2][KS];
#pragma unroll
    for (int s = 0; s < KS; ++s) {
        const int q = 2 * s + (t >> 1);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
            const int row = i0 + 4 * (r >> 1) + 2 * tile + (r & 1);
            double2 kk = (q < M) ? Kall[q * lds + row] : make_double2(0.0, 0.0);
            kf[tile][s] = -((t & 1) ? kk.y : kk.x);
        }
    }
    double* p = P + (int64_t)(i0 + 4 * t) + (int64_t)(c0 + r) * ld;
    const double* bs = Bs + r * SROW + t;
#pragma unroll 1
    for (int g = 0; g < NG; g += UNROLL) {
        double4a v[UNROLL];
        bool ok[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            ok[u] = (g + u < NG) && (c0 + (g + u) * 8 + r < ncols);
            if (ok[u]) v[u] = ld_stream4(p + (int64_t)(g + u) * 8 * ld);
            else { v[u].x = v[u].y = v[u].z = v[u].w = 0.0; }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int gg = (g + u < NG) ? (g + u) : (NG - 1);
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                const double a = bs[gg * 8 * SROW + 4 * s];
                dmma8x8x4(v[u].x, v[u].y, a, kf[0][s]);
                dmma8x8x4(v[u].z, v[u].w, a, kf[1][s]);
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (ok[u]) st_stream4(p + (int64_t)(g + u) * 8 * ld, v[u]);
    }
}

template <int KS, int TC, int UNROLL>
static void launch_dd_mma(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                          const double2* Ball, int64_t lds, int M, int col_off) {
    dim3 grid((n + 63) / 64, (ncols + TC - 1) / TC);
    k_downdate_mma<KS, TC, UNROLL><<<grid, 128, 0, st>>>(P, ld, n, ncols, sym, Kall, Ball, lds, M, col_off);
}

// ------------------------------------------------------------------ K3c rank-2M flush, TMA-staged (bulk async copies)
// Same arithmetic as k_downdate_mma (same fragments, same term order -> bit-identical), but P never travels
// through registers on its way from/to HBM: a producer warp moves 128-row column segments (1 KB contiguous each)
// global -> shared with cp.async.bulk + mbarrier complete_tx, 8 consumer warps update the stage in place with DMMA,
// and the producer writes it back shared -> global with cp.async.bulk (bulk_group).  NSTAGE stages of SCOLS columns
// ring through shared memory, so every SM keeps NSTAGE*SCOLS KB of reads in flight with one issuing warp and the
// DRAM sees long contiguous bursts.  Columns sit CS = 1040 B apart in shared memory (1 KB + 16 B pad) so the
// 8-column x 128-B fragment reads (LDS.128) are bank-conflict free.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Live-tile enumeration for the TMA flush.  Tiles are TR rows x TC columns with TR = RPC * TC (RPC = 1 or 2).  In SYM
// storage column tile y keeps row tiles 0 .. min(R, f(y)) - 1, f(y) = floor((A + y) / RPC) + 1 with
// A = floor((col_off + TC - 1) / TC): the tiles that touch the upper triangle.  Launching exactly those (1-D grid, column
// tile major) instead of the R x CT rectangle removes ~half the CTAs of a SYM launch, which only existed to exit.
struct TileMap { int R, CT, A, RPC, ystar, order, zero; };   // ystar: first column tile whose live count is clipped at R; order 1: two fronts (k_downdate_tma)
__host__ __device__ inline long long tm_G(long long m, int r) {          // sum_{t<m} floor(t / r)
    const long long q = m / r, rem = m % r;
    return (long long)r * q * (q - 1) / 2 + q * rem;
}
__host__ __device__ inline long long tm_prefix(const TileMap& tm, int Y) {   // live tiles in column tiles [0, Y)
    const int y1 = Y < tm.ystar ? Y : tm.ystar;
    return tm_G((long long)tm.A + y1, tm.RPC) - tm_G(tm.A, tm.RPC) + y1 + (long long)tm.R * (Y - y1);
}
static inline TileMap tm_make(int R, int CT, int TC, int RPC, int col_off, int sym) {
    TileMap tm;
    tm.R = R; tm.CT = CT; tm.RPC = RPC; tm.order = 0; tm.zero = 0;   // zero: a 0 the compiler cannot see (k_downdate_hyb's slot release)
    tm.A = (col_off + TC - 1) / TC;
    long long ys = sym ? (long long)RPC * (R - 1) - tm.A : 0;            // f(y) >= R  <=>  y >= RPC (R - 1) - A
    tm.ystar = (int)(ys < 0 ? 0 : (ys > CT ? CT : ys));
    return tm;
}
// Device-side decode runs once per CTA before anything else can start, so it is 32-bit and division-free (RPC is a
// template constant: / and % become shifts); launchers guarantee the live-tile count fits 31 bits.
template <int RPC> __device__ __forceinline__ int tm_G32(int m) {
    const int q = m / RPC, rem = m % RPC;
    return RPC * (q * (q - 1) / 2) + q * rem;
}
template <int RPC> __device__ __forceinline__ int tm_prefix32(const TileMap& tm, int gA, int Y) {
    const int y1 = Y < tm.ystar ? Y : tm.ystar;
    return tm_G32<RPC>(tm.A + y1) - gA + y1 + tm.R * (Y - y1);
}
template <int RPC> __device__ __forceinline__ void tm_decode(const TileMap& tm, int t, int& bx, int& by) {
    const int gA = tm_G32<RPC>(tm.A);
    int lo = 0, hi = tm.CT - 1;                                          // largest by with prefix(by) <= t
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (tm_prefix32<RPC>(tm, gA, mid) <= t) lo = mid; else hi = mid - 1;
    }
    by = lo;
    bx = t - tm_prefix32<RPC>(tm, gA, lo);
}

template <int KS, int TC, int SCOLS, int NSTAGE, int CG = 1>
struct TmaCfg {
    static constexpr int TR = 128;                       // rows per CTA tile: 8 consumer warps x 16 rows
    static constexpr int CS = TR * 8 + 16;               // bytes between columns of a stage in shared memory
    static constexpr int SROW = 4 * KS + 4;              // doubles per column of the B operand
    static constexpr int STAGE_BYTES = SCOLS * CS;
    static constexpr int SMEM = NSTAGE * STAGE_BYTES + TC * SROW * 8 + 2 * NSTAGE * 8;
    static constexpr int CWARPS = 8 * CG;                // consumer warps: CG sets of 8 (set c takes column groups c, c+CG, ..)
    static constexpr int THREADS = (CWARPS + 1) * 32;    // + 1 producer warp
};

// One tile (rows r0.., columns c0..) of the flush.  TRIM (SYM storage, column 0 of this P is global column 0): column
// j of a diagonal tile is copied only down to the diagonal (rows <= j, rounded to a row pair), so the bytes moved are
// exactly the live ones -- it matters for the small matrices of the Monte-Carlo batch, where diagonal tiles are a
// large share; consumers then work on stale shared memory below the diagonal, which is never stored.
template <int KS, int TC, int SCOLS, int NSTAGE, int CG, bool TRIM>
__device__ __forceinline__ void dd_tma_tile(double* __restrict__ P, int64_t ld, int n, int ncols,
                                            const double2* __restrict__ Kall, const double2* __restrict__ Ball,
                                            int64_t lds, int M, int r0, int c0) {
    typedef TmaCfg<KS, TC, SCOLS, NSTAGE, CG> C;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* stage = smem;
    double* Bs = (double*)(smem + NSTAGE * C::STAGE_BYTES);
    uint64_t* bars = (uint64_t*)(smem + NSTAGE * C::STAGE_BYTES + TC * C::SROW * 8);
    const uint32_t full0 = smem_u32(bars), done0 = smem_u32(bars + NSTAGE);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile_cols = min(TC, ncols - c0);
    const int ns = (tile_cols + SCOLS - 1) / SCOLS;      // stages of this tile
    const int rows_here = (int)min((int64_t)C::TR, ld - r0);     // ld is a multiple of 16 rows -> 128-byte multiple
    const uint32_t col_bytes = (uint32_t)rows_here * 8u;

    if (tid == 0) {
        for (int i = 0; i < NSTAGE; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(done0 + 8 * i, C::CWARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();                                             // barriers visible; nothing has touched memory yet

    if (warp != C::CWARPS) {
        // consumers stage the B operand of the tile (zero beyond M / ncols) while the producer already streams P
        for (int e = tid; e < TC * 2 * KS; e += C::CWARPS * 32) {
            int j = e / (2 * KS), q = e - j * (2 * KS);
            double2 b = (j < tile_cols && q < M) ? Ball[q * lds + c0 + j] : make_double2(0.0, 0.0);
            Bs[j * C::SROW + 2 * q] = b.x;
            Bs[j * C::SROW + 2 * q + 1] = b.y;
        }
    }

    if (warp == C::CWARPS) {
        // ------------------------------------------------ producer: lane c owns column c of every stage
        const bool col_lane = lane < SCOLS;
        auto bytes_of = [&](int s, int nc) -> uint32_t {                // bytes this lane moves for stage s
            if (!(col_lane && lane < nc)) return 0u;
            if (!TRIM) return col_bytes;
            const int live = (c0 + s * SCOLS + lane) - r0 + 1;           // rows r0 .. j of column j
            return live <= 0 ? 0u : min(col_bytes, (uint32_t)((live + 1) & ~1) * 8u);
        };
        auto load_stage = [&](int s) {
            const int slot = s % NSTAGE;
            const int nc = min(SCOLS, tile_cols - s * SCOLS);
            const uint32_t mine = bytes_of(s, nc);
            const uint32_t total = TRIM ? __reduce_add_sync(0xffffffffu, mine) : (uint32_t)nc * col_bytes;
            if (lane == 0) mbar_arrive_expect_tx(full0 + 8 * slot, total);
            __syncwarp();
            if (mine)
                bulk_g2s(smem_u32(stage + slot * C::STAGE_BYTES + lane * C::CS),
                         P + (int64_t)r0 + (int64_t)(c0 + s * SCOLS + lane) * ld, mine, full0 + 8 * slot);
        };
        for (int s = 0; s < NSTAGE - 1 && s < ns; ++s) load_stage(s);
        for (int s = 0; s < ns; ++s) {
            const int slot = s % NSTAGE;
            mbar_wait(done0 + 8 * slot, (s / NSTAGE) & 1);           // consumers finished stage s (writes fenced)
            const int nc = min(SCOLS, tile_cols - s * SCOLS);
            const uint32_t mine = bytes_of(s, nc);
            if (mine)
                bulk_s2g(P + (int64_t)r0 + (int64_t)(c0 + s * SCOLS + lane) * ld,
                         smem_u32(stage + slot * C::STAGE_BYTES + lane * C::CS), mine);
            bulk_commit();
            const int s2 = s + NSTAGE - 1;                           // refill the slot stage s-1 was stored from
            if (s2 < ns) {
                if (s >= 1) bulk_wait_read<1>();                     // every group but the newest (stage s) has read its smem
                __syncwarp();
                load_stage(s2);
            }
        }
        bulk_wait_read<0>();                                         // shared memory must outlive the last store's read
        return;
    }

    // ---------------------------------------------------- consumers: warp w owns rows r0 + 16w .. +15
    const int r = lane >> 2, t = lane & 3;
    const int rw = warp & 7, cg = warp >> 3;                         // row-warp within the tile, consumer set
    const int i0 = r0 + rw * 16;
    const bool active = i0 < n;                                      // rows [n, ld) are padding: K is zero there
    double kf[2][KS];
#pragma unroll
    for (int s = 0; s < KS; ++s) {
        const int q = 2 * s + (t >> 1);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
            const int row = i0 + 4 * (r >> 1) + 2 * tile + (r & 1);
            double2 kk = (active && q < M) ? Kall[q * lds + row] : make_double2(0.0, 0.0);
            kf[tile][s] = -((t & 1) ? kk.y : kk.x);
        }
    }
    const int row_off = (rw * 16 + 4 * t) * 8;                     // byte offset of this thread's 4 rows in a column
    asm volatile("bar.sync 1, %0;" :: "n"(C::CWARPS * 32) : "memory");   // B operand staged (consumer warps only)
    for (int s = 0; s < ns; ++s) {
        const int slot = s % NSTAGE;
        mbar_wait(full0 + 8 * slot, (s / NSTAGE) & 1);
        if (active) {
            unsigned char* sb = stage + slot * C::STAGE_BYTES + row_off;
#pragma unroll
            for (int g = cg; g < SCOLS / 8; g += CG) {
                const int cs = g * 8 + r;                            // column within the stage
                const int ct = s * SCOLS + cs;                       // column within the tile
                if (s * SCOLS + g * 8 >= tile_cols) break;           // warp-uniform: whole group beyond the tile
                const bool ok = ct < tile_cols;                      // per-lane; mma.sync itself must stay warp-wide
                double2* pv = (double2*)(sb + cs * C::CS);
                double2 lo = make_double2(0.0, 0.0), hi = lo;
                if (ok) { lo = pv[0]; hi = pv[1]; }
                const double* bs = Bs + (ok ? ct : 0) * C::SROW + t;
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) {
                    const double a = bs[4 * ks];
                    dmma8x8x4(lo.x, lo.y, a, kf[0][ks]);
                    dmma8x8x4(hi.x, hi.y, a, kf[1][ks]);
                }
                if (ok) { pv[0] = lo; pv[1] = hi; }
            }
        }
        fence_async_smem();                                          // generic-proxy writes -> visible to the bulk store
        __syncwarp();
        if (lane == 0) mbar_arrive(done0 + 8 * slot);
    }
}

template <int KS, int TC, int SCOLS, int NSTAGE, int CG, int MINB = 1>
__global__ void __launch_bounds__((8 * CG + 1) * 32, MINB)
k_downdate_tma(double* __restrict__ P, int64_t ld, int n, int ncols, int sym,
               const double2* __restrict__ Kall, const double2* __restrict__ Ball, int64_t lds, int M, TileMap tm) {
    typedef TmaCfg<KS, TC, SCOLS, NSTAGE, CG> C;
    int bx, by;
    // Launch order of the live tiles.  Natural = column-tile major (neighbouring CTAs stream neighbouring 1 KB segments of
    // the same columns: DRAM page locality; a scattered order measures 12 % slower).  For the triangle the natural order
    // starts with the short column tiles, where that locality is weakest, and ends far from where it started; walking
    // the list from both ends at once (even CTAs ascending, odd CTAs descending, meeting in the middle) measured 2.5 %
    // faster at 10k for the rank-16 flush (0.527 -> 0.514 ms, profiles/README.md) and is its default in SYM storage.
    const int T = (int)gridDim.x, t0 = (int)blockIdx.x;
    const int t = tm.order == 1 ? ((t0 & 1) ? T - 1 - (t0 >> 1) : (t0 >> 1)) : t0;
    tm_decode<C::TR / TC>(tm, t, bx, by);   // only live tiles are launched
    dd_tma_tile<KS, TC, SCOLS, NSTAGE, CG, false>(P, ld, n, ncols, Kall, Ball, lds, M, bx * C::TR, by * TC);
}

// The same tile pipeline over a batch of small matrices (Monte-Carlo filters, ekf_mc.cu): blockIdx.y = filter,
// blockIdx.x = live tile of that filter's n x n matrix (n and the pending count come from device arrays).
template <int KS, int TC, int SCOLS, int NSTAGE, int CG>
__global__ void __launch_bounds__((8 * CG + 1) * 32)
k_mc_flush_tma(McArgs a) {
    typedef TmaCfg<KS, TC, SCOLS, NSTAGE, CG> C;
    const int f = blockIdx.y;
    const int n = 3 + 2 * a.N[f], M = a.Mpend[f];
    if (M == 0) return;
    TileMap tm;
    tm.R = (n + C::TR - 1) / C::TR; tm.CT = (n + TC - 1) / TC; tm.RPC = C::TR / TC; tm.A = 0;
    const int ys = a.sym ? tm.RPC * (tm.R - 1) : 0;
    tm.ystar = ys > tm.CT ? tm.CT : ys;
    if ((long long)blockIdx.x >= tm_prefix(tm, tm.CT)) return;
    int bx, by;
    tm_decode<C::TR / TC>(tm, (int)blockIdx.x, bx, by);
    double* P = a.P + (int64_t)f * a.pstride;
    const double2* Kg = a.Kg + (int64_t)f * a.maxd * a.ldp;
    const double2* Bg = a.Bg + (int64_t)f * a.maxd * a.ldp;
    if (a.sym) dd_tma_tile<KS, TC, SCOLS, NSTAGE, CG, true>(P, a.ldp, n, n, Kg, Bg, a.ldp, M, bx * C::TR, by * TC);
    else dd_tma_tile<KS, TC, SCOLS, NSTAGE, CG, false>(P, a.ldp, n, n, Kg, Bg, a.ldp, M, bx * C::TR, by * TC);
}

template <int KS>
static void launch_mc_flush_tma_ks(cudaStream_t st, const McArgs& a, int n_filters, int cap_n) {
    constexpr int TC = 64, SCOLS = 8, NSTAGE = 6, CG = 1;
    typedef TmaCfg<KS, TC, SCOLS, NSTAGE, CG> C;
    smem_optin<k_mc_flush_tma<KS, TC, SCOLS, NSTAGE, CG>>(C::SMEM);
    TileMap tm = tm_make((cap_n + C::TR - 1) / C::TR, (cap_n + TC - 1) / TC, TC, C::TR / TC, 0, a.sym);
    dim3 grid((unsigned)tm_prefix(tm, tm.CT), n_filters);            // live tiles at capacity; smaller filters exit early
    k_mc_flush_tma<KS, TC, SCOLS, NSTAGE, CG><<<grid, C::THREADS, C::SMEM, st>>>(a);
}

void launch_mc_flush_tma(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max) {
    switch ((m_max + 1) / 2) {
        case 1: launch_mc_flush_tma_ks<1>(st, a, n_filters, cap_n); break;
        case 2: launch_mc_flush_tma_ks<2>(st, a, n_filters, cap_n); break;
        case 3: launch_mc_flush_tma_ks<3>(st, a, n_filters, cap_n); break;
        default: launch_mc_flush_tma_ks<4>(st, a, n_filters, cap_n); break;
    }
}

template <int KS, int TC, int SCOLS, int NSTAGE, int CG = 1, int MINB = 1>
static void launch_dd_tma(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                          const double2* Ball, int64_t lds, int M, int col_off) {
    typedef TmaCfg<KS, TC, SCOLS, NSTAGE, CG> C;
    smem_optin<k_downdate_tma<KS, TC, SCOLS, NSTAGE, CG, MINB>>(C::SMEM);
    static_assert(C::TR % TC == 0 && (C::TR / TC == 1 || C::TR / TC == 2), "tile map needs TR = TC or 2 TC");
    const int R = (n + C::TR - 1) / C::TR, CT = (ncols + TC - 1) / TC;   // row tiles cover the active rows only
    TileMap tm = tm_make(R, CT, TC, C::TR / TC, col_off, sym);
    { static int ord = -3; if (ord == -3) { const char* v = getenv("EKF_TILE_ORDER"); ord = v ? atoi(v) : -2; }   // -3: not read yet, -2: no override
      // default: two fronts for the 128-column tiles of the rank >= 10 flush; the 64-column tile shapes of the lower ranks
      // measure 1-2 % faster in natural order (profiles/README.md)
      tm.order = !sym ? 0 : (ord == -2 ? (TC == 128 ? 1 : 0) : ord); }
    const long long live = tm_prefix(tm, CT);
    if (live <= 0 || live > 0x3fffffffLL) return;        // device decode is 32-bit
    k_downdate_tma<KS, TC, SCOLS, NSTAGE, CG, MINB><<<(unsigned)live, C::THREADS, C::SMEM, st>>>(P, ld, n, ncols, sym, Kall, Ball, lds, M, tm);
}

// ------------------------------------------------------------------ K3d rank-2M flush, hybrid: bulk-async in, registers out
// k_downdate_tma couples its loads to its arithmetic: a stage's slot can only be refilled after the consumers have
// finished the DMMA chain, written the stage back to shared memory, fenced, and the producer has stored it out -- so a
// CTA's loads run only ~2 stages ahead of its math, and at rank 32 (16 dependent-pair DMMAs per warp and stage) neither
// HBM (0.75) nor the FP64 tensor pipe (0.53) is kept busy.  Here the way IN is the same (one producer warp, 1 KB column
// segments global -> shared with cp.async.bulk + mbarrier complete_tx, NSTAGE ring), but the consumers read their
// fragment out of the stage into the accumulator registers, RELEASE THE SLOT once the last fragment read has been
// consumed by an issued DMMA, and store the updated 4 consecutive rows of a column straight from registers with one
// 256-bit st.global (four lanes = one full 128-byte line, like k_downdate_mma): no STS, no proxy fence, no store
// handshake, and all NSTAGE slots can be in flight while the DMMA chains run.
// BST = 1: the B operand of a stage (SCOLS columns x 4 KS doubles, from the column-major copy Bt[col][SROW] that
// k_pack_bt writes) rides in the stage itself with one extra bulk copy, instead of a per-CTA 37 KB operand tile:
// shared memory per stage 10.6 KB -> deeper rings at the same CTAs/SM.
// CH = 4 (SCOLS = 16): a warp runs the two 8-column groups of a stage as four interleaved DMMA chains.
// Arithmetic, fragments and term order are those of k_downdate_mma / k_downdate_tma: bit-identical results.
template <int KS, int SCOLS, int NSTAGE, int BST>
struct HybCfg {
    static constexpr int TR = 128, TC = 128;
    static constexpr int CS = TR * 8 + 16;               // bytes between columns of a stage (16 B pad: conflict-free LDS.128)
    static constexpr int SROW = 4 * KS + 4;              // doubles per column of the B operand
    static constexpr int PBYTES = SCOLS * CS;
    static constexpr int BBYTES = BST ? SCOLS * SROW * 8 : 0;
    static constexpr int STAGE_BYTES = (PBYTES + BBYTES + 127) / 128 * 128;
    static constexpr int BS_BYTES = BST ? 0 : TC * SROW * 8;
    static constexpr int SMEM = NSTAGE * STAGE_BYTES + BS_BYTES + 2 * NSTAGE * 8;
    static constexpr int THREADS = 9 * 32;               // 8 consumer warps x 16 rows + 1 producer warp
};

// Bt[col][SROW]: column-major copy of the pending B slabs, Bt[col * SROW + 2 q + c] = B_q(c, col), zero for q >= M.
template <int KS>
__global__ void k_pack_bt(const double2* __restrict__ Ball, int64_t lds, int M, int ncols, double* __restrict__ Bt) {
    constexpr int SROW = 4 * KS + 4;
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= ncols) return;
    double* o = Bt + (int64_t)col * SROW;
#pragma unroll
    for (int q = 0; q < 2 * KS; ++q) {
        const double2 b = q < M ? Ball[q * lds + col] : make_double2(0.0, 0.0);
        *(double2*)(o + 2 * q) = b;
    }
    *(double2*)(o + 4 * KS) = make_double2(0.0, 0.0);
    *(double2*)(o + 4 * KS + 2) = make_double2(0.0, 0.0);
}

template <int KS, int SCOLS, int NSTAGE, int BST, int MINB>
__global__ void __launch_bounds__(288, MINB)
k_downdate_hyb(const double* P, double* Pout, int64_t ld, int n, int ncols, const double2* __restrict__ Kall,
               const double2* __restrict__ Ball, const double* __restrict__ Bt, int64_t lds, int M, TileMap tm) {
    typedef HybCfg<KS, SCOLS, NSTAGE, BST> C;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* stage = smem;
    double* Bs = (double*)(smem + NSTAGE * C::STAGE_BYTES);
    uint64_t* bars = (uint64_t*)(smem + NSTAGE * C::STAGE_BYTES + C::BS_BYTES);
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + NSTAGE);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    int bx, by;
    {
        const int T = (int)gridDim.x, t0 = (int)blockIdx.x;
        const int t = tm.order == 1 ? ((t0 & 1) ? T - 1 - (t0 >> 1) : (t0 >> 1)) : t0;   // two fronts (see k_downdate_tma)
        tm_decode<1>(tm, t, bx, by);
    }
    const int r0 = bx * C::TR, c0 = by * C::TC;
    const int tile_cols = min(C::TC, ncols - c0);
    const int ns = (tile_cols + SCOLS - 1) / SCOLS;
    const int rows_here = (int)min((int64_t)C::TR, ld - r0);
    const uint32_t col_bytes = (uint32_t)rows_here * 8u;

    if (tid == 0) {
        for (int i = 0; i < NSTAGE; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 8) {
        // ------------------------------------------------ producer: lane c moves column c of every stage, lane 31 its B operand
        for (int s = 0; s < ns; ++s) {
            const int slot = s % NSTAGE;
            if (s >= NSTAGE) mbar_wait(empty0 + 8 * slot, ((s / NSTAGE) - 1) & 1);     // every consumer warp has read the slot
            const int nc = min(SCOLS, tile_cols - s * SCOLS);
            const uint32_t bbytes = BST ? (uint32_t)nc * C::SROW * 8u : 0u;
            if (lane == 0) mbar_arrive_expect_tx(full0 + 8 * slot, (uint32_t)nc * col_bytes + bbytes);
            __syncwarp();
            unsigned char* sb = stage + slot * C::STAGE_BYTES;
            if (lane < nc)
                bulk_g2s(smem_u32(sb + lane * C::CS), P + (int64_t)r0 + (int64_t)(c0 + s * SCOLS + lane) * ld, col_bytes,
                         full0 + 8 * slot);
            if (BST && lane == 31)
                bulk_g2s(smem_u32(sb + C::PBYTES), Bt + (int64_t)(c0 + s * SCOLS) * C::SROW, bbytes, full0 + 8 * slot);
        }
        return;
    }

    // ---------------------------------------------------- consumers: warp w owns rows r0 + 16 w .. + 15
    if (!BST) {
        for (int e = tid; e < C::TC * 2 * KS; e += 256) {
            const int j = e / (2 * KS), q = e - j * (2 * KS);
            const double2 b = (j < tile_cols && q < M) ? Ball[q * lds + c0 + j] : make_double2(0.0, 0.0);
            Bs[j * C::SROW + 2 * q] = b.x;
            Bs[j * C::SROW + 2 * q + 1] = b.y;
        }
    }
    const int r = lane >> 2, t = lane & 3;
    const int i0 = r0 + warp * 16;
    const bool active = i0 < n;                                      // rows [n, ld) are padding: never stored
    double kf[2][KS];
#pragma unroll
    for (int s = 0; s < KS; ++s) {
        const int q = 2 * s + (t >> 1);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
            const int row = i0 + 4 * (r >> 1) + 2 * tile + (r & 1);
            const double2 kk = (active && q < M) ? Kall[q * lds + row] : make_double2(0.0, 0.0);
            kf[tile][s] = -((t & 1) ? kk.y : kk.x);
        }
    }
    const int row_off = (warp * 16 + 4 * t) * 8;                     // byte offset of this lane's 4 rows within a column
    double* gp = Pout + (int64_t)(i0 + 4 * t) + (int64_t)(c0 + r) * ld;  // this lane's 4 rows of column r of group 0 (Pout == P: in place)
    if (!BST) asm volatile("bar.sync 1, 256;" ::: "memory");         // B operand staged (consumer warps only)
    constexpr int NG = SCOLS / 8;
    for (int s = 0; s < ns; ++s) {
        const int slot = s % NSTAGE;
        mbar_wait(full0 + 8 * slot, (s / NSTAGE) & 1);
        const unsigned char* sb = stage + slot * C::STAGE_BYTES;
        const double* bop = BST ? (const double*)(sb + C::PBYTES) : Bs + (int64_t)s * SCOLS * C::SROW;
        double2 lo[NG], hi[NG];
        bool ok[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) {
            const int cs = g * 8 + r;
            ok[g] = active && (s * SCOLS + cs < tile_cols);
            lo[g] = make_double2(0.0, 0.0); hi[g] = lo[g];
            if (ok[g]) {
                const double2* pv = (const double2*)(sb + cs * C::CS + row_off);
                lo[g] = pv[0]; hi[g] = pv[1];
            }
        }
        int dep = 0;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                // columns beyond the tile read stale operand rows: they only reach accumulator rows that are never stored
                const double a = bop[(g * 8 + r) * C::SROW + 4 * ks + t];
                if (BST && ks + 3 >= KS) dep ^= __double2hiint(a);
                dmma8x8x4(lo[g].x, lo[g].y, a, kf[0][ks]);
                dmma8x8x4(hi[g].x, hi[g].y, a, kf[1][ks]);
            }
        }
        // The slot may be refilled once every shared-memory read of this stage has RETURNED.  With the B operand in the stage
        // (BST) the last reads happen late, and ptxas schedules the arrive between them and the DMMAs that consume them (an
        // LDS that has only been issued can still be waiting in the memory pipe -- beside kernels that flood it with
        // uncoalesced loads, long enough for the producer's refill to land first: seen as wrong elements when this kernel
        // ran next to the gather / gain kernels of the continuous flush).  So the arrive's address is made to depend on the
        // values of the last reads: dep is compared with a pattern, and the offset it would select is a zero the compiler
        // cannot see (tm.zero) -- the arrive cannot issue before those loads have written their registers.
        __syncwarp();
        if (lane == 0) mbar_arrive(empty0 + 8 * slot + (BST && dep == 0x7ff4dead ? (uint32_t)tm.zero : 0u));
#pragma unroll
        for (int g = 0; g < NG; ++g)
            if (ok[g]) {
                double4a v; v.x = lo[g].x; v.y = lo[g].y; v.z = hi[g].x; v.w = hi[g].y;
                st_stream4(gp + (int64_t)(s * SCOLS + g * 8) * ld, v);
            }
    }
}

template <int KS, int SCOLS, int NSTAGE, int BST, int MINB>
static void launch_dd_hyb(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                          const double2* Ball, double* Bt, int64_t lds, int M, int col_off, double* Pout, bool prepacked) {
    typedef HybCfg<KS, SCOLS, NSTAGE, BST> C;
    smem_optin<k_downdate_hyb<KS, SCOLS, NSTAGE, BST, MINB>>(C::SMEM);
    const int R = (n + C::TR - 1) / C::TR, CT = (ncols + C::TC - 1) / C::TC;
    TileMap tm = tm_make(R, CT, C::TC, 1, col_off, sym);
    { static int ord = -3; if (ord == -3) { const char* v = getenv("EKF_TILE_ORDER"); ord = v ? atoi(v) : -2; }
      tm.order = !sym ? 0 : (ord == -2 ? 1 : ord); }
    const long long live = tm_prefix(tm, CT);
    if (live <= 0 || live > 0x3fffffffLL) return;
    if (BST && !prepacked) k_pack_bt<KS><<<cdiv(ncols, 128), 128, 0, st>>>(Ball, lds, M, ncols, Bt);
    k_downdate_hyb<KS, SCOLS, NSTAGE, BST, MINB><<<(unsigned)live, C::THREADS, C::SMEM, st>>>(P, Pout ? Pout : P, ld, n, ncols, Kall, Ball, Bt, lds, M, tm);
}

// hybrid variants by number (EKF_DOWNDATE_VARIANT >= 30); returns false when `v` names none
template <int KS>
static bool launch_dd_hyb_v(int v, cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                            const double2* Ball, double* Bt, int64_t lds, int M, int col_off, double* Pout = nullptr,
                            bool prepacked = false) {
#define HY(SC, NS, BS, MB) launch_dd_hyb<KS, SC, NS, BS, MB>(st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, M, col_off, Pout, prepacked); return true
    if (Bt == nullptr && v != 30 && v != 37) return false;       // the in-stage operand needs the Bt scratch buffer
    switch (v) {
        case 30: HY(8, 4, 0, 3);       // operand tile per CTA as in k_downdate_tma, 4 x 8-col stages
        case 31: HY(8, 6, 1, 3);       // operand in the stage: 6 x 10.6 KB = 64 KB, 3 CTAs/SM
        case 32: HY(8, 4, 1, 3);
        case 33: HY(8, 8, 1, 2);       // 8 stages, 2 CTAs/SM
        case 34: HY(16, 3, 1, 3);      // 16-col stages (four DMMA chains per warp), 3 x 21.2 KB
        case 35: HY(16, 4, 1, 2);
        case 36: HY(8, 5, 1, 4);       // 4 CTAs/SM if the registers allow (<= 56)
        case 37: HY(8, 6, 0, 2);
        case 38: HY(16, 5, 1, 2);
        default: return false;
    }
#undef HY
}

template <int M, int VEC, int BLOCK, int TC, int UNROLL>
static void launch_dd(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall, const double2* Ball,
                      int64_t lds, int col_off) {
    dim3 grid(cdiv(n, BLOCK * VEC), cdiv(ncols, TC));
    k_downdate<M, VEC, BLOCK, TC, UNROLL><<<grid, BLOCK, 0, st>>>(P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
}


// The B operand of the hybrid kernels, packed ahead of the pass (continuous flush: on the chain's stream when the set is
// closed, into the buffer of that pass, so that nothing stands between two passes on the pass stream).
void launch_pack_bt(cudaStream_t st, const double2* Ball, int64_t lds, int m, int ncols, double* Bt) {
    if (m < 1 || ncols < 1) return;
    switch ((m + 1) / 2) {
#define PB(KS_) case KS_: k_pack_bt<KS_><<<cdiv(ncols, 128), 128, 0, st>>>(Ball, lds, m, ncols, Bt); break;
        PB(1) PB(2) PB(3) PB(4) PB(5) PB(6) PB(7) PB(8)
#undef PB
        default: break;
    }
}

// m pending rank-2 terms (1 <= m <= EKF_MAX_DEFER) applied in one pass over P.
void launch_downdate(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall, const double2* Ball,
                     int64_t lds, int m, int col_off, int variant, double* Bt, double* Pout, bool prepacked) {
    if (m < 1) return;
    if (Pout && Pout != P) {
        // out of place (continuous flush): the hybrid kernels store straight from registers, so the destination is just
        // another pointer; any other kernel choice gets a copy of the matrix first and runs in place on it
        const bool hyb = (variant == 30 || variant == 37) || (Bt != nullptr && ((variant >= 31 && variant <= 38) || (variant < 30 && m > 8 && variant != 13 && variant != 21)));
        if (!hyb) {
            cudaMemcpyAsync(Pout, P, sizeof(double) * (size_t)ld * (size_t)ncols, cudaMemcpyDeviceToDevice, st);
            P = Pout;
            Pout = nullptr;
        }
    }
    if (variant >= 30) {    // hybrid kernels (bulk-async in, registers out), any rank
        bool done = false;
        switch ((m + 1) / 2) {
#define HV(KS_) case KS_: done = launch_dd_hyb_v<KS_>(variant, st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, m, col_off, Pout, prepacked); break;
            HV(1) HV(2) HV(3) HV(4) HV(5) HV(6) HV(7) HV(8)
#undef HV
            default: break;
        }
        if (done) return;
        variant = -1;
    }
    // variants: 0..3 CUDA-core DFMA (LDG/STG), 4..7 DMMA with LDG/STG fragments, >= 8 DMMA with bulk-async (TMA) staging
    // default (variant < 0), from the B200 sweep in profiles/README.md: TMA staging everywhere; tile shape by rank
    //   m == 1: 64 x 16-col stages, 2 consumer sets (v18)   m = 2..4: 64-col tiles, 6 x 8-col stages (v17)
    //   m >= 5: 128-col tiles, 4 x 8-col stages, 4 CTAs/SM (v13)
    const int v = variant >= 0 ? variant : (m == 1 ? 18 : (m <= 4 ? 17 : 13));
    if (m > 8) {
        // cross-step deferral (9..16 pending terms, rank <= 32).  Default: the hybrid kernel with the B operand in its stages
        // (k_downdate_hyb<KS, 8-col stages, 4 stages, BST, 3 CTAs/SM>): 0.520 ms at 10k SYM = 0.94 of the measured HBM peak,
        // FULL storage 0.983 ms = 0.996 (profiles/r02/sweep_flush_call1.jsonl); the round-1 TMA-in/TMA-out kernel (0.660 ms,
        // 0.74) stays selectable (variant 13; 21 = its 64-column 4-CTA/SM shape) and is the fallback without a Bt scratch buffer.
        bool done = false;
        if (Bt != nullptr && variant != 13 && variant != 21) {
            switch ((m + 1) / 2) {
                case 5: done = launch_dd_hyb_v<5>(32, st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, m, col_off, Pout, prepacked); break;
                case 6: done = launch_dd_hyb_v<6>(32, st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, m, col_off, Pout, prepacked); break;
                case 7: done = launch_dd_hyb_v<7>(32, st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, m, col_off, Pout, prepacked); break;
                default: done = launch_dd_hyb_v<8>(32, st, P, ld, n, ncols, sym, Kall, Ball, Bt, lds, m, col_off, Pout, prepacked); break;
            }
        }
        if (done) return;
        switch ((m + 1) / 2) {
            // MINB = 3: 72 registers x 288 threads and 70 KB of shared memory both allow 3 CTAs/SM
            case 5: launch_dd_tma<5, 128, 8, 4, 1, 3>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); break;
            case 6: launch_dd_tma<6, 128, 8, 4, 1, 3>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); break;
            case 7: launch_dd_tma<7, 128, 8, 4, 1, 3>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); break;
            default:
                if (variant == 21) launch_dd_tma<8, 64, 8, 4, 1, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off);
                else launch_dd_tma<8, 128, 8, 4, 1, 3>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off);
                break;
        }
        return;
    }
    if (m >= 1 && v >= 8) {
        const int ks = (m + 1) / 2;
#define DT(KS_) \
        if (v == 9) launch_dd_tma<KS_, 128, 16, 3>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 11) launch_dd_tma<KS_, 128, 8, 6>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 13) launch_dd_tma<KS_, 128, 8, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 14) launch_dd_tma<KS_, 64, 8, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 15) launch_dd_tma<KS_, 128, 8, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 16) launch_dd_tma<KS_, 128, 16, 3, 2>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 17) launch_dd_tma<KS_, 64, 8, 6>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 18) launch_dd_tma<KS_, 64, 16, 4, 2>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else launch_dd_tma<KS_, 128, 16, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off);
        switch (ks) { case 1: { DT(1) } break; case 2: { DT(2) } break; case 3: { DT(3) } break; default: { DT(4) } break; }
#undef DT
        return;
    }
    if (m >= 2 && v >= 4) {
        const int ks = (m + 1) / 2;
#define DM(KS_) \
        if (v == 5) launch_dd_mma<KS_, 128, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 6) launch_dd_mma<KS_, 64, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else if (v == 7) launch_dd_mma<KS_, 256, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off); \
        else launch_dd_mma<KS_, 128, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, m, col_off);
        switch (ks) { case 1: { DM(1) } break; case 2: { DM(2) } break; case 3: { DM(3) } break; default: { DM(4) } break; }
#undef DM
        return;
    }
#define DD(M_) \
    if (v == 1) launch_dd<M_, 2, 128, 64, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off); \
    else if (v == 2) launch_dd<M_, 2, 256, 32, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off); \
    else if (v == 3) launch_dd<M_, 2, 128, 32, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off); \
    else launch_dd<M_, 2, 128, 32, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
    switch (m) {
        case 1:
            if (v == 1) launch_dd<1, 2, 128, 64, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
            else if (v == 2) launch_dd<1, 4, 64, 64, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
            else if (v == 3) launch_dd<1, 4, 128, 64, 4>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
            else launch_dd<1, 4, 128, 32, 8>(st, P, ld, n, ncols, sym, Kall, Ball, lds, col_off);
            break;
        case 2: { DD(2) } break;
        case 3: { DD(3) } break;
        case 4: { DD(4) } break;
        case 5: { DD(5) } break;
        case 6: { DD(6) } break;
        case 7: { DD(7) } break;
        case 8: { DD(8) } break;
        default: break;
    }
#undef DD
}

}  // namespace ekf
