// tools/tri_stream.cu -- does the HBM care whether the live (upper-triangular) part of many small column-major
// matrices is strided (ld apart, dead gap after every column) or packed?  Streams nf matrices of n x n doubles
// read-modify-write three ways with the same thread mapping (one warp per column, lanes over row pairs).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ long long off_packed(int j) {   // doubles before column j when column c holds (c+2)&~1 entries
    const long long m = j >> 1;
    return 2 * m * (m + 1) + ((j & 1) ? (2 * m + 2) : 0);
}
template <int MODE>   // 0 full columns, 1 upper part at stride ld, 2 packed upper
__global__ void k_stream(double* P, int n, long long ld, long long pstride) {
    const int f = blockIdx.y;
    double* A = P + (long long)f * pstride;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int j = blockIdx.x * nw + warp; j < n; j += gridDim.x * nw) {
        const int len = MODE == 0 ? ((n + 1) & ~1) : ((j + 2) & ~1);
        double* col = MODE == 2 ? A + off_packed(j) : A + (long long)j * ld;
        for (int r = 2 * lane; r < len; r += 64) {
            double2 v = *(double2*)(col + r);
            v.x = v.x * 1.0000001 - 1e-9; v.y = v.y * 0.9999999 + 1e-9;
            *(double2*)(col + r) = v;
        }
    }
}

// ---- replica of k_mc_flush_w (ekf_mc.cu) with its parts switchable, to find which one costs bandwidth ----
// STAGE: stage the K slab in shared memory (+ barrier); COMP: 0 trivial arithmetic, 1 rank-2M update from shared memory;
// TC columns per CTA; PF: prefetch one item ahead
template <int M, int TC, int STAGE, int COMP, int PF, int MINB, int NC>
__global__ void __launch_bounds__(256, MINB) k_like(double* Pall, const double2* Kall, const double2* Ball, int n, long long ld, long long pstride, int cap) {
    extern __shared__ __align__(16) unsigned char dyn[];
    const int f = blockIdx.y, c0 = blockIdx.x * TC, tid = threadIdx.x;
    if (c0 >= n) return;
    double* P = Pall + (long long)f * pstride;
    const double2* Kg = Kall + (long long)f * M * ld;
    const double2* Bg = Ball + (long long)f * M * ld;
    const int nc = min(TC, n - c0);
    const int nrows = min(n + 1, ((c0 + nc + 1) & ~1));
    const int NRH = cap + 2;
    double2* Ksh = (double2*)dyn;
    double2* Bsh = Ksh + (size_t)2 * NRH * M;
    const int warp = tid >> 5, lane = tid & 31;
    auto grp_rend = [&](int g) { const int jmax = c0 + min(g + NC - 1, nc - 1); return min(n, jmax + 1); };
    auto load = [&](double2 (&v)[NC], int g, int rb) {
        const int r = rb + 2 * lane;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int col = c0 + g + c;
            if (g + c < nc && r <= col) v[c] = *(const double2*)(P + r + (long long)col * ld);
        }
    };
    int g = warp * NC, rb = 0;
    bool have = g < nc;
    double2 vn[NC];
    if (PF && have) load(vn, g, rb);
    if (STAGE) {
        for (int e = tid; e < nrows * M; e += 256) {
            const int q = e / nrows, i = e - q * nrows;
            Ksh[(q * 2 + (i & 1)) * NRH + (i >> 1)] = Kg[q * ld + i];
        }
        for (int e = tid; e < nc * M; e += 256) {
            const int j = e / M, q = e - j * M;
            Bsh[e] = Bg[q * ld + c0 + j];
        }
        __syncthreads();
    }
    while (have) {
        double2 v[NC];
        if (PF) {
#pragma unroll
            for (int c = 0; c < NC; ++c) v[c] = vn[c];
        } else load(v, g, rb);
        const int cg = g, r = rb + 2 * lane;
        rb += 64;
        if (rb >= grp_rend(g)) { g += 8 * NC; rb = 0; }
        have = g < nc;
        if (PF && have) load(vn, g, rb);
        if (r <= n) {
            if (COMP) {
#pragma unroll
                for (int q = 0; q < M; ++q) {
                    const double2 k0 = Ksh[(q * 2) * NRH + (r >> 1)], k1 = Ksh[(q * 2 + 1) * NRH + (r >> 1)];
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const double2 b = Bsh[min(cg + c, nc - 1) * M + q];
                        v[c].x = fma(-k0.y, b.y, fma(-k0.x, b.x, v[c].x));
                        v[c].y = fma(-k1.y, b.y, fma(-k1.x, b.x, v[c].y));
                    }
                }
            } else {
#pragma unroll
                for (int c = 0; c < NC; ++c) { v[c].x = v[c].x * 1.0000001 - 1e-9; v[c].y = v[c].y * 0.9999999 + 1e-9; }
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int col = c0 + cg + c;
                if (cg + c < nc && r <= col) *(double2*)(P + r + (long long)col * ld) = v[c];
            }
        }
    }
}
template <int TC, int STAGE, int COMP, int PF, int MINB, int NC = 4>
static void run_like(double* P, double2* K, double2* B, int n, long long ld, long long pstride, int nf, cudaEvent_t e0, cudaEvent_t e1, int smem_mode = 2) {
    constexpr int M = 5; const int cap = (n - 3) / 2;
    // smem_mode 0: no dynamic shared memory (only valid with STAGE 0), 1: allocate it, default carve-out, 2: allocate + max-shared carve-out
    const size_t smem = smem_mode == 0 ? 0 : ((size_t)(cap + 2) * 2 * M + (size_t)TC * M) * sizeof(double2);
    auto kf = k_like<M, TC, STAGE, COMP, PF, MINB, NC>;
    cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kf, cudaFuncAttributePreferredSharedMemoryCarveout, smem_mode == 2 ? (int)cudaSharedmemCarveoutMaxShared : smem_mode > 2 ? smem_mode : (int)cudaSharedmemCarveoutDefault);
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
        dim3 grid((n + TC - 1) / TC, nf);
        cudaEventRecord(e0);
        kf<<<grid, 256, smem>>>(P, K, B, n, ld, pstride, cap);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
    }
    printf("like smem_mode %d NC %d TC %3d stage %d comp %d pf %d minb %d: %.3f ms  %.0f GB/s  (%s)\n", smem_mode, NC, TC, STAGE, COMP, PF, MINB, best, 8.0 * n * (n + 1) * nf / best / 1e6,
           cudaGetErrorString(cudaGetLastError()));
}

template <int UNR, int CONTIG>
__global__ void k_stream_v(double* P, int n, long long ld, long long pstride) {
    const int f = blockIdx.y;
    double* A = P + (long long)f * pstride;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    const int per = (n + gridDim.x - 1) / gridDim.x;   // CONTIG: CTA owns a contiguous column range, warps interleave inside it
    const int jb = CONTIG ? blockIdx.x * per + warp : blockIdx.x * nw + warp;
    const int je = CONTIG ? min(n, (blockIdx.x + 1) * per) : n;
    const int js = CONTIG ? nw : gridDim.x * nw;
    for (int j = jb; j < je; j += js) {
        const int len = (j + 2) & ~1;
        double* col = A + (long long)j * ld;
#pragma unroll UNR
        for (int r = 2 * lane; r < len; r += 64) {
            double2 v = *(double2*)(col + r);
            v.x = v.x * 1.0000001 - 1e-9; v.y = v.y * 0.9999999 + 1e-9;
            *(double2*)(col + r) = v;
        }
    }
}
template <int UNR, int CONTIG>
static void run_v(double* P, int n, long long ld, long long pstride, int nf, int gx, int threads, cudaEvent_t e0, cudaEvent_t e1) {
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
        dim3 grid(gx, nf);
        cudaEventRecord(e0);
        k_stream_v<UNR, CONTIG><<<grid, threads>>>(P, n, ld, pstride);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
    }
    printf("stream_v unroll %d contig %d grid.x %d threads %d: %.3f ms  %.0f GB/s\n", UNR, CONTIG, gx, threads, best, 8.0 * n * (n + 1) * nf / best / 1e6);
}
int main() {
    const int n = 403, nf = 4096; const long long ld = 416, pstride = ld * n;
    double* P; cudaMalloc(&P, sizeof(double) * pstride * nf); cudaMemset(P, 0, sizeof(double) * pstride * nf);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double live = 8.0 * n * (n + 1) * nf, full = 16.0 * n * 404 * nf;
    for (int mode = 0; mode < 3; ++mode)
        for (int gx : {1, 4, 13}) {
            float best = 1e9f;
            for (int rep = 0; rep < 5; ++rep) {
                dim3 grid(gx, nf);
                cudaEventRecord(e0);
                if (mode == 0) k_stream<0><<<grid, 256>>>(P, n, ld, pstride);
                else if (mode == 1) k_stream<1><<<grid, 256>>>(P, n, ld, pstride);
                else k_stream<2><<<grid, 256>>>(P, n, ld, pstride / 2 + 1024);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
            }
            printf("mode %d (%s) grid.x %2d: %.3f ms  %.0f GB/s\n", mode, mode == 0 ? "full" : mode == 1 ? "upper, strided" : "upper, packed",
                   gx, best, (mode == 0 ? full : live) / best / 1e6);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    double2 *K, *B; cudaMalloc(&K, sizeof(double2) * 5 * ld * nf); cudaMalloc(&B, sizeof(double2) * 5 * ld * nf);
    cudaMemset(K, 0, sizeof(double2) * 5 * ld * nf); cudaMemset(B, 0, sizeof(double2) * 5 * ld * nf);
    for (int sm : {1, 40, 50, 60, 70, 80}) {     // > 2: explicit carve-out percentage
        run_like<64, 1, 1, 1, 4, 4>(P, K, B, n, ld, pstride, nf, e0, e1, sm);
        run_like<64, 1, 1, 1, 5, 4>(P, K, B, n, ld, pstride, nf, e0, e1, sm);
        run_like<64, 1, 1, 1, 6, 4>(P, K, B, n, ld, pstride, nf, e0, e1, sm);
    }
    run_like<64, 1, 1, 1, 4, 2>(P, K, B, n, ld, pstride, nf, e0, e1, 1);
    run_like<64, 1, 1, 1, 6, 2>(P, K, B, n, ld, pstride, nf, e0, e1, 1);
    run_like<32, 1, 1, 1, 4, 4>(P, K, B, n, ld, pstride, nf, e0, e1, 1);
    run_like<128, 1, 1, 1, 4, 4>(P, K, B, n, ld, pstride, nf, e0, e1, 1);
    return 0;
}
