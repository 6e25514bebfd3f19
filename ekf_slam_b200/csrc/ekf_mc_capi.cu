// ekf_mc_capi.cu -- C ABI of the Monte-Carlo filter batch (ekfb_* in include/ekfslam_b200.h) over k_mc_step.
#include "../../include/ekfslam_b200.h"
#include "ekf_kernels.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace ekf;

struct ekfb_handle {
    ekf_params p;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int nf = 0, max_obs = 0, max_table = 0, cap_n = 0, maxd = 1;
    int64_t ldp = 0, pstride = 0, sstride = 0;
    double *P = nullptr, *x = nullptr, *s = nullptr;
    int* N = nullptr;
    int* err = nullptr;
    double2 *Kg = nullptr, *Bg = nullptr;    // split step: gain slabs [nf][maxd][ldp]
    int* Mpend = nullptr;
    int dev_smem = 0;                        // opt-in shared memory per block of the device
    int front_kind = 1;                      // 1: k_mc_step<.,1>, one observation at a time (default, measured faster); EKF_MC_FRONT=run -> 0 (run-based k_mc_front)
    int flush_kind = 0;                      // 0: warp-per-column-group flush (default); EKF_MC_FLUSH=ldg -> 1 (tiled, gains in registers), =tma -> 2 (TMA-staged DMMA)
    int flush_tc = 64;                       // columns of one filter per CTA of the warp flush (EKF_MC_TC = 32 | 64 | 128)
    bool split = true;                       // front (latency) + flush (bandwidth) kernels; EKF_MC_FUSED=1 selects the one-kernel step
    double flush_ms = 0.0;
    int* h_err = nullptr;                    // pinned
    // early pose read-back (EKF_MC_EARLY_POSES=0 disables): the poses of a split step are final when the front kernel
    // ends, so ekfb_get_poses packs and copies them on a side stream behind `ev_front` while the flush still streams P
    cudaStream_t sx = nullptr;
    cudaEvent_t ev_front = nullptr;
    bool early = true, front_ev_valid = false;
    double* d_poses = nullptr;               // [nf][3] packed + 2 ints of the device error word behind them
    double* h_poses = nullptr;               // pinned mirror of d_poses
    // staging for host inputs (mem = 0)
    double *d_u = nullptr, *d_obs = nullptr, *d_loc = nullptr;
    int *d_counts = nullptr, *d_idx = nullptr, *d_ntab = nullptr, *d_act = nullptr;
    size_t smem = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    bool prof = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> evs;
    size_t ev_used = 0;
    double step_ms = 0.0;
    int64_t steps = 0;
    std::string errmsg;
};

static int32_t failb(ekfb_handle* h, int32_t code, const std::string& msg) {
    if (h) h->errmsg = msg;
    return code;
}
#define CKB(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return failb(h, EKF_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));       \
    } while (0)

extern "C" const char* ekfb_last_error(ekfb_handle* h) { return h ? h->errmsg.c_str() : "null handle"; }

extern "C" int32_t ekfb_destroy(ekfb_handle* h) {
    if (!h) return EKF_OK;
    cudaSetDevice(h->p.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (auto& e : h->evs) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    cudaFree(h->P); cudaFree(h->x); cudaFree(h->s); cudaFree(h->N); cudaFree(h->err);
    cudaFree(h->Kg); cudaFree(h->Bg); cudaFree(h->Mpend);
    cudaFree(h->d_u); cudaFree(h->d_obs); cudaFree(h->d_loc); cudaFree(h->d_counts); cudaFree(h->d_idx);
    cudaFree(h->d_ntab); cudaFree(h->d_act);
    if (h->h_err) cudaFreeHost(h->h_err);
    if (h->sx) { cudaStreamSynchronize(h->sx); cudaStreamDestroy(h->sx); }
    if (h->ev_front) cudaEventDestroy(h->ev_front);
    cudaFree(h->d_poses);
    if (h->h_poses) cudaFreeHost(h->h_poses);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return EKF_OK;
}

extern "C" int32_t ekfb_create(const ekf_params* p, int32_t n_filters, int32_t max_obs, int32_t max_table,
                               ekfb_handle** out) {
    if (!p || !out || n_filters < 1 || max_obs < 1 || max_table < 1 || p->capacity_landmarks < 1) return EKF_EINVAL;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= p->device) {
        fprintf(stderr, "ekfslam_b200: no CUDA device %d (this library has no CPU fallback)\n", p->device);
        return EKF_ECUDA;
    }
    ekfb_handle* h = new ekfb_handle();
    h->p = *p;
    h->nf = n_filters; h->max_obs = max_obs; h->max_table = max_table;
    int32_t rc = [&]() -> int32_t {
        CKB(cudaSetDevice(p->device));
        if (p->stream) h->stream = (cudaStream_t)(uintptr_t)p->stream;
        else { CKB(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)); h->own_stream = true; }
        h->cap_n = 3 + 2 * p->capacity_landmarks;
        h->ldp = ((int64_t)h->cap_n + 15) / 16 * 16;
        h->pstride = h->ldp * h->cap_n;
        h->sstride = ((int64_t)p->capacity_landmarks + 1 + 15) / 16 * 16;
        // gain slabs of <= maxd pending observations live in shared memory: 2 slabs x maxd x ldp x 16 B per CTA
        int maxd = p->defer_m < 1 ? 1 : std::min(p->defer_m, EKF_MAX_DEFER);
        if (!h->split) maxd = std::min(maxd, max_obs);     // fused: slabs cost shared memory; split: global memory
        if (const char* v = getenv("EKF_MC_FUSED")) h->split = atoi(v) == 0;
        if (const char* v = getenv("EKF_MC_FLUSH")) h->flush_kind = std::string(v) == "ldg" ? 1 : (std::string(v) == "tma" ? 2 : 0);
        if (const char* v = getenv("EKF_MC_TC")) h->flush_tc = atoi(v);
        if (const char* v = getenv("EKF_MC_FRONT")) h->front_kind = std::string(v) == "run" ? 0 : 1;
        int dev_smem = 0;
        CKB(cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device));
        h->dev_smem = dev_smem;
        if (!h->split) {
            while (maxd > 1 && (size_t)2 * maxd * h->ldp * 16 + 4096 > (size_t)dev_smem) --maxd;
            if ((size_t)2 * maxd * h->ldp * 16 + 4096 > (size_t)dev_smem)
                return failb(h, EKF_ECAPACITY, "ekfb_create: capacity_landmarks too large for the fused one-CTA-per-filter kernel "
                                                "(gain slabs exceed shared memory); unset EKF_MC_FUSED");
        }
        h->maxd = maxd;
        h->smem = (size_t)2 * maxd * h->ldp * 16;
        if (h->split) {
            CKB(cudaMalloc(&h->Kg, sizeof(double2) * (size_t)n_filters * maxd * h->ldp));
            CKB(cudaMalloc(&h->Bg, sizeof(double2) * (size_t)n_filters * maxd * h->ldp));
            CKB(cudaMalloc(&h->Mpend, sizeof(int) * n_filters));
            CKB(cudaMemset(h->Mpend, 0, sizeof(int) * n_filters));
            CKB(cudaMemset(h->Kg, 0, sizeof(double2) * (size_t)n_filters * maxd * h->ldp));   // rows >= n stay zero: padding rows of P never change
            CKB(cudaMemset(h->Bg, 0, sizeof(double2) * (size_t)n_filters * maxd * h->ldp));
        }
        CKB(cudaMalloc(&h->P, sizeof(double) * h->pstride * n_filters));
        CKB(cudaMalloc(&h->x, sizeof(double) * h->ldp * n_filters));
        CKB(cudaMalloc(&h->s, sizeof(double) * h->sstride * n_filters));
        CKB(cudaMalloc(&h->N, sizeof(int) * n_filters));
        CKB(cudaMalloc(&h->err, sizeof(int) * 2));
        CKB(cudaMallocHost(&h->h_err, sizeof(int) * 2));
        if (const char* v = getenv("EKF_MC_EARLY_POSES")) h->early = atoi(v) != 0;
        {   // highest priority, so that the pack kernel is scheduled beside a flush that still has CTAs waiting
            int pr_lo = 0, pr_hi = 0;
            CKB(cudaDeviceGetStreamPriorityRange(&pr_lo, &pr_hi));
            CKB(cudaStreamCreateWithPriority(&h->sx, cudaStreamNonBlocking, pr_hi));
        }
        CKB(cudaEventCreateWithFlags(&h->ev_front, cudaEventDisableTiming));
        CKB(cudaMalloc(&h->d_poses, sizeof(double) * (3 * (size_t)n_filters + 1)));
        CKB(cudaMallocHost(&h->h_poses, sizeof(double) * (3 * (size_t)n_filters + 1)));
        CKB(cudaMalloc(&h->d_u, sizeof(double) * 2 * n_filters));
        CKB(cudaMalloc(&h->d_obs, sizeof(double) * 3 * max_obs * n_filters));
        CKB(cudaMalloc(&h->d_loc, sizeof(double) * 2 * max_table * n_filters));
        CKB(cudaMalloc(&h->d_counts, sizeof(int) * n_filters));
        CKB(cudaMalloc(&h->d_idx, sizeof(int) * max_table * n_filters));
        CKB(cudaMalloc(&h->d_ntab, sizeof(int) * n_filters));
        CKB(cudaMalloc(&h->d_act, sizeof(int) * max_obs * n_filters));
        // x = [0 0 0], P = 0.1*eye(3), s = [] for every filter (EKF_SLAM.m:28-31)
        CKB(cudaMemsetAsync(h->P, 0, sizeof(double) * h->pstride * n_filters, h->stream));
        CKB(cudaMemsetAsync(h->x, 0, sizeof(double) * h->ldp * n_filters, h->stream));
        CKB(cudaMemsetAsync(h->s, 0, sizeof(double) * h->sstride * n_filters, h->stream));
        CKB(cudaMemsetAsync(h->N, 0, sizeof(int) * n_filters, h->stream));
        CKB(cudaMemsetAsync(h->err, 0, sizeof(int) * 2, h->stream));
        // diagonal of the 3x3 corner of every filter: one strided 2-D copy per diagonal element
        const std::vector<double> diag(n_filters, 0.1);
        for (int i = 0; i < 3; ++i)
            CKB(cudaMemcpy2DAsync(h->P + i + i * h->ldp, sizeof(double) * h->pstride, diag.data(), sizeof(double),
                                  sizeof(double), n_filters, cudaMemcpyHostToDevice, h->stream));
        CKB(cudaStreamSynchronize(h->stream));
        return EKF_OK;
    }();
    if (rc != EKF_OK) {
        fprintf(stderr, "ekfslam_b200: ekfb_create failed: %s\n", h->errmsg.c_str());
        ekfb_destroy(h);
        return rc;
    }
    *out = h;
    return EKF_OK;
}

static int32_t decode_err(ekfb_handle* h, int code, int f);
static int32_t check_err(ekfb_handle* h) {
    CKB(cudaMemcpyAsync(h->h_err, h->err, sizeof(int) * 2, cudaMemcpyDeviceToHost, h->stream));
    CKB(cudaStreamSynchronize(h->stream));
    if (h->h_err[0] == 0) return EKF_OK;
    const int code = h->h_err[0], f = h->h_err[1];
    CKB(cudaMemsetAsync(h->err, 0, sizeof(int) * 2, h->stream));
    return decode_err(h, code, f);
}

extern "C" int32_t ekfb_sync(ekfb_handle* h) {
    if (!h) return EKF_EINVAL;
    CKB(cudaSetDevice(h->p.device));
    return check_err(h);
}

extern "C" int32_t ekfb_step(ekfb_handle* h, int32_t mem, const double* u, const double* obs, const int32_t* counts,
                             const double* loc, const int32_t* idx, const int32_t* ntab, int32_t shared_table,
                             int32_t do_predict, int32_t do_measure, int32_t* actions_out) {
    if (!h) return EKF_EINVAL;
    if (!u || (do_measure && (!obs || !counts || !loc || !idx || !ntab))) return failb(h, EKF_EINVAL, "ekfb_step: null argument");
    CKB(cudaSetDevice(h->p.device));
    const int nf = h->nf, nt = shared_table ? 1 : nf;
    if (mem == 0) {                              // host buffers: stage them (inside the caller's timed region)
        CKB(cudaMemcpyAsync(h->d_u, u, sizeof(double) * 2 * nf, cudaMemcpyHostToDevice, h->stream));
        if (do_measure) {
            CKB(cudaMemcpyAsync(h->d_obs, obs, sizeof(double) * 3 * h->max_obs * nf, cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->d_counts, counts, sizeof(int) * nf, cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->d_loc, loc, sizeof(double) * 2 * h->max_table * nt, cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->d_idx, idx, sizeof(int) * h->max_table * nt, cudaMemcpyHostToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->d_ntab, ntab, sizeof(int) * nt, cudaMemcpyHostToDevice, h->stream));
        }
        u = h->d_u; obs = h->d_obs; counts = h->d_counts; loc = h->d_loc; idx = h->d_idx; ntab = h->d_ntab;
    }
    McArgs a;
    a.P = h->P; a.ldp = h->ldp; a.pstride = h->pstride;
    a.x = h->x; a.xstride = h->ldp;
    a.s = h->s; a.sstride = h->sstride;
    a.N = h->N; a.cap = h->p.capacity_landmarks;
    a.u = u; a.obs = obs; a.max_obs = h->max_obs; a.counts = counts;
    a.loc = loc; a.loc_stride = shared_table ? 0 : 2 * (int64_t)h->max_table;
    a.idx = idx; a.idx_stride = shared_table ? 0 : h->max_table;
    a.ntab = ntab; a.ntab_stride = shared_table ? 0 : 1;
    a.max_table = h->max_table;
    a.actions = actions_out ? h->d_act : nullptr;
    a.err = h->err;
    a.C = h->p.C; a.Rc0 = h->p.Rc[0]; a.Rc1 = h->p.Rc[1]; a.s_cost = h->p.s_cost; a.s_thresh = h->p.s_thresh;
    a.variant = h->p.variant; a.assoc_mode = h->p.assoc_mode; a.sym = h->p.storage == EKF_STORAGE_SYM ? 1 : 0;
    a.do_predict = do_predict; a.do_measure = do_measure; a.maxd = h->maxd;
    if (actions_out) CKB(cudaMemsetAsync(h->d_act, 0, sizeof(int) * h->max_obs * nf, h->stream));
    a.Kg = h->Kg; a.Bg = h->Bg; a.Mpend = h->Mpend;
    const size_t epp = 2;                                // event pairs per step: whole step, flush launch
    auto ev = [&](size_t i) -> cudaEvent_t* {            // event i of this step (0,1: whole step; 2,3: flush kernel)
        while (h->evs.size() < epp * (h->ev_used + 1)) {
            cudaEvent_t a0 = nullptr, a1 = nullptr;
            cudaEventCreate(&a0); cudaEventCreate(&a1);
            h->evs.push_back({a0, a1});
        }
        auto& pr = h->evs[epp * h->ev_used + i / 2];
        return (i & 1) ? &pr.second : &pr.first;
    };
    auto flush = [&](cudaStream_t st, const McArgs& s, int cnt_f) {
        const int mm = std::min(h->maxd, h->max_obs);
        if (h->flush_kind == 1) launch_mc_flush(st, s, cnt_f, h->cap_n, mm);
        else if (h->flush_kind == 2) launch_mc_flush_tma(st, s, cnt_f, h->cap_n, mm);
        else launch_mc_flush_w(st, s, cnt_f, h->cap_n, mm, h->flush_tc);
    };
    if (h->prof) CKB(cudaEventRecord(*ev(0), h->stream));
    if (h->split) {
        launch_mc_front(h->stream, a, nf, (size_t)h->dev_smem, h->front_kind);
        if (h->early) { CKB(cudaEventRecord(h->ev_front, h->stream)); h->front_ev_valid = true; }   // x, N, err are final here; the flush only touches P
        if (h->prof) CKB(cudaEventRecord(*ev(2), h->stream));
        if (do_measure) flush(h->stream, a, nf);
        if (h->prof) CKB(cudaEventRecord(*ev(3), h->stream));
    } else {
        launch_mc_step(h->stream, a, nf, 256, h->smem);
        h->front_ev_valid = false;
        if (h->prof) { CKB(cudaEventRecord(*ev(2), h->stream)); CKB(cudaEventRecord(*ev(3), h->stream)); }
    }
    if (h->prof) { CKB(cudaEventRecord(*ev(1), h->stream)); h->ev_used++; }
    h->steps += 1;
    CKB(cudaGetLastError());
    if (actions_out) {
        CKB(cudaMemcpyAsync(actions_out, h->d_act, sizeof(int) * h->max_obs * nf, cudaMemcpyDeviceToHost, h->stream));
        return check_err(h);
    }
    return EKF_OK;
}

extern "C" int32_t ekfb_profile(ekfb_handle* h, int32_t enable, double* step_ms_total, int64_t* steps,
                                double* flush_ms_total) {
    if (!h) return EKF_EINVAL;
    CKB(cudaStreamSynchronize(h->stream));
    const size_t epp = 2;
    for (size_t i = 0; i < h->ev_used; ++i) {
        float ms = 0.f;
        CKB(cudaEventElapsedTime(&ms, h->evs[epp * i].first, h->evs[epp * i].second));
        h->step_ms += ms;
        for (size_t c = 1; c < epp; ++c) {
            CKB(cudaEventElapsedTime(&ms, h->evs[epp * i + c].first, h->evs[epp * i + c].second));
            h->flush_ms += ms;
        }
    }
    h->ev_used = 0;
    if (step_ms_total) *step_ms_total = h->step_ms;
    if (flush_ms_total) *flush_ms_total = h->flush_ms;
    if (steps) *steps = h->steps;
    if (enable >= 0) { h->prof = enable != 0; h->step_ms = 0.0; h->flush_ms = 0.0; h->steps = 0; }
    if (h->prof)            // event pool created here, outside anybody's timed region
        while (h->evs.size() < 4096) {
            cudaEvent_t a0 = nullptr, a1 = nullptr;
            CKB(cudaEventCreate(&a0)); CKB(cudaEventCreate(&a1));
            h->evs.push_back({a0, a1});
        }
    return EKF_OK;
}

extern "C" int32_t ekfb_n_landmarks(ekfb_handle* h, int32_t* N) {
    if (!h || !N) return EKF_EINVAL;
    CKB(cudaSetDevice(h->p.device));
    CKB(cudaMemcpyAsync(N, h->N, sizeof(int) * h->nf, cudaMemcpyDeviceToHost, h->stream));
    return check_err(h);
}

static int32_t filter_N(ekfb_handle* h, int f, int* N) {
    if (f < 0 || f >= h->nf) return failb(h, EKF_EINVAL, "filter index out of range");
    CKB(cudaSetDevice(h->p.device));
    CKB(cudaMemcpyAsync(N, h->N + f, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    return check_err(h);
}

extern "C" int32_t ekfb_get_x(ekfb_handle* h, int32_t f, double* x, int32_t len) {
    if (!h || !x) return EKF_EINVAL;
    int N = 0;
    int32_t rc = filter_N(h, f, &N);
    if (rc != EKF_OK) return rc;
    if (len < 3 + 2 * N) return failb(h, EKF_EINVAL, "ekfb_get_x: buffer too small");
    CKB(cudaMemcpyAsync(x, h->x + (int64_t)f * h->ldp, sizeof(double) * (3 + 2 * N), cudaMemcpyDeviceToHost, h->stream));
    CKB(cudaStreamSynchronize(h->stream));
    return EKF_OK;
}

extern "C" int32_t ekfb_get_s(ekfb_handle* h, int32_t f, double* s, int32_t len) {
    if (!h) return EKF_EINVAL;
    int N = 0;
    int32_t rc = filter_N(h, f, &N);
    if (rc != EKF_OK) return rc;
    if (N == 0) return EKF_OK;
    if (!s || len < N) return failb(h, EKF_EINVAL, "ekfb_get_s: buffer too small");
    CKB(cudaMemcpyAsync(s, h->s + (int64_t)f * h->sstride, sizeof(double) * N, cudaMemcpyDeviceToHost, h->stream));
    CKB(cudaStreamSynchronize(h->stream));
    return EKF_OK;
}

extern "C" int32_t ekfb_get_P(ekfb_handle* h, int32_t f, double* P, int32_t lda) {
    if (!h || !P) return EKF_EINVAL;
    int N = 0;
    int32_t rc = filter_N(h, f, &N);
    if (rc != EKF_OK) return rc;
    const int n = 3 + 2 * N;
    if (lda < n) return failb(h, EKF_EINVAL, "ekfb_get_P: lda < n");
    CKB(cudaMemcpy2DAsync(P, sizeof(double) * lda, h->P + (int64_t)f * h->pstride, sizeof(double) * h->ldp,
                          sizeof(double) * n, n, cudaMemcpyDeviceToHost, h->stream));
    CKB(cudaStreamSynchronize(h->stream));
    if (h->p.storage == EKF_STORAGE_SYM)        // only i <= j is live on the device: mirror the caller's copy
        for (int j = 0; j < n; ++j)
            for (int i = j + 1; i < n; ++i) P[i + (size_t)j * lda] = P[j + (size_t)i * lda];
    return EKF_OK;
}

extern "C" int32_t ekfb_set_state(ekfb_handle* h, int32_t f, const double* x, const double* P, int32_t lda,
                                  const double* s, int32_t N) {
    if (h) h->front_ev_valid = false;      // x changes on the handle's stream from here on
    if (!h) return EKF_EINVAL;
    if (f < 0) {                                  // every filter gets the same state: upload once, replicate on the device
        int32_t rc = ekfb_set_state(h, 0, x, P, lda, s, N);
        if (rc != EKF_OK) return rc;
        for (int g = 1; g < h->nf; ++g) {
            CKB(cudaMemcpyAsync(h->P + (int64_t)g * h->pstride, h->P, sizeof(double) * h->pstride, cudaMemcpyDeviceToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->x + (int64_t)g * h->ldp, h->x, sizeof(double) * h->ldp, cudaMemcpyDeviceToDevice, h->stream));
            CKB(cudaMemcpyAsync(h->s + (int64_t)g * h->sstride, h->s, sizeof(double) * h->sstride, cudaMemcpyDeviceToDevice, h->stream));
        }
        const std::vector<int> Ns(h->nf, N);
        CKB(cudaMemcpyAsync(h->N, Ns.data(), sizeof(int) * h->nf, cudaMemcpyHostToDevice, h->stream));
        CKB(cudaStreamSynchronize(h->stream));
        return EKF_OK;
    }
    if (f >= h->nf) return failb(h, EKF_EINVAL, "filter index out of range");
    if (N < 0 || N > h->p.capacity_landmarks) return failb(h, EKF_ECAPACITY, "ekfb_set_state: N beyond capacity");
    const int n = 3 + 2 * N;
    if (!x || !P || lda < n || (N > 0 && !s)) return failb(h, EKF_EINVAL, "ekfb_set_state: bad argument");
    CKB(cudaSetDevice(h->p.device));
    CKB(cudaMemsetAsync(h->x + (int64_t)f * h->ldp, 0, sizeof(double) * h->ldp, h->stream));
    CKB(cudaMemcpyAsync(h->x + (int64_t)f * h->ldp, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    if (N > 0) CKB(cudaMemcpyAsync(h->s + (int64_t)f * h->sstride, s, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
    CKB(cudaMemcpy2DAsync(h->P + (int64_t)f * h->pstride, sizeof(double) * h->ldp, P, sizeof(double) * lda,
                          sizeof(double) * n, n, cudaMemcpyHostToDevice, h->stream));
    CKB(cudaMemcpyAsync(h->N + f, &N, sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CKB(cudaStreamSynchronize(h->stream));
    return EKF_OK;
}

// Robot poses x(1:3) of every filter, the per-step result a Monte-Carlo driver reads back: one small kernel packs the
// nf strided triples (and the device error word) into a contiguous buffer, one copy brings them to pinned memory.
// After a split step this runs on the side stream behind the front kernel's event, so the caller gets the poses while
// the flush of that step is still streaming P (it does not touch x); otherwise on the handle's stream.
static int32_t decode_err(ekfb_handle* h, int code, int f) {
    if (code == MC_ERR_CAPACITY)
        return failb(h, EKF_ECAPACITY, "filter " + std::to_string(f) + ": append beyond capacity_landmarks");
    if (code == MC_ERR_TABLE)
        return failb(h, EKF_EINVAL, "filter " + std::to_string(f) + ": measure: landmark index not in the table (EKF_SLAM.m:111,120 / EKF_SLAM_UC.m:123)");
    return failb(h, EKF_EINVAL, "filter " + std::to_string(f) + ": update slot outside 1..N (EKF_SLAM.m:125)");
}

extern "C" int32_t ekfb_get_poses(ekfb_handle* h, double* poses) {
    if (!h || !poses) return EKF_EINVAL;
    CKB(cudaSetDevice(h->p.device));
    cudaStream_t st = h->stream;
    if (h->early && h->front_ev_valid) {
        st = h->sx;
        CKB(cudaStreamWaitEvent(st, h->ev_front, 0));
    }
    launch_mc_pack_poses(st, h->x, h->ldp, h->nf, h->d_poses, h->err);
    CKB(cudaGetLastError());
    CKB(cudaMemcpyAsync(h->h_poses, h->d_poses, sizeof(double) * (3 * (size_t)h->nf + 1), cudaMemcpyDeviceToHost, st));
    CKB(cudaStreamSynchronize(st));
    memcpy(poses, h->h_poses, sizeof(double) * 3 * (size_t)h->nf);
    int e2[2];
    memcpy(e2, h->h_poses + 3 * (size_t)h->nf, sizeof(e2));
    if (e2[0] == 0) return EKF_OK;
    CKB(cudaMemsetAsync(h->err, 0, sizeof(int) * 2, h->stream));      // stream-ordered before the next step's front
    return decode_err(h, e2[0], e2[1]);
}

extern "C" int32_t ekfb_downdate_bytes(ekfb_handle* h, double* bytes) {
    if (!h || !bytes) return EKF_EINVAL;
    std::vector<int> N(h->nf);
    int32_t rc = ekfb_n_landmarks(h, N.data());
    if (rc != EKF_OK) return rc;
    double b = 0.0;
    for (int f = 0; f < h->nf; ++f) {
        const double n = 3.0 + 2.0 * N[f];
        b += h->p.storage == EKF_STORAGE_SYM ? 8.0 * n * (n + 1.0) : 16.0 * n * n;
    }
    *bytes = b;
    return EKF_OK;
}
