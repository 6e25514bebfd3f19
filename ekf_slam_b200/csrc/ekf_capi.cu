// ekf_capi.cu -- C ABI (include/ekfslam_b200.h) over the sm_100a kernels.  No CPU fallback: every entry point
// that computes enqueues CUDA kernels on the handle's stream; without a device ekf_create fails.
#include "../../include/ekfslam_b200.h"
#include "ekf_kernels.h"
#include "ekf_device.cuh"

#include <nccl.h>
#include <dlfcn.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <string>
#include <vector>
#include <unordered_map>

using namespace ekf;

struct ekf_handle {
    ekf_params p;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int N = 0;                 // landmarks in the state (host-tracked; appends are host-driven)
    int cap_n = 0;             // 3 + 2*capacity
    int64_t ld = 0;            // leading dimension of P (multiple of 16)
    double* P = nullptr;
    double* x = nullptr;
    double* s = nullptr;
    double2* Kd = nullptr;     // EKF_MAX_DEFER slabs of ld double2: gains K_q of the pending observations
    double2* Bd = nullptr;     // EKF_MAX_DEFER slabs: B_q = H_q P
    double* Bt = nullptr;      // (local columns) x EKF_BT_SROW: column-major copy of the pending B slabs for the hybrid flush kernels
    double* Bt2 = nullptr;     // continuous flush: the operand of the other pass of the queue (packed when its set is closed)
    // ---- column-slab sharding (world > 1): this rank stores columns [c_lo, c_hi) of P, FULL storage ----
    int rank = 0, world = 1;
    int c_lo = 0, c_hi = 0;
    // ---- single-process multi-device filter (ekf_create_multi): the FRONT handle owns one sub-handle per device and fans
    // every call out to them; a sub-handle is an ordinary slab handle whose peers live in the same process (direct peer
    // pointers after cudaDeviceEnablePeerAccess, no cudaIpc, no NCCL)
    std::vector<ekf_handle*> subs;
    bool in_process = false;   // sub-handle of a multi front: peers are plain pointers
    ncclComm_t comm = nullptr;
    double* Pc = nullptr;      // 2 x EKF_MAX_SLOTS x ld: exchanged columns of the current batch (double-buffered by batch parity)
    double* Pc_cur = nullptr;  // buffer holding the batch being consumed
    bool p2p = false;          // exchange by direct stores into cudaIpc-mapped peer memory (default) instead of NCCL
    PeerSet peers;             // every rank's Pc / flag words as seen from this process
    int* xflags = nullptr;     // [world] batch number last published by each rank
    unsigned int* xticket = nullptr;
    int epoch = 0;
    int aepoch = 0;            // epoch of the sharded mode-B association (candidate exchange)
    void* d_cand = nullptr;    // this rank's (cost, index) candidate + (NCCL fallback) the all-gathered ones: (1 + EKF_MAX_PEERS) x 16 B
    double xch_ms = 0.0;       // profile: exchange time (CUDA events)
    std::vector<cudaEvent_t> xev;
    size_t xev_used = 0;
    double2* Bc = nullptr;     // EKF_MAX_DEFER x EKF_MAX_SLOTS: B_q(:, col(slot))
    struct QObs { double z0, z1, R0, R1; int a; };
    std::vector<QObs> q;       // sharded: observations queued since the last flush (executed as one batch)
    SlotSet ss_cur;            // slot columns of the batch between exec_send and exec_finish
    bool lazy_now = false;     // that batch takes the lazy-flush path
    int m_pend = 0;            // pending (not yet flushed) rank-2 terms
    // ---- decoupled state chain (EKF_ASYNC_X=0 disables) ----
    // x, s and the mode-A association do not depend on the O(n^2) flush of P, so the state half of predict, the batched
    // signature association and the x read-back run on a side stream `sx`, ordered against the handle's stream by two
    // events: ev_mx = "every kernel enqueued so far on the main stream that reads or writes x / s", ev_pp = "the covariance
    // half of the last predict" (it reads xprev, which the next state half overwrites), ev_px = "the state half of the
    // last predict".  A caller that reads x after measure() gets it when the gains are
    // done, while the flush of that measure still streams P; its next predict/measure calls queue up behind the flush.
    bool async_x = false;
    cudaStream_t sx = nullptr;
    cudaEvent_t ev_mx = nullptr, ev_px = nullptr, ev_pp = nullptr;
    double* xprev = nullptr;   // pre-motion pose saved by the state half of predict for the covariance half
    double* x_pin = nullptr;   // pinned staging of x for ekf_get_x (a copy into pageable memory goes through the driver's bounce buffers)
    bool x_host = false;       // x_pin holds the device's x, except (x_head_dirty) the pose x(1:3) that a predict has moved since:
    bool x_head_dirty = false; // the read-back after predict then needs 24 bytes instead of 8 n (single-GPU handles) -- and those the
    double* pose_host = nullptr;   // state half of predict has already written into mapped pinned memory (pose_dev = its device alias),
    double* pose_dev = nullptr;    // so the read waits for that kernel's event and copies nothing
    bool pose_fresh = false;
    double* x_pin_dev = nullptr;   // device alias of x_pin (mapped): k_batch_gain writes x there as well
    bool xpin_fresh = false;       // x_pin is the device's x once ev_mx has fired (the last writer of x was such a gain kernel)
    int defer = 1;             // flush when m_pend reaches this (1 = immediate); also the chunk size of the fused batch (<= 8)
    // cross-step deferral (single GPU, defer_m = 9..EKF_MAX_PEND): terms stay pending across measure() and predict();
    // the flush runs when the next chunk would not fit (or at a getter / append / mode-B association).  predict then
    // transforms the pending slabs (K_q <- F K_q, B_q <- B_q F') instead of flushing.  Not bit-identical to the immediate
    // path (F is applied to P0 and to the terms separately), but the same estimator to rounding: parity <= 1e-9.
    int pend_cap = 1;          // = defer unless cross-step
    bool xstep = false;
    // lazy flush (cross-step deferral, SYM storage; EKF_LAZY_FLUSH=0 disables): a full set of pending terms is not flushed
    // when it fills up but when the NEXT batch arrives: that batch's slot columns are copied out of P0 first (local gather /
    // sharded exchange), then the K3 pass runs on the main stream while the batch is linearised on the side stream `sg`
    // from those copies + the pending terms, writing its gains into the other pair of slabs (Kd2, Bd2).  Same operands,
    // same order as flush-then-linearise: bit-identical, but k_batch_lin / k_batch_gain leave the critical path.
    bool lazy = false;
    double2* Kd2 = nullptr;
    double2* Bd2 = nullptr;
    cudaStream_t sg = nullptr;
    cudaEvent_t ev_g1 = nullptr, ev_g2 = nullptr;
    // continuous flush ("shadow"; single GPU, SYM storage, cross-step deferral, n >= shadow_min_n; EKF_SHADOW=0 disables):
    // the K3 pass of a full set of terms runs OUT OF PLACE on its own low-priority stream `sf` (P -> P2, from frozen copies
    // Kf / Bf of the set), while predict and the following batches carry on from the untouched P on the handle's stream: the
    // gather of a batch's slot columns reads P + the robot strip (rows 0..2, which predict rewrites; ekf_kernels.cu) and the
    // pending-aware linearisation / gain kernels correct them by the whole window [0, m_pend) = the mA terms of the pass in
    // flight + the mQ terms of the pass queued behind it (their copies in the window keep following predict's F) + the set
    // that is filling up.  A set is closed the moment it is full: its pass is enqueued on `sf` at once -- behind the running
    // one, from that one's output back into the buffer the chain is still reading, which is why the chain then has to retire
    // the running pass (wait for it, let the strip drop its terms, swap P and P2) before it may gather again.  The K3 kernel
    // so runs back to back, and every latency-bound kernel, and the host's turnaround between calls, hides beside it.
    bool shadow = false;
    int shadow_min_n = 0;      // state size from which the mode is used (below it the pass is too short to hide anything)
    double* P2 = nullptr;
    double* strip = nullptr;   // 4 x ld doubles
    bool strip_valid = false;
    // The window of pending terms is [Kd, Kd + m_pend) with Kd / Bd moving through a larger allocation (Kd0 / Bd0, EKF_WIN_SLOTS
    // slabs): retiring a pass advances the base past its terms instead of moving the rest down, and a pass reads its set
    // where it lies -- predict keeps rewriting rows 0, 1 of those slabs (K_q <- F K_q, B_q <- B_q F'), which only ever reach
    // rows / columns 0-1 of the pass's output, the part of the matrix the strip stands in for.  Compacted to the start
    // (with nothing in flight) when the end of the allocation comes near: once in ~13 passes.
    double2* Kd0 = nullptr;
    double2* Bd0 = nullptr;
    int kw0 = 0;               // slabs between Kd0 and Kd
    cudaStream_t sf = nullptr;
    cudaEvent_t ev_fz = nullptr, ev_fd[2] = {nullptr, nullptr};
    int nfl = 0;               // passes outstanding on sf: 0, 1 (running) or 2 (running + queued)
    int fcur = 0;              // ev_fd / Kf index of the running pass
    int mA = 0;                // terms of the running pass (window slots [0, mA))
    int mQ = 0;                // terms of the queued pass (window slots [mA, mA + mQ))
    int dd_variant = -1;       // downdate kernel variant (-1 = library default; EKF_DOWNDATE_VARIANT env at create)
    Lin* lin = nullptr;
    BatchLin* blin = nullptr;  // scratch of the fused batch linearisation (ekf_update_batch)
    void* partial = nullptr;
    unsigned int* ticket = nullptr;
    AssocResult* d_res = nullptr;
    AssocResult* h_res = nullptr;   // pinned
    // ---- measure() decided on the device (EKF_SLAM_UC, mode A; EKF_ASYNC_MEASURE=0 disables): ekf_measure_batch_async
    // uploads observed_LL, lets k_measure_decide pick the slots, enqueues the update kernels behind it and returns without
    // waiting; the actions reach the caller at the next call on the handle (or ekf_measure_wait).  The host takes this path
    // only when every row's signature is already in the map -- `sigs` mirrors the signatures it has appended / uploaded --
    // so that no row can need the landmark table (an append), which only exists for the duration of the call.
    std::unordered_map<uint64_t, int> sigs;
    bool am_enable = false, am_pending = false;
    int am_k = 0;
    int32_t* am_act_out = nullptr;
    double* am_obs_h = nullptr;      // pinned staging of observed_LL (k x 3)
    double* am_obs_d = nullptr;
    BatchObs* am_chunks = nullptr;   // EKF_ASSOC_BATCH / EKF_BATCH_MAX batch descriptions
    int* am_skip = nullptr;
    int* am_status = nullptr;        // mapped pinned: [0] = 1 if the device found a row it could not serve, [1 + i] = action of row i
    int* am_status_d = nullptr;
    int* am_scratch = nullptr;
    unsigned int* am_ticket = nullptr;
    cudaEvent_t ev_dec = nullptr;
    // sticky device->host status words in mapped pinned memory (kernels write them only when something is wrong; the host
    // looks at them after every synchronisation): [0] non-finite innovation covariance, [1] peer exchange timed out
    int* hflags = nullptr;
    int* dflags = nullptr;     // device alias of hflags
    long long xch_timeout = 0; // clock cycles k_wait_flags spins before it gives up (EKF_XCHG_TIMEOUT_MS, default 20 s)
    bool prof = false;
    std::vector<cudaEvent_t> cev;      // EKF_CHAIN_TIMING (debug)
    size_t cev_used = 0;
    std::vector<cudaEvent_t> ev;
    size_t ev_used = 0;
    double dd_ms = 0.0;
    int64_t dd_launches = 0;
    int64_t dd_terms = 0;      // rank-2 terms (observations) applied by those launches
    int64_t launches = 0;
    std::string err;
};

static int32_t fail(ekf_handle* h, int32_t code, const std::string& msg) {
    if (h) h->err = msg;
    return code;
}
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(h, EKF_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e__));        \
    } while (0)
#define CKH(h)                                                                                     \
    do { if (!(h)) return EKF_EINVAL; } while (0)

static int32_t flush(ekf_handle* h);
static int32_t shadow_strip_ready(ekf_handle* h);
static int32_t shadow_retire(ekf_handle* h);
static int32_t shadow_close(ekf_handle* h);
static int32_t shadow_settle(ekf_handle* h);
static int32_t fused_chunk(ekf_handle* h, const BatchObs& ob, ObSrc src);
static int32_t exec_sharded(ekf_handle* h);
static int32_t downdate_sharded(ekf_handle* h);
static int32_t front_exec(ekf_handle* f);
static int32_t front_flush(ekf_handle* f);
#define FLUSH(h) do { int32_t rc__ = flush(h); if (rc__ != EKF_OK) return rc__; } while (0)

// Called after every stream synchronisation: surface what the kernels flagged since the last one.
static int32_t after_sync(ekf_handle* h) {
    if (!h->hflags) return EKF_OK;
    if (h->hflags[1])       // sticky: the handle is unusable once a peer stopped answering
        return fail(h, EKF_ENCCL, "sharded exchange: a peer rank did not publish its columns in time (EKF_XCHG_TIMEOUT_MS)");
    if (h->hflags[0]) {
        h->hflags[0] = 0;   // reported once; x and P keep the non-finite values, as the reference's would
        return fail(h, EKF_ENONFINITE, "non-finite innovation covariance in an update since the last synchronisation "
                                       "(EKF_SLAM.m:137-143: q = 0 when the robot stands on the observed landmark); x and P are poisoned");
    }
    return EKF_OK;
}

static inline uint64_t sig_key(double v) { uint64_t k; if (v == 0.0) v = 0.0; memcpy(&k, &v, 8); return k; }
static void sigs_reset(ekf_handle* h, const double* s, int N) {
    h->sigs.clear();
    for (int i = 0; i < N; ++i) h->sigs[sig_key(s[i])] += 1;
}

// Decisions of the last ekf_measure_batch_async: wait for the decision kernel (NOT for the updates or the flush behind
// it), hand the actions to the caller's buffer.  Called at the top of every entry point; usually the kernel is long done.
static int32_t resolve_measure(ekf_handle* h) {
    if (!h->am_pending) return EKF_OK;
    h->am_pending = false;
    CK(cudaEventSynchronize(h->ev_dec));
    if (h->am_act_out)
        for (int i = 0; i < h->am_k; ++i) h->am_act_out[i] = h->am_status[1 + i];
    h->am_act_out = nullptr;
    if (h->am_status[0])        // cannot happen while the signature mirror is right; the batch contributed nothing
        return fail(h, EKF_EINVAL, "measure: the device found a row whose signature is not in the map although the host's "
                                   "signature mirror said otherwise; that measure() was not applied");
    return EKF_OK;
}
#define RESOLVE(h) do { if ((h)->am_pending) { int32_t rc__ = resolve_measure(h); if (rc__ != EKF_OK) return rc__; } } while (0)

static inline int32_t mark_x(ekf_handle* h) {         // call after enqueuing main-stream work that touches x / s
    h->x_host = false;
    h->pose_fresh = false;
    h->xpin_fresh = false;
    if (h->async_x) CK(cudaEventRecord(h->ev_mx, h->stream));
    return EKF_OK;
}
#define MARK_X(h) do { int32_t rc__ = mark_x(h); if (rc__ != EKF_OK) return rc__; } while (0)

static inline int n_of(const ekf_handle* h) { return 3 + 2 * h->N; }
static inline int sym_of(const ekf_handle* h) { return h->p.storage == EKF_STORAGE_SYM ? 1 : 0; }
static inline bool shadow_on(const ekf_handle* h) { return h->shadow && n_of(h) >= h->shadow_min_n; }   // continuous flush in use at this size

// ---- NCCL is bound lazily (dlopen) so that the single-GPU library loads anywhere libcudart does ----
struct NcclApi {
    void* so = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static bool nccl_load(std::string& err) {
    if (g_nccl.so) return true;
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!so) { err = std::string("dlopen libnccl.so.2: ") + dlerror(); return false; }
#define SYM(field, name) \
    *(void**)(&g_nccl.field) = dlsym(so, name); \
    if (!g_nccl.field) { err = std::string("dlsym ") + name; return false; }
    SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
    SYM(Broadcast, "ncclBroadcast") SYM(AllReduce, "ncclAllReduce") SYM(AllGather, "ncclAllGather") SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.so = so;
    return true;
}
#define CKN(call)                                                                                  \
    do {                                                                                           \
        ncclResult_t r__ = (call);                                                                 \
        if (r__ != ncclSuccess)                                                                    \
            return fail(h, EKF_ENCCL, std::string(#call) + ": " + g_nccl.GetErrorString(r__));     \
    } while (0)

// Slab partition: landmark-aligned contiguous column ranges, balanced on capacity.  Rank 0 also owns the robot
// columns 0..2; a landmark's two columns never straddle a boundary.
// FULL storage: equal column counts.  SYM storage: column j has j+1 live rows, so equal live area puts the
// boundaries at capN * sqrt(g / world).
static inline int slab_first_landmark(int capN, int world, int g, int storage) {
    if (g <= 0) return 0;
    if (g >= world) return capN;
    if (storage == EKF_STORAGE_SYM) {
        int l = (int)llround((double)capN * sqrt((double)g / (double)world));
        return std::min(std::max(l, g), capN - (world - g));      // every rank keeps at least one landmark
    }
    return (int)((int64_t)g * capN / world);
}
extern "C" int32_t ekf_slab_range(int32_t capacity_landmarks, int32_t world, int32_t rank, int32_t storage,
                                  int32_t* c_lo, int32_t* c_hi) {
    if (capacity_landmarks < 1 || world < 1 || rank < 0 || rank >= world || !c_lo || !c_hi) return EKF_EINVAL;
    if (capacity_landmarks < world) return EKF_EINVAL;
    *c_lo = rank == 0 ? 0 : 3 + 2 * slab_first_landmark(capacity_landmarks, world, rank, storage);
    *c_hi = 3 + 2 * slab_first_landmark(capacity_landmarks, world, rank + 1, storage);
    return EKF_OK;
}
static inline int owner_of_col(const ekf_handle* h, int col) {
    if (col < 3) return 0;
    const int l = (col - 3) / 2;
    for (int g = h->world - 1; g > 0; --g)
        if (l >= slab_first_landmark(h->p.capacity_landmarks, h->world, g, h->p.storage)) return g;
    return 0;
}
static inline int ncols_loc(const ekf_handle* h) {      // active columns of the local slab
    int hi = std::min(h->c_hi, 3 + 2 * h->N);
    return hi > h->c_lo ? hi - h->c_lo : 0;
}

extern "C" int32_t ekf_nccl_unique_id(uint8_t* id128) {
    std::string err;
    if (!id128) return EKF_EINVAL;
    if (!nccl_load(err)) { fprintf(stderr, "ekfslam_b200: %s\n", err.c_str()); return EKF_ENCCL; }
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != ncclSuccess) return EKF_ENCCL;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    memcpy(id128, &id, 128);
    return EKF_OK;
}

extern "C" const char* ekf_version(void) { return "ekfslam_b200 0.1 (sm_100a)"; }

extern "C" int32_t ekf_params_default(ekf_params* p, int32_t variant) {
    if (!p) return EKF_EINVAL;
    memset(p, 0, sizeof(*p));
    p->C = 0.2;                                         // EKF_SLAM.m:12
    p->Rc[0] = (variant == EKF_VARIANT_UC) ? 0.1 : 0.01;   // EKF_SLAM_UC.m:13 / EKF_SLAM.m:13
    p->Rc[1] = 5.0;
    p->s_cost = 0.00000000001;                          // EKF_SLAM.m:14, EKF_SLAM_UC.m:16
    p->s_thresh = 1000000000.0;                         // EKF_SLAM.m:16
    p->variant = variant;
    p->assoc_mode = EKF_ASSOC_SIGNATURE;                // Correspondence.m:75
    p->storage = EKF_STORAGE_FULL;
    p->capacity_landmarks = 1024;
    p->device = 0;
    p->defer_m = 0;
    p->stream = 0;
    return EKF_OK;
}

extern "C" int32_t ekf_destroy(ekf_handle* h) {
    if (!h) return EKF_OK;
    if (!h->subs.empty()) {
        for (ekf_handle* sh : h->subs) { cudaSetDevice(sh->p.device); if (sh->stream) cudaStreamSynchronize(sh->stream); }
        for (ekf_handle* sh : h->subs) ekf_destroy(sh);
        delete h;
        return EKF_OK;
    }
    cudaSetDevice(h->p.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->sf) cudaStreamSynchronize(h->sf);          // a K3 pass of the continuous flush may still be running
    for (auto e : h->ev) cudaEventDestroy(e);
    for (auto e : h->cev) cudaEventDestroy(e);
    cudaFree(h->P); cudaFree(h->x); cudaFree(h->s); cudaFree(h->Kd0 ? h->Kd0 : h->Kd); cudaFree(h->Bd0 ? h->Bd0 : h->Bd); cudaFree(h->Bt); cudaFree(h->Bt2);
    cudaFree(h->Kd2); cudaFree(h->Bd2);
    if (h->sf) { cudaStreamSynchronize(h->sf); cudaStreamDestroy(h->sf); }
    if (h->ev_fz) cudaEventDestroy(h->ev_fz);
    for (int i = 0; i < 2; ++i) if (h->ev_fd[i]) cudaEventDestroy(h->ev_fd[i]);
    cudaFree(h->P2); cudaFree(h->strip);
    if (h->sg) { cudaStreamSynchronize(h->sg); cudaStreamDestroy(h->sg); }
    if (h->ev_g1) cudaEventDestroy(h->ev_g1);
    if (h->ev_g2) cudaEventDestroy(h->ev_g2);
    for (int g = 0; g < h->world && h->p2p && !h->in_process; ++g)
        if (g != h->rank) { if (h->peers.pc[g]) cudaIpcCloseMemHandle(h->peers.pc[g]); if (h->peers.flags[g]) cudaIpcCloseMemHandle(h->peers.flags[g]); }
    for (auto e : h->xev) cudaEventDestroy(e);
    cudaFree(h->Pc); cudaFree(h->Bc); cudaFree(h->xflags); cudaFree(h->xticket); cudaFree(h->d_cand);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    cudaFree(h->lin); cudaFree(h->blin); cudaFree(h->partial); cudaFree(h->ticket); cudaFree(h->d_res);
    if (h->h_res) cudaFreeHost(h->h_res);
    if (h->x_pin) cudaFreeHost(h->x_pin);
    if (h->pose_host) cudaFreeHost(h->pose_host);
    if (h->hflags) cudaFreeHost(h->hflags);
    if (h->am_obs_h) cudaFreeHost(h->am_obs_h);
    if (h->am_status) cudaFreeHost(h->am_status);
    cudaFree(h->am_obs_d); cudaFree(h->am_chunks); cudaFree(h->am_skip); cudaFree(h->am_scratch); cudaFree(h->am_ticket);
    if (h->ev_dec) cudaEventDestroy(h->ev_dec);
    if (h->sx) { cudaStreamSynchronize(h->sx); cudaStreamDestroy(h->sx); }
    if (h->ev_mx) cudaEventDestroy(h->ev_mx);
    if (h->ev_px) cudaEventDestroy(h->ev_px);
    if (h->ev_pp) cudaEventDestroy(h->ev_pp);
    cudaFree(h->xprev);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return EKF_OK;
}

static int32_t init_state(ekf_handle* h) {
    if (h->Kd0) { h->Kd = h->Kd0; h->Bd = h->Bd0; h->kw0 = 0; }
    // x = [0 0 0], P = 0.1*eye(3), s = []   (EKF_SLAM.m:28-31)
    CK(cudaMemsetAsync(h->P, 0, sizeof(double) * h->ld * (h->c_hi - h->c_lo), h->stream));
    CK(cudaMemsetAsync(h->x, 0, sizeof(double) * h->ld, h->stream));
    CK(cudaMemsetAsync(h->s, 0, sizeof(double) * (h->p.capacity_landmarks + 1), h->stream));
    CK(cudaMemsetAsync(h->Kd, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
    CK(cudaMemsetAsync(h->Bd, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
    const double d = 0.1;
    for (int i = 0; i < 3 && h->rank == 0; ++i)
        CK(cudaMemcpyAsync(h->P + i + i * h->ld, &d, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    h->N = 0;
    h->m_pend = 0;
    h->mA = h->mQ = 0;
    h->nfl = 0;
    h->strip_valid = false;
    h->sigs.clear();
    return EKF_OK;
}

// Map every rank's Pc buffer and flag words into this process (cudaIpc), handles all-gathered over NCCL once at
// creation.  If any rank cannot map a peer (no P2P path), every rank falls back to the NCCL exchange together.
static int32_t setup_p2p(ekf_handle* h) {
    const int W = h->world;
    CK(cudaMalloc(&h->xflags, sizeof(int) * EKF_XF_INTS));
    CK(cudaMemset(h->xflags, 0, sizeof(int) * EKF_XF_INTS));
    CK(cudaMalloc(&h->xticket, sizeof(unsigned int)));
    CK(cudaMemset(h->xticket, 0, sizeof(unsigned int)));
    struct Pack { cudaIpcMemHandle_t pc, fl; int ok; int pad[15]; };
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    Pack mine;
    memset(&mine, 0, sizeof(mine));
    mine.ok = cudaIpcGetMemHandle(&mine.pc, h->Pc) == cudaSuccess && cudaIpcGetMemHandle(&mine.fl, h->xflags) == cudaSuccess;
    cudaGetLastError();
    Pack* d_all = nullptr;
    CK(cudaMalloc(&d_all, sizeof(Pack) * (W + 1)));
    CK(cudaMemcpy(d_all + W, &mine, sizeof(Pack), cudaMemcpyHostToDevice));
    CKN(g_nccl.AllGather(d_all + W, d_all, sizeof(Pack), ncclChar, h->comm, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<Pack> all(W);
    CK(cudaMemcpy(all.data(), d_all, sizeof(Pack) * W, cudaMemcpyDeviceToHost));
    int ok = 1;
    for (int g = 0; g < W; ++g) ok &= all[g].ok;
    memset(&h->peers, 0, sizeof(h->peers));
    h->peers.world = W; h->peers.rank = h->rank;
    for (int g = 0; g < W && ok; ++g) {
        if (g == h->rank) { h->peers.pc[g] = h->Pc; h->peers.flags[g] = h->xflags; continue; }
        void *ppc = nullptr, *pfl = nullptr;
        if (cudaIpcOpenMemHandle(&ppc, all[g].pc, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess ||
            cudaIpcOpenMemHandle(&pfl, all[g].fl, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); break; }
        h->peers.pc[g] = (double*)ppc; h->peers.flags[g] = (int*)pfl;
    }
    // agree: P2P only if every rank mapped every peer
    int* d_ok = (int*)d_all;
    CK(cudaMemcpy(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice));
    CKN(g_nccl.AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, h->comm, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(d_all);
    h->p2p = ok != 0;
    return EKF_OK;
}

static int32_t create_common(const ekf_params* p, int rank, int world, const uint8_t* id128, ekf_handle** out,
                             bool in_process = false) {
    if (!p || !out || p->capacity_landmarks < 1) return EKF_EINVAL;
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world || (world > 1 && !id128 && !in_process)) return EKF_EINVAL;
    if (world > 1 && p->capacity_landmarks < world) {
        fprintf(stderr, "ekfslam_b200: sharded filters need capacity_landmarks >= world\n");
        return EKF_EINVAL;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= p->device) {
        fprintf(stderr, "ekfslam_b200: no CUDA device %d (this library has no CPU fallback)\n", p->device);
        return EKF_ECUDA;
    }
    ekf_handle* h = new ekf_handle();
    h->p = *p;
    h->rank = rank;
    h->world = world;
    h->in_process = in_process;
    h->defer = p->defer_m < 1 ? 1 : (p->defer_m > EKF_MAX_DEFER ? EKF_MAX_DEFER : p->defer_m);
    h->pend_cap = h->defer;
    // cross-step deferral: single GPU, or SYM column slabs (every rank holds all of K_q and B_q = (P H')', so the exchanged
    // P0 columns can be corrected by the pending terms on arrival); FULL slabs keep B_q only where its column lives: <= 8
    if ((world == 1 || p->storage == EKF_STORAGE_SYM) && p->defer_m > EKF_MAX_DEFER) {
        h->pend_cap = p->defer_m > EKF_MAX_PEND ? EKF_MAX_PEND : p->defer_m;
        h->xstep = true;
    }
    if (const char* v = getenv("EKF_DOWNDATE_VARIANT")) h->dd_variant = atoi(v);
    int32_t rc = [&]() -> int32_t {
        CK(cudaSetDevice(p->device));
        h->cap_n = 3 + 2 * p->capacity_landmarks;
        h->ld = ((int64_t)h->cap_n + 15) / 16 * 16;
        {
            // continuous flush: needs a second P (checked against the free memory), a stream of its own making (the chain
            // must outrank the K3 pass in the block scheduler) and the decoupled state chain
            const char* sh = getenv("EKF_SHADOW");
            const char* ax = getenv("EKF_ASYNC_X");
            const char* lz = getenv("EKF_LAZY_FLUSH");
            // On slabs the mode pays only while a slab's pass is longer than the per-batch chain it hides (exchange, linearise,
            // gains: ~0.06 ms alone, more beside a pass).  Measured at 10k landmarks: 2 GPUs 40.5k -> 44.3k updates/s, 8 GPUs
            // 63.4k -> 57.6k.  Default: on when a rank's live share of P is >= 0.6 GB (10k: 2 GPUs yes, 4 and 8 no) -- a number
            // every rank computes alike; EKF_SHADOW_SLABS=0 / 1 overrides.
            const char* ss = getenv("EKF_SHADOW_SLABS");
            const double live_per_rank = 4.0 * (double)h->cap_n * (double)(h->cap_n + 1) / (double)world;
            const bool slabs_ok = world == 1 || (ss ? atoi(ss) != 0 : live_per_rank >= 0.6e9);
            h->shadow = h->xstep && p->storage == EKF_STORAGE_SYM && slabs_ok &&
                        !(sh && atoi(sh) == 0) && !(ax && atoi(ax) == 0) && !(lz && atoi(lz) != 0);
            if (h->shadow) {
                size_t fr = 0, tot = 0;
                CK(cudaMemGetInfo(&fr, &tot));
                if (world == 1) {
                    const size_t need = 2 * sizeof(double) * (size_t)h->ld * (size_t)h->cap_n + ((size_t)2 << 30);
                    if (fr < need) h->shadow = false;
                } else {
                    // slabs: every rank must take the same decision (x and the gains are replicated: the modes differ in
                    // rounding), so it is made from numbers all ranks share: the largest slab (rank 0's) twice, plus
                    // margin, within 60 % of the device's total memory.  10k landmarks: yes; 80k: no (stays with the lazy flush)
                    int lo0 = 0, hi0 = 0;
                    ekf_slab_range(p->capacity_landmarks, world, 0, p->storage, &lo0, &hi0);
                    const double need = 2.0 * sizeof(double) * (double)h->ld * (double)(hi0 - lo0) + 4.0 * (double)((size_t)1 << 30);
                    if (need > 0.6 * (double)tot) h->shadow = false;
                }
            }
        }
        int pr_lo = 0, pr_hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&pr_lo, &pr_hi));
        if (p->stream) { h->stream = (cudaStream_t)(uintptr_t)p->stream; }      // continuous flush: give it a priority above the lowest, or nothing overlaps
        else if (h->shadow) { CK(cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, pr_hi)); h->own_stream = true; }
        else { CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)); h->own_stream = true; }
        if (h->shadow) {
            CK(cudaStreamCreateWithPriority(&h->sf, cudaStreamNonBlocking, pr_lo));
            CK(cudaEventCreateWithFlags(&h->ev_fz, cudaEventDisableTiming));
            for (int i = 0; i < 2; ++i) CK(cudaEventCreateWithFlags(&h->ev_fd[i], cudaEventDisableTiming));
            const char* mn = getenv("EKF_SHADOW_MIN_N");
            h->shadow_min_n = mn ? atoi(mn) : 8000;       // below ~4000 landmarks a K3 pass is shorter than the kernels it would hide
        }
        ekf_slab_range(p->capacity_landmarks, world, rank, p->storage, &h->c_lo, &h->c_hi);
        CK(cudaMalloc(&h->P, sizeof(double) * h->ld * (h->c_hi - h->c_lo)));
        if (h->shadow) {
            CK(cudaMalloc(&h->P2, sizeof(double) * h->ld * (h->c_hi - h->c_lo)));
            CK(cudaMemset(h->P2, 0, sizeof(double) * h->ld * (h->c_hi - h->c_lo)));
            CK(cudaMalloc(&h->strip, sizeof(double) * 4 * h->ld));
            CK(cudaMemset(h->strip, 0, sizeof(double) * 4 * h->ld));
            if (world == 1) { CK(cudaMalloc(&h->Pc, sizeof(double) * h->ld * EKF_MAX_SLOTS)); h->Pc_cur = h->Pc; }
        }
        if (world > 1 && in_process) {       // peers are wired by ekf_create_multi once every sub-handle exists
            CK(cudaMalloc(&h->Pc, sizeof(double) * h->ld * EKF_MAX_SLOTS * 2));
            h->Pc_cur = h->Pc;
            CK(cudaMalloc(&h->Bc, sizeof(double2) * EKF_MAX_DEFER * EKF_MAX_SLOTS));
            CK(cudaMalloc(&h->xflags, sizeof(int) * EKF_XF_INTS));
            CK(cudaMemset(h->xflags, 0, sizeof(int) * EKF_XF_INTS));
            CK(cudaMalloc(&h->xticket, sizeof(unsigned int)));
            CK(cudaMemset(h->xticket, 0, sizeof(unsigned int)));
        } else if (world > 1) {
            if (!nccl_load(h->err)) return EKF_ENCCL;
            CK(cudaMalloc(&h->Pc, sizeof(double) * h->ld * EKF_MAX_SLOTS * 2));
            h->Pc_cur = h->Pc;
            CK(cudaMalloc(&h->Bc, sizeof(double2) * EKF_MAX_DEFER * EKF_MAX_SLOTS));
            ncclUniqueId id;
            memcpy(&id, id128, 128);
            CKN(g_nccl.CommInitRank(&h->comm, world, id, rank));
            const char* xe = getenv("EKF_SHARD_EXCHANGE");
            if (!(xe && std::string(xe) == "nccl") && world <= EKF_MAX_PEERS) {
                int32_t rp = setup_p2p(h);
                if (rp != EKF_OK) return rp;
            }
        }
        CK(cudaMalloc(&h->x, sizeof(double) * h->ld));
        CK(cudaHostAlloc(&h->x_pin, sizeof(double) * h->ld, cudaHostAllocMapped | cudaHostAllocPortable));
        CK(cudaHostGetDevicePointer(&h->x_pin_dev, h->x_pin, 0));
        CK(cudaHostAlloc(&h->pose_host, 4 * sizeof(double), cudaHostAllocMapped | cudaHostAllocPortable));
        CK(cudaHostGetDevicePointer(&h->pose_dev, h->pose_host, 0));
        CK(cudaMalloc(&h->s, sizeof(double) * (p->capacity_landmarks + 1)));
        CK(cudaMalloc(&h->Kd, sizeof(double2) * h->ld * (h->shadow ? EKF_WIN_SLOTS : EKF_MAX_PEND)));
        CK(cudaMalloc(&h->Bd, sizeof(double2) * h->ld * (h->shadow ? EKF_WIN_SLOTS : EKF_MAX_PEND)));
        if (h->shadow) {
            h->Kd0 = h->Kd; h->Bd0 = h->Bd;
            CK(cudaMemset(h->Kd0, 0, sizeof(double2) * h->ld * EKF_WIN_SLOTS));
            CK(cudaMemset(h->Bd0, 0, sizeof(double2) * h->ld * EKF_WIN_SLOTS));
        }
        CK(cudaMalloc(&h->Bt, sizeof(double) * EKF_BT_SROW * ((size_t)(h->c_hi - h->c_lo) + 16)));
        if (h->shadow) CK(cudaMalloc(&h->Bt2, sizeof(double) * EKF_BT_SROW * ((size_t)(h->c_hi - h->c_lo) + 16)));
        CK(cudaMalloc(&h->lin, sizeof(Lin)));
        CK(cudaMalloc(&h->blin, sizeof(BatchLin)));
        CK(cudaMemset(h->lin, 0, sizeof(Lin)));
        CK(cudaMemset(h->blin, 0, sizeof(BatchLin)));
        CK(cudaHostAlloc(&h->hflags, 4 * sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
        memset(h->hflags, 0, 4 * sizeof(int));
        {
            int* dflags = nullptr;
            CK(cudaHostGetDevicePointer(&dflags, h->hflags, 0));
            h->dflags = dflags;
            CK(cudaMemcpy(&h->lin->nf, &dflags, sizeof(int*), cudaMemcpyHostToDevice));
            CK(cudaMemcpy(&h->blin->nf, &dflags, sizeof(int*), cudaMemcpyHostToDevice));
            int khz = 0;
            CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, p->device));
            const char* tv = getenv("EKF_XCHG_TIMEOUT_MS");
            const double tms = tv ? atof(tv) : 20000.0;
            h->xch_timeout = (long long)(tms * (double)(khz > 0 ? khz : 1965000));
        }
        CK(cudaMalloc(&h->partial, 16 * (size_t)(assoc_blocks(p->capacity_landmarks) + 1)));
        CK(cudaMalloc(&h->ticket, sizeof(unsigned int)));
        CK(cudaMemset(h->ticket, 0, sizeof(unsigned int)));
        CK(cudaMalloc(&h->d_res, sizeof(AssocResult) * EKF_ASSOC_BATCH));
        if (world > 1) CK(cudaMalloc(&h->d_cand, 16 * (1 + EKF_MAX_PEERS)));
        CK(cudaMallocHost(&h->h_res, sizeof(AssocResult) * EKF_ASSOC_BATCH));
        {
            const char* ax = getenv("EKF_ASYNC_X");
            h->async_x = !(ax && atoi(ax) == 0);
        }
        {
            const char* lz = getenv("EKF_LAZY_FLUSH");
            // default: on for slabs, off on one GPU.  Measured (profiles/r02): one GPU 25 784 -> 25 864 updates/s while K3
            // slows from 0.520 to 0.530 ms beside the side-stream kernels (the batch after a flush is the cheap one, m0 = 0,
            // so only ~28 us per two steps can hide and the gather + the slower flush eat them); at 1k landmarks the extra
            // launches cost 15 %.  EKF_LAZY_FLUSH=1 / 0 forces it.
            h->lazy = !h->shadow && h->xstep && p->storage == EKF_STORAGE_SYM && (lz ? atoi(lz) != 0 : world > 1);
        }
        if (h->lazy) {
            int pr_lo = 0, pr_hi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&pr_lo, &pr_hi));
            CK(cudaStreamCreateWithPriority(&h->sg, cudaStreamNonBlocking, pr_hi));
            CK(cudaEventCreateWithFlags(&h->ev_g1, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_g2, cudaEventDisableTiming));
            CK(cudaMalloc(&h->Kd2, sizeof(double2) * h->ld * EKF_MAX_PEND));
            CK(cudaMalloc(&h->Bd2, sizeof(double2) * h->ld * EKF_MAX_PEND));
            CK(cudaMemset(h->Kd2, 0, sizeof(double2) * h->ld * EKF_MAX_PEND));
            CK(cudaMemset(h->Bd2, 0, sizeof(double2) * h->ld * EKF_MAX_PEND));
            if (world == 1) { CK(cudaMalloc(&h->Pc, sizeof(double) * h->ld * EKF_MAX_SLOTS)); h->Pc_cur = h->Pc; }
        }
        {
            const char* am = getenv("EKF_ASYNC_MEASURE");
            h->am_enable = world == 1 && h->async_x && h->defer >= 2 && p->variant == EKF_VARIANT_UC &&
                           p->assoc_mode == EKF_ASSOC_SIGNATURE && !(am && atoi(am) == 0);
        }
        if (h->am_enable) {
            CK(cudaMallocHost(&h->am_obs_h, sizeof(double) * 3 * EKF_ASSOC_BATCH));
            CK(cudaMalloc(&h->am_obs_d, sizeof(double) * 3 * EKF_ASSOC_BATCH));
            CK(cudaMalloc(&h->am_chunks, sizeof(BatchObs) * (EKF_ASSOC_BATCH / EKF_BATCH_MAX)));
            CK(cudaMalloc(&h->am_skip, sizeof(int)));
            CK(cudaMemset(h->am_skip, 0, sizeof(int)));
            CK(cudaHostAlloc(&h->am_status, sizeof(int) * (1 + EKF_ASSOC_BATCH), cudaHostAllocMapped | cudaHostAllocPortable));
            memset(h->am_status, 0, sizeof(int) * (1 + EKF_ASSOC_BATCH));
            CK(cudaHostGetDevicePointer(&h->am_status_d, h->am_status, 0));
            CK(cudaMalloc(&h->am_scratch, sizeof(int) * EKF_ASSOC_BATCH));
            CK(cudaMalloc(&h->am_ticket, sizeof(unsigned int)));
            CK(cudaMemset(h->am_ticket, 0, sizeof(unsigned int)));
            CK(cudaEventCreateWithFlags(&h->ev_dec, cudaEventDisableTiming));
        }
        if (h->async_x) {
            // highest priority: its tiny kernels must be scheduled at once beside a flush that still has ~20 waves of
            // CTAs waiting (at equal priority the block scheduler drains the flush's grid first)
            int pr_lo = 0, pr_hi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&pr_lo, &pr_hi));
            CK(cudaStreamCreateWithPriority(&h->sx, cudaStreamNonBlocking, pr_hi));
            CK(cudaEventCreateWithFlags(&h->ev_mx, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_px, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_pp, cudaEventDisableTiming));
            CK(cudaMalloc(&h->xprev, sizeof(double) * 4));
        }
        return init_state(h);
    }();
    if (rc != EKF_OK) {
        fprintf(stderr, "ekfslam_b200: ekf_create failed: %s\n", h->err.c_str());
        ekf_destroy(h);
        return rc;
    }
    *out = h;
    return EKF_OK;
}

extern "C" int32_t ekf_create(const ekf_params* p, ekf_handle** out) { return create_common(p, 0, 1, nullptr, out); }

extern "C" int32_t ekf_create_sharded(const ekf_params* p, int32_t rank, int32_t world, const uint8_t* id128,
                                      ekf_handle** out) {
    return create_common(p, rank, world, id128, out);
}

// One process, G devices (SURVEY 8(b) "Threading": MATLAB calls MEX from its single interpreter thread, so a sharded
// filter under MATLAB cannot be one process per GPU).  The front handle owns G ordinary slab handles and makes every
// call on each of them in turn; all work is asynchronous on the devices' streams, so the G slabs run concurrently.  The
// per-batch column exchange is the same peer-memory scatter as in the multi-process path, with the peers' buffers
// addressed directly (cudaDeviceEnablePeerAccess) instead of through cudaIpc handles; no NCCL is involved.
extern "C" int32_t ekf_create_multi(const ekf_params* p, int32_t n_devices, const int32_t* devices, ekf_handle** out) {
    if (!p || !out || n_devices < 1 || n_devices > EKF_MAX_PEERS) return EKF_EINVAL;
    *out = nullptr;
    if (n_devices == 1) {
        ekf_params q = *p;
        if (devices) q.device = devices[0];
        return ekf_create(&q, out);
    }
    if (p->stream != 0) { fprintf(stderr, "ekfslam_b200: ekf_create_multi creates its own streams (params.stream must be 0)\n"); return EKF_EINVAL; }
    for (int g = 0; g < n_devices; ++g)
        for (int q = 0; q < g; ++q)
            if ((devices ? devices[g] : g) == (devices ? devices[q] : q)) {
                fprintf(stderr, "ekfslam_b200: ekf_create_multi needs %d distinct devices (slabs that wait on each other must not share a GPU)\n", n_devices);
                return EKF_EINVAL;
            }
    ekf_handle* f = new ekf_handle();
    f->p = *p;
    f->world = n_devices;
    int32_t rc = EKF_OK;
    for (int g = 0; g < n_devices && rc == EKF_OK; ++g) {
        ekf_params q = *p;
        q.device = devices ? devices[g] : g;
        ekf_handle* sh = nullptr;
        rc = create_common(&q, g, n_devices, nullptr, &sh, true);
        if (rc == EKF_OK) f->subs.push_back(sh);
    }
    for (int g = 0; g < n_devices && rc == EKF_OK; ++g) {      // wire the peers: every slab sees every slab's Pc and flag words
        ekf_handle* sh = f->subs[g];
        cudaSetDevice(sh->p.device);
        memset(&sh->peers, 0, sizeof(sh->peers));
        sh->peers.world = n_devices; sh->peers.rank = g;
        for (int q = 0; q < n_devices; ++q) {
            if (q != g) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, sh->p.device, f->subs[q]->p.device);
                cudaError_t e = can ? cudaDeviceEnablePeerAccess(f->subs[q]->p.device, 0) : cudaErrorInvalidDevice;
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                    fprintf(stderr, "ekfslam_b200: ekf_create_multi: device %d cannot map device %d (peer access is required)\n",
                            sh->p.device, f->subs[q]->p.device);
                    rc = EKF_ECUDA;
                }
                cudaGetLastError();
            }
            sh->peers.pc[q] = f->subs[q]->Pc;
            sh->peers.flags[q] = f->subs[q]->xflags;
        }
        sh->p2p = true;
    }
    if (rc != EKF_OK) { ekf_destroy(f); return rc; }
    ekf_handle* s0 = f->subs[0];
    f->defer = s0->defer; f->pend_cap = s0->pend_cap; f->xstep = s0->xstep;
    f->cap_n = s0->cap_n; f->ld = s0->ld; f->c_lo = 0; f->c_hi = s0->cap_n;
    *out = f;
    return EKF_OK;
}

// Multi front: make `call` on every sub-handle; the first failure is returned with that sub-handle's message.
#define FAN_OUT(h, sh, call)                                                                       \
    do {                                                                                           \
        for (ekf_handle* sh : (h)->subs) {                                                         \
            int32_t rc__ = (call);                                                                 \
            if (rc__ != EKF_OK) { (h)->err = "device " + std::to_string(sh->p.device) + ": " + sh->err; return rc__; }   \
        }                                                                                          \
    } while (0)
static inline bool is_front(const ekf_handle* h) { return !h->subs.empty(); }

extern "C" int32_t ekf_shard_info(ekf_handle* h, int32_t* rank, int32_t* world, int32_t* c_lo, int32_t* c_hi) {
    CKH(h);
    if (rank) *rank = h->rank;
    if (world) *world = h->world;
    if (c_lo) *c_lo = h->c_lo;
    if (c_hi) *c_hi = h->c_hi;
    return EKF_OK;
}

extern "C" const char* ekf_last_error(ekf_handle* h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int32_t ekf_sync(ekf_handle* h) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) { FAN_OUT(h, sh, ekf_sync(sh)); return EKF_OK; }
    CK(cudaStreamSynchronize(h->stream));
    if (h->sx) CK(cudaStreamSynchronize(h->sx));
    if (h->sf) CK(cudaStreamSynchronize(h->sf));
    return after_sync(h);
}

extern "C" int32_t ekf_n_landmarks(ekf_handle* h, int32_t* N) {
    CKH(h);
    if (!N) return fail(h, EKF_EINVAL, "N is null");
    *N = h->N;
    return EKF_OK;
}

extern "C" int32_t ekf_defer_info(ekf_handle* h, int32_t* batch, int32_t* pend_cap) {
    CKH(h);
    if (batch) *batch = h->defer;          // (a multi front mirrors its sub-handles' values)
    if (pend_cap) *pend_cap = h->pend_cap;
    return EKF_OK;
}

extern "C" int32_t ekf_set_noise(ekf_handle* h, double C, const double Rc[2], double s_cost, double s_thresh) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) FAN_OUT(h, sh, ekf_set_noise(sh, C, Rc, s_cost, s_thresh));
    h->p.C = C; h->p.Rc[0] = Rc[0]; h->p.Rc[1] = Rc[1]; h->p.s_cost = s_cost; h->p.s_thresh = s_thresh;
    return EKF_OK;
}

// ---------------------------------------------------------------- predict / update / append / associate
extern "C" int32_t ekf_predict(ekf_handle* h, const double u[2]) {
    CKH(h);
    RESOLVE(h);
    if (!u) return fail(h, EKF_EINVAL, "u is null");
    if (is_front(h)) {                       // x must be final on every slab before the motion step
        int32_t rcf = h->xstep ? front_exec(h) : front_flush(h);
        if (rcf != EKF_OK) return rcf;
        FAN_OUT(h, sh, ekf_predict(sh, u));
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    h->x_head_dirty = true;                  // predict moves x(1:3) only (EKF_SLAM.m:57-60)
    if (!h->xstep) FLUSH(h);                 // cross-step deferral keeps the terms pending and transforms them below
    else if (h->world > 1) { int32_t rcq = exec_sharded(h); if (rcq != EKF_OK) return rcq; }   // x must be final before the motion step
    if (h->async_x) {
        // state half first, on the side stream, as soon as the last writer of x is done (not the flush behind it);
        // covariance half on the main stream with the heading the state half saved (EKF_SLAM.m:42,63-64 use theta before f())
        CK(cudaStreamWaitEvent(h->sx, h->ev_mx, 0));
        CK(cudaStreamWaitEvent(h->sx, h->ev_pp, 0));
        launch_predict_x_save(h->sx, h->x, u[0], u[1], h->xprev, h->world == 1 ? h->pose_dev : nullptr);
        CK(cudaEventRecord(h->ev_px, h->sx));
        h->pose_fresh = h->world == 1 && h->pose_dev != nullptr;
        CK(cudaStreamWaitEvent(h->stream, h->ev_px, 0));
        if (shadow_on(h)) {                  // covariance half on the robot strip: P itself may be under a K3 pass
            int32_t rcs = shadow_strip_ready(h);
            if (rcs != EKF_OK) return rcs;
            launch_predict_strip(h->stream, h->strip, h->xprev, h->c_lo, ncols_loc(h), u[0], u[1], h->p.C, h->Kd, h->Bd, h->ld, h->m_pend);
        } else if (h->world > 1)
            launch_predict_shard_P(h->stream, h->P, h->ld, h->c_lo, h->c_hi, h->xprev, n_of(h), u[0], u[1], h->p.C, sym_of(h),
                                   h->Kd, h->Bd, h->ld, h->xstep ? h->m_pend : 0);
        else
            launch_predict_P(h->stream, h->P, h->ld, h->xprev, n_of(h), u[0], u[1], h->p.C, sym_of(h),
                             h->Kd, h->Bd, h->ld, h->xstep ? h->m_pend : 0);      // the pending slabs ride in the same launch
        CK(cudaGetLastError());
        CK(cudaEventRecord(h->ev_pp, h->stream));   // NOT ev_mx: a reader of x must not wait for this kernel (it sits behind the flush)
    } else if (h->world > 1 && h->xstep && h->m_pend > 0) {   // single stream, slab: covariance half, slab transform, state half
        launch_predict_shard_P(h->stream, h->P, h->ld, h->c_lo, h->c_hi, h->x, n_of(h), u[0], u[1], h->p.C, sym_of(h));
        launch_predict_slabs(h->stream, h->Kd, h->Bd, h->ld, h->m_pend, h->x, u[0]);
        launch_predict_x(h->stream, h->x, u[0], u[1]);
        h->launches += 1;
    } else if (h->world > 1)
        launch_predict_shard(h->stream, h->P, h->ld, h->c_lo, h->c_hi, h->x, n_of(h), u[0], u[1], h->p.C, sym_of(h));
    else if (h->xstep && h->m_pend > 0) {    // single stream: covariance half, slab transform, then the state half (all see the pre-motion heading)
        launch_predict_P(h->stream, h->P, h->ld, h->x, n_of(h), u[0], u[1], h->p.C, sym_of(h));
        launch_predict_slabs(h->stream, h->Kd, h->Bd, h->ld, h->m_pend, h->x, u[0]);
        launch_predict_x(h->stream, h->x, u[0], u[1]);
        h->launches += 1;
    } else
        launch_predict(h->stream, h->P, h->ld, h->x, n_of(h), u[0], u[1], h->p.C, sym_of(h));
    h->launches += 2;
    CK(cudaGetLastError());
    return EKF_OK;
}

// Apply the m_pend pending rank-2 terms to P in one pass (the K3 kernel).
static int32_t prof_begin(ekf_handle* h) {
    if (!h->prof) return EKF_OK;
    if (h->ev_used + 2 > h->ev.size()) {
        size_t old = h->ev.size();
        h->ev.resize(old + 512);
        for (size_t i = old; i < h->ev.size(); ++i) CK(cudaEventCreate(&h->ev[i]));
    }
    CK(cudaEventRecord(h->ev[h->ev_used], h->stream));
    return EKF_OK;
}
static int32_t prof_end(ekf_handle* h) {
    if (!h->prof) return EKF_OK;
    CK(cudaEventRecord(h->ev[h->ev_used + 1], h->stream));
    h->ev_used += 2;
    return EKF_OK;
}

// The one data-path collective: make the current columns of the slot set available on every rank.
//   FULL slabs: broadcast each column from its owner (robot columns from rank 0).
//   SYM  slabs: each rank contributes the part of every symmetric column it stores, zero elsewhere; all-reduce.
// phase: 0 = whole exchange; 1 = send only, 2 = wait only (a multi-device front sends on EVERY device before it lets any
// of them wait: with one host thread, a spinning waiter whose peers' sends are not enqueued yet would stall the first
// call that has to synchronise that context -- lazy module loading of a not-yet-used kernel is enough)
static int32_t exchange_cols(ekf_handle* h, const SlotSet& ss, int n, int phase = 0) {
    const int sym = sym_of(h);
    if (h->prof && phase != 2) {
        if (h->xev_used + 2 > h->xev.size()) {
            size_t old = h->xev.size();
            h->xev.resize(old + 256);
            for (size_t i = old; i < h->xev.size(); ++i) CK(cudaEventCreate(&h->xev[i]));
        }
        CK(cudaEventRecord(h->xev[h->xev_used], h->stream));
    }
    if (h->p2p) {
        // peer-memory scatter: every rank stores the elements it owns into every rank's buffer, then publishes the batch
        if (phase != 2) {
            h->epoch += 1;
            const int64_t buf_off = (int64_t)(h->epoch & 1) * EKF_MAX_SLOTS * h->ld;
            launch_scatter_cols(h->stream, h->P, h->ld, h->c_lo, h->c_hi, ss, n, h->peers, buf_off, sym, h->epoch, h->xticket,
                                h->strip_valid ? h->strip : nullptr);
            h->Pc_cur = h->Pc + buf_off;
            h->launches += 1;
        }
        if (phase != 1) {
            launch_wait_flags(h->stream, h->xflags, h->world, h->epoch, h->xch_timeout, h->dflags + 1);
            h->launches += 1;
        }
        CK(cudaGetLastError());
    } else if (phase != 2) {
        h->Pc_cur = h->Pc;
        launch_pack_cols(h->stream, h->P, h->ld, h->c_lo, h->c_hi, ss, n, h->Pc, h->ld, sym, h->strip_valid ? h->strip : nullptr);
        CK(cudaGetLastError());
        if (sym) {
            CKN(g_nccl.AllReduce(h->Pc, h->Pc, (size_t)ss.n * h->ld, ncclDouble, ncclSum, h->comm, h->stream));
        } else {
            CKN(g_nccl.GroupStart());
            CKN(g_nccl.Broadcast(h->Pc, h->Pc, 3 * (size_t)h->ld, ncclDouble, 0, h->comm, h->stream));
            for (int sl = 3; sl + 1 < ss.n; sl += 2) {
                double* p = h->Pc + (size_t)sl * h->ld;
                CKN(g_nccl.Broadcast(p, p, 2 * (size_t)h->ld, ncclDouble, owner_of_col(h, ss.col[sl]), h->comm, h->stream));
            }
            CKN(g_nccl.GroupEnd());
        }
        h->launches += 1;
    }
    if (h->prof && phase != 1) {
        CK(cudaEventRecord(h->xev[h->xev_used + 1], h->stream));
        h->xev_used += 2;
    }
    return EKF_OK;
}

// Lazy flush (see ekf_handle::lazy).  On entry the raw slot columns of the batch `ob` sit in h->Pc_cur (enqueued on the
// main stream) and h->m_pend terms are pending in (Kd, Bd).  Launches the K3 pass over P on the main stream and, beside it
// on `sg`, the batch's linearisation + gains from the copies corrected by those terms, into (Kd2, Bd2); then swaps.
static int32_t lazy_flush_and_batch(ekf_handle* h, const BatchObs& ob, ObSrc src = ObSrc{nullptr, nullptr}) {
    const int m_old = h->m_pend;
    CK(cudaEventRecord(h->ev_g1, h->stream));                  // the copies are complete here; nothing has touched P yet
    int32_t rc = (h->world > 1) ? downdate_sharded(h) : flush(h);              // K3 on the main stream; m_pend = 0
    if (rc != EKF_OK) return rc;
    CK(cudaStreamWaitEvent(h->sg, h->ev_g1, 0));
    launch_batch_lin_gain(h->sg, h->P, h->ld, h->x, n_of(h), sym_of(h), ob, h->blin, h->Kd, h->Bd, h->ld, h->Pc_cur, h->ld,
                          h->c_lo, std::min(h->c_hi, h->cap_n), m_old, h->Kd2, h->Bd2, 0, src);
    h->launches += 2;
    CK(cudaGetLastError());
    h->x_host = false; h->xpin_fresh = false; h->pose_fresh = false;
    if (h->async_x) CK(cudaEventRecord(h->ev_mx, h->sg));      // x of the batch is final after the gain kernel on sg
    CK(cudaEventRecord(h->ev_g2, h->sg));
    CK(cudaStreamWaitEvent(h->stream, h->ev_g2, 0));           // later main-stream work sees x and the new terms
    std::swap(h->Kd, h->Kd2);
    std::swap(h->Bd, h->Bd2);
    h->m_pend = ob.k;
    return EKF_OK;
}

// Sharded filters queue their observations (h->q) and execute them as batches of <= defer_m:
//   exec_sharded      one exchange of the 3 + 2k slot columns of P0, then k communication-free linearise/gain steps
//                     replicated on every rank (x, K_q, B_q final; m_pend grows by k)
//   downdate_sharded  one rank-2 m_pend pass over the local slab
// Without cross-step deferral the two always run back to back (flush_sharded).  With it (SYM slabs, defer_m = 9..16) the
// terms of a batch stay pending across predict exactly as on one GPU: the exchanged columns are raw P0 and every rank
// corrects them by the pending terms it already holds (k_batch_lin / k_batch_gain with m0), so the slab is streamed
// once per pend_cap observations.
static int32_t downdate_sharded(ekf_handle* h) {
    if (h->shadow) return shadow_settle(h);
    if (h->m_pend == 0) return EKF_OK;
    const int ncl = ncols_loc(h);
    if (ncl > 0) {
        int32_t rc = prof_begin(h);
        if (rc != EKF_OK) return rc;
        launch_downdate(h->stream, h->P, h->ld, n_of(h), ncl, sym_of(h), h->Kd, h->Bd + h->c_lo, h->ld, h->m_pend, h->c_lo,
                        h->dd_variant, h->Bt);
        rc = prof_end(h);
        if (rc != EKF_OK) return rc;
        h->dd_launches += 1;
        h->dd_terms += h->m_pend;
        h->launches += 1;
        CK(cudaGetLastError());
    }
    h->m_pend = 0;
    return EKF_OK;
}

// exec in two halves (see exchange_cols): exec_send = make room + send this slab's share of the slot columns;
// exec_finish = wait for everybody's share, linearise, gains.
static int32_t exec_send(ekf_handle* h) {
    const int k = (int)h->q.size();
    if (k == 0) return EKF_OK;
    h->lazy_now = h->lazy && h->m_pend > 0 && h->m_pend + k > h->pend_cap && k >= 2 && !getenv("EKF_SHARD_PEROBS");
    if (shadow_on(h)) {                                     // continuous flush on the slab: same state machine as fused_chunk
        int32_t rc = shadow_strip_ready(h);
        if (rc != EKF_OK) return rc;
        if (h->mQ > 0 && (rc = shadow_retire(h)) != EKF_OK) return rc;
        if (h->m_pend - h->mA + k > h->pend_cap) {
            if ((rc = shadow_close(h)) != EKF_OK) return rc;
            if (h->mQ > 0 && (rc = shadow_retire(h)) != EKF_OK) return rc;
        }
    } else if (h->m_pend + k > h->pend_cap && !h->lazy_now) {      // make room: one pass over the slab for everything pending
        int32_t rc = downdate_sharded(h);
        if (rc != EKF_OK) return rc;
    }
    SlotSet& ss = h->ss_cur;
    ss.n = 3 + 2 * k;
    ss.col[0] = 0; ss.col[1] = 1; ss.col[2] = 2;
    for (int m = 0; m < k; ++m) { ss.col[3 + 2 * m] = h->q[m].a; ss.col[4 + 2 * m] = h->q[m].a + 1; }
    return exchange_cols(h, ss, n_of(h), 1);
}

static int32_t exec_finish(ekf_handle* h) {
    const int k = (int)h->q.size();
    if (k == 0) return EKF_OK;
    const int m0 = h->m_pend;
    const int n = n_of(h);
    const SlotSet& ss = h->ss_cur;
    int32_t rcx = exchange_cols(h, ss, n, 2);
    if (rcx != EKF_OK) return rcx;
    if (h->lazy_now) {                      // the exchanged columns are raw P0: flush on the main stream, this batch beside it
        BatchObs ob;
        ob.k = k;
        for (int m = 0; m < k; ++m) {
            const ekf_handle::QObs& o = h->q[m];
            ob.z0[m] = o.z0; ob.z1[m] = o.z1; ob.R0[m] = o.R0; ob.R1[m] = o.R1; ob.a[m] = o.a;
        }
        h->q.clear();
        h->lazy_now = false;
        return lazy_flush_and_batch(h, ob);
    }
    if (k >= 2 && !getenv("EKF_SHARD_PEROBS")) {
        // fused: one CTA walks the k linearisations on the exchanged (3+2k)^2 block, one grid-wide kernel builds all
        // k gain slabs (replicated on every rank from identical inputs) -- 2 launches per batch
        BatchObs ob;
        ob.k = k;
        for (int m = 0; m < k; ++m) {
            const ekf_handle::QObs& o = h->q[m];
            ob.z0[m] = o.z0; ob.z1[m] = o.z1; ob.R0[m] = o.R0; ob.R1[m] = o.R1; ob.a[m] = o.a;
        }
        static int fr = -1;
        if (fr < 0) { const char* e = getenv("EKF_BATCH_FRONT"); fr = e ? atoi(e) : 1; }
        if (fr && sym_of(h) && m0 > 0)      // pending terms: their correction of the exchanged columns runs beside the linearisation chain
            launch_batch_front_gain(h->stream, h->P, h->ld, nullptr, h->x, n, ob, h->blin, h->Pc_cur, h->ld, 0, h->Kd, h->Bd, h->ld, m0);
        else
            launch_batch_lin_gain(h->stream, h->P, h->ld, h->x, n, sym_of(h), ob, h->blin, h->Kd, h->Bd, h->ld, h->Pc_cur, h->ld,
                                  h->c_lo, h->c_hi, m0);
        h->launches += 2;
    } else {
        for (int m = 0; m < k; ++m) {       // slab slot m0 + m; cur_col subtracts the m0 + m terms before it
            const ekf_handle::QObs& o = h->q[m];
            launch_lin_gain_shard(h->stream, h->P, h->ld, h->c_lo, h->c_hi, h->x, n, o.a, 3 + 2 * m, o.z0, o.z1, o.R0, o.R1,
                                  ss, h->Pc_cur, h->ld, h->lin, h->Kd, h->Bd, h->ld, h->Bc, m0 + m, sym_of(h));
        }
        h->launches += 2 * k;
    }
    CK(cudaGetLastError());
    MARK_X(h);                              // x of the batch is final here, before any slab flush
    h->q.clear();
    h->m_pend = m0 + k;
    if (shadow_on(h)) {
        if (h->m_pend - h->mA - h->mQ + std::min(h->defer, EKF_BATCH_MAX) > h->pend_cap) return shadow_close(h);
        return EKF_OK;
    }
    if (h->m_pend >= h->pend_cap && !h->lazy) return downdate_sharded(h);    // full: same flush points as the single-GPU path (bitwise equal)
    return EKF_OK;
}

static int32_t exec_sharded(ekf_handle* h) {
    int32_t rc = exec_send(h);
    if (rc != EKF_OK) return rc;
    return exec_finish(h);
}

// Multi-device front: the same batch on every slab, all sends before any wait.
static int32_t front_exec(ekf_handle* f) {
    for (ekf_handle* sh : f->subs) { cudaSetDevice(sh->p.device); int32_t rc = exec_send(sh); if (rc != EKF_OK) { f->err = sh->err; return rc; } }
    for (ekf_handle* sh : f->subs) { cudaSetDevice(sh->p.device); int32_t rc = exec_finish(sh); if (rc != EKF_OK) { f->err = sh->err; return rc; } }
    return EKF_OK;
}
static int32_t front_flush(ekf_handle* f) {
    int32_t rc = front_exec(f);
    if (rc != EKF_OK) return rc;
    for (ekf_handle* sh : f->subs) { cudaSetDevice(sh->p.device); rc = downdate_sharded(sh); if (rc != EKF_OK) { f->err = sh->err; return rc; } }
    return EKF_OK;
}

static int32_t flush_sharded(ekf_handle* h) {
    int32_t rc = exec_sharded(h);
    if (rc != EKF_OK) return rc;
    return downdate_sharded(h);
}

// ---- continuous flush (see ekf_handle::shadow) ----
static int32_t prof_slot(ekf_handle* h) {
    if (h->ev_used + 2 > h->ev.size()) {
        size_t old = h->ev.size();
        h->ev.resize(old + 512);
        for (size_t i = old; i < h->ev.size(); ++i) CK(cudaEventCreate(&h->ev[i]));
    }
    return EKF_OK;
}
static int32_t shadow_strip_ready(ekf_handle* h) {
    if (h->strip_valid) return EKF_OK;
    launch_strip_load(h->stream, h->P, h->ld, h->c_lo, ncols_loc(h), h->strip);
    h->launches += 1;
    CK(cudaGetLastError());
    h->strip_valid = true;
    return EKF_OK;
}
// The running pass is done (the chain waits for it here): its output buffer becomes the matrix, the strip drops the
// retired terms, the window keeps what came after them; a queued pass becomes the running one.
static int32_t shadow_retire(ekf_handle* h) {
    if (h->nfl == 0) return EKF_OK;
    CK(cudaStreamWaitEvent(h->stream, h->ev_fd[h->fcur], 0));
    launch_strip_fix(h->stream, h->strip, h->c_lo, ncols_loc(h), h->Kd, h->Bd, h->ld, h->mA);
    std::swap(h->P, h->P2);
    h->Kd += (int64_t)h->mA * h->ld;              // the window starts behind the retired terms now
    h->Bd += (int64_t)h->mA * h->ld;
    h->kw0 += h->mA;
    h->launches += 1;
    CK(cudaGetLastError());
    h->m_pend -= h->mA;
    h->mA = h->mQ;
    h->mQ = 0;
    h->nfl -= 1;
    h->fcur ^= 1;
    return EKF_OK;
}
// Close the set that is filling up (window slots [mA + mQ, m_pend)) and enqueue its K3 pass on `sf`, out of place: as the
// running pass (P -> P2) when nothing is outstanding, else queued behind the running one (its output P2 -> back into P;
// the chain must retire the running pass before it reads P again).  Needs a free place in the queue (nfl < 2, mQ == 0).
static int32_t shadow_close(ekf_handle* h) {
    if (h->m_pend - h->mA - h->mQ <= 0) return EKF_OK;
    if (h->nfl >= 2 || h->mQ != 0) return fail(h, EKF_EINVAL, "internal: continuous flush queue is full");
    if (h->kw0 + 2 * EKF_MAX_WINDOW > EKF_WIN_SLOTS) {
        // the window has wandered to the end of its allocation: bring it back to the start.  Passes in flight read their
        // sets where they lie, so they are retired first (a wait of at most one pass, once in ~13 passes).
        int32_t rc = EKF_OK;
        while (h->nfl > 0) if ((rc = shadow_retire(h)) != EKF_OK) return rc;
        launch_rotate_slabs(h->stream, h->Kd0, h->Bd0, h->ld, n_of(h), h->kw0, h->m_pend, nullptr, nullptr);
        h->Kd = h->Kd0; h->Bd = h->Bd0; h->kw0 = 0;
        h->launches += 1;
    }
    const int off = h->mA + h->mQ, m = h->m_pend - off;
    const int slot = h->nfl == 0 ? h->fcur : (h->fcur ^ 1);
    const double2* Kset = h->Kd + (int64_t)off * h->ld;
    const double2* Bset = h->Bd + (int64_t)off * h->ld;
    const int ncl = ncols_loc(h);                 // a slab passes its own columns (and B_q from its first column on)
    double* bt = slot ? h->Bt2 : h->Bt;           // the pass two places back used this buffer: it has been retired
    const bool pre = m > 8 && h->dd_variant < 0;  // the default (hybrid, operand in the stage) kernel: pack its B operand here, on the chain's stream
    if (pre && ncl > 0) { launch_pack_bt(h->stream, Bset + h->c_lo, h->ld, m, ncl, bt); h->launches += 1; }
    CK(cudaEventRecord(h->ev_fz, h->stream));     // the set is complete; every gather enqueued so far has read the buffer a queued pass overwrites
    CK(cudaStreamWaitEvent(h->sf, h->ev_fz, 0));
    if (h->prof) { int32_t rc = prof_slot(h); if (rc != EKF_OK) return rc; CK(cudaEventRecord(h->ev[h->ev_used], h->sf)); }
    double* src = h->nfl == 0 ? h->P : h->P2;
    double* dst = h->nfl == 0 ? h->P2 : h->P;
    // EKF_SHADOW_VARIANT: hybrid kernel shape of the overlapped pass.  Default = the library's (4-stage ring, 3 CTAs/SM).  The
    // 2-CTA/SM shapes (33, 35, 38: as fast when alone) leave a third of every SM free for the chain's CTAs -- measured at 10k:
    // end to end 23.6k -> 24.6k updates/s, but the pass itself slows from 0.563 to 0.603 ms beside them (resident 27.4k -> 26.2k).
    static int sv = -2;
    if (sv == -2) { const char* e = getenv("EKF_SHADOW_VARIANT"); sv = (e && *e) ? atoi(e) : -3; }
    const int svh = sv != -3 ? sv : (h->world > 1 ? 33 : -1);     // slabs: the 2-CTA/SM shape (2 GPUs, 10k: 41.2k -> 44.3k updates/s)
    if (ncl > 0)
        launch_downdate(h->sf, src, h->ld, n_of(h), ncl, 1, Kset, Bset + h->c_lo, h->ld, m, h->c_lo,
                        h->dd_variant >= 0 ? h->dd_variant : (m > 8 ? svh : -1), bt, dst, pre);
    if (h->prof) { CK(cudaEventRecord(h->ev[h->ev_used + 1], h->sf)); h->ev_used += 2; }
    CK(cudaEventRecord(h->ev_fd[slot], h->sf));
    CK(cudaGetLastError());
    h->dd_launches += 1;
    h->dd_terms += m;
    h->launches += 1;
    if (h->nfl == 0) h->mA = m; else h->mQ = m;
    h->nfl += 1;
    return EKF_OK;
}
// Forget everything pending (the state is about to be replaced).
static int32_t shadow_reset(ekf_handle* h) {
    if (!h->shadow) return EKF_OK;
    if (h->nfl > 0) { CK(cudaStreamSynchronize(h->sf)); }
    if (h->nfl == 1) { /* the running pass wrote P2; P is still the matrix the caller replaces */ }
    h->nfl = 0;
    h->mA = h->mQ = 0;
    h->strip_valid = false;
    if (h->Kd0) { h->Kd = h->Kd0; h->Bd = h->Bd0; h->kw0 = 0; }
    return EKF_OK;
}

// P (the local slab) must be whole: retire the passes in flight, apply what is left of the window in place, put the strip back.
static int32_t shadow_settle(ekf_handle* h) {
    int32_t rc = EKF_OK;
    while (h->nfl > 0) if ((rc = shadow_retire(h)) != EKF_OK) return rc;
    const int ncl = ncols_loc(h);
    if (h->m_pend > 0) {
        if (ncl > 0) {
            if (h->prof) { rc = prof_slot(h); if (rc != EKF_OK) return rc; CK(cudaEventRecord(h->ev[h->ev_used], h->stream)); }
            launch_downdate(h->stream, h->P, h->ld, n_of(h), ncl, 1, h->Kd, h->Bd + h->c_lo, h->ld, h->m_pend, h->c_lo, h->dd_variant, h->Bt);
            if (h->prof) { CK(cudaEventRecord(h->ev[h->ev_used + 1], h->stream)); h->ev_used += 2; }
            if (h->strip_valid) launch_strip_fix(h->stream, h->strip, h->c_lo, ncl, h->Kd, h->Bd, h->ld, h->m_pend);
            h->dd_launches += 1;
            h->dd_terms += h->m_pend;
            h->launches += 2;
        }
        h->m_pend = 0;
    }
    if (h->strip_valid) {
        launch_strip_store(h->stream, h->P, h->ld, h->c_lo, ncl, h->strip);
        h->launches += 1;
        h->strip_valid = false;
    }
    if (h->Kd0) { h->Kd = h->Kd0; h->Bd = h->Bd0; h->kw0 = 0; }      // nothing pending: the window starts over
    CK(cudaGetLastError());
    return EKF_OK;
}

static int32_t flush(ekf_handle* h) {
    if (is_front(h)) return front_flush(h);
    if (h->world > 1) return flush_sharded(h);
    if (h->shadow) return shadow_settle(h);
    if (h->m_pend == 0) return EKF_OK;
    if (h->prof) {
        if (h->ev_used + 2 > h->ev.size()) {
            size_t old = h->ev.size();
            h->ev.resize(old + 512);
            for (size_t i = old; i < h->ev.size(); ++i) CK(cudaEventCreate(&h->ev[i]));
        }
        CK(cudaEventRecord(h->ev[h->ev_used], h->stream));
    }
    launch_downdate(h->stream, h->P, h->ld, n_of(h), n_of(h), sym_of(h), h->Kd, h->Bd, h->ld, h->m_pend, 0, h->dd_variant, h->Bt);
    if (h->prof) {
        CK(cudaEventRecord(h->ev[h->ev_used + 1], h->stream));
        h->ev_used += 2;
    }
    h->dd_launches += 1;
    h->dd_terms += h->m_pend;
    h->launches += 1;
    h->m_pend = 0;
    CK(cudaGetLastError());
    return EKF_OK;
}

extern "C" int32_t ekf_flush_info(ekf_handle* h, int32_t* continuous, int32_t* passes_outstanding) {
    CKH(h);
    const ekf_handle* q = is_front(h) ? h->subs[0] : h;
    if (continuous) *continuous = shadow_on(q) ? 1 : 0;
    if (passes_outstanding) *passes_outstanding = q->nfl;
    return EKF_OK;
}

extern "C" int32_t ekf_flush(ekf_handle* h) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) return flush(h);
    CK(cudaSetDevice(h->p.device));
    return flush(h);
}

extern "C" int32_t ekf_update(ekf_handle* h, const double z[2], const double Rdiag[2], int32_t idx1) {
    CKH(h);
    RESOLVE(h);
    if (!z || !Rdiag) return fail(h, EKF_EINVAL, "null argument");
    if (idx1 < 1 || idx1 > h->N)      // MATLAB would raise "index exceeds array bounds" at EKF_SLAM.m:125
        return fail(h, EKF_EINVAL, "ekf_update: landmark index " + std::to_string(idx1) + " outside 1.." + std::to_string(h->N));
    if (is_front(h)) {                       // queue on every slab; the front executes full batches in lockstep
        FAN_OUT(h, sh, ekf_update(sh, z, Rdiag, idx1));
        if ((int)h->subs[0]->q.size() >= h->defer) return h->xstep ? front_exec(h) : front_flush(h);
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    const int a = 3 + 2 * (idx1 - 1);
    if (h->world > 1) {                   // queue; the batch runs at the next flush point
        h->q.push_back({z[0], z[1], Rdiag[0], Rdiag[1], a});
        if (h->in_process) return EKF_OK;     // the multi-device front decides when its slabs execute
        if ((int)h->q.size() >= h->defer) return h->xstep ? exec_sharded(h) : flush_sharded(h);
        return EKF_OK;
    }
    if (shadow_on(h)) {                   // one-observation batch through the strip-aware path
        BatchObs ob;
        memset(&ob, 0, sizeof(ob));
        ob.k = 1; ob.z0[0] = z[0]; ob.z1[0] = z[1]; ob.R0[0] = Rdiag[0]; ob.R1[0] = Rdiag[1]; ob.a[0] = a;
        return fused_chunk(h, ob, ObSrc{nullptr, nullptr});
    }
    if (h->m_pend >= h->pend_cap) FLUSH(h);      // lazy flush leaves a full set pending: make room first
    launch_linearize_gain(h->stream, h->P, h->ld, h->x, n_of(h), a, z[0], z[1], Rdiag[0], Rdiag[1], sym_of(h),
                          h->lin, h->Kd, h->Bd, h->ld, h->m_pend);
    h->launches += 2;
    CK(cudaGetLastError());
    MARK_X(h);
    h->m_pend += 1;
    if (h->m_pend >= h->pend_cap && !h->lazy) FLUSH(h);     // lazy: a full set waits for the next observation / batch
    return EKF_OK;
}

// append on a slab: the new rows need P(:,1:3), one 3-column exchange from rank 0 (phase as in exchange_cols)
static int32_t append_sharded(ekf_handle* h, const double u[2], const double Rdiag[2], const double pos[2], double signature,
                              int phase) {
    SlotSet ss; ss.n = 3; ss.col[0] = 0; ss.col[1] = 1; ss.col[2] = 2;
    int32_t rcx = exchange_cols(h, ss, n_of(h), phase);
    if (rcx != EKF_OK || phase == 1) return rcx;
    launch_append_shard(h->stream, h->P, h->ld, h->c_lo, h->c_hi, h->x, h->s, h->N, u[0], u[1], Rdiag[0], Rdiag[1],
                        pos[0], pos[1], signature, h->Pc_cur, h->ld);
    h->launches += 1;
    CK(cudaGetLastError());
    MARK_X(h);
    h->N += 1;
    return EKF_OK;
}

extern "C" int32_t ekf_append(ekf_handle* h, const double u[2], const double Rdiag[2], const double pos[2],
                              double signature) {
    CKH(h);
    RESOLVE(h);
    if (!u || !Rdiag || !pos) return fail(h, EKF_EINVAL, "null argument");
    if (h->N >= h->p.capacity_landmarks)
        return fail(h, EKF_ECAPACITY, "ekf_append: capacity_landmarks=" + std::to_string(h->p.capacity_landmarks) + " reached");
    if (is_front(h)) {                    // flush everywhere, then the 3-column exchange: all sends before any wait
        int32_t rcq = front_flush(h);
        if (rcq != EKF_OK) return rcq;
        for (int phase = 1; phase <= 2; ++phase)
            for (ekf_handle* sh : h->subs) {
                cudaSetDevice(sh->p.device);
                int32_t rcf = append_sharded(sh, u, Rdiag, pos, signature, phase);
                if (rcf != EKF_OK) { h->err = sh->err; return rcf; }
            }
        h->N += 1;
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);
    if (h->world > 1) return append_sharded(h, u, Rdiag, pos, signature, 0);
    launch_append(h->stream, h->P, h->ld, h->x, h->s, h->N, u[0], u[1], Rdiag[0], Rdiag[1], pos[0], pos[1],
                  signature, sym_of(h));
    h->launches += 1;
    CK(cudaGetLastError());
    MARK_X(h);
    h->N += 1;
    h->sigs[sig_key(signature)] += 1;
    return EKF_OK;
}

// Mode-B association on a slab (Correspondence.m:49-87 with :74): every rank scores the landmarks whose columns it stores
// -- their 2x2 diagonal block and robot cross-columns are local, the robot 3x3 block (and, FULL storage, the robot
// columns) comes from one 3-column exchange --, reduces them to one lexicographic (cost, index) candidate, and the
// candidates cross NVLink as 16-byte stores into the peers' flag area (all-gather over NCCL without peer memory).
// stage 1: send the robot columns; 2: wait for them, score, send the candidate; 3: fold.  0: all three.
static int32_t associate_b_sharded(ekf_handle* h, const double z[3], const double Rdiag[2], int stage) {
    SlotSet ss; ss.n = 3; ss.col[0] = 0; ss.col[1] = 1; ss.col[2] = 2;
    const int n = n_of(h);
    if (stage == 0 || stage == 1) {
        int32_t rc = exchange_cols(h, ss, n, stage == 0 ? 0 : 1);
        if (rc != EKF_OK || stage == 1) return rc;
    }
    if (stage == 0 || stage == 2) {
        if (stage == 2) { int32_t rc = exchange_cols(h, ss, n, 2); if (rc != EKF_OK) return rc; }
        const int lo = std::max(h->c_lo, 3), hi = std::min(h->c_hi, n);
        const int lm_lo = (lo - 3) / 2, lm_cnt = hi > lo ? (hi - lo) / 2 : 0;
        launch_associate_slab(h->stream, h->P, h->ld, h->c_lo, lm_lo, lm_cnt, h->x, h->s, h->N, z[0], z[1], z[2], Rdiag[0], Rdiag[1],
                              h->p.s_cost, h->p.s_thresh, sym_of(h), h->Pc_cur, h->ld, h->partial, h->ticket, h->d_cand);
        h->launches += 1;
        CK(cudaGetLastError());
        if (h->p2p) {
            h->aepoch += 1;
            launch_cand_send(h->stream, h->d_cand, h->peers, h->aepoch);
            h->launches += 1;
            CK(cudaGetLastError());
        }
        if (stage == 2) return EKF_OK;
    }
    if (h->p2p) launch_cand_fold(h->stream, h->xflags, h->world, h->aepoch, h->xch_timeout, h->dflags + 1, h->N, h->d_res);
    else {
        char* all = (char*)h->d_cand + 16;
        CKN(g_nccl.AllGather(h->d_cand, all, 16, ncclChar, h->comm, h->stream));
        launch_cand_fold_gathered(h->stream, all, h->world, h->N, h->d_res);
    }
    h->launches += 1;
    CK(cudaGetLastError());
    return EKF_OK;
}

extern "C" int32_t ekf_associate(ekf_handle* h, const double z[3], const double Rdiag[2], int32_t* newLL,
                                 int32_t* idx1) {
    CKH(h);
    RESOLVE(h);
    if (!z || !Rdiag || !newLL || !idx1) return fail(h, EKF_EINVAL, "null argument");
    if (h->N == 0) { *newLL = 1; *idx1 = 1; return EKF_OK; }   // Correspondence.m:33,40 with an empty loop
    if (is_front(h)) {
        if (h->p.assoc_mode == EKF_ASSOC_POSITION) {       // every slab scores its landmarks; all sends before any wait
            int32_t rcq = front_flush(h);
            if (rcq != EKF_OK) return rcq;
            for (int stage = 1; stage <= 3; ++stage)
                for (ekf_handle* sh : h->subs) {
                    cudaSetDevice(sh->p.device);
                    int32_t rcf = associate_b_sharded(sh, z, Rdiag, stage);
                    if (rcf != EKF_OK) { h->err = sh->err; return rcf; }
                }
            ekf_handle* s0 = h->subs[0];
            cudaSetDevice(s0->p.device);
            if (cudaMemcpyAsync(s0->h_res, s0->d_res, sizeof(AssocResult), cudaMemcpyDeviceToHost, s0->stream) != cudaSuccess ||
                cudaStreamSynchronize(s0->stream) != cudaSuccess) return fail(h, EKF_ECUDA, "ekf_associate: copy of the result failed");
            *newLL = s0->h_res->newLL;
            *idx1 = s0->h_res->index1;
            int32_t rca = after_sync(s0);
            if (rca != EKF_OK) h->err = s0->err;
            return rca;
        }
        int32_t rcf = ekf_associate(h->subs[0], z, Rdiag, newLL, idx1);   // mode A reads only s, which every slab holds
        if (rcf != EKF_OK) h->err = h->subs[0]->err;
        return rcf;
    }
    CK(cudaSetDevice(h->p.device));
    if (h->world > 1 && h->p.assoc_mode == EKF_ASSOC_POSITION) {
        FLUSH(h);
        int32_t rcb = associate_b_sharded(h, z, Rdiag, 0);
        if (rcb != EKF_OK) return rcb;
        CK(cudaMemcpyAsync(h->h_res, h->d_res, sizeof(AssocResult), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        *newLL = h->h_res->newLL;
        *idx1 = h->h_res->index1;
        return after_sync(h);
    }
    if (h->p.assoc_mode == EKF_ASSOC_POSITION) FLUSH(h);      // mode B reads P blocks; mode A only s
    cudaStream_t st = h->stream;
    if (h->async_x && h->p.assoc_mode == EKF_ASSOC_SIGNATURE) {   // mode A: behind the last writer of s, not behind a flush
        st = h->sx;
        CK(cudaStreamWaitEvent(st, h->ev_mx, 0));
    }
    launch_associate(st, h->P, h->ld, h->x, h->s, h->N, z[0], z[1], z[2], Rdiag[0], Rdiag[1],
                     h->p.s_cost, h->p.s_thresh, h->p.assoc_mode, sym_of(h), h->partial, h->ticket, h->d_res);
    h->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(h->h_res, h->d_res, sizeof(AssocResult), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *newLL = h->h_res->newLL;
    *idx1 = h->h_res->index1;
    return after_sync(h);
}

// k updates with explicit slots: the update branch of measure() (EKF_SLAM.m:124-145) for a batch whose
// correspondences are already decided.  z is k x 2 and Rdiag k x 2, column-major like observed_LL.
// One fused chunk (<= 8 observations with decided slots) on a single-GPU handle: make room / lazy flush, k_batch_lin +
// k_batch_gain, flush policy.  `ob` carries the batch by value, or only its size when `src` names a device-side
// description (slots decided by k_measure_decide).
static int32_t fused_chunk(ekf_handle* h, const BatchObs& ob, ObSrc src) {
    if (shadow_on(h)) {
        int32_t rc = shadow_strip_ready(h);
        if (rc != EKF_OK) return rc;
        if (h->mQ > 0 && (rc = shadow_retire(h)) != EKF_OK) return rc;     // a queued pass writes into P: move on to the new P first
        if (h->m_pend - h->mA + ob.k > h->pend_cap) {                        // does not fit the filling set: close it now
            if ((rc = shadow_close(h)) != EKF_OK) return rc;
            if (h->mQ > 0 && (rc = shadow_retire(h)) != EKF_OK) return rc;
        }
        static int fr = -1;                                                 // EKF_BATCH_FRONT=0: gather, k_batch_lin, k_batch_gain(m0) one after another
        if (fr < 0) { const char* e = getenv("EKF_BATCH_FRONT"); fr = e ? atoi(e) : 1; }
        double* xh = h->async_x ? h->x_pin_dev : nullptr;                   // the gains also write x into mapped host memory
        static int ct = -1;                                                 // EKF_CHAIN_TIMING=1 (debug): CUDA-event times of front and gains
        if (ct < 0) ct = getenv("EKF_CHAIN_TIMING") ? 1 : 0;
        if (fr && ct) {
            if (h->cev.empty()) { h->cev.resize(3 * 4096); for (auto& e : h->cev) CK(cudaEventCreate(&e)); }
            if (h->cev_used + 3 <= h->cev.size()) {
                CK(cudaEventRecord(h->cev[h->cev_used], h->stream));
                launch_batch_front_only(h->stream, h->P, h->ld, h->strip, h->x, n_of(h), ob, h->blin, h->Pc, h->ld, 1, h->Kd, h->Bd, h->ld, h->m_pend, src);
                CK(cudaEventRecord(h->cev[h->cev_used + 1], h->stream));
                launch_batch_gain_only(h->stream, h->P, h->ld, h->x, n_of(h), ob, h->blin, h->Pc, h->ld, h->Kd, h->Bd, h->ld, h->m_pend, src, xh);
                CK(cudaEventRecord(h->cev[h->cev_used + 2], h->stream));
                h->cev_used += 3;
            } else launch_batch_front_gain(h->stream, h->P, h->ld, h->strip, h->x, n_of(h), ob, h->blin, h->Pc, h->ld, 1, h->Kd, h->Bd, h->ld, h->m_pend, src, xh);
            h->launches += 2;
        } else if (fr) {
            launch_batch_front_gain(h->stream, h->P, h->ld, h->strip, h->x, n_of(h), ob, h->blin, h->Pc, h->ld, 1, h->Kd, h->Bd, h->ld,
                                    h->m_pend, src, xh);
            h->launches += 2;
        } else {
            launch_gather_cols_strip(h->stream, h->P, h->ld, n_of(h), ob, h->Pc, h->ld, src, h->strip);
            launch_batch_lin_gain(h->stream, h->P, h->ld, h->x, n_of(h), 1, ob, h->blin, h->Kd, h->Bd, h->ld, h->Pc, h->ld,
                                  0, n_of(h), h->m_pend, nullptr, nullptr, 0, src, xh);
            h->launches += 3;
        }
        CK(cudaGetLastError());
        MARK_X(h);
        if (xh) { h->xpin_fresh = true; h->x_head_dirty = false; }          // x_pin is x once ev_mx (just recorded) has fired
        h->m_pend += ob.k;
        if (h->m_pend - h->mA - h->mQ + std::min(h->defer, EKF_BATCH_MAX) > h->pend_cap)   // no further full chunk fits: its pass can start
            return shadow_close(h);
        return EKF_OK;
    }
    if (h->lazy && h->m_pend > 0 && h->m_pend + ob.k > h->pend_cap) {
        // lazy flush: copy the batch's slot columns out of P0, then K3 on the main stream || lin + gain on sg
        launch_gather_cols(h->stream, h->P, h->ld, n_of(h), sym_of(h), ob, h->Pc, h->ld, src);
        h->Pc_cur = h->Pc;
        h->launches += 1;
        CK(cudaGetLastError());
        return lazy_flush_and_batch(h, ob, src);
    }
    if (h->xstep && h->m_pend + ob.k > h->pend_cap) FLUSH(h);       // make room: one pass over P for everything pending
    const int m0 = h->xstep ? h->m_pend : 0;
    static int fr = -1;
    if (fr < 0) { const char* e = getenv("EKF_BATCH_FRONT"); fr = e ? atoi(e) : 1; }
    double* xh = h->async_x ? h->x_pin_dev : nullptr;               // the gains also write x into mapped host memory
    if (fr && h->shadow && h->Pc && m0 > 0)         // in-place deferral on a handle that has the copies buffer: same front kernel, no strip
        launch_batch_front_gain(h->stream, h->P, h->ld, nullptr, h->x, n_of(h), ob, h->blin, h->Pc, h->ld, 1, h->Kd, h->Bd, h->ld, m0, src, xh);
    else
        launch_batch_lin_gain(h->stream, h->P, h->ld, h->x, n_of(h), sym_of(h), ob, h->blin, h->Kd, h->Bd, h->ld,
                              nullptr, 0, 0, 0, m0, nullptr, nullptr, 0, src, xh);
    h->launches += 2;
    CK(cudaGetLastError());
    MARK_X(h);                      // x is final for this chunk here, before the flush
    if (xh) { h->xpin_fresh = true; h->x_head_dirty = false; }      // x_pin is x once ev_mx (just recorded) has fired
    h->m_pend = m0 + ob.k;
    if (!h->xstep || (h->m_pend >= h->pend_cap && !h->lazy)) FLUSH(h);   // lazy: a full set waits for the next batch
    return EKF_OK;
}

extern "C" int32_t ekf_update_batch(ekf_handle* h, int32_t k, const double* z, const double* Rdiag, const int32_t* idx1) {
    CKH(h);
    RESOLVE(h);
    if (k < 0 || (k > 0 && (!z || !Rdiag || !idx1))) return fail(h, EKF_EINVAL, "null argument");
    if (is_front(h)) {
        for (int j = 0; j < k; ++j)
            if (idx1[j] < 1 || idx1[j] > h->N)
                return fail(h, EKF_EINVAL, "ekf_update_batch: landmark index " + std::to_string(idx1[j]) + " outside 1.." + std::to_string(h->N));
        for (int j = 0; j < k; ++j) {
            const double zj[2] = {z[j], z[j + k]}, Rj[2] = {Rdiag[j], Rdiag[j + k]};
            int32_t rcf = ekf_update(h, zj, Rj, idx1[j]);
            if (rcf != EKF_OK) return rcf;
        }
        return h->xstep ? front_exec(h) : front_flush(h);
    }
    CK(cudaSetDevice(h->p.device));
    int i = 0;
    if (h->world == 1 && h->defer >= 2 && k >= 2) {
        // fused path: chunks of <= defer_m observations, 2 launches (k_batch_lin + k_batch_gain) + 1 flush per chunk
        for (int j = 0; j < k; ++j)
            if (idx1[j] < 1 || idx1[j] > h->N)
                return fail(h, EKF_EINVAL, "ekf_update_batch: landmark index " + std::to_string(idx1[j]) + " outside 1.." + std::to_string(h->N));
        if (!h->xstep) FLUSH(h);
        while (k - i >= 2) {
            BatchObs ob;
            ob.k = std::min(k - i, std::min(h->defer, EKF_BATCH_MAX));
            for (int q = 0; q < ob.k; ++q) {
                ob.z0[q] = z[i + q]; ob.z1[q] = z[i + q + k];
                ob.R0[q] = Rdiag[i + q]; ob.R1[q] = Rdiag[i + q + k];
                ob.a[q] = 3 + 2 * (idx1[i + q] - 1);
            }
            int32_t rcc = fused_chunk(h, ob, ObSrc{nullptr, nullptr});
            if (rcc != EKF_OK) return rcc;
            i += ob.k;
        }
    }
    for (; i < k; ++i) {
        const double zi[2] = {z[i], z[i + k]}, Ri[2] = {Rdiag[i], Rdiag[i + k]};
        int32_t rc = ekf_update(h, zi, Ri, idx1[i]);
        if (rc != EKF_OK) return rc;
    }
    if (h->xstep) return h->world > 1 ? exec_sharded(h) : EKF_OK;      // x final; the terms stay pending
    return flush(h);
}

// The whole measure() on the device when it can be (see ekf_handle::sigs): returns 1 when it enqueued the call, 0 when the
// host loop has to run (an append may be needed, a variant / mode the device path does not serve), < 0 on error.
static int32_t measure_on_device(ekf_handle* h, const double* obs, int32_t k, int32_t* action_out) {
    if (!h->am_enable || h->N == 0 || k < 1 || k > EKF_ASSOC_BATCH) return 0;
    for (int i = 0; i < k; ++i)
        if (h->sigs.find(sig_key(obs[i + 2 * k])) == h->sigs.end()) return 0;     // unseen signature: an append may be needed
    CK(cudaSetDevice(h->p.device));
    memcpy(h->am_obs_h, obs, sizeof(double) * 3 * k);
    // decisions on the side stream, behind the last writer of s (like the batched association), not behind a flush
    CK(cudaStreamWaitEvent(h->sx, h->ev_mx, 0));
    CK(cudaMemcpyAsync(h->am_obs_d, h->am_obs_h, sizeof(double) * 3 * k, cudaMemcpyHostToDevice, h->sx));
    launch_measure_decide(h->sx, h->s, h->N, h->am_obs_d, k, h->p.Rc[0], h->p.Rc[1], h->p.s_cost, h->p.s_thresh, h->am_chunks,
                          h->am_skip, h->am_status_d, h->am_scratch, h->am_ticket);
    h->launches += 1;
    CK(cudaGetLastError());
    CK(cudaEventRecord(h->ev_dec, h->sx));
    CK(cudaStreamWaitEvent(h->stream, h->ev_dec, 0));
    if (h->sg) CK(cudaStreamWaitEvent(h->sg, h->ev_dec, 0));
    if (!h->xstep) FLUSH(h);
    for (int c = 0; c * EKF_BATCH_MAX < k; ++c) {                      // rows 8c .. 8c+7, in order (EKF_SLAM_UC.m:109)
        BatchObs ob;
        memset(&ob, 0, sizeof(ob));
        ob.k = std::min(EKF_BATCH_MAX, k - c * EKF_BATCH_MAX);
        int32_t rc = fused_chunk(h, ob, ObSrc{h->am_chunks + c, h->am_skip});
        if (rc != EKF_OK) return rc;
    }
    h->am_pending = true;
    h->am_k = k;
    h->am_act_out = action_out;
    return 1;
}

extern "C" int32_t ekf_measure_wait(ekf_handle* h) {
    CKH(h);
    if (is_front(h)) return EKF_OK;
    return resolve_measure(h);
}

static int32_t measure_host_loop(ekf_handle* h, const double* obs, int32_t k, const double u[2], const double* loc_table,
                                 const int32_t* index_table, int32_t n_table, int32_t* action_out);

// measure(): EKF_SLAM.m:105-150 / EKF_SLAM_UC.m:107-151.  _async: the actions (and a possible error of the device path)
// arrive at the next call on the handle or at ekf_measure_wait; action_out must stay valid until then.
extern "C" int32_t ekf_measure_batch_async(ekf_handle* h, const double* obs, int32_t k, const double u[2],
                                           const double* loc_table, const int32_t* index_table, int32_t n_table,
                                           int32_t* action_out) {
    CKH(h);
    if (k < 0 || (k > 0 && !obs) || !u) return fail(h, EKF_EINVAL, "null argument");
    if (!is_front(h)) {
        RESOLVE(h);
        int32_t rd = measure_on_device(h, obs, k, action_out);
        if (rd != 0) {
            if (rd < 0) return rd;
            return h->xstep ? EKF_OK : flush(h);
        }
    }
    return measure_host_loop(h, obs, k, u, loc_table, index_table, n_table, action_out);
}

extern "C" int32_t ekf_measure_batch(ekf_handle* h, const double* obs, int32_t k, const double u[2],
                                     const double* loc_table, const int32_t* index_table, int32_t n_table,
                                     int32_t* action_out) {
    int32_t rc = ekf_measure_batch_async(h, obs, k, u, loc_table, index_table, n_table, action_out);
    if (rc != EKF_OK) return rc;
    return ekf_measure_wait(h);
}

// measure() control flow on the host (appends, mode-B association, the known-correspondence variant, sharded handles).
static int32_t measure_host_loop(ekf_handle* h, const double* obs, int32_t k, const double u[2], const double* loc_table,
                                 const int32_t* index_table, int32_t n_table, int32_t* action_out) {
    auto loc_by_index = [&](double want, double pos[2]) -> bool {          // find([..index]==want), first hit
        for (int i = 0; i < n_table; ++i)
            if ((double)index_table[i] == want) { pos[0] = loc_table[i]; pos[1] = loc_table[i + n_table]; return true; }
        return false;
    };
    // Updates whose slot is already decided are queued and applied through ekf_update_batch (fused linearise/gain,
    // one rank-2m pass over P per defer_m of them); an append drains the queue first, so the order of EKF_SLAM.m:107 holds.
    std::vector<double> qz0, qz1, qr0, qr1;
    std::vector<int32_t> qidx;
    auto drain = [&]() -> int32_t {
        const int m = (int)qidx.size();
        if (m == 0) return EKF_OK;
        std::vector<double> zz(2 * m), rr(2 * m);
        for (int i = 0; i < m; ++i) { zz[i] = qz0[i]; zz[i + m] = qz1[i]; rr[i] = qr0[i]; rr[i + m] = qr1[i]; }
        int32_t rc = ekf_update_batch(h, m, zz.data(), rr.data(), qidx.data());
        qz0.clear(); qz1.clear(); qr0.clear(); qr1.clear(); qidx.clear();
        return rc;
    };
    auto queue_update = [&](const double z[3], const double Rd[2], int32_t idx) -> int32_t {
        if (idx < 1 || idx > h->N)
            return fail(h, EKF_EINVAL, "ekf_update: landmark index " + std::to_string(idx) + " outside 1.." + std::to_string(h->N));
        qz0.push_back(z[0]); qz1.push_back(z[1]); qr0.push_back(Rd[0]); qr1.push_back(Rd[1]); qidx.push_back(idx);
        return EKF_OK;
    };
    // The reference applies every earlier row before MATLAB raises on a later one: an early error return first applies
    // the updates queued so far, so that the state after a failed measure() equals "rows 1..ii-1 done" on every path.
    auto bail = [&](int32_t code, const std::string& msg) -> int32_t {
        drain();
        if (!h->xstep) flush(h);
        return fail(h, code, msg);
    };
    // Mode-A associations of rows [from, k) against the current s, <= EKF_ASSOC_BATCH per launch + one D2H copy.
    const bool batch_assoc = h->p.variant == EKF_VARIANT_UC && h->p.assoc_mode == EKF_ASSOC_SIGNATURE;
    int assoc_from = 0, assoc_to = 0;                                      // h_res[j - assoc_from] valid for j in [from, to)
    ekf_handle* dv = is_front(h) ? h->subs[0] : h;                         // device-side resources (a multi front: s is replicated on every slab)
    auto associate_ahead = [&](int from) -> int32_t {
        AssocBatchZ zb;
        const int cnt = std::min(k - from, EKF_ASSOC_BATCH);
        for (int j = 0; j < cnt; ++j) zb.z2[j] = obs[from + j + 2 * k];
        CK(cudaSetDevice(dv->p.device));
        cudaStream_t st = dv->stream;
        if (dv->async_x) {                  // mode A reads only s (written by append): no need to wait for a flush
            st = dv->sx;
            CK(cudaStreamWaitEvent(st, dv->ev_mx, 0));
        }
        launch_associate_sig_batch(st, dv->s, dv->N, zb, cnt, h->p.s_cost, h->p.s_thresh, dv->d_res);
        dv->launches += 1;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(dv->h_res, dv->d_res, sizeof(AssocResult) * cnt, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        assoc_from = from; assoc_to = from + cnt;
        return EKF_OK;
    };
    for (int ii = 1; ii <= k; ++ii) {                                      // :107
        const double z[3] = {obs[ii - 1], obs[ii - 1 + k], obs[ii - 1 + 2 * k]};
        const double Rd[2] = {z[0] * h->p.Rc[0], z[1] * h->p.Rc[1]};       // :108
        double pos[2];
        int32_t rc = EKF_OK, act;
        if (h->N == 0) {                                                   // :110  length(x) < 4
            bool ok = false;                                               // find([..index]) -> first non-zero index (F15)
            for (int i = 0; i < n_table && !ok; ++i)
                if (index_table[i] != 0) { pos[0] = loc_table[i]; pos[1] = loc_table[i + n_table]; ok = true; }
            if (!ok) return bail(EKF_EINVAL, "measure: no indexed landmark in the table (EKF_SLAM.m:111)");
            rc = ekf_append(h, u, Rd, pos, 1.0);                           // :111
            assoc_to = 0;                                                  // s changed
            act = -1;
        } else if (h->p.variant == EKF_VARIANT_KNOWN) {
            if (z[2] > (double)h->N) {                                     // :118
                if (!loc_by_index(z[2], pos)) return bail(EKF_EINVAL, "measure: landmark index not in table (EKF_SLAM.m:120)");
                if ((rc = drain()) != EKF_OK) return rc;
                rc = ekf_append(h, u, Rd, pos, z[2]);                      // :120
                act = -(int32_t)z[2];
            } else {
                rc = queue_update(z, Rd, ii);                              // :123  idx = ii (row number)
                act = ii;
            }
        } else {
            int32_t newLL = 0, idx = 0;
            if (batch_assoc) {
                if (ii - 1 >= assoc_to || ii - 1 < assoc_from) {
                    if ((rc = associate_ahead(ii - 1)) != EKF_OK) return rc;
                }
                newLL = dv->h_res[ii - 1 - assoc_from].newLL;              // EKF_SLAM_UC.m:119
                idx = dv->h_res[ii - 1 - assoc_from].index1;
            } else {
                if ((rc = drain()) != EKF_OK) return rc;                   // mode B scores against the current x, P
                rc = ekf_associate(h, z, Rd, &newLL, &idx);                // EKF_SLAM_UC.m:119
                if (rc != EKF_OK) return rc;
            }
            if (newLL) {                                                   // :121
                if (!loc_by_index((double)idx, pos)) return bail(EKF_EINVAL, "measure: landmark index not in table (EKF_SLAM_UC.m:123)");
                if ((rc = drain()) != EKF_OK) return rc;
                rc = ekf_append(h, u, Rd, pos, (double)idx);               // :123
                assoc_to = 0;                                              // s changed: later rows are re-associated
                act = -idx;
            } else {
                rc = queue_update(z, Rd, idx);                             // :125-146
                act = idx;
            }
        }
        if (rc != EKF_OK) {                                                // rows before this one are applied, like the reference
            const std::string msg = h->err;
            drain();
            if (!h->xstep) flush(h);
            h->err = msg;
            return rc;
        }
        if (action_out) action_out[ii - 1] = act;
    }
    {
        int32_t rc = drain();
        if (rc != EKF_OK) return rc;
    }
    if (h->xstep) return (h->world > 1 && !is_front(h)) ? exec_sharded(h) : EKF_OK;   // cross-step deferral: P is whole again at the next getter
    return flush(h);                                                       // P is whole again when measure() returns
}

// ---------------------------------------------------------------- state access
extern "C" int32_t ekf_get_x(ekf_handle* h, double* x, int32_t len) {
    CKH(h);
    RESOLVE(h);
    if (!x || len < n_of(h)) return fail(h, EKF_EINVAL, "ekf_get_x: buffer too small");
    if (is_front(h)) {                    // every slab executes its queued batch (they wait for each other's columns), then slab 0 answers
        int32_t rcq = h->xstep ? front_exec(h) : front_flush(h);
        if (rcq != EKF_OK) return rcq;
        for (size_t g = 1; g < h->subs.size(); ++g) {            // a full pending set: see below (slab 0 does its own inside get_x)
            ekf_handle* sh = h->subs[g];
            if (sh->lazy && sh->m_pend >= sh->pend_cap) { cudaSetDevice(sh->p.device); rcq = downdate_sharded(sh); if (rcq != EKF_OK) { h->err = sh->err; return rcq; } }
        }
        int32_t rcf = ekf_get_x(h->subs[0], x, len);
        if (rcf != EKF_OK) h->err = h->subs[0]->err;
        return rcf;
    }
    CK(cudaSetDevice(h->p.device));
    if (h->world > 1) {                   // sharded: x of queued observations is produced by the batch
        int32_t rcq = h->xstep ? exec_sharded(h) : flush_sharded(h);
        if (rcq != EKF_OK) return rcq;
    }
    if (h->async_x && h->x_pin && h->world == 1 && (h->x_host || h->xpin_fresh) && (!h->x_head_dirty || h->pose_fresh)) {
        // No copy at all: the body of x is in pinned memory already -- from the last read (x_host), or because the gain
        // kernel of the last batch wrote it there beside the device copy (xpin_fresh: wait for that kernel's event) -- and
        // if a predict has moved the pose since, its state kernel wrote the new pose into mapped memory too.
        if (h->xpin_fresh) CK(cudaEventSynchronize(h->ev_mx));
        if (h->x_head_dirty) {
            CK(cudaEventSynchronize(h->ev_px));
            h->x_pin[0] = h->pose_host[0]; h->x_pin[1] = h->pose_host[1]; h->x_pin[2] = h->pose_host[2];
        }
        memcpy(x, h->x_pin, sizeof(double) * n_of(h));
        h->x_host = true;
        h->xpin_fresh = false;
        h->x_head_dirty = false;
        return after_sync(h);
    }
    cudaStream_t st = h->stream;
    if (h->async_x) {                     // behind the last writer of x, not behind the flush that follows it
        st = h->sx;
        CK(cudaStreamWaitEvent(st, h->ev_mx, 0));
    }
    double* dst = h->x_pin ? h->x_pin : x;
    const bool head_only = h->x_pin && h->x_host && h->world == 1;         // only predict has touched x since the last read
    if (!head_only) CK(cudaMemcpyAsync(dst, h->x, sizeof(double) * n_of(h), cudaMemcpyDeviceToHost, st));
    else if (h->x_head_dirty) CK(cudaMemcpyAsync(dst, h->x, sizeof(double) * 3, cudaMemcpyDeviceToHost, st));
    if (h->lazy && h->m_pend >= h->pend_cap) {
        // The caller is about to block on x while a full set of terms is pending: start its K3 pass now, so that it runs
        // during the host's turnaround instead of waiting (lazily) for the next batch.  A loop that never reads x keeps the
        // lazy overlap; a loop that reads x every step (the reference's, SLAM.m:110-116) keeps the device busy this way.
        int32_t rcl = (h->world > 1) ? downdate_sharded(h) : flush(h);
        if (rcl != EKF_OK) return rcl;
    }
    CK(cudaStreamSynchronize(st));
    if (h->x_pin) memcpy(x, h->x_pin, sizeof(double) * n_of(h));
    h->x_host = h->x_pin != nullptr && h->world == 1;
    h->x_head_dirty = false;
    return after_sync(h);
}

extern "C" int32_t ekf_get_s(ekf_handle* h, double* s, int32_t len) {
    CKH(h);
    RESOLVE(h);
    if (h->N == 0) return EKF_OK;
    if (!s || len < h->N) return fail(h, EKF_EINVAL, "ekf_get_s: buffer too small");
    if (is_front(h)) { int32_t rcf = ekf_get_s(h->subs[0], s, len); if (rcf != EKF_OK) h->err = h->subs[0]->err; return rcf; }
    CK(cudaSetDevice(h->p.device));
    CK(cudaMemcpyAsync(s, h->s, sizeof(double) * h->N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return after_sync(h);
}

extern "C" int32_t ekf_get_P(ekf_handle* h, double* P, int32_t lda) {
    CKH(h);
    RESOLVE(h);
    const int n = n_of(h);
    if (!P || lda < n) return fail(h, EKF_EINVAL, "ekf_get_P: lda < n");
    if (is_front(h)) {                    // assemble the slabs; SYM slabs hold rows <= column: mirror on the host
        int32_t rcf = ekf_get_P_block(h, 0, 0, n, n, P, lda);
        if (rcf != EKF_OK) return rcf;
        if (sym_of(h))
            for (int j = 0; j < n; ++j)
                for (int i = j + 1; i < n; ++i) P[(size_t)i + (size_t)j * lda] = P[(size_t)j + (size_t)i * lda];
        return EKF_OK;
    }
    if (h->world > 1) return fail(h, EKF_EINVAL, "ekf_get_P: sharded filter, read the local slab with ekf_get_P_block");
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);
    if (sym_of(h)) { launch_mirror_upper(h->stream, h->P, h->ld, n); h->launches += 1; CK(cudaGetLastError()); }
    CK(cudaMemcpy2DAsync(P, sizeof(double) * lda, h->P, sizeof(double) * h->ld, sizeof(double) * n, n,
                         cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return after_sync(h);
}

extern "C" int32_t ekf_get_P_block(ekf_handle* h, int32_t r0, int32_t c0, int32_t nr, int32_t nc, double* out,
                                   int32_t ldo) {
    CKH(h);
    RESOLVE(h);
    const int n = n_of(h);
    if (!out || r0 < 0 || c0 < 0 || nr < 1 || nc < 1 || r0 + nr > n || c0 + nc > n || ldo < nr)
        return fail(h, EKF_EINVAL, "ekf_get_P_block: bad block");
    if (is_front(h)) {                    // every slab delivers the columns it stores
        { int32_t rcq = front_flush(h); if (rcq != EKF_OK) return rcq; }
        for (ekf_handle* sh : h->subs) {
            const int lo = std::max(c0, sh->c_lo), hi = std::min(c0 + nc, std::min(sh->c_hi, n));
            if (hi <= lo) continue;
            int32_t rcf = ekf_get_P_block(sh, r0, lo, nr, hi - lo, out + (size_t)(lo - c0) * ldo, ldo);
            if (rcf != EKF_OK) { h->err = sh->err; return rcf; }
        }
        return EKF_OK;
    }
    if (c0 < h->c_lo || c0 + nc > h->c_hi)
        return fail(h, EKF_EINVAL, "ekf_get_P_block: columns outside this rank's slab [" + std::to_string(h->c_lo) + "," +
                                       std::to_string(h->c_hi) + ")");
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);
    if (sym_of(h) && h->world == 1) { launch_mirror_upper(h->stream, h->P, h->ld, n); h->launches += 1; CK(cudaGetLastError()); }
    CK(cudaMemcpy2DAsync(out, sizeof(double) * ldo, h->P + r0 + (int64_t)(c0 - h->c_lo) * h->ld, sizeof(double) * h->ld,
                         sizeof(double) * nr, nc, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return after_sync(h);
}

// Headless stand-in for the ellipse part of plot() (EKF_SLAM.m:180-229): numbers instead of figure calls.
extern "C" int32_t ekf_get_ellipses(ekf_handle* h, double* out, int32_t ldo) {
    CKH(h);
    RESOLVE(h);
    const int nb = h->N + 1;
    if (!out || ldo < nb) return fail(h, EKF_EINVAL, "ekf_get_ellipses: buffer too small");
    if (h->world > 1 || is_front(h)) return fail(h, EKF_EINVAL, "ekf_get_ellipses: not available on a sharded filter (diagonal blocks live on different ranks)");
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);
    double* d = nullptr;
    CK(cudaMallocAsync(&d, sizeof(double) * 5 * nb, h->stream));
    launch_ellipses(h->stream, h->P, h->ld, h->x, h->N, d, nb);
    h->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpy2DAsync(out, sizeof(double) * ldo, d, sizeof(double) * nb, sizeof(double) * nb, 5, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaFreeAsync(d, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return after_sync(h);
}

extern "C" int32_t ekf_set_state(ekf_handle* h, const double* x, const double* P, int32_t lda, const double* s,
                                 int32_t N) {
    CKH(h);
    RESOLVE(h);
    if (N < 0 || N > h->p.capacity_landmarks) return fail(h, EKF_ECAPACITY, "ekf_set_state: N beyond capacity");
    const int n = 3 + 2 * N;
    if (!x || !P || lda < n || (N > 0 && !s)) return fail(h, EKF_EINVAL, "ekf_set_state: bad argument");
    if (is_front(h)) { FAN_OUT(h, sh, ekf_set_state(sh, x, P, lda, s, N)); h->N = N; return EKF_OK; }
    CK(cudaSetDevice(h->p.device));
    h->m_pend = 0;
    { int32_t rcs = shadow_reset(h); if (rcs != EKF_OK) return rcs; }
    CK(cudaMemsetAsync(h->Kd, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
    if (h->Kd2) CK(cudaMemsetAsync(h->Kd2, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
    CK(cudaMemsetAsync(h->x, 0, sizeof(double) * h->ld, h->stream));
    CK(cudaMemcpyAsync(h->x, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    if (N > 0) CK(cudaMemcpyAsync(h->s, s, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
    {
        const int hi = std::min(h->c_hi, n);
        if (hi > h->c_lo)
            CK(cudaMemcpy2DAsync(h->P, sizeof(double) * h->ld, P + (size_t)h->c_lo * lda, sizeof(double) * lda,
                                 sizeof(double) * n, hi - h->c_lo, cudaMemcpyHostToDevice, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    MARK_X(h);
    h->q.clear();
    h->N = N;
    if (h->hflags) h->hflags[0] = 0;
    sigs_reset(h, s, N);
    return EKF_OK;
}

// Property writes h.x = v / h.P = M / h.s = v of the reference classes (public read/write, EKF_SLAM.m:6-9): one of the three
// at the current number of landmarks, the other two untouched.  Pending rank-2 terms are applied first (they belong to
// the old P); x and s are final as soon as their last kernel has run.
extern "C" int32_t ekf_set_x(ekf_handle* h, const double* x, int32_t len) {
    CKH(h);
    RESOLVE(h);
    if (!x || len != n_of(h)) return fail(h, EKF_EINVAL, "ekf_set_x: expected " + std::to_string(n_of(h)) + " elements (3 + 2 N); use ekf_set_state to change the number of landmarks");
    if (is_front(h)) {                    // enqueue every slab's flush first: a flush exchanges columns, and set_x synchronises
        { int32_t rcq = front_flush(h); if (rcq != EKF_OK) return rcq; }
        FAN_OUT(h, sh, ekf_set_x(sh, x, len));
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);                             // sharded handles: queued observations still have to read the old x
    CK(cudaStreamSynchronize(h->stream));
    if (h->sx) CK(cudaStreamSynchronize(h->sx));
    CK(cudaMemcpyAsync(h->x, x, sizeof(double) * len, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    MARK_X(h);
    return EKF_OK;
}

extern "C" int32_t ekf_set_s(ekf_handle* h, const double* s, int32_t len) {
    CKH(h);
    RESOLVE(h);
    if (len != h->N || (len > 0 && !s)) return fail(h, EKF_EINVAL, "ekf_set_s: expected " + std::to_string(h->N) + " signatures");
    if (len == 0) return EKF_OK;
    if (is_front(h)) { FAN_OUT(h, sh, ekf_set_s(sh, s, len)); return EKF_OK; }
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->sx) CK(cudaStreamSynchronize(h->sx));
    CK(cudaMemcpyAsync(h->s, s, sizeof(double) * len, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    MARK_X(h);
    sigs_reset(h, s, len);
    return EKF_OK;
}

extern "C" int32_t ekf_set_P(ekf_handle* h, const double* P, int32_t lda) {
    CKH(h);
    RESOLVE(h);
    const int n = n_of(h);
    if (!P || lda < n) return fail(h, EKF_EINVAL, "ekf_set_P: lda < n");
    if (is_front(h)) { int32_t rcq = front_flush(h); if (rcq != EKF_OK) return rcq; FAN_OUT(h, sh, ekf_set_P(sh, P, lda)); return EKF_OK; }
    CK(cudaSetDevice(h->p.device));
    FLUSH(h);
    const int hi = std::min(h->c_hi, n);  // a sharded handle keeps its own columns of the caller's full matrix
    if (hi > h->c_lo)
        CK(cudaMemcpy2DAsync(h->P, sizeof(double) * h->ld, P + (size_t)h->c_lo * lda, sizeof(double) * lda,
                             sizeof(double) * n, hi - h->c_lo, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (h->hflags) h->hflags[0] = 0;
    return EKF_OK;
}

extern "C" int32_t ekf_set_state_lowrank(ekf_handle* h, const double* x, const double* s, int32_t N, const double* A,
                                         int32_t r, const double* d) {
    CKH(h);
    RESOLVE(h);
    if (N < 0 || N > h->p.capacity_landmarks) return fail(h, EKF_ECAPACITY, "ekf_set_state_lowrank: N beyond capacity");
    const int n = 3 + 2 * N;
    if (!x || !A || !d || r < 1 || (N > 0 && !s)) return fail(h, EKF_EINVAL, "ekf_set_state_lowrank: bad argument");
    if (is_front(h)) { FAN_OUT(h, sh, ekf_set_state_lowrank(sh, x, s, N, A, r, d)); h->N = N; return EKF_OK; }
    CK(cudaSetDevice(h->p.device));
    h->m_pend = 0;
    { int32_t rcs = shadow_reset(h); if (rcs != EKF_OK) return rcs; }
    double *dA = nullptr, *dd = nullptr;
    int32_t rc = [&]() -> int32_t {               // scratch freed on every path
        CK(cudaMalloc(&dA, sizeof(double) * (size_t)n * r));
        CK(cudaMalloc(&dd, sizeof(double) * n));
        CK(cudaMemcpyAsync(dA, A, sizeof(double) * (size_t)n * r, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(dd, d, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemsetAsync(h->Kd, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
        if (h->Kd2) CK(cudaMemsetAsync(h->Kd2, 0, sizeof(double2) * h->ld * EKF_MAX_PEND, h->stream));
        CK(cudaMemsetAsync(h->x, 0, sizeof(double) * h->ld, h->stream));
        CK(cudaMemcpyAsync(h->x, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
        if (N > 0) CK(cudaMemcpyAsync(h->s, s, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
        launch_init_lowrank(h->stream, h->P, h->ld, n, h->c_lo, std::min(h->c_hi, n) - h->c_lo, dA, r, dd);
        h->q.clear();
        h->launches += 1;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
        return mark_x(h);
    }();
    cudaFree(dA); cudaFree(dd);
    if (rc != EKF_OK) return rc;
    h->N = N;
    if (h->hflags) h->hflags[0] = 0;
    sigs_reset(h, s, N);
    return EKF_OK;
}

// ---------------------------------------------------------------- measurement hooks
extern "C" int32_t ekf_profile_enable(ekf_handle* h, int32_t on) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) { FAN_OUT(h, sh, ekf_profile_enable(sh, on)); return EKF_OK; }
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    h->prof = on != 0;
    if (h->prof) {          // create the event pool here, outside anybody's timed region (cudaEventCreate costs microseconds)
        CK(cudaSetDevice(h->p.device));
        while (h->ev.size() < 8192) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->ev.push_back(e); }
        while (h->world > 1 && h->xev.size() < 2048) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->xev.push_back(e); }
    }
    h->ev_used = 0;
    h->xev_used = 0;
    h->xch_ms = 0.0;
    h->dd_ms = 0.0;
    h->dd_launches = 0;
    h->dd_terms = 0;
    h->launches = 0;
    return EKF_OK;
}

extern "C" int32_t ekf_profile_read(ekf_handle* h, double* downdate_ms_total, int64_t* downdate_launches,
                                    int64_t* kernel_launches_total) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) {                    // the slowest slab's downdate time, launches summed over the slabs
        double worst = -1.0; int64_t nd = 0, nl = 0;
        for (size_t g = 0; g < h->subs.size(); ++g) {
            double ms = 0.0; int64_t d1 = 0, l1 = 0;
            int32_t rcf = ekf_profile_read(h->subs[g], &ms, &d1, &l1);
            if (rcf != EKF_OK) { h->err = h->subs[g]->err; return rcf; }
            if (ms > worst) { worst = ms; nd = d1; h->rank = (int)g; }      // front->rank: slab the roofline numbers refer to
            nl += l1;
        }
        if (downdate_ms_total) *downdate_ms_total = worst;
        if (downdate_launches) *downdate_launches = nd;
        if (kernel_launches_total) *kernel_launches_total = nl;
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->sf) CK(cudaStreamSynchronize(h->sf));
    for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
        h->dd_ms += ms;
    }
    if (h->cev_used >= 3) {
        double tf = 0.0, tg = 0.0, tf_max = 0.0, tg_max = 0.0; int nc = 0;
        for (size_t i = 0; i + 2 < h->cev_used; i += 3) {
            float a = 0.f, b = 0.f;
            if (cudaEventElapsedTime(&a, h->cev[i], h->cev[i + 1]) == cudaSuccess && cudaEventElapsedTime(&b, h->cev[i + 1], h->cev[i + 2]) == cudaSuccess) {
                tf += a; tg += b; tf_max = a > tf_max ? a : tf_max; tg_max = b > tg_max ? b : tg_max; ++nc;
            }
        }
        fprintf(stderr, "ekfslam_b200: %d batches: k_batch_front mean %.1f us (max %.1f), k_batch_gain mean %.1f us (max %.1f)\n", nc,
                nc ? 1e3 * tf / nc : 0.0, 1e3 * tf_max, nc ? 1e3 * tg / nc : 0.0, 1e3 * tg_max);
        h->cev_used = 0;
    }
    if (getenv("EKF_PROFILE_GAPS") && h->ev_used >= 4) {       // debug aid: idle time between consecutive K3 passes
        double gap = 0.0, gmax = 0.0; int ng = 0;
        for (size_t i = 2; i + 1 < h->ev_used; i += 2) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, h->ev[i - 1], h->ev[i]) == cudaSuccess) { gap += ms; gmax = ms > gmax ? ms : gmax; ++ng; }
        }
        fprintf(stderr, "ekfslam_b200: %d K3 passes, mean gap between passes %.4f ms (max %.4f)\n", ng + 1, ng ? gap / ng : 0.0, gmax);
    }
    h->ev_used = 0;
    if (downdate_ms_total) *downdate_ms_total = h->dd_ms;
    if (downdate_launches) *downdate_launches = h->dd_launches;
    if (kernel_launches_total) *kernel_launches_total = h->launches;
    return EKF_OK;
}

extern "C" int32_t ekf_profile_exchange(ekf_handle* h, double* exchange_ms_total, int64_t* exchanges, int32_t* p2p) {
    CKH(h);
    RESOLVE(h);
    if (is_front(h)) {
        double worst = 0.0; int64_t cnt = 0; int32_t pp = 1;
        for (ekf_handle* sh : h->subs) {
            double ms = 0.0; int64_t c1 = 0; int32_t p1 = 0;
            int32_t rcf = ekf_profile_exchange(sh, &ms, &c1, &p1);
            if (rcf != EKF_OK) { h->err = sh->err; return rcf; }
            if (ms > worst) worst = ms;
            cnt = c1; pp &= p1;
        }
        if (exchange_ms_total) *exchange_ms_total = worst;
        if (exchanges) *exchanges = cnt;
        if (p2p) *p2p = pp;
        return EKF_OK;
    }
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    const int64_t cnt = (int64_t)(h->xev_used / 2);
    for (size_t i = 0; i + 1 < h->xev_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->xev[i], h->xev[i + 1]));
        h->xch_ms += ms;
    }
    h->xev_used = 0;
    if (exchange_ms_total) *exchange_ms_total = h->xch_ms;
    if (exchanges) *exchanges = cnt;
    if (p2p) *p2p = h->p2p ? 1 : 0;
    h->xch_ms = 0.0;
    return EKF_OK;
}

extern "C" int32_t ekf_downdate_bytes(ekf_handle* h, double* bytes) {
    CKH(h);
    RESOLVE(h);
    if (!bytes) return fail(h, EKF_EINVAL, "bytes is null");
    if (is_front(h)) return ekf_downdate_bytes(h->subs[(size_t)h->rank < h->subs.size() ? h->rank : 0], bytes);   // slab of the last ekf_profile_read
    const double n = (double)n_of(h);
    if (h->world > 1) {                                         // local slab
        const double lo = (double)h->c_lo, hi = lo + (double)ncols_loc(h);
        *bytes = sym_of(h) ? 8.0 * (hi * (hi + 1.0) - lo * (lo + 1.0)) : 16.0 * n * (hi - lo);
        return EKF_OK;
    }
    *bytes = sym_of(h) ? 8.0 * n * (n + 1.0) : 16.0 * n * n;   // read + write of every live element
    return EKF_OK;
}
