/*
 * ekfslam_b200.h -- C ABI of the B200-native EKF-SLAM filter core.
 *
 * This is the drop-in boundary under the reference's MATLAB class interface
 * (SamShue/EKF_SLAM).  The reference has no FFI of its own; the seam is the
 * class API (SLAM.m:110 `h.slam.predict(h.u)`, SLAM.m:116
 * `h.slam.measure(laserData,h.u,h.LM)`, EKF_SLAM_UC.m:119
 * `correspondence.estimateCorrespondence(z,R,x,P,s)`).  Each entry point below
 * names the reference method whose body it replaces.
 *
 * Conventions
 *   - every function returns an int32 status (EKF_OK == 0, negative on error);
 *     ekf_last_error() returns the message for the last failure on a handle.
 *   - host buffers are caller-owned, fp64, MATLAB column-major; indices that
 *     cross the ABI are 1-based int32 like MATLAB's.
 *   - the library owns all device memory.  One handle = one CUDA device and one
 *     stream; a handle is NOT thread-safe, distinct handles are independent.
 *   - calls enqueue work on the handle's stream and return; getters and
 *     ekf_sync() synchronise.  There is no CPU fallback: without a CUDA device
 *     ekf_create fails with EKF_ECUDA.
 */
#ifndef EKFSLAM_B200_H
#define EKFSLAM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EKF_OK          0
#define EKF_EINVAL     (-1)   /* bad argument / bad state (e.g. idx out of range) */
#define EKF_ECAPACITY  (-2)   /* append beyond capacity_landmarks */
#define EKF_ECUDA      (-3)   /* CUDA runtime error (message in ekf_last_error) */
#define EKF_ENCCL      (-4)   /* NCCL error, or a peer rank of a sharded filter did not answer within EKF_XCHG_TIMEOUT_MS (sticky) */
#define EKF_ENONFINITE (-5)   /* an update since the last synchronising call had a non-finite innovation covariance (q = 0 at
                                 EKF_SLAM.m:137: robot on the landmark); reported once by the next getter / ekf_sync, x and P stay poisoned */

#define EKF_VARIANT_KNOWN  0  /* EKF_SLAM.m    */
#define EKF_VARIANT_UC     1  /* EKF_SLAM_UC.m */

#define EKF_ASSOC_SIGNATURE 0 /* Correspondence.m:75 as shipped (signature cost only) */
#define EKF_ASSOC_POSITION  1 /* Correspondence.m:74 (commented out upstream): position + signature */

#define EKF_STORAGE_FULL 0    /* P <- P - K (H P): every element read+written, non-symmetric form of EKF_SLAM.m:145 */
#define EKF_STORAGE_SYM  1    /* only tiles on/above the diagonal are kept live (half the HBM traffic) */

typedef struct ekf_handle ekf_handle;

/* Property defaults of the reference classes (EKF_SLAM.m:12-16, EKF_SLAM_UC.m:12-16). */
typedef struct ekf_params {
    double   C;                  /* process-noise constant, EKF_SLAM.m:12 */
    double   Rc[2];              /* measurement-noise constants, EKF_SLAM.m:13 / EKF_SLAM_UC.m:13 */
    double   s_cost;             /* Correspondence.m:16 */
    double   s_thresh;           /* Correspondence.m:17 */
    int32_t  variant;            /* EKF_VARIANT_* */
    int32_t  assoc_mode;         /* EKF_ASSOC_* */
    int32_t  storage;            /* EKF_STORAGE_* */
    int32_t  capacity_landmarks; /* P is allocated once at 3+2*capacity (reference reallocates per append, EKF_SLAM.m:91) */
    int32_t  device;             /* CUDA device ordinal */
    int32_t  defer_m;            /* <=1: one pass over P per observation; 2..8: deferred, one rank-2m pass per m observations (same sequence of linearisations, bit-identical); 9..16 (single GPU, or SYM slabs of a sharded filter): terms also stay pending across ekf_predict, which transforms them (K_q <- F K_q, B_q <- B_q F'): equal to rounding, not bit for bit; on a large map (n >= 8000, SYM storage) the rank-32 pass then runs out of place on a stream of its own beside predict and the next batches (continuous flush: ekf_flush_info, EKF_SHADOW=0 disables; needs room for a second copy of P) */
    uint64_t stream;             /* cudaStream_t to enqueue on, 0 = library creates one */
} ekf_params;

/* Fill *p with the reference's defaults for EKF_SLAM (variant 0) or EKF_SLAM_UC (variant 1). */
int32_t ekf_params_default(ekf_params* p, int32_t variant);

/* Constructor: x=[0 0 0], P=0.1*eye(3), s=[]  (EKF_SLAM.m:26-34, EKF_SLAM_UC.m:27-36). */
int32_t ekf_create(const ekf_params* p, ekf_handle** out);
int32_t ekf_destroy(ekf_handle* h);

/* One large filter across `world` GPUs (one process per GPU): rank g stores the column slab
 * [c_lo, c_hi) of the column-major P (by symmetry the "row slab" of a row-major view), x/s/gains replicated.
 * Every rank makes the same calls with the same arguments; per batch of <= defer_m observations the only
 * exchange is that of the 3 robot columns and the 2 columns of each observed landmark: by default every rank stores the
 * elements it owns straight into every peer's buffer (cudaIpc-mapped peer memory over NVLink, flag-word handshake);
 * EKF_SHARD_EXCHANGE=nccl (or no P2P path) uses NCCL instead (broadcast from the owners / all-reduce below).
 * id128 = 128 bytes from ekf_nccl_unique_id() on rank 0, distributed by the host (MPI, torch.distributed, file).
 * FULL slabs hold equal column counts and exchange by broadcast from the owners; SYM slabs hold equal live
 * area (boundaries at capacity*sqrt(g/world)) and assemble symmetric columns with one all-reduce of zero-padded
 * contributions (exact).  ekf_get_P_block reads local columns only (SYM: rows <= column are live, no
 * mirroring); association mode B is not available.  (The reference has no counterpart: EKF_SLAM.m is single-process.) */
int32_t ekf_nccl_unique_id(uint8_t* id128);
int32_t ekf_create_sharded(const ekf_params* p, int32_t rank, int32_t world, const uint8_t* id128, ekf_handle** out);
/* The same filter from ONE process over n_devices GPUs (what a MATLAB host needs: MEX runs on the interpreter's single
 * thread, SLAM.m:21-40 constructs one object in one process).  The returned handle is used exactly like a single-GPU one
 * -- every entry point of this header accepts it -- and internally owns one column-slab handle per device; the per-batch
 * column exchange goes through directly mapped peer memory (cudaDeviceEnablePeerAccess), no NCCL, no launcher.  Devices
 * must be distinct and peer-accessible (EKF_ECUDA otherwise); params.stream must be 0; devices == NULL means 0..n-1.
 * ekf_get_P / ekf_get_P_block assemble the slabs on the host; association mode B and ekf_get_ellipses are not available. */
int32_t ekf_create_multi(const ekf_params* p, int32_t n_devices, const int32_t* devices, ekf_handle** out);
int32_t ekf_shard_info(ekf_handle* h, int32_t* rank, int32_t* world, int32_t* c_lo, int32_t* c_hi);
/* Pure host helper: column range of `rank` for a given capacity (landmark-aligned, rank 0 holds the robot). */
int32_t ekf_slab_range(int32_t capacity_landmarks, int32_t world, int32_t rank, int32_t storage,
                       int32_t* c_lo, int32_t* c_hi);

/* predict(h,u): EKF_SLAM.m:40-51 (+ f, :56-65).  u = [dD, dtheta_deg]. */
int32_t ekf_predict(ekf_handle* h, const double u[2]);

/* Update branch of measure(): EKF_SLAM.m:124-145 == EKF_SLAM_UC.m:125-146.
 * z = [range, bearing_deg]; Rdiag = diag of R (EKF_SLAM.m:108); idx1 = 1-based landmark slot. */
int32_t ekf_update(ekf_handle* h, const double z[2], const double Rdiag[2], int32_t idx1);

/* The same update for k observations whose slots are known (z, Rdiag: k x 2 column-major); ends with a flush. */
int32_t ekf_update_batch(ekf_handle* h, int32_t k, const double* z, const double* Rdiag, const int32_t* idx1);

/* append(h,u,R,landmarkPos,signature): EKF_SLAM.m:67-98 (append.m:1-27). */
int32_t ekf_append(ekf_handle* h, const double u[2], const double Rdiag[2],
                   const double pos[2], double signature);

/* [newLL,index] = estimateCorrespondence(z,R,x,P,s): Correspondence.m:28-88.
 * z = [range, bearing_deg, signature].  Synchronises to return the two scalars. */
int32_t ekf_associate(ekf_handle* h, const double z[3], const double Rdiag[2],
                      int32_t* newLL, int32_t* idx1);

/* Association (K4, Correspondence.m:28-88 as ekf_associate computes it) for callers that may not allocate.
 * NO COUNTERPART IN THE REFERENCE: this entry point was added for code under CUDA
 * graph capture, tight serving loops and harnesses that forbid cudaMalloc.  It takes no handle and calls no cudaMalloc /
 * cudaFree / allocating library routine: the kernel's scratch (per-block candidates + the fold's ticket word) lives in a
 * file-scope `static __device__` array of 65 792 bytes that every GPU loading the library holds until exit, which bounds
 * the map at EKF_ASSOC_STATIC_MAX_LANDMARKS landmarks (EKF_EINVAL beyond, message in errbuf).
 *   d_x (3+2N), d_s (N), d_P (ld x (3+2N) column-major; read in mode B only, may be NULL in mode A): DEVICE pointers on the
 *   current device, the caller's; d_result: DEVICE pointer to 16 bytes, written as {int32 newLL; int32 idx1; double cost}
 *   (cost = min_log_likelihood, Inf when newLL); z = [range, bearing_deg, signature], Rdiag: host; s_cost, s_thresh,
 *   assoc_mode (EKF_ASSOC_*), storage (EKF_STORAGE_*: SYM reads only elements on/above the diagonal of P) as in ekf_params;
 *   stream: cudaStream_t on the current device (0 = default stream).  The call enqueues one 4-byte memset and one kernel
 *   and returns; it does not synchronise.  errbuf (may be NULL) receives the message of a failure.
 * Every call on a GPU shares the one scratch array: calls that overlap on the GPU (different streams or threads) race;
 * calls on one stream, or otherwise ordered, are safe back to back. */
#define EKF_ASSOC_STATIC_MAX_LANDMARKS 1048576
int32_t ekf_associate_static(const double* d_x, const double* d_P, int64_t ld, const double* d_s, int32_t N,
                             const double z[3], const double Rdiag[2], double s_cost, double s_thresh,
                             int32_t assoc_mode, int32_t storage, void* d_result, uint64_t stream,
                             char* errbuf, int32_t errlen);

/* Whole measure() loop for k observations: EKF_SLAM.m:105-150 / EKF_SLAM_UC.m:107-151.
 * observed_LL is k x 3 column-major ([range | bearing | index], ld = k).
 * loc_table (n_table x 2 column-major) and index_table (n_table) are the fields of
 * landmark_list.landmarkObj.landmark(:) the reference reads (.loc, .index).
 * action_out[i] receives +idx for an update of slot idx, -idx for an append with signature idx
 * (may be NULL). */
int32_t ekf_measure_batch(ekf_handle* h, const double* observed_LL, int32_t k,
                          const double u[2], const double* loc_table,
                          const int32_t* index_table, int32_t n_table,
                          int32_t* action_out);

/* The same call without the wait: EKF_SLAM_UC with the shipped (signature) association decides append / associate /
 * update per row ON THE DEVICE (k_measure_decide), the update kernels read the chosen slots from device memory, and the
 * call returns as soon as everything is enqueued -- no device-to-host copy inside it.  action_out (may be NULL) is filled
 * at the next call on the handle or by ekf_measure_wait, and must stay valid until then.  The library takes the device
 * path when every row's signature is already in the map (it mirrors the signatures it was given, so it knows that no row
 * can need the landmark table, which only exists during the call); otherwise -- an unseen signature, the first landmark,
 * the known-correspondence class (which needs no association and was never blocking), mode B, sharded handles -- the
 * host loop runs as in ekf_measure_batch.  ekf_measure_wait blocks for the decision kernel only, not for the updates or
 * the covariance flush behind it.  ekf_measure_batch == ekf_measure_batch_async + ekf_measure_wait.
 * EKF_ASYNC_MEASURE=0 forces the host loop. */
int32_t ekf_measure_batch_async(ekf_handle* h, const double* observed_LL, int32_t k,
                                const double u[2], const double* loc_table,
                                const int32_t* index_table, int32_t n_table,
                                int32_t* action_out);
int32_t ekf_measure_wait(ekf_handle* h);

/* State access (public properties x, P, s of the reference classes, EKF_SLAM.m:6-9). */
int32_t ekf_n_landmarks(ekf_handle* h, int32_t* N);
/* ekf_get_x returns x as of every call made before it.  It waits for the last kernel that writes x, not for the
 * O(n^2) covariance flush queued behind that kernel (single-GPU handles; EKF_ASYNC_X=0 makes it wait for the stream). */
int32_t ekf_get_x(ekf_handle* h, double* x, int32_t len);            /* len >= 3+2N */
int32_t ekf_get_s(ekf_handle* h, double* s, int32_t len);            /* len >= N */
int32_t ekf_get_P(ekf_handle* h, double* P, int32_t lda);            /* n x n column-major, lda >= n; SYM storage is mirrored */
int32_t ekf_get_P_block(ekf_handle* h, int32_t r0, int32_t c0, int32_t nr, int32_t nc,
                        double* out, int32_t ldo);                   /* 0-based sub-block (sampled parity at large n) */
int32_t ekf_set_state(ekf_handle* h, const double* x, const double* P, int32_t lda,
                      const double* s, int32_t N);
/* Property writes `h.x = v`, `h.P = M`, `h.s = v` (x, P, s are public read/write properties upstream, EKF_SLAM.m:6-9):
 * one of the three at the current number of landmarks (len = 3+2N resp. N; P is n x n column-major, lda >= n), the
 * other two untouched.  ekf_set_state replaces all three and may change N.  A sharded handle keeps its own columns of P. */
int32_t ekf_set_x(ekf_handle* h, const double* x, int32_t len);
int32_t ekf_set_P(ekf_handle* h, const double* P, int32_t lda);
int32_t ekf_set_s(ekf_handle* h, const double* s, int32_t len);
/* Populated synthetic map without a 3.2 GB H2D copy: P = A A^T / r + diag(d), A is n x r column-major. */
int32_t ekf_set_state_lowrank(ekf_handle* h, const double* x, const double* s, int32_t N,
                              const double* A, int32_t r, const double* d);
/* plot(h, landmark_list), EKF_SLAM.m:154-234, minus the drawing: the covariance ellipses it would draw (:180-229), as
 * (N+1) x 5 column-major rows [cx, cy, semi-axis along the major eigenvector, semi-axis across, arc_rad]; row 0 is the
 * robot (scale .25, :198), row b the landmark b (scale .50, :221); chi_square = 2.2788 as upstream. */
int32_t ekf_get_ellipses(ekf_handle* h, double* out, int32_t ldo);
/* Property writes (public, mutable after construction in the reference). */
int32_t ekf_set_noise(ekf_handle* h, double C, const double Rc[2], double s_cost, double s_thresh);

/* What the handle does with defer_m: *batch = observations linearised per fused batch (<= 8), *pend_cap = rank-2 terms
 * that may be pending before a pass over P is forced (> 8: cross-step deferral is on). */
int32_t ekf_defer_info(ekf_handle* h, int32_t* batch, int32_t* pend_cap);
/* Whether the handle runs the continuous flush at its current size (the K3 pass of a full set of terms out of place on a
 * stream of its own, beside predict and the next batches: SYM storage, defer_m = 9..16, n >= EKF_SHADOW_MIN_N (8000), room for
 * a second copy of P / of the largest slab; EKF_SHADOW=0 disables), and how many passes are outstanding on that stream right
 * now (0..2).  No counterpart in the reference (its P = (I - K H) P is one dense statement, EKF_SLAM.m:145); bench.py prints it. */
int32_t ekf_flush_info(ekf_handle* h, int32_t* continuous, int32_t* passes_outstanding);
/* Apply any deferred (defer_m > 1) rank-2 terms to P now.  predict, append, getters of P and the end of
 * ekf_measure_batch flush implicitly; ekf_update flushes when defer_m terms are pending.  With defer_m = 9..16 predict and
 * the end of ekf_measure_batch / ekf_update_batch do NOT flush (append, mode-B association and the getters of P still do). */
int32_t ekf_flush(ekf_handle* h);
int32_t ekf_sync(ekf_handle* h);
const char* ekf_last_error(ekf_handle* h);

/* Measurement hooks used by bench.py: CUDA-event timing of the covariance downdate kernel. */
int32_t ekf_profile_enable(ekf_handle* h, int32_t on);
int32_t ekf_profile_read(ekf_handle* h, double* downdate_ms_total, int64_t* downdate_launches,
                         int64_t* kernel_launches_total);
/* Sharded filters: CUDA-event time of the column exchanges since profiling was enabled (and whether the peer-memory
 * path is in use, 0 = NCCL collectives).  Resets the accumulated time. */
int32_t ekf_profile_exchange(ekf_handle* h, double* exchange_ms_total, int64_t* exchanges, int32_t* p2p);
/* Bytes the downdate kernel must move per launch at the current n and storage. */
int32_t ekf_downdate_bytes(ekf_handle* h, double* bytes);

/* ---------------------------------------------------------------------------------------------------------------
 * Monte-Carlo batch: n_filters independent filters of one class (BASELINE config 5: 4096 x EKF_SLAM_UC x 200
 * landmarks).  Replaces a host loop over n_filters reference objects, each doing
 *     slam.predict(u); slam.measure(laserData, u, LM)        (SLAM.m:110,116)
 * One CUDA block owns one filter and runs the whole iteration on the device (association, append / update decision,
 * deferred covariance downdate: one pass over that filter's P per measure()).  Filters never communicate; across GPUs
 * the caller gives each rank its own share of the filters (filter f on rank f mod G) -- no collective.
 * Per-filter inputs are dense arrays, filter-major:
 *   u            n_filters x 2                          [dD, dtheta_deg] per filter
 *   observed_LL  n_filters x (max_obs x 3 column-major) rows [range | bearing_deg | index], counts[f] rows valid
 *   loc_table    (n_filters or 1) x (max_table x 2 column-major), index_table (n_filters or 1) x max_table,
 *   n_table      (n_filters or 1): landmark_list.landmarkObj.landmark(:).{loc,index};  shared_table != 0: one table for all
 * mem = 0: host pointers (copied to the device inside the call);  mem = 1: device pointers (inputs already resident).
 * actions_out (host, n_filters x max_obs, may be NULL) receives +idx for an update, -idx for an append; asking for it
 * synchronises.  Errors inside a filter (landmark not in the table, capacity, slot out of range) surface at the next
 * ekfb_sync / getter as EKF_EINVAL / EKF_ECAPACITY with the filter number in ekfb_last_error. */
typedef struct ekfb_handle ekfb_handle;
int32_t ekfb_create(const ekf_params* p, int32_t n_filters, int32_t max_obs, int32_t max_table, ekfb_handle** out);
int32_t ekfb_destroy(ekfb_handle* h);
int32_t ekfb_step(ekfb_handle* h, int32_t mem, const double* u, const double* observed_LL, const int32_t* counts,
                  const double* loc_table, const int32_t* index_table, const int32_t* n_table, int32_t shared_table,
                  int32_t do_predict, int32_t do_measure, int32_t* actions_out);
int32_t ekfb_sync(ekfb_handle* h);
int32_t ekfb_n_landmarks(ekfb_handle* h, int32_t* N /* n_filters */);
int32_t ekfb_get_x(ekfb_handle* h, int32_t filter, double* x, int32_t len);
int32_t ekfb_get_s(ekfb_handle* h, int32_t filter, double* s, int32_t len);
int32_t ekfb_get_P(ekfb_handle* h, int32_t filter, double* P, int32_t lda);      /* SYM storage is mirrored on the host side copy */
int32_t ekfb_set_state(ekfb_handle* h, int32_t filter /* < 0: every filter */, const double* x, const double* P, int32_t lda, const double* s, int32_t N);
/* Poses as of every step made before the call; after a split step it returns when the front kernel of that step is done
 * (the flush does not touch x) and reports that step's device-side errors.  EKF_MC_EARLY_POSES=0: waits for the stream. */
int32_t ekfb_get_poses(ekfb_handle* h, double* poses /* n_filters x 3, filter-major: x(1:3) of every filter */);
/* Algorithmic HBM bytes of the covariance pass of one step over all filters at their current sizes. */
int32_t ekfb_downdate_bytes(ekfb_handle* h, double* bytes);
/* CUDA-event time of the step kernels (front + flush) and of the bandwidth-bound flush kernel alone; enable < 0 only reads. */
int32_t ekfb_profile(ekfb_handle* h, int32_t enable, double* step_ms_total, int64_t* steps, double* flush_ms_total);
const char* ekfb_last_error(ekfb_handle* h);

/* Library identification. */
const char* ekf_version(void);

#ifdef __cplusplus
}
#endif
#endif /* EKFSLAM_B200_H */
