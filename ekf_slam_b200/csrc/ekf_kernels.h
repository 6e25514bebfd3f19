// ekf_kernels.h -- host-side launchers of the sm_100a kernels (definitions in ekf_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

struct Lin;
struct AssocResult;
struct BatchObs;
struct BatchLin;

void launch_predict(cudaStream_t st, double* P, int64_t ld, double* x, int n, double u1, double u2, double C, int sym);
#define EKF_MAX_DEFER 8
#define EKF_MAX_PEND 16                             // single-GPU handles: rank-2 terms that may stay pending across predict (defer_m = 9..16)
#define EKF_MAX_WINDOW (2 * EKF_MAX_PEND)            // continuous flush: the set a K3 pass is applying + the set that is filling up
#define EKF_WIN_SLOTS 256                             // slabs the window's allocation holds (the window moves through it, see ekf_capi.cu)
#define EKF_MAX_SLOTS (3 + 2 * EKF_MAX_DEFER)      // columns exchanged per sharded batch: robot + 2 per observation
struct SlotSet { int n; int col[EKF_MAX_SLOTS]; }; // global column index of each slot (passed by value)
struct ObSrc { const BatchObs* ob; const int* skip; };   // device-side batch description + skip word (both null: by-value BatchObs)
void launch_linearize_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int a, double z0, double z1,
                           double R0, double R1, int sym, Lin* lin, double2* Kall, double2* Ball, int64_t lds, int m);
// n rows x ncols columns of P (square: ncols == n; a column slab passes its own width and Ball + first column)
void launch_downdate(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                     const double2* Ball, int64_t lds, int m, int col_off = 0, int variant = -1,
                     double* Bt = nullptr, double* Pout = nullptr, bool prepacked = false);   // prepacked: Bt already holds the B operand (launch_pack_bt)
void launch_pack_bt(cudaStream_t st, const double2* Ball, int64_t lds, int m, int ncols, double* Bt);   // col_off: global index of column 0 (SYM slabs); variant < 0: default; Bt: scratch of
                                              // ncols x EKF_BT_SROW doubles for the kernels that carry the B operand in their stages
                                              // Pout != null: out of place, Pout = P - sum K_q B_q on every live tile, P untouched (continuous flush)
#define EKF_BT_SROW (4 * 8 + 4)
// fused batch (<= 8 observations with known slots; nothing may be pending): fills gain slabs 0..k-1 and updates x
// Pc != null (sharded): slot columns come from the exchanged Pc[slot][0..n), P is the local slab of columns [c_lo, c_hi)
void launch_batch_lin_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int sym, const BatchObs& ob,
                           BatchLin* lin, double2* Kall, double2* Ball, int64_t lds, const double* Pc = nullptr,
                           int64_t ldc = 0, int c_lo = 0, int c_hi = 0, int m0 = 0,    // m0: terms already pending in slabs 0..m0-1; this batch fills m0..m0+k-1
                           double2* Kout = nullptr, double2* Bout = nullptr, int o0 = 0,    // ... or slots o0.. of another pair of slabs (lazy flush)
                           ObSrc src = ObSrc{nullptr, nullptr}, double* xhost = nullptr);   // xhost: mapped host copy of x written beside x
// SYM storage: linearisation chain (block 0) beside the gather / pending-term correction of every row's slot entries (into Pc),
// then the gains with nothing pending, written to slots m0..: bit-identical to launch_batch_lin_gain(..., m0), one launch less
void launch_batch_front_gain(cudaStream_t st, const double* P, int64_t ld, const double* S, double* x, int n, const BatchObs& ob,
                             BatchLin* lin, double* Pc, int64_t ldc, int gather, double2* Kall, double2* Ball, int64_t lds, int m0,
                             ObSrc src = ObSrc{nullptr, nullptr}, double* xhost = nullptr);
void launch_batch_front_only(cudaStream_t st, const double* P, int64_t ld, const double* S, double* x, int n, const BatchObs& ob,
                             BatchLin* lin, double* Pc, int64_t ldc, int gather, double2* Kall, double2* Ball, int64_t lds, int m0,
                             ObSrc src = ObSrc{nullptr, nullptr});
void launch_batch_gain_only(cudaStream_t st, const double* P, int64_t ld, double* x, int n, const BatchObs& ob, BatchLin* lin,
                            double* Pc, int64_t ldc, double2* Kall, double2* Ball, int64_t lds, int m0, ObSrc src = ObSrc{nullptr, nullptr},
                            double* xhost = nullptr);
void launch_gather_cols(cudaStream_t st, const double* P, int64_t ld, int n, int sym, const BatchObs& ob, double* Pc, int64_t ldc,
                        ObSrc src = ObSrc{nullptr, nullptr});
// ---- continuous flush ("shadow", single GPU, SYM storage): robot strip S[4 j + r] = P0(r, j), r < 3; see ekf_kernels.cu ----
void launch_strip_load(cudaStream_t st, const double* P, int64_t ld, int c_lo, int ncols, double* S);       // column range: (0, n) on one GPU,
void launch_strip_store(cudaStream_t st, double* P, int64_t ld, int c_lo, int ncols, const double* S);      // the local columns of a slab otherwise
void launch_predict_strip(cudaStream_t st, double* S, const double* xprev, int c_lo, int ncols, double u1, double u2, double C,
                          double2* Kall, double2* Ball, int64_t lds, int m);
void launch_strip_fix(cudaStream_t st, double* S, int c_lo, int ncols, const double2* Kall, const double2* Ball, int64_t lds, int m);
void launch_gather_cols_strip(cudaStream_t st, const double* P, int64_t ld, int n, const BatchObs& ob, double* Pc, int64_t ldc,
                              ObSrc src, const double* S);
void launch_rotate_slabs(cudaStream_t st, double2* Kall, double2* Ball, int64_t lds, int rows, int mA, int mB, double2* Kf,
                         double2* Bf);
// measure() decisions on the device (EKF_SLAM_UC.m:107-124, mode A): per row of observed_LL the slot it updates; see ekf_kernels.cu
void launch_measure_decide(cudaStream_t st, const double* s, int N, const double* obs, int k, double Rc0, double Rc1,
                           double s_cost, double s_thresh, BatchObs* ob_chunks, int* skip, int* status_host, int* scratch,
                           unsigned int* ticket);
// predict with m pending terms: P0 takes F . F' + Q as usual, the slabs take K_q <- F K_q, B_q <- B_q F' (rows / columns 0, 1 only)
void launch_predict_slabs(cudaStream_t st, double2* Kall, double2* Ball, int64_t lds, int m, const double* xpre, double u1);
void launch_append(cudaStream_t st, double* P, int64_t ld, double* x, double* s, int N, double u1, double u2,
                   double R0, double R1, double px, double py, double sig, int sym);
#define EKF_ASSOC_BATCH 32
struct AssocBatchZ { double z2[EKF_ASSOC_BATCH]; };   // signatures z(3) of up to 32 observations (by value)
void launch_associate_sig_batch(cudaStream_t st, const double* s, int N, const AssocBatchZ& zb, int k, double s_cost,
                                double s_thresh, AssocResult* out);
int  assoc_blocks(int N);
void launch_associate(cudaStream_t st, const double* P, int64_t ld, const double* x, const double* s, int N,
                      double z0, double z1, double z2, double R0, double R1, double s_cost, double s_thresh,
                      int mode, int sym, void* partial, unsigned int* ticket, AssocResult* out);
// columns [c_lo, c_lo+ncols) of P = A A'/r + diag(d) into a slab whose first column is c_lo
void launch_init_lowrank(cudaStream_t st, double* P, int64_t ld, int n, int c_lo, int ncols, const double* A, int r,
                         const double* d);
void launch_predict_x(cudaStream_t st, double* x, double u1, double u2);
void launch_predict_x_save(cudaStream_t st, double* x, double u1, double u2, double* xprev, double* pose_host = nullptr);      // decoupled predict, state half (side stream)
void launch_predict_P(cudaStream_t st, double* P, int64_t ld, const double* xprev, int n, double u1, double u2, double C, int sym,
                      double2* Kall = nullptr, double2* Ball = nullptr, int64_t lds = 0, int m = 0);   // covariance half (+ the m pending slabs, one extra block)
// ---- column-slab sharded filter (ekf_shard.cu) ----
#define EKF_MAX_PEERS 8
// layout of a rank's flag allocation (ints): [0, 8) column-exchange epochs per source rank, [EKF_XF_ASSOC, +8) association
// epochs, [EKF_XF_CAND, ...) 2 x 8 candidates of 16 bytes (cost, index) for the sharded mode-B association
#define EKF_XF_ASSOC 8
#define EKF_XF_CAND 64
#define EKF_XF_INTS 256
struct PeerSet { double* pc[EKF_MAX_PEERS]; int* flags[EKF_MAX_PEERS]; int world, rank; };   // cudaIpc-mapped Pc buffers / flag words of every rank
void launch_scatter_cols(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, const SlotSet& ss, int n,
                         const PeerSet& ps, int64_t buf_off, int sym, int epoch, unsigned int* ticket,
                         const double* strip = nullptr);     // strip (continuous flush): rows 0..2 of the local columns come from it
// wait (bounded) until every rank has published `epoch`; err_word: mapped host memory, set to 1 when a peer does not answer in time
void launch_wait_flags(cudaStream_t st, const int* flags, int world, int epoch, long long timeout_cycles, int* err_word);
void launch_pack_cols(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, const SlotSet& ss, int n,
                      double* Pc, int64_t ldc, int sym, const double* strip = nullptr);
void launch_lin_gain_shard(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, int a,
                           int sa, double z0, double z1, double R0, double R1, const SlotSet& ss, const double* Pc,
                           int64_t ldc, Lin* lin, double2* Kall, double2* Ball, int64_t lds, double2* Bc, int m, int sym);
void launch_predict_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, double u1,
                          double u2, double C, int sym);
void launch_predict_shard_P(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, const double* xprev, int n,
                            double u1, double u2, double C, int sym, double2* Kall = nullptr, double2* Ball = nullptr,
                            int64_t lds = 0, int m = 0);   // covariance half of the decoupled predict (+ the m pending slabs)
void launch_append_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, double* s, int N,
                         double u1, double u2, double R0, double R1, double px, double py, double sig, const double* Pc,
                         int64_t ldc);
void launch_associate_slab(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int lm_lo, int lm_cnt, const double* x,
                           const double* s, int N, double z0, double z1, double z2, double R0, double R1, double s_cost,
                           double s_thresh, int sym, const double* Pc, int64_t ldc, void* partial, unsigned int* ticket,
                           void* cand_out);                      // mode B over the landmarks of this slab -> one 16-byte candidate
void launch_cand_send(cudaStream_t st, const void* mine, const PeerSet& ps, int epoch);
void launch_cand_fold(cudaStream_t st, const int* flags, int world, int epoch, long long timeout_cycles, int* err_word, int N,
                      AssocResult* out);
void launch_cand_fold_gathered(cudaStream_t st, const void* all, int world, int N, AssocResult* out);
void launch_mirror_upper(cudaStream_t st, double* P, int64_t ld, int n);
void launch_ellipses(cudaStream_t st, const double* P, int64_t ld, const double* x, int N, double* out, int64_t ldo);

// downdate kernel variant: < 0 = library default (TMA-staged DMMA flush, tile shape by rank; see launch_downdate);
// the EKF_DOWNDATE_VARIANT environment variable overrides it per handle at ekf_create time.

// ---- Monte-Carlo batch of independent filters (ekf_mc.cu): one CTA per filter, whole predict+measure on device ----
struct McArgs {
    double* P; int64_t ldp; int64_t pstride;     // per filter: ldp x cap_n column-major, pstride doubles apart
    double* x; int64_t xstride;
    double* s; int64_t sstride;
    int* N;                                      // landmarks per filter
    int cap;                                     // capacity_landmarks
    const double* u;                             // [nf][2]
    const double* obs; int max_obs;              // [nf][3][max_obs]  (k x 3 column-major per filter, ld = max_obs)
    const int* counts;                           // [nf] observations this step
    const double* loc; int64_t loc_stride;       // [.][2][max_table] column-major; stride 0 = one table for all filters
    const int* idx; int64_t idx_stride;          // [.][max_table]
    const int* ntab; int ntab_stride;            // [.] valid table entries
    int max_table;
    int* actions;                                // [nf][max_obs] or null: +idx update, -idx append
    int* err;                                    // [0] = first error code (0 ok), [1] = filter
    double2* Kg; double2* Bg;                    // split step: gain slabs in global memory, [nf][maxd][ldp]
    int* Mpend;                                  // split step: pending rank-2 terms per filter
    double C, Rc0, Rc1, s_cost, s_thresh;
    int variant, assoc_mode, sym, do_predict, do_measure, maxd;
    int robot_cache;                             // split front: keep the 3 robot rows of P0 in shared memory for the whole step
};

enum { MC_ERR_TABLE = 1, MC_ERR_CAPACITY = 2, MC_ERR_INDEX = 3 };
void launch_mc_step(cudaStream_t st, const McArgs& a, int n_filters, int threads, size_t smem);   // fused (one kernel)
void launch_mc_pack_poses(cudaStream_t st, const double* x, int64_t xstride, int n_filters, double* out, const int* err);   // out[3 nf] = x(1:3) of every filter, then the 2-int error word
void launch_mc_front(cudaStream_t st, const McArgs& a, int n_filters, size_t smem_limit, int front_kind);   // split: latency half (run-based k_mc_front, or k_mc_step<.,1> when front_kind = 1 / mode B / the slot cache does not fit)
void launch_mc_flush_tma(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max);  // split: bandwidth half, TMA-staged DMMA (default)
void launch_mc_flush(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max);
void launch_mc_flush_w(cudaStream_t st, const McArgs& a, int n_filters, int cap_n, int m_max, int tc = 64);   // tc: columns of a filter per CTA (32 | 64 | 128)    // split: bandwidth half, warp-per-column-group (default)      // split: bandwidth half (m_max >= every Mpend[f])

}  // namespace ekf
