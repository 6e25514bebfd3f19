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
#define EKF_MAX_SLOTS (3 + 2 * EKF_MAX_DEFER)      // columns exchanged per sharded batch: robot + 2 per observation
struct SlotSet { int n; int col[EKF_MAX_SLOTS]; }; // global column index of each slot (passed by value)
void launch_linearize_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int a, double z0, double z1,
                           double R0, double R1, int sym, Lin* lin, double2* Kall, double2* Ball, int64_t lds, int m);
// n rows x ncols columns of P (square: ncols == n; a column slab passes its own width and Ball + first column)
void launch_downdate(cudaStream_t st, double* P, int64_t ld, int n, int ncols, int sym, const double2* Kall,
                     const double2* Ball, int64_t lds, int m, int col_off = 0, int variant = -1);   // col_off: global index of column 0 (SYM slabs); variant < 0: default
// fused batch (<= 8 observations with known slots; nothing may be pending): fills gain slabs 0..k-1 and updates x
void launch_batch_lin_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int sym, const BatchObs& ob,
                           BatchLin* lin, double2* Kall, double2* Ball, int64_t lds);
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
// ---- column-slab sharded filter (ekf_shard.cu) ----
void launch_pack_cols(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, const SlotSet& ss, int n,
                      double* Pc, int64_t ldc, int sym);
void launch_lin_gain_shard(cudaStream_t st, const double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, int a,
                           int sa, double z0, double z1, double R0, double R1, const SlotSet& ss, const double* Pc,
                           int64_t ldc, Lin* lin, double2* Kall, double2* Ball, int64_t lds, double2* Bc, int m, int sym);
void launch_predict_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, int n, double u1,
                          double u2, double C, int sym);
void launch_append_shard(cudaStream_t st, double* Ploc, int64_t ld, int c_lo, int c_hi, double* x, double* s, int N,
                         double u1, double u2, double R0, double R1, double px, double py, double sig, const double* Pc,
                         int64_t ldc);
void launch_mirror_upper(cudaStream_t st, double* P, int64_t ld, int n);

// downdate kernel variant: < 0 = library default (TMA-staged DMMA flush, tile shape by rank; see launch_downdate);
// the EKF_DOWNDATE_VARIANT environment variable overrides it per handle at ekf_create time.

}  // namespace ekf
