// ekf_kernels.h -- host-side launchers of the sm_100a kernels (definitions in ekf_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ekf {

struct Lin;
struct AssocResult;

void launch_predict(cudaStream_t st, double* P, int64_t ld, double* x, int n, double u1, double u2, double C, int sym);
#define EKF_MAX_DEFER 8
void launch_linearize_gain(cudaStream_t st, const double* P, int64_t ld, double* x, int n, int a, double z0, double z1,
                           double R0, double R1, int sym, Lin* lin, double2* Kall, double2* Ball, int64_t lds, int m,
                           double* Bpk);
void launch_downdate(cudaStream_t st, double* P, int64_t ld, int n, int sym, const double2* Kall, const double2* Ball,
                     int64_t lds, int m, const double* Bpk);
void launch_append(cudaStream_t st, double* P, int64_t ld, double* x, double* s, int N, double u1, double u2,
                   double R0, double R1, double px, double py, double sig, int sym);
int  assoc_blocks(int N);
void launch_associate(cudaStream_t st, const double* P, int64_t ld, const double* x, const double* s, int N,
                      double z0, double z1, double z2, double R0, double R1, double s_cost, double s_thresh,
                      int mode, int sym, void* partial, unsigned int* ticket, AssocResult* out);
void launch_init_lowrank(cudaStream_t st, double* P, int64_t ld, int n, const double* A, int r, const double* d);
void launch_mirror_upper(cudaStream_t st, double* P, int64_t ld, int n);

extern int g_downdate_cch;       // columns per work unit of the tensor-core flush (EKF_DOWNDATE_CCH)
extern int g_downdate_variant;   // tuning knob (EKF_DOWNDATE_VARIANT env), see launch_downdate

}  // namespace ekf
