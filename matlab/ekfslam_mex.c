/*
 * ekfslam_mex.c -- MEX gateway: MATLAB / GNU Octave  ->  C ABI of libekfslam_b200.so (include/ekfslam_b200.h).
 *
 *   out = ekfslam_mex('verb', uint64 handle, args...)
 *
 * Pure marshalling: every verb unpacks mxArrays into plain pointers/sizes and makes exactly one C-ABI call.
 * No filter arithmetic lives here.  A non-zero status becomes mexErrMsgIdAndTxt('ekfslam:<code>', message),
 * which is how the MATLAB classes in this directory (EKF_SLAM.m, EKF_SLAM_UC.m, Correspondence.m) surface
 * errors; the reference itself has no error convention.
 *
 * Build (where MATLAB or Octave exists; neither is in this image):
 *   MATLAB : mex -I../include ekfslam_mex.c -L../ekf_slam_b200 -lekfslam_b200
 *   Octave : mkoctfile --mex -I../include ekfslam_mex.c -L../ekf_slam_b200 -lekfslam_b200
 * In this repo it is compile-checked against tests/mex_stub/mex.h (a minimal stand-in declaring only the
 * mx/mex functions used below) by tests/test_matlab_shim.py, and exercised through that stand-in.
 */
#include <string.h>
#include <stdint.h>
#include "mex.h"
#include "ekfslam_b200.h"

static ekf_handle* get_handle(const mxArray* a) {
    if (!mxIsUint64(a) || mxGetNumberOfElements(a) != 1)
        mexErrMsgIdAndTxt("ekfslam:EINVAL", "handle must be a uint64 scalar");
    return (ekf_handle*)(uintptr_t)(*(uint64_t*)mxGetData(a));
}

static const char* code_name(int32_t rc) {
    switch (rc) {
        case EKF_EINVAL: return "ekfslam:EINVAL";
        case EKF_ECAPACITY: return "ekfslam:ECAPACITY";
        case EKF_ECUDA: return "ekfslam:ECUDA";
        case EKF_ENCCL: return "ekfslam:ENCCL";
        case EKF_ENONFINITE: return "ekfslam:ENONFINITE";
        default: return "ekfslam:EUNKNOWN";
    }
}

static void check(ekf_handle* h, int32_t rc) {
    if (rc != EKF_OK) mexErrMsgIdAndTxt(code_name(rc), "%s", h ? ekf_last_error(h) : "ekfslam_b200 call failed");
}

static const double* dbl(const mxArray* a, size_t at_least, const char* what) {
    if (!mxIsDouble(a) || mxIsComplex(a) || mxGetNumberOfElements(a) < at_least)
        mexErrMsgIdAndTxt("ekfslam:EINVAL", "%s: expected a real double array with >= %d elements", what, (int)at_least);
    return mxGetPr(a);
}

static double scalar_field(const mxArray* s, const char* name, double dflt) {
    const mxArray* f = mxIsStruct(s) ? mxGetField(s, 0, name) : NULL;
    return f ? mxGetScalar(f) : dflt;
}

/* R arrives either as the 2x2 diag matrix the reference builds (EKF_SLAM.m:108) or as its diagonal. */
static void rdiag(const mxArray* a, double out[2]) {
    const double* r = dbl(a, 2, "R");
    if (mxGetM(a) == 2 && mxGetN(a) == 2) { out[0] = r[0]; out[1] = r[3]; }
    else { out[0] = r[0]; out[1] = r[1]; }
}

static mxArray* handle_out(ekf_handle* h) {
    mxArray* o = mxCreateNumericMatrix(1, 1, mxUINT64_CLASS, mxREAL);
    *(uint64_t*)mxGetData(o) = (uint64_t)(uintptr_t)h;
    return o;
}

void mexFunction(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    char verb[32];
    (void)nlhs;
    if (nrhs < 1 || mxGetString(prhs[0], verb, sizeof verb) != 0)
        mexErrMsgIdAndTxt("ekfslam:EINVAL", "usage: ekfslam_mex('verb', handle, ...)");

    if (!strcmp(verb, "version")) { plhs[0] = mxCreateString(ekf_version()); return; }

    if (!strcmp(verb, "create")) {            /* h = ekfslam_mex('create', variant, opts-struct) */
        ekf_params p;
        ekf_handle* h = NULL;
        int32_t variant = nrhs > 1 ? (int32_t)mxGetScalar(prhs[1]) : EKF_VARIANT_KNOWN;
        check(NULL, ekf_params_default(&p, variant));
        if (nrhs > 2) {
            const mxArray* o = prhs[2];
            p.capacity_landmarks = (int32_t)scalar_field(o, "capacity_landmarks", p.capacity_landmarks);
            p.storage = (int32_t)scalar_field(o, "storage", p.storage);
            p.assoc_mode = (int32_t)scalar_field(o, "assoc_mode", p.assoc_mode);
            p.device = (int32_t)scalar_field(o, "device", p.device);
            p.defer_m = (int32_t)scalar_field(o, "defer_m", p.defer_m);
        }
        {
            /* opts.devices = [0 1 2 3]: ONE filter in column slabs over several GPUs, driven from this one process */
            const mxArray* dv = (nrhs > 2 && mxIsStruct(prhs[2])) ? mxGetField(prhs[2], 0, "devices") : NULL;
            size_t nd = dv ? mxGetNumberOfElements(dv) : 0;
            int32_t rc;
            if (nd > 1) {
                int32_t devs[8];
                size_t g;
                const double* dd = dbl(dv, nd, "devices");
                if (nd > 8) mexErrMsgIdAndTxt("ekfslam:EINVAL", "at most 8 devices");
                for (g = 0; g < nd; ++g) devs[g] = (int32_t)dd[g];
                rc = ekf_create_multi(&p, (int32_t)nd, devs, &h);
            } else {
                if (nd == 1) p.device = (int32_t)mxGetScalar(dv);
                rc = ekf_create(&p, &h);
            }
            if (rc != EKF_OK) mexErrMsgIdAndTxt(code_name(rc), "ekf_create failed (a CUDA device is required; there is no CPU fallback)");
        }
        plhs[0] = handle_out(h);
        return;
    }

    if (nrhs < 2) mexErrMsgIdAndTxt("ekfslam:EINVAL", "verb '%s' needs a handle", verb);
    {   /* right-hand-side count per verb (incl. verb and handle): a short call is an error, not an out-of-bounds read */
        static const struct { const char* verb; int nrhs; } need[] = {
            {"destroy", 2}, {"predict", 3}, {"update", 5}, {"append", 6}, {"associate", 4}, {"measure", 6},
            {"n_landmarks", 2}, {"get_x", 2}, {"get_s", 2}, {"get_P", 2}, {"get_P_block", 6}, {"get_ellipses", 2},
            {"set_state", 5}, {"set_x", 3}, {"set_P", 3}, {"set_s", 3}, {"set_noise", 6}, {"flush", 2}, {"sync", 2}};
        size_t v;
        for (v = 0; v < sizeof need / sizeof need[0]; ++v)
            if (!strcmp(verb, need[v].verb) && nrhs < need[v].nrhs)
                mexErrMsgIdAndTxt("ekfslam:EINVAL", "verb '%s' needs %d arguments, got %d", verb, need[v].nrhs, nrhs);
    }
    ekf_handle* h = get_handle(prhs[1]);

    if (!strcmp(verb, "destroy")) { check(NULL, ekf_destroy(h)); return; }

    if (!strcmp(verb, "predict")) {           /* ekfslam_mex('predict', h, u)  -- EKF_SLAM.m:40 */
        check(h, ekf_predict(h, dbl(prhs[2], 2, "u")));
        return;
    }
    if (!strcmp(verb, "update")) {            /* ekfslam_mex('update', h, z, R, idx) -- EKF_SLAM.m:124-145 */
        double Rd[2];
        rdiag(prhs[3], Rd);
        check(h, ekf_update(h, dbl(prhs[2], 2, "z"), Rd, (int32_t)mxGetScalar(prhs[4])));
        return;
    }
    if (!strcmp(verb, "append")) {            /* ekfslam_mex('append', h, u, R, pos, signature) -- EKF_SLAM.m:67 */
        double Rd[2];
        rdiag(prhs[3], Rd);
        check(h, ekf_append(h, dbl(prhs[2], 2, "u"), Rd, dbl(prhs[4], 2, "landmarkPos"), mxGetScalar(prhs[5])));
        return;
    }
    if (!strcmp(verb, "associate")) {         /* [newLL, index] = ekfslam_mex('associate', h, z, R) -- Correspondence.m:28 */
        double Rd[2];
        int32_t newLL = 1, idx = 0;
        rdiag(prhs[3], Rd);
        check(h, ekf_associate(h, dbl(prhs[2], 3, "z"), Rd, &newLL, &idx));
        plhs[0] = mxCreateLogicalScalar(newLL != 0);
        if (nlhs > 1) plhs[1] = mxCreateDoubleScalar((double)idx);
        return;
    }
    if (!strcmp(verb, "measure")) {           /* actions = ekfslam_mex('measure', h, observed_LL(kx3), u, loc(Tx2), index(Tx1)) */
        const mxArray* obs = prhs[2];
        size_t k = mxGetM(obs), T = mxGetNumberOfElements(prhs[5]);
        size_t i;
        int32_t* idx_tab = (int32_t*)mxCalloc(T ? T : 1, sizeof(int32_t));
        int32_t* act = (int32_t*)mxCalloc(k ? k : 1, sizeof(int32_t));
        const double* idx_d = T ? dbl(prhs[5], T, "index table") : NULL;
        int32_t rc;
        if (k && mxGetN(obs) != 3) mexErrMsgIdAndTxt("ekfslam:EINVAL", "observed_LL must be k x 3");
        for (i = 0; i < T; ++i) idx_tab[i] = (int32_t)idx_d[i];
        rc = ekf_measure_batch(h, k ? dbl(obs, 3 * k, "observed_LL") : NULL, (int32_t)k, dbl(prhs[3], 2, "u"),
                               T ? dbl(prhs[4], 2 * T, "loc table") : NULL, idx_tab, (int32_t)T, act);
        plhs[0] = mxCreateDoubleMatrix(k, 1, mxREAL);
        for (i = 0; i < k; ++i) mxGetPr(plhs[0])[i] = (double)act[i];
        mxFree(idx_tab);
        mxFree(act);
        check(h, rc);
        return;
    }
    if (!strcmp(verb, "n_landmarks")) {
        int32_t N = 0;
        check(h, ekf_n_landmarks(h, &N));
        plhs[0] = mxCreateDoubleScalar((double)N);
        return;
    }
    if (!strcmp(verb, "get_x")) {             /* 1 x n row vector, EKF_SLAM.m:28 */
        int32_t N = 0;
        check(h, ekf_n_landmarks(h, &N));
        plhs[0] = mxCreateDoubleMatrix(1, 3 + 2 * (size_t)N, mxREAL);
        check(h, ekf_get_x(h, mxGetPr(plhs[0]), 3 + 2 * N));
        return;
    }
    if (!strcmp(verb, "get_s")) {             /* N x 1 column, EKF_SLAM.m:70 */
        int32_t N = 0;
        check(h, ekf_n_landmarks(h, &N));
        plhs[0] = mxCreateDoubleMatrix((size_t)N, 1, mxREAL);
        if (N > 0) check(h, ekf_get_s(h, mxGetPr(plhs[0]), N));
        return;
    }
    if (!strcmp(verb, "get_P")) {             /* n x n, column-major both sides: one 2-D copy */
        int32_t N = 0, n;
        check(h, ekf_n_landmarks(h, &N));
        n = 3 + 2 * N;
        plhs[0] = mxCreateDoubleMatrix((size_t)n, (size_t)n, mxREAL);
        check(h, ekf_get_P(h, mxGetPr(plhs[0]), n));
        return;
    }
    if (!strcmp(verb, "get_P_block")) {       /* ekfslam_mex('get_P_block', h, r1, c1, nr, nc), 1-based like MATLAB */
        int32_t r1 = (int32_t)mxGetScalar(prhs[2]), c1 = (int32_t)mxGetScalar(prhs[3]);
        int32_t nr = (int32_t)mxGetScalar(prhs[4]), nc = (int32_t)mxGetScalar(prhs[5]);
        plhs[0] = mxCreateDoubleMatrix((size_t)nr, (size_t)nc, mxREAL);
        check(h, ekf_get_P_block(h, r1 - 1, c1 - 1, nr, nc, mxGetPr(plhs[0]), nr));
        return;
    }
    if (!strcmp(verb, "get_ellipses")) {      /* (N+1) x 5: what plot() would draw, EKF_SLAM.m:180-229 */
        int32_t N = 0;
        check(h, ekf_n_landmarks(h, &N));
        plhs[0] = mxCreateDoubleMatrix((size_t)N + 1, 5, mxREAL);
        check(h, ekf_get_ellipses(h, mxGetPr(plhs[0]), N + 1));
        return;
    }
    if (!strcmp(verb, "set_state")) {         /* ekfslam_mex('set_state', h, x, P, s) */
        size_t n = mxGetNumberOfElements(prhs[2]);
        int32_t N = (int32_t)((n - 3) / 2);
        double dummy = 0.0;
        if (n < 3 || ((n - 3) & 1) || mxGetM(prhs[3]) != n || mxGetN(prhs[3]) != n)
            mexErrMsgIdAndTxt("ekfslam:EINVAL", "set_state: x must have 3+2N elements and P must be n x n");
        check(h, ekf_set_state(h, dbl(prhs[2], n, "x"), dbl(prhs[3], n * n, "P"), (int32_t)n,
                               N ? dbl(prhs[4], (size_t)N, "s") : &dummy, N));
        return;
    }
    if (!strcmp(verb, "set_x")) {             /* h.x = v, EKF_SLAM.m:6 */
        size_t n = mxGetNumberOfElements(prhs[2]);
        check(h, ekf_set_x(h, dbl(prhs[2], n, "x"), (int32_t)n));
        return;
    }
    if (!strcmp(verb, "set_s")) {             /* h.s = v, EKF_SLAM.m:9 */
        size_t N = mxGetNumberOfElements(prhs[2]);
        double dummy = 0.0;
        check(h, ekf_set_s(h, N ? dbl(prhs[2], N, "s") : &dummy, (int32_t)N));
        return;
    }
    if (!strcmp(verb, "set_P")) {             /* h.P = M, EKF_SLAM.m:7 */
        size_t n = mxGetM(prhs[2]);
        int32_t N = 0;
        check(h, ekf_n_landmarks(h, &N));
        if (n != (size_t)(3 + 2 * N) || mxGetN(prhs[2]) != n)
            mexErrMsgIdAndTxt("ekfslam:EINVAL", "set_P: P must be n x n with n = 3 + 2 N = %d", 3 + 2 * N);
        check(h, ekf_set_P(h, dbl(prhs[2], n * n, "P"), (int32_t)n));
        return;
    }
    if (!strcmp(verb, "set_noise")) {         /* ekfslam_mex('set_noise', h, C, Rc, s_cost, s_thresh) */
        check(h, ekf_set_noise(h, mxGetScalar(prhs[2]), dbl(prhs[3], 2, "Rc"), mxGetScalar(prhs[4]), mxGetScalar(prhs[5])));
        return;
    }
    if (!strcmp(verb, "flush")) { check(h, ekf_flush(h)); return; }
    if (!strcmp(verb, "sync")) { check(h, ekf_sync(h)); return; }

    mexErrMsgIdAndTxt("ekfslam:EINVAL", "unknown verb '%s'", verb);
}
