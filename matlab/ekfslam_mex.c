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
#include <stdlib.h>
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

/* Monte-Carlo batch (ekfb_*): the uint64 handed to MATLAB points at this record -- the C ABI has no getters for the
 * batch's dimensions, and every verb needs them to size its outputs. */
typedef struct { ekfb_handle* h; int32_t nf, max_obs, max_table; } mex_batch;

static mex_batch* get_batch(const mxArray* a) {
    mex_batch* b;
    if (!mxIsUint64(a) || mxGetNumberOfElements(a) != 1)
        mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch handle must be a uint64 scalar");
    b = (mex_batch*)(uintptr_t)(*(uint64_t*)mxGetData(a));
    if (!b || !b->h) mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch handle is null or destroyed");
    return b;
}

static void bcheck(mex_batch* b, int32_t rc) {
    if (rc != EKF_OK) mexErrMsgIdAndTxt(code_name(rc), "%s", b && b->h ? ekfb_last_error(b->h) : "ekfslam_b200 batch call failed");
}

static int32_t filter_index(const mex_batch* b, const mxArray* a, int allow_all) {      /* 1-based like MATLAB; 0 = every filter */
    int32_t f = (int32_t)mxGetScalar(a);
    if (f < (allow_all ? 0 : 1) || f > b->nf) mexErrMsgIdAndTxt("ekfslam:EINVAL", "filter index %d outside 1..%d", (int)f, (int)b->nf);
    return f - 1;
}

/* The batch verbs: n_filters objects of one reference class driven in lock step (SLAM.m:110,116 per object).
 *   hb      = ekfslam_mex('batch_create', variant, n_filters, max_obs, max_table, opts)
 *   actions = ekfslam_mex('batch_step', hb, u(2 x nf), observed_LL(max_obs x 3 x nf), counts(nf), loc(T x 2), index(T), do_predict, do_measure)
 *             one landmark table shared by all filters; actions: max_obs x nf, +idx update / -idx append
 *   N = ekfslam_mex('batch_n_landmarks', hb);  x = ('batch_get_x', hb, f);  s = ('batch_get_s', hb, f);  P = ('batch_get_P', hb, f)
 *   poses = ekfslam_mex('batch_poses', hb)  (3 x nf);  ('batch_set_state', hb, f | 0 = all, x, P, s);  ('batch_sync', hb);  ('batch_destroy', hb) */
static int batch_verb(const char* verb, int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]) {
    (void)nlhs;
    if (strncmp(verb, "batch_", 6)) return 0;
    if (!strcmp(verb, "batch_create")) {
        ekf_params p;
        mex_batch* b;
        int32_t variant, rc;
        if (nrhs < 5) mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch_create needs variant, n_filters, max_obs, max_table");
        variant = (int32_t)mxGetScalar(prhs[1]);
        check(NULL, ekf_params_default(&p, variant));
        if (nrhs > 5) {
            const mxArray* o = prhs[5];
            p.capacity_landmarks = (int32_t)scalar_field(o, "capacity_landmarks", p.capacity_landmarks);
            p.storage = (int32_t)scalar_field(o, "storage", p.storage);
            p.assoc_mode = (int32_t)scalar_field(o, "assoc_mode", p.assoc_mode);
            p.device = (int32_t)scalar_field(o, "device", p.device);
            p.defer_m = (int32_t)scalar_field(o, "defer_m", p.defer_m);
        }
        b = (mex_batch*)calloc(1, sizeof(mex_batch));      /* outlives the call: not mxCalloc (MATLAB frees that on return) */
        b->nf = (int32_t)mxGetScalar(prhs[2]); b->max_obs = (int32_t)mxGetScalar(prhs[3]); b->max_table = (int32_t)mxGetScalar(prhs[4]);
        rc = ekfb_create(&p, b->nf, b->max_obs, b->max_table, &b->h);
        if (rc != EKF_OK) { free(b); mexErrMsgIdAndTxt(code_name(rc), "ekfb_create failed (a CUDA device is required; there is no CPU fallback)"); }
        plhs[0] = mxCreateNumericMatrix(1, 1, mxUINT64_CLASS, mxREAL);
        *(uint64_t*)mxGetData(plhs[0]) = (uint64_t)(uintptr_t)b;
        return 1;
    }
    if (nrhs < 2) mexErrMsgIdAndTxt("ekfslam:EINVAL", "verb '%s' needs a batch handle", verb);
    {
        mex_batch* b = get_batch(prhs[1]);
        if (!strcmp(verb, "batch_destroy")) { ekfb_destroy(b->h); b->h = NULL; free(b); return 1; }
        if (!strcmp(verb, "batch_sync")) { bcheck(b, ekfb_sync(b->h)); return 1; }
        if (!strcmp(verb, "batch_step")) {
            size_t nf = (size_t)b->nf, mo = (size_t)b->max_obs, mt = (size_t)b->max_table, T, i, f;
            int32_t *counts, *idx_tab, *act, ntab, rc;
            double* loc;
            const double *cd, *ld_, *id_;
            int do_predict, do_measure;
            if (nrhs < 9) mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch_step needs u, observed_LL, counts, loc, index, do_predict, do_measure");
            do_predict = mxGetScalar(prhs[7]) != 0.0; do_measure = mxGetScalar(prhs[8]) != 0.0;
            T = mxGetNumberOfElements(prhs[6]);
            if (T > mt) mexErrMsgIdAndTxt("ekfslam:EINVAL", "landmark table has %d entries, max_table is %d", (int)T, (int)mt);
            counts = (int32_t*)mxCalloc(nf, sizeof(int32_t));
            idx_tab = (int32_t*)mxCalloc(mt ? mt : 1, sizeof(int32_t));
            loc = (double*)mxCalloc(2 * (mt ? mt : 1), sizeof(double));
            act = (int32_t*)mxCalloc((nf * mo) > 0 ? nf * mo : 1, sizeof(int32_t));
            cd = dbl(prhs[4], nf, "counts");
            for (f = 0; f < nf; ++f) counts[f] = (int32_t)cd[f];
            if (T) {
                ld_ = dbl(prhs[5], 2 * T, "loc table"); id_ = dbl(prhs[6], T, "index table");
                for (i = 0; i < T; ++i) { idx_tab[i] = (int32_t)id_[i]; loc[i] = ld_[i]; loc[mt + i] = ld_[T + i]; }   /* T x 2 -> max_table x 2 */
            }
            ntab = (int32_t)T;
            rc = ekfb_step(b->h, 0, dbl(prhs[2], 2 * nf, "u"), do_measure ? dbl(prhs[3], 3 * mo * nf, "observed_LL") : NULL,
                           do_measure ? counts : NULL, do_measure ? loc : NULL, do_measure ? idx_tab : NULL, do_measure ? &ntab : NULL,
                           1, do_predict, do_measure, act);                   /* asking for the actions synchronises: the buffers may go */
            plhs[0] = mxCreateDoubleMatrix(mo, nf, mxREAL);
            for (i = 0; i < mo * nf; ++i) mxGetPr(plhs[0])[i] = (double)act[i];
            mxFree(counts); mxFree(idx_tab); mxFree(loc); mxFree(act);
            bcheck(b, rc);
            return 1;
        }
        if (!strcmp(verb, "batch_n_landmarks")) {
            size_t f;
            int32_t* N = (int32_t*)mxCalloc((size_t)b->nf, sizeof(int32_t));
            int32_t rc = ekfb_n_landmarks(b->h, N);
            plhs[0] = mxCreateDoubleMatrix((size_t)b->nf, 1, mxREAL);
            for (f = 0; f < (size_t)b->nf; ++f) mxGetPr(plhs[0])[f] = (double)N[f];
            mxFree(N);
            bcheck(b, rc);
            return 1;
        }
        if (!strcmp(verb, "batch_poses")) {
            plhs[0] = mxCreateDoubleMatrix(3, (size_t)b->nf, mxREAL);          /* filter-major triples == 3 x nf column-major */
            bcheck(b, ekfb_get_poses(b->h, mxGetPr(plhs[0])));
            return 1;
        }
        if (!strcmp(verb, "batch_get_x") || !strcmp(verb, "batch_get_s") || !strcmp(verb, "batch_get_P")) {
            int32_t f, n, rc, *N;
            if (nrhs < 3) mexErrMsgIdAndTxt("ekfslam:EINVAL", "verb '%s' needs a filter index", verb);
            f = filter_index(b, prhs[2], 0);
            N = (int32_t*)mxCalloc((size_t)b->nf, sizeof(int32_t));
            rc = ekfb_n_landmarks(b->h, N);
            n = 3 + 2 * N[f];
            if (rc == EKF_OK) {
                if (verb[10] == 'x') { plhs[0] = mxCreateDoubleMatrix(1, (size_t)n, mxREAL); rc = ekfb_get_x(b->h, f, mxGetPr(plhs[0]), n); }
                else if (verb[10] == 's') { plhs[0] = mxCreateDoubleMatrix((size_t)N[f], 1, mxREAL); if (N[f] > 0) rc = ekfb_get_s(b->h, f, mxGetPr(plhs[0]), N[f]); }
                else { plhs[0] = mxCreateDoubleMatrix((size_t)n, (size_t)n, mxREAL); rc = ekfb_get_P(b->h, f, mxGetPr(plhs[0]), n); }
            }
            mxFree(N);
            bcheck(b, rc);
            return 1;
        }
        if (!strcmp(verb, "batch_set_state")) {           /* ('batch_set_state', hb, f | 0 = every filter, x, P, s) */
            size_t n;
            int32_t f, N;
            double dummy = 0.0;
            if (nrhs < 6) mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch_set_state needs f, x, P, s");
            f = filter_index(b, prhs[2], 1);
            n = mxGetNumberOfElements(prhs[3]);
            N = (int32_t)((n - 3) / 2);
            if (n < 3 || ((n - 3) & 1) || mxGetM(prhs[4]) != n || mxGetN(prhs[4]) != n)
                mexErrMsgIdAndTxt("ekfslam:EINVAL", "batch_set_state: x must have 3+2N elements and P must be n x n");
            bcheck(b, ekfb_set_state(b->h, f, dbl(prhs[3], n, "x"), dbl(prhs[4], n * n, "P"), (int32_t)n,
                                     N ? dbl(prhs[5], (size_t)N, "s") : &dummy, N));
            return 1;
        }
    }
    mexErrMsgIdAndTxt("ekfslam:EINVAL", "unknown verb '%s'", verb);
    return 1;
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
    if (batch_verb(verb, nlhs, plhs, nrhs, prhs)) return;

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
            {"set_state", 5}, {"set_x", 3}, {"set_P", 3}, {"set_s", 3}, {"set_noise", 6}, {"flush", 2}, {"sync", 2},
            {"update_batch", 5}, {"flush_info", 2}, {"defer_info", 2}};
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
    if (!strcmp(verb, "update_batch")) {      /* ekfslam_mex('update_batch', h, z(k x 2), Rdiag(k x 2), idx(k)): EKF_SLAM.m:124-145 for k known slots */
        size_t k = mxGetM(prhs[2]), i;
        int32_t* idx = (int32_t*)mxCalloc(k ? k : 1, sizeof(int32_t));
        const double* id = k ? dbl(prhs[4], k, "idx") : NULL;
        int32_t rc;
        if (k && (mxGetN(prhs[2]) != 2 || mxGetM(prhs[3]) != k || mxGetN(prhs[3]) != 2))
            mexErrMsgIdAndTxt("ekfslam:EINVAL", "update_batch: z and Rdiag must be k x 2");
        for (i = 0; i < k; ++i) idx[i] = (int32_t)id[i];
        rc = ekf_update_batch(h, (int32_t)k, k ? dbl(prhs[2], 2 * k, "z") : NULL, k ? dbl(prhs[3], 2 * k, "Rdiag") : NULL, idx);
        mxFree(idx);
        check(h, rc);
        return;
    }
    if (!strcmp(verb, "flush_info")) {        /* [continuous, passes_outstanding] = ekfslam_mex('flush_info', h) */
        int32_t c = 0, q = 0;
        check(h, ekf_flush_info(h, &c, &q));
        plhs[0] = mxCreateLogicalScalar(c != 0);
        if (nlhs > 1) plhs[1] = mxCreateDoubleScalar((double)q);
        return;
    }
    if (!strcmp(verb, "defer_info")) {        /* [batch, pend_cap] = ekfslam_mex('defer_info', h) */
        int32_t bsz = 0, cap = 0;
        check(h, ekf_defer_info(h, &bsz, &cap));
        plhs[0] = mxCreateDoubleScalar((double)bsz);
        if (nlhs > 1) plhs[1] = mxCreateDoubleScalar((double)cap);
        return;
    }
    if (!strcmp(verb, "flush")) { check(h, ekf_flush(h)); return; }
    if (!strcmp(verb, "sync")) { check(h, ekf_sync(h)); return; }

    mexErrMsgIdAndTxt("ekfslam:EINVAL", "unknown verb '%s'", verb);
}
