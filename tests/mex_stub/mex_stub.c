/* Tiny mxArray runtime behind tests/mex_stub/mex.h -- TEST INFRASTRUCTURE ONLY.  mexErrMsgIdAndTxt longjmps
 * back to stub_mex(), which is what tests call instead of the MATLAB interpreter. */
#include "mex.h"
#include <setjmp.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct mxArray_tag {
    mxClassID cls;
    size_t m, n;
    void* data;
    int nfields;
    char fname[16][32];
    mxArray* fval[16];
};

static jmp_buf g_jmp;
static char g_err_id[64], g_err_msg[512];

static mxArray* mk(mxClassID cls, size_t m, size_t n, size_t elem) {
    mxArray* a = (mxArray*)calloc(1, sizeof(mxArray));
    a->cls = cls; a->m = m; a->n = n;
    a->data = calloc((m * n) > 0 ? m * n : 1, elem);
    return a;
}
bool mxIsUint64(const mxArray* a) { return a && a->cls == mxUINT64_CLASS; }
bool mxIsDouble(const mxArray* a) { return a && a->cls == mxDOUBLE_CLASS; }
bool mxIsComplex(const mxArray* a) { (void)a; return false; }
bool mxIsStruct(const mxArray* a) { return a && a->cls == mxSTRUCT_CLASS; }
size_t mxGetNumberOfElements(const mxArray* a) { return a->m * a->n; }
size_t mxGetM(const mxArray* a) { return a->m; }
size_t mxGetN(const mxArray* a) { return a->n; }
void* mxGetData(const mxArray* a) { return a->data; }
double* mxGetPr(const mxArray* a) { return (double*)a->data; }
double mxGetScalar(const mxArray* a) {
    if (a->cls == mxUINT64_CLASS) return (double)*(uint64_t*)a->data;
    if (a->cls == mxLOGICAL_CLASS) return (double)*(unsigned char*)a->data;
    return *(double*)a->data;
}
mxArray* mxGetField(const mxArray* a, size_t index, const char* name) {
    (void)index;
    for (int i = 0; i < a->nfields; ++i)
        if (!strcmp(a->fname[i], name)) return a->fval[i];
    return NULL;
}
int mxGetString(const mxArray* a, char* buf, size_t buflen) {
    if (!a || a->cls != mxCHAR_CLASS || a->m * a->n + 1 > buflen) return 1;
    memcpy(buf, a->data, a->m * a->n);
    buf[a->m * a->n] = 0;
    return 0;
}
mxArray* mxCreateNumericMatrix(size_t m, size_t n, mxClassID cls, mxComplexity c) { (void)c; return mk(cls, m, n, 8); }
mxArray* mxCreateDoubleMatrix(size_t m, size_t n, mxComplexity c) { (void)c; return mk(mxDOUBLE_CLASS, m, n, 8); }
mxArray* mxCreateDoubleScalar(double v) { mxArray* a = mk(mxDOUBLE_CLASS, 1, 1, 8); *(double*)a->data = v; return a; }
mxArray* mxCreateLogicalScalar(bool v) { mxArray* a = mk(mxLOGICAL_CLASS, 1, 1, 1); *(unsigned char*)a->data = v; return a; }
mxArray* mxCreateString(const char* s) {
    size_t n = strlen(s);
    mxArray* a = mk(mxCHAR_CLASS, 1, n, 1);
    memcpy(a->data, s, n);
    return a;
}
void* mxCalloc(size_t n, size_t size) { return calloc(n, size); }
void mxFree(void* p) { free(p); }
void mexErrMsgIdAndTxt(const char* id, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    snprintf(g_err_id, sizeof g_err_id, "%s", id);
    vsnprintf(g_err_msg, sizeof g_err_msg, fmt, ap);
    va_end(ap);
    longjmp(g_jmp, 1);
}

/* ---- harness entry points used from Python (ctypes) ---- */
mxArray* stub_double(size_t m, size_t n, const double* src) {
    mxArray* a = mxCreateDoubleMatrix(m, n, mxREAL);
    if (src && (m * n) > 0) memcpy(a->data, src, m * n * 8);
    return a;
}
mxArray* stub_uint64(uint64_t v) { mxArray* a = mk(mxUINT64_CLASS, 1, 1, 8); *(uint64_t*)a->data = v; return a; }
mxArray* stub_string(const char* s) { return mxCreateString(s); }
mxArray* stub_struct(void) { return mk(mxSTRUCT_CLASS, 1, 1, 8); }
void stub_set_field(mxArray* s, const char* name, mxArray* v) {
    if (s->nfields < 16) { snprintf(s->fname[s->nfields], 32, "%s", name); s->fval[s->nfields++] = v; }
}
void stub_free(mxArray* a) { if (a) { free(a->data); free(a); } }
int stub_class(const mxArray* a) { return (int)a->cls; }
const char* stub_error_id(void) { return g_err_id; }
const char* stub_error_msg(void) { return g_err_msg; }
/* returns 0 on success, 1 when the gateway raised (id/message via stub_error_*) */
int stub_mex(int nlhs, mxArray** plhs, int nrhs, const mxArray** prhs) {
    g_err_id[0] = g_err_msg[0] = 0;
    if (setjmp(g_jmp)) return 1;
    mexFunction(nlhs, plhs, nrhs, prhs);
    return 0;
}
