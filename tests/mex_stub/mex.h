/* Minimal stand-in for MATLAB's mex.h -- TEST INFRASTRUCTURE ONLY (no MATLAB/Octave in this image).
 * Declares just the mx/mex functions matlab/ekfslam_mex.c uses, with MATLAB's signatures, so that the gateway
 * is compile-checked and can be driven from tests/test_matlab_shim.py through mex_stub.c. */
#ifndef EKF_TEST_MEX_H
#define EKF_TEST_MEX_H
#include <stddef.h>
#include <stdbool.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct mxArray_tag mxArray;
typedef enum { mxUNKNOWN_CLASS = 0, mxSTRUCT_CLASS = 2, mxLOGICAL_CLASS = 3, mxCHAR_CLASS = 4,
               mxDOUBLE_CLASS = 6, mxUINT64_CLASS = 13 } mxClassID;
typedef enum { mxREAL = 0, mxCOMPLEX = 1 } mxComplexity;
bool mxIsUint64(const mxArray*);
bool mxIsDouble(const mxArray*);
bool mxIsComplex(const mxArray*);
bool mxIsStruct(const mxArray*);
size_t mxGetNumberOfElements(const mxArray*);
size_t mxGetM(const mxArray*);
size_t mxGetN(const mxArray*);
void* mxGetData(const mxArray*);
double* mxGetPr(const mxArray*);
double mxGetScalar(const mxArray*);
mxArray* mxGetField(const mxArray*, size_t index, const char* name);
int mxGetString(const mxArray*, char* buf, size_t buflen);
mxArray* mxCreateNumericMatrix(size_t m, size_t n, mxClassID cls, mxComplexity c);
mxArray* mxCreateDoubleMatrix(size_t m, size_t n, mxComplexity c);
mxArray* mxCreateDoubleScalar(double v);
mxArray* mxCreateLogicalScalar(bool v);
mxArray* mxCreateString(const char* s);
void* mxCalloc(size_t n, size_t size);
void mxFree(void* p);
void mexErrMsgIdAndTxt(const char* id, const char* fmt, ...);
void mexFunction(int nlhs, mxArray* plhs[], int nrhs, const mxArray* prhs[]);
#ifdef __cplusplus
}
#endif
#endif
