/* Minimal stand-in for MATLAB's mex.h, only so that the reference's own CPU sources
 * (src/C/mex/HenyeyGreenstein.cc, src/C/mex/timestamp.cpp, src/C/vr/volumeRender.cpp) compile
 * UNMODIFIED where they lie under /root/reference.  Test infrastructure only (oracle/_ref).
 * The functions are implemented in ref_driver.cpp. */
#ifndef STUB_MEX_H_
#define STUB_MEX_H_
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

typedef size_t mwSize;
typedef size_t mwIndex;
typedef enum { mxUNKNOWN_CLASS = 0, mxLOGICAL_CLASS = 3, mxDOUBLE_CLASS = 6, mxSINGLE_CLASS = 7, mxUINT64_CLASS = 15 } mxClassID;
typedef enum { mxREAL = 0, mxCOMPLEX = 1 } mxComplexity;
typedef struct mxArray_tag mxArray;

#ifdef __cplusplus
extern "C" {
#endif
void mexErrMsgTxt(const char *msg);
void mexWarnMsgTxt(const char *msg);
int mexPrintf(const char *fmt, ...);
void mexLock(void);
void mexUnlock(void);
double mxGetScalar(const mxArray *a);
mxArray *mxCreateNumericArray(mwSize ndim, const mwSize *dims, mxClassID cls, mxComplexity c);
mxArray *mxCreateNumericMatrix(mwSize m, mwSize n, mxClassID cls, mxComplexity c);
double *mxGetPr(const mxArray *a);
void *mxGetData(const mxArray *a);
const mwSize *mxGetDimensions(const mxArray *a);
mwSize mxGetNumberOfDimensions(const mxArray *a);
size_t mxGetNumberOfElements(const mxArray *a);
size_t mxGetN(const mxArray *a);
size_t mxGetM(const mxArray *a);
mxClassID mxGetClassID(const mxArray *a);
bool mxIsComplex(const mxArray *a);
bool mxIsClass(const mxArray *a, const char *name);
int mxGetString(const mxArray *a, char *buf, mwSize buflen);
mxArray *mxGetProperty(const mxArray *a, mwIndex idx, const char *name);
mxArray *mxGetPropertyShared(const mxArray *a, mwIndex idx, const char *name);
#ifdef __cplusplus
}
#endif
#endif
