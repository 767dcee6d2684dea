/* Compile-check stand-in for MATLAB's mex.h (this image has no MATLAB).  Declarations only: the
 * gateway objects are compiled and their undefined mx/mex symbols are left for MATLAB's libmex /
 * libmx to resolve (`mex -R2018a volumeRender_mex.cpp -lvr_b200`).  Not shipped to users. */
#ifndef VR_STUB_MEX_H_
#define VR_STUB_MEX_H_
#include <stddef.h>
#include <stdint.h>
typedef size_t mwSize;
typedef size_t mwIndex;
typedef enum { mxUNKNOWN_CLASS = 0, mxLOGICAL_CLASS = 3, mxDOUBLE_CLASS = 6, mxSINGLE_CLASS = 7, mxINT32_CLASS = 12, mxUINT64_CLASS = 15 } mxClassID;
typedef enum { mxREAL = 0, mxCOMPLEX = 1 } mxComplexity;
typedef struct mxArray_tag mxArray;
#ifdef __cplusplus
extern "C" {
#endif
void mexErrMsgTxt(const char *msg);
void mexErrMsgIdAndTxt(const char *id, const char *fmt, ...);
int mexPrintf(const char *fmt, ...);
void mexLock(void);
void mexUnlock(void);
double mxGetScalar(const mxArray *a);
mxArray *mxCreateNumericArray(mwSize ndim, const mwSize *dims, mxClassID cls, mxComplexity c);
mxArray *mxCreateNumericMatrix(mwSize m, mwSize n, mxClassID cls, mxComplexity c);
void *mxGetData(const mxArray *a);
const mwSize *mxGetDimensions(const mxArray *a);
mwSize mxGetNumberOfDimensions(const mxArray *a);
size_t mxGetNumberOfElements(const mxArray *a);
mxClassID mxGetClassID(const mxArray *a);
bool mxIsClass(const mxArray *a, const char *name);
bool mxIsLogical(const mxArray *a);
int mxGetString(const mxArray *a, char *buf, mwSize buflen);
mxArray *mxGetProperty(const mxArray *a, mwIndex idx, const char *name);
mxArray *mxGetPropertyShared(const mxArray *a, mwIndex idx, const char *name);
#ifdef __cplusplus
}
#endif
#endif
