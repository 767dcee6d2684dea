// HenyeyGreenstein(n[,g]) mex - drop-in for src/C/mex/HenyeyGreenstein.cc:29-96, evaluated on the
// GPU through vr_henyey_greenstein.  Output: n x n x n single, index k = c*n*n + a*n + b.
#include <mex.h>

#include "../include/vr_abi.h"

extern "C" void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]) {
  if (nrhs < 1 || nrhs > 2) mexErrMsgTxt("insufficient parameter!");  // HenyeyGreenstein.cc:30-33
  if (nlhs > 1) mexErrMsgTxt("too many outputs");
  const double n = mxGetScalar(prhs[0]);
  const float g = nrhs == 2 ? (float)mxGetScalar(prhs[1]) : 0.8f;  // HenyeyGreenstein.cc:39
  if (g > 1.0f || g < -1.0f) mexErrMsgTxt("g must be in interval [-1,1]");  // :44-45
  const mwSize dims[3] = {(mwSize)n, (mwSize)n, (mwSize)n};
  plhs[0] = mxCreateNumericArray(3, dims, mxSINGLE_CLASS, mxREAL);
  if (vr_henyey_greenstein((uint32_t)n, g, static_cast<float *>(mxGetData(plhs[0]))) != VR_OK) mexErrMsgTxt(vr_last_error(nullptr));
}
