// timestamp() mex - drop-in for src/C/mex/timestamp.cpp:13-33: ms since the epoch as a uint64
// scalar.  The reference stores through an int* (low 32 bits only); the full value is returned
// here, which only makes the dirty test `last_update > timeLastMemSync` more robust.
#include <mex.h>

#include "../include/vr_abi.h"

extern "C" void mexFunction(int, mxArray *plhs[], int, const mxArray *[]) {
  plhs[0] = mxCreateNumericMatrix(1, 1, mxUINT64_CLASS, mxREAL);
  *static_cast<uint64_t *>(mxGetData(plhs[0])) = vr_timestamp_ms(0);
}
