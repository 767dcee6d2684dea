// volumeRender_mex.cpp - the `volumeRender` mex of the MATLAB classes, as a thin gateway over the
// C ABI (include/vr_abi.h).  Drop-in for the reference's src/C/mex/render.cpp:50-277: same
// command strings, same argument positions, same return values, so src/matlab/VolumeRender/*.m
// run unmodified.  All re-ordering the reference does while unpacking (axis reversal of light
// positions, element sizes and the rotation rows, render.cpp:162-168,195,211-219) happens behind
// the ABI; this file only turns mxArrays into plain pointers.
//
// Build (on a machine with MATLAB):  mex -R2018a -I../include volumeRender_mex.cpp -L../volume_renderer_b200/csrc -lvr_b200 -output volumeRender
// Here: compile-checked against mex/stub/mex.h (mex/Makefile).
#include <mex.h>

#include <cstdint>
#include <cstring>
#include <vector>

#include "../include/vr_abi.h"

namespace {

[[noreturn]] void fail(vr_ctx *ctx, const char *fallback) {
  const char *msg = vr_last_error(ctx);
  mexErrMsgTxt(msg && *msg ? msg : fallback);  // longjmps out of the mex like the reference
  for (;;) {
  }
}

// uint64 scalar handle <-> context (class_handle.hpp:56-70 convertPtr2Mat / convertMat2Ptr)
mxArray *handle_to_mat(vr_ctx *ctx) {
  mexLock();
  mxArray *out = mxCreateNumericMatrix(1, 1, mxUINT64_CLASS, mxREAL);
  *static_cast<uint64_t *>(mxGetData(out)) = reinterpret_cast<uint64_t>(ctx);
  return out;
}
vr_ctx *mat_to_handle(const mxArray *in) {
  if (mxGetNumberOfElements(in) != 1 || mxGetClassID(in) != mxUINT64_CLASS) mexErrMsgTxt("Input must be a real uint64 scalar.");
  return reinterpret_cast<vr_ctx *>(*static_cast<uint64_t *>(mxGetData(in)));
}

// vr::mxMake_volume (volumeRender.cpp:307-342): a MATLAB `Volume` object -> vr_volume.  The data
// pointer is borrowed through mxGetPropertyShared (no copy) for the duration of the call.
vr_volume make_volume(const mxArray *obj) {
  vr_volume v;
  memset(&v, 0, sizeof v);
  const mxArray *data = mxGetPropertyShared(obj, 0, "Data");
  const mxArray *stamp = mxGetProperty(obj, 0, "TimeLastUpdate");
  if (!data || mxGetClassID(data) != mxSINGLE_CLASS) mexErrMsgTxt("Volume.Data must be single");
  const mwSize nd = mxGetNumberOfDimensions(data);
  const mwSize *d = mxGetDimensions(data);
  v.data = mxGetData(data);
  v.dims[0] = d[0], v.dims[1] = nd > 1 ? d[1] : 1, v.dims[2] = nd > 2 ? d[2] : 1;  // volumeRender.cpp:330-339
  v.last_update = stamp ? *static_cast<uint64_t *>(mxGetData(stamp)) : 0;
  v.dtype = VR_DTYPE_F32;
  v.memspace = VR_MEM_HOST;
  return v;
}

const float *single_vec(const mxArray *a, size_t n, const char *what) {
  if (mxGetClassID(a) != mxSINGLE_CLASS || mxGetNumberOfElements(a) < n) mexErrMsgTxt(what);
  return static_cast<const float *>(mxGetData(a));
}

}  // namespace

extern "C" void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]) {
  char cmd[64];
  if (nrhs < 1 || mxGetString(prhs[0], cmd, sizeof cmd)) mexErrMsgTxt("First input should be a command string less than 64 characters long.");  // render.cpp:54-56

  if (!strcmp("new", cmd)) {  // render.cpp:58-65
    if (nlhs != 1) mexErrMsgTxt("New: One output expected.");
    vr_ctx *ctx = nullptr;
    if (vr_create(&ctx) != VR_OK) fail(nullptr, "New: could not create the render context.");
    plhs[0] = handle_to_mat(ctx);
    return;
  }
  if (nrhs < 2) mexErrMsgTxt("Second input should be a class instance handle.");  // render.cpp:68
  vr_ctx *ctx = mat_to_handle(prhs[1]);

  if (!strcmp("delete", cmd)) {  // render.cpp:72-79; memClear + destructor double-delete is tolerated
    if (vr_destroy(ctx) == VR_OK) mexUnlock();
    return;
  }
  if (!strcmp("mem_info", cmd)) {  // render.cpp:85-87
    size_t need = 0;
    if (vr_mem_info(ctx, nullptr, 0, &need) != VR_OK) fail(ctx, "mem_info failed");
    std::vector<char> buf(need + 1);
    if (vr_mem_info(ctx, buf.data(), buf.size(), &need) != VR_OK) fail(ctx, "mem_info failed");
    mexPrintf("%s", buf.data());
    return;
  }
  if (!strcmp("sync_volumes", cmd)) {  // render.cpp:93-129
    if (nrhs != 6 && nrhs != 9) mexErrMsgTxt("insufficient parameter!");
    const uint64_t T = (uint64_t)mxGetScalar(prhs[2]);  // render.cpp:94
    const vr_volume em = make_volume(prhs[3]), re = make_volume(prhs[4]), ab = make_volume(prhs[5]);
    vr_volume g[3];
    if (nrhs == 9)
      for (int i = 0; i < 3; ++i) g[i] = make_volume(prhs[6 + i]);
    const vr_status s = vr_sync_volumes(ctx, T, &em, &re, &ab, nrhs == 9 ? &g[0] : nullptr, nrhs == 9 ? &g[1] : nullptr,
                                        nrhs == 9 ? &g[2] : nullptr);
    if (s != VR_OK) fail(ctx, "sync_volumes failed");
    return;
  }
  if (!strcmp("update_box", cmd)) {  // extension, no reference counterpart: volumeRender('update_box', h, role, volume, lo, hi)
    if (nrhs != 6) mexErrMsgTxt("insufficient parameter!");
    if (mxGetClassID(prhs[4]) != mxUINT64_CLASS || mxGetClassID(prhs[5]) != mxUINT64_CLASS || mxGetNumberOfElements(prhs[4]) < 3 ||
        mxGetNumberOfElements(prhs[5]) < 3)
      mexErrMsgTxt("update_box: lo and hi must be uint64[3] (zero-based, hi exclusive)");
    const vr_volume v = make_volume(prhs[3]);
    const vr_status s = vr_update_box(ctx, (int)mxGetScalar(prhs[2]), &v, static_cast<const uint64_t *>(mxGetData(prhs[4])),
                                      static_cast<const uint64_t *>(mxGetData(prhs[5])));
    if (s != VR_OK) fail(ctx, "update_box failed");
    return;
  }
  if (!strcmp("set_option", cmd)) {  // extension, no reference counterpart: volumeRender('set_option', h, option, value)  (vr_option, include/vr_abi.h)
    if (nrhs != 4) mexErrMsgTxt("insufficient parameter!");
    if (vr_set_option(ctx, (int)mxGetScalar(prhs[2]), (int64_t)mxGetScalar(prhs[3])) != VR_OK) fail(ctx, "set_option failed");
    return;
  }
  if (!strcmp("set_devices", cmd)) {  // extension, no reference counterpart: volumeRender('set_devices', h, int32([0 1 2 3]))  (vr_set_devices)
    if (nrhs != 3 || mxGetClassID(prhs[2]) != mxINT32_CLASS) mexErrMsgTxt("set_devices: device ids must be an int32 vector");
    if (vr_set_devices(ctx, (uint32_t)mxGetNumberOfElements(prhs[2]), static_cast<const int *>(mxGetData(prhs[2]))) != VR_OK) fail(ctx, "set_devices failed");
    return;
  }
  if (!strcmp("render", cmd)) {  // render.cpp:134-277
    if (nrhs < 11) mexErrMsgTxt("insufficient parameter!");  // render.cpp:136-140
    // lights + illumination: skipped when either is a logical placeholder (render.cpp:145-146)
    std::vector<vr_light> lights;
    vr_volume lut;
    const bool lit = !mxIsLogical(prhs[2]) && !mxIsLogical(prhs[3]);
    if (lit) {
      const size_t n = mxGetNumberOfElements(prhs[2]);
      lights.resize(n);
      for (size_t i = 0; i < n; ++i) {
        const float *p = single_vec(mxGetProperty(prhs[2], i, "Position"), 3, "LightSource.Position must be single[3]");
        const float *c = single_vec(mxGetProperty(prhs[2], i, "Color"), 3, "LightSource.Color must be single[3]");
        memcpy(lights[i].position, p, 12);  // as stored in MATLAB; reversed behind the ABI
        memcpy(lights[i].color, c, 12);
      }
      lut = make_volume(prhs[3]);
    }
    const float *factors = single_vec(prhs[4], 3, "factors must be single[3]");
    const float *es = single_vec(prhs[5], 3, "ElementSizeUm must be single[3]");
    if (mxGetClassID(prhs[6]) != mxUINT64_CLASS || mxGetNumberOfElements(prhs[6]) < 2) mexErrMsgTxt("resolution must be uint64[2]");
    const uint64_t *res = static_cast<const uint64_t *>(mxGetData(prhs[6]));  // [H, W]
    const float *rot = single_vec(prhs[7], 9, "rotation matrix must be single 3x3");
    const float *props = single_vec(prhs[8], 3, "properties must be single[3]");
    const float thr = (float)mxGetScalar(prhs[9]);
    const float *color = single_vec(prhs[10], 3, "color must be single[3]");
    // prhs[11..13] (gradient volumes) are ignored by the reference's 'render' as well
    const mwSize dims[3] = {(mwSize)res[0], (mwSize)res[1], 3};
    plhs[0] = mxCreateNumericArray(3, dims, mxSINGLE_CLASS, mxREAL);  // H x W x 3 single, render.cpp:248
    const vr_status s = vr_render(ctx, lit ? lights.data() : nullptr, lights.size(), lit ? &lut : nullptr, factors, es, res, rot, props,
                                  thr, color, static_cast<float *>(mxGetData(plhs[0])));  // written in place: no second host copy
    if (s != VR_OK) fail(ctx, "render failed");
    return;
  }
  mexErrMsgTxt("Command not recognized.");
}
