/* ref_driver.cpp - drives the reference's OWN CPU sources, compiled unmodified from
 * /root/reference into oracle/_ref/libref_cpu.so (see oracle/Makefile: target ref).
 * TEST INFRASTRUCTURE ONLY.  It implements the few mx* / mex* functions of stub/mex.h and
 * exposes C entry points that call
 *   - HenyeyGreenstein.cc's mexFunction                    (renamed hg_mexFunction by -D)
 *   - timestamp.cpp's mexFunction                          (renamed ts_mexFunction by -D)
 *   - vr::initRender / vr::mxMake_volume / operator==      (volumeRender.cpp)
 * so that the oracle's restatements can be pinned bit-for-bit against the reference.
 */
#include "stub/mex.h"

#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

struct mxArray_tag {
  mxClassID cls = mxUNKNOWN_CLASS;
  std::vector<mwSize> dims;
  std::vector<unsigned char> own;
  void *data = nullptr; /* own.data() or borrowed */
  std::string str;      /* char arrays */
  std::string class_name;
  std::map<std::string, mxArray *> props;
};

static std::vector<mxArray *> g_pool;
static mxArray *track(mxArray *a) {
  g_pool.push_back(a);
  return a;
}
static void free_pool() {
  for (auto *a : g_pool) delete a;
  g_pool.clear();
}
static size_t elsize(mxClassID c) {
  switch (c) {
  case mxSINGLE_CLASS: return 4;
  case mxDOUBLE_CLASS: return 8;
  case mxUINT64_CLASS: return 8;
  case mxLOGICAL_CLASS: return 1;
  default: return 1;
  }
}
static size_t numel(const mxArray *a) {
  size_t n = 1;
  for (auto d : a->dims) n *= d;
  return n;
}

extern "C" {
void mexErrMsgTxt(const char *msg) { throw std::runtime_error(msg); }
void mexWarnMsgTxt(const char *) {}
int mexPrintf(const char *, ...) { return 0; }
void mexLock(void) {}
void mexUnlock(void) {}
double mxGetScalar(const mxArray *a) {
  switch (a->cls) {
  case mxSINGLE_CLASS: return *(const float *)a->data;
  case mxDOUBLE_CLASS: return *(const double *)a->data;
  case mxUINT64_CLASS: return (double)*(const uint64_t *)a->data;
  case mxLOGICAL_CLASS: return *(const unsigned char *)a->data;
  default: return 0;
  }
}
mxArray *mxCreateNumericArray(mwSize ndim, const mwSize *dims, mxClassID cls, mxComplexity) {
  mxArray *a = track(new mxArray);
  a->cls = cls;
  a->dims.assign(dims, dims + ndim);
  a->own.assign(numel(a) * elsize(cls), 0);
  a->data = a->own.data();
  return a;
}
mxArray *mxCreateNumericMatrix(mwSize m, mwSize n, mxClassID cls, mxComplexity c) {
  mwSize d[2] = {m, n};
  return mxCreateNumericArray(2, d, cls, c);
}
double *mxGetPr(const mxArray *a) { return (double *)a->data; }
void *mxGetData(const mxArray *a) { return a->data; }
const mwSize *mxGetDimensions(const mxArray *a) { return a->dims.data(); }
mwSize mxGetNumberOfDimensions(const mxArray *a) { return a->dims.size(); }
size_t mxGetNumberOfElements(const mxArray *a) { return numel(a); }
size_t mxGetM(const mxArray *a) { return a->dims.empty() ? 0 : a->dims[0]; }
size_t mxGetN(const mxArray *a) { return a->dims.empty() ? 0 : numel(a) / a->dims[0]; }
mxClassID mxGetClassID(const mxArray *a) { return a->cls; }
bool mxIsComplex(const mxArray *) { return false; }
bool mxIsClass(const mxArray *a, const char *name) {
  if (!strcmp(name, "logical")) return a->cls == mxLOGICAL_CLASS;
  return a->class_name == name;
}
int mxGetString(const mxArray *a, char *buf, mwSize buflen) {
  if (a->str.size() + 1 > buflen) return 1;
  strcpy(buf, a->str.c_str());
  return 0;
}
mxArray *mxGetProperty(const mxArray *a, mwIndex, const char *name) {
  auto it = a->props.find(name);
  return it == a->props.end() ? nullptr : it->second;
}
mxArray *mxGetPropertyShared(const mxArray *a, mwIndex i, const char *name) { return mxGetProperty(a, i, name); }
} // extern "C"

/* the reference's entry points (compiled from /root/reference, see Makefile) */
void hg_mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]);
void ts_mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]);
#include <vr/volumeRender.h>

/* vr::render (volumeRender.cpp:233-300) references the CUDA launcher that lives in the .cu file,
 * which cannot be compiled (texture references).  It is never called from here. */
#ifndef REF_WITH_CUDA_KERNEL /* the shimmed CUDA build (libref_cuda.so) links the reference's real launcher */
namespace vr {
void render_kernel(float *, const dim3 &, const dim3 &, const RenderOptions &, const float3 &, const float3 &) {
  throw std::runtime_error("render_kernel is not available in the CPU-only reference build");
}
} // namespace vr
#endif

static mxArray *scalar_double(double v) {
  mxArray *a = mxCreateNumericMatrix(1, 1, mxDOUBLE_CLASS, mxREAL);
  *(double *)a->data = v;
  return a;
}
/* a MATLAB `Volume` object: properties Data (single, borrowed) and TimeLastUpdate (uint64) */
static mxArray *volume_object(const float *data, const uint64_t *dims, int ndim, uint64_t last_update) {
  mxArray *d = track(new mxArray);
  d->cls = mxSINGLE_CLASS;
  d->dims.assign(dims, dims + ndim);
  d->data = (void *)data;
  mxArray *t = mxCreateNumericMatrix(1, 1, mxUINT64_CLASS, mxREAL);
  *(uint64_t *)t->data = last_update;
  mxArray *o = track(new mxArray);
  o->class_name = "Volume";
  o->dims = {1, 1};
  o->props["Data"] = d;
  o->props["TimeLastUpdate"] = t;
  return o;
}

extern "C" {

/* HenyeyGreenstein(n) / HenyeyGreenstein(n, g): returns 0 on success, fills out[n^3] */
int ref_henyey_greenstein(uint32_t n, int have_g, float g, float *out, char *err, size_t errcap) {
  int rc = 0;
  try {
    const mxArray *prhs[2] = {scalar_double((double)n), scalar_double((double)g)};
    mxArray *plhs[1] = {nullptr};
    hg_mexFunction(1, plhs, have_g ? 2 : 1, prhs);
    memcpy(out, plhs[0]->data, (size_t)n * n * n * sizeof(float));
  } catch (const std::exception &e) {
    if (err && errcap) snprintf(err, errcap, "%s", e.what());
    rc = 1;
  }
  free_pool();
  return rc;
}

/* timestamp(): the uint64 scalar exactly as the reference fills it */
uint64_t ref_timestamp(void) {
  mxArray *plhs[1] = {nullptr};
  ts_mexFunction(1, plhs, 0, nullptr);
  uint64_t v = *(uint64_t *)plhs[0]->data;
  free_pool();
  return v;
}

/* vr::mxMake_volume on a fake Volume object, then vr::initRender with that extent.
 * es_xyz is the element size AFTER the gateway's reversal (render.cpp:195). */
void ref_init_render(const float *data, const uint64_t *dims, int ndim, uint64_t last_update, const float es_xyz[3],
                     uint64_t W, uint64_t H, const float factors_em_re_ab[3], float thr, uint64_t extent_out[3],
                     uint64_t *memory_size, float boxmin[3], float boxmax[3], float *tstep,
                     float factors_out_em_ab_re[3]) {
  vr::Volume v = vr::mxMake_volume(volume_object(data, dims, ndim, last_update));
  extent_out[0] = v.extent.width, extent_out[1] = v.extent.height, extent_out[2] = v.extent.depth;
  *memory_size = v.memory_size;
  vr::float4x3 m;
  memset(&m, 0, sizeof(m));
  vr::RenderOptions o = vr::initRender(W, H, factors_em_re_ab[0], factors_em_re_ab[1], factors_em_re_ab[2],
                                       make_float3(es_xyz[0], es_xyz[1], es_xyz[2]), m, thr, v.extent);
  boxmin[0] = o.boxmin.x, boxmin[1] = o.boxmin.y, boxmin[2] = o.boxmin.z;
  boxmax[0] = o.boxmax.x, boxmax[1] = o.boxmax.y, boxmax[2] = o.boxmax.z;
  *tstep = o.tstep;
  factors_out_em_ab_re[0] = o.factor_emission;
  factors_out_em_ab_re[1] = o.factor_absorption;
  factors_out_em_ab_re[2] = o.factor_reflection;
  free_pool();
}

/* vr::operator== (volumeRender.cpp:24-26) */
int ref_volume_equal(const float *a, uint64_t ta, const uint64_t da[3], const float *b, uint64_t tb,
                     const uint64_t db[3]) {
  cudaExtent ea = make_cudaExtent(da[0], da[1], da[2]), eb = make_cudaExtent(db[0], db[1], db[2]);
  vr::Volume va = vr::make_volume((float *)a, ta, ea), vb = vr::make_volume((float *)b, tb, eb);
  return va == vb ? 1 : 0;
}

size_t ref_idivup(size_t a, size_t b) { return vr::iDivUp(a, b); }

} // extern "C"
