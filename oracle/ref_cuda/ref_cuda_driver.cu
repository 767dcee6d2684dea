/* ref_cuda_driver.cu - TEST INFRASTRUCTURE (oracle/_ref/libref_cuda.so, "shimmed reference").
 * Drives the reference's OWN CUDA path - its d_render kernel, its upload / texture binding code
 * (src/C/vr/volumeRender_kernel.cu, compiled through texref_shim.h) and its host setup
 * (src/C/vr/volumeRender.cpp: initRender, render) - with the call sequence of the reference's mex
 * gateway, which itself is MSVC-only (src/C/mex/render.cpp:93-129 'sync_volumes' incl.
 * MManager::sync, mmanager.hxx:178-201; :134-277 'render').  The argument marshalling below
 * (axis reversals) restates render.cpp:162-168,195,211-221; everything that computes a pixel is
 * the reference's code.  Used to produce the golden images under tests/golden/ on a B200 and to
 * check the oracle and the product against them; never part of the product. */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/vr_abi.h"
#include "vr/volumeRender.h"

using namespace vr;

namespace {
Volume to_volume(const vr_volume *v) { /* mxMake_volume (volumeRender.cpp:307-342) */
  cudaExtent e = make_cudaExtent((size_t)v->dims[0], (size_t)v->dims[1], (size_t)v->dims[2]);
  return make_volume((float *)v->data, v->last_update, e);
}
/* process-wide device state, like mm::MManager's members */
cudaArray *g_em = 0, *g_ab = 0, *g_re = 0, *g_dx = 0, *g_dy = 0, *g_dz = 0, *g_light = 0;
Volume g_emission;
uint64_t g_T = 0;
bool g_have_grad = false;
}  // namespace

extern "C" {

/* 'sync_volumes' (render.cpp:93-129 + MManager::sync) */
int refcuda_sync_volumes(uint64_t time_last_mem_sync, const vr_volume *em, const vr_volume *re, const vr_volume *ab,
                         const vr_volume *gx, const vr_volume *gy, const vr_volume *gz) {
  g_T = time_last_mem_sync;
  g_emission = to_volume(em);
  const Volume vab = to_volume(ab), vre = to_volume(re);
  const bool grad = gx && gy && gz;
  setGradientMethod(grad ? gradientLookup : gradientCompute); /* render.cpp:105-121 */
  syncWithDevice(g_emission, vab, vre, g_T, g_em, g_ab, g_re);
  if (grad) {
    const Volume vx = to_volume(gx), vy = to_volume(gy), vz = to_volume(gz);
    setGradientTextures(vx, vy, vz, g_dx, g_dy, g_dz, g_T);
  }
  g_have_grad = grad;
  return (int)cudaGetLastError();
}

/* 'render' (render.cpp:134-277); out = H*W*3 floats */
int refcuda_render(const vr_light *lights, size_t n_lights, const vr_volume *illumination, const float factors[3],
                   const float es_matlab[3], const uint64_t res[2], const float R[9], const float props[3], float thr,
                   const float color_[3], float *out) {
  if (lights && illumination) {
    LightSource *ls = new LightSource[n_lights];
    for (size_t l = 0; l < n_lights; ++l) /* make_float3Inv on Position (render.cpp:167) */
      ls[l] = make_lightSource(make_float3(lights[l].position[2], lights[l].position[1], lights[l].position[0]),
                               make_float3(lights[l].color[0], lights[l].color[1], lights[l].color[2]));
    copyLightSources(ls, n_lights);
    delete[] ls;
    g_light = setIlluminationTexture(to_volume(illumination), g_light, g_T);
  }
  const float3 es = make_float3(es_matlab[2], es_matlab[1], es_matlab[0]); /* render.cpp:195 */
  float4x3 rot;
  rot.m[0] = make_float3(R[2], R[1], R[0]); /* render.cpp:211-219 */
  rot.m[1] = make_float3(R[5], R[4], R[3]);
  rot.m[2] = make_float3(R[8], R[7], R[6]);
  rot.m[3] = make_float3(props[0], props[1], props[2]);
  const size_t H = (size_t)res[0], W = (size_t)res[1];
  const dim3 block(16, 16), grid((unsigned)iDivUp(W, 16), (unsigned)iDivUp(H, 16)); /* render.cpp:236-238 */
  const float3 color = make_float3(color_[0], color_[1], color_[2]);
  RenderOptions o = initRender(W, H, factors[0], factors[1], factors[2], es, rot, thr, g_emission.extent);
  float *img = render(block, grid, o, g_emission.extent, color);
  memcpy(out, img, H * W * 3 * sizeof(float));
  free(img);
  return (int)cudaGetLastError();
}

/* a probe of the hardware rule the oracle restates: tex3D with a NaN coordinate reads as coordinate 0 */
int refcuda_device_count(void) {
  int n = 0;
  return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

}  /* extern "C" */
