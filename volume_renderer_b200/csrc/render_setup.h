// render_setup.h - per-frame setup shared by the CUDA context and the host-side emulation used
// in tests: marshals the 'render' arguments exactly as the mex gateway lays them out
// (src/C/mex/render.cpp:194-245) and derives box + step like vr::initRender
// (src/C/vr/volumeRender.cpp:112-156) and the gradient step of vr::render (:273-275).
#ifndef RENDER_SETUP_H_
#define RENDER_SETUP_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#include "vr_types.h"

namespace vrb {

// rotation rows, axis-reversed (render.cpp:211-219), and the property row (render.cpp:220-221)
inline void fill_camera(Camera &C, const float R[9], const float props[3]) {
  for (int k = 0; k < 3; ++k) C.m[k][0] = R[3 * k + 2], C.m[k][1] = R[3 * k + 1], C.m[k][2] = R[3 * k + 0];
  C.m[3][0] = props[0], C.m[3][1] = props[1], C.m[3][2] = props[2];
}

inline void setup_frame(KernelParams &P, const float factors[3], const float es[3], uint64_t W, uint64_t H,
                        const float R[9], const float props[3], float thr, const float color[3], uint64_t w,
                        uint64_t h, uint64_t d) {
  P.W = (int)W, P.H = (int)H;
  P.fEm = factors[0], P.fRe = factors[1], P.fAb = factors[2];  // VolumeRender.m:558 + volumeRender.cpp:112-117
  // element size arrives as in MATLAB and is axis-reversed (render.cpp:195)
  const float esx = es[2], esy = es[1], esz = es[0];
  P.boxmax[0] = 1.f;  // volumeRender.cpp:125-131
  P.boxmax[1] = (esy * h) / ((float)w * esx);
  P.boxmax[2] = (esz * d) / ((float)w * esx);
  for (int i = 0; i < 3; ++i) P.boxmin[i] = -1.f * P.boxmax[i];
  const float dxy = sqrtf((float)(w * w + h * h)), dyz = sqrtf((float)(h * h + d * d)), dxz = sqrtf((float)(w * w + d * d));
  const float minDiagonal = fminf(dxy, fminf(dyz, dxz));  // named maxDiagonal in the reference, volumeRender.cpp:141
  P.tstep = 1.f / (2.2f * minDiagonal);                    // volumeRender.cpp:144-145
  fill_camera(P.cam, R, props);
  P.thr = thr;
  memcpy(P.color, color, sizeof P.color);
  P.gstep[0] = 1.f / w, P.gstep[1] = 1.f / h, P.gstep[2] = 1.f / d;  // volumeRender.cpp:273-275
}

}  // namespace vrb
#endif
