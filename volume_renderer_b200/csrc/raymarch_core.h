// raymarch_core.h - per-ray device functions of the ray march (one lane = one pixel).
//
// B200-native restatement of the reference's d_render and the device functions it calls
// (src/C/vr/volumeRender_kernel.cu:155-199 intersectBox, :212-276 gradients, :284-287 angle,
// :308-353 shade, :365-507 d_render).  Nothing is translated line by line: volumes are padded
// (optionally bricked) arrays in HBM sampled with explicit trilinear arithmetic instead of
// texture references, role->texture indirection is resolved on the host, the gradient method
// is a template parameter instead of a device function-pointer table (kernel.cu:49-50), and the
// march is split into a cheap empty-space phase and a shading phase (see raymarch_kernel.cu).
// Every function is VR_HD so that tests/host_emul can run the very same per-lane code on the
// CPU against the checker before a GPU is spent on it.
#ifndef RAYMARCH_CORE_H_
#define RAYMARCH_CORE_H_

#include <math.h>
#include <stdint.h>

#include "vr_types.h"

#if defined(__CUDA_ARCH__)
#define VR_LDG(ptr) __ldg(ptr)
#define VR_EXPF(x) __expf(x) /* the reference's intrinsic, kernel.cu:456 */
#define VR_RSQRT(x) (1.0f / sqrtf(x)) /* IEEE divide + sqrt: the normal must not depend on MUFU.RSQ's last bits */
#else
#define VR_LDG(ptr) (*(ptr))
#define VR_EXPF(x) exp2f((x) * 1.4426950408889634f)
#define VR_RSQRT(x) (1.0f / sqrtf(x))
#endif

namespace vrb {

struct F3 {
  float x, y, z;
};
VR_HD F3 mk3(float x, float y, float z) {
  F3 r;
  r.x = x, r.y = y, r.z = z;
  return r;
}
VR_HD F3 operator+(F3 a, F3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
VR_HD F3 operator-(F3 a, F3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
VR_HD F3 operator*(F3 a, F3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
VR_HD F3 operator*(F3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
VR_HD F3 operator*(float s, F3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
VR_HD float dot3(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
VR_HD float length3(F3 v) { return sqrtf(dot3(v, v)); }
// helper_math.h:1325-1328: v * rsqrtf(dot(v,v)); normalize(0) = NaN on purpose
VR_HD F3 normalize3(F3 v) { return v * VR_RSQRT(dot3(v, v)); }
VR_HD F3 ld3(const float *p) { return mk3(p[0], p[1], p[2]); }

#define VR_PI_F 3.14159265358979323846f /* kernel.cu:20 */

// ---- storage element types -------------------------------------------------------------
struct ElemF32 {
  typedef float T;
  static VR_HD float load(const void *p, uint32_t off) { return VR_LDG(static_cast<const float *>(p) + off); }
};
struct ElemBF16 {
  typedef uint16_t T;
  static VR_HD float load(const void *p, uint32_t off) {
    uint32_t u = (uint32_t)VR_LDG(static_cast<const uint16_t *>(p) + off) << 16;
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
  }
};

// ---- addressing of the padded storage --------------------------------------------------
template <int LAYOUT>
struct Addr;
template <>
struct Addr<LAYOUT_LINEAR> {
  static VR_HD uint32_t x(int X) { return (uint32_t)X; }
  static VR_HD uint32_t y(int Y, uint32_t sy) { return (uint32_t)Y * sy; }
  static VR_HD uint32_t z(int Z, uint32_t sz) { return (uint32_t)Z * sz; }
};
template <>
struct Addr<LAYOUT_BRICK> {
  static VR_HD uint32_t x(int X) { return (((uint32_t)X >> BRICK_SHIFT) << 9) + ((uint32_t)X & BRICK_MASK); }
  static VR_HD uint32_t y(int Y, uint32_t sy) { return ((uint32_t)Y >> BRICK_SHIFT) * sy + (((uint32_t)Y & BRICK_MASK) << 3); }
  static VR_HD uint32_t z(int Z, uint32_t sz) { return ((uint32_t)Z >> BRICK_SHIFT) * sz + (((uint32_t)Z & BRICK_MASK) << 6); }
};

// One axis of a normalized-coordinate linear-filter fetch (CUDA Programming Guide, "Linear
// Filtering"; configured at kernel.cu:544-548): xB = x*N - 0.5, i = floor(xB), a = frac(xB).
// x is clamped to [0,1] in float first: that IS clamp-to-edge addressing (any x<0 filters to
// the edge voxel, as does x=0), it maps a NaN coordinate to 0 (fmaxf returns the non-NaN
// operand - the texture unit's rule), and it bounds i to [-1, N-1] so that with the padded
// storage no integer clamps are needed.  X0 = i+1 is the padded index of the low tap.
VR_HD void axis_setup(float x, float fN, int frac_bits, int &X0, float &f) {
  x = fminf(fmaxf(x, 0.0f), 1.0f);
  float xb = x * fN - 0.5f;
  float fl = floorf(xb);
  f = xb - fl;
  if (frac_bits) f = floorf(f * 256.0f + 0.5f) * (1.0f / 256.0f);  // 1.8 fixed-point weights
  X0 = (int)fl + 1;
}
VR_HD float lerpf(float a, float b, float f) { return a + f * (b - a); }

// tex3D on a resident volume (kind == TEX_VOLUME).
template <int LAYOUT, typename E>
VR_HD float tex3d_volume(const TexDesc &t, float x, float y, float z, int frac_bits) {
  int X, Y, Z;
  float fx, fy, fz;
  axis_setup(x, t.fw, frac_bits, X, fx);
  axis_setup(y, t.fh, frac_bits, Y, fy);
  axis_setup(z, t.fd, frac_bits, Z, fz);
  typedef Addr<LAYOUT> A;
  const uint32_t x0 = A::x(X), x1 = A::x(X + 1);
  const uint32_t y0 = A::y(Y, t.sy), y1 = A::y(Y + 1, t.sy);
  const uint32_t z0 = A::z(Z, t.sz), z1 = A::z(Z + 1, t.sz);
  const uint32_t b00 = y0 + z0, b10 = y1 + z0, b01 = y0 + z1, b11 = y1 + z1;
  const float c000 = E::load(t.p, b00 + x0), c100 = E::load(t.p, b00 + x1);
  const float c010 = E::load(t.p, b10 + x0), c110 = E::load(t.p, b10 + x1);
  const float c001 = E::load(t.p, b01 + x0), c101 = E::load(t.p, b01 + x1);
  const float c011 = E::load(t.p, b11 + x0), c111 = E::load(t.p, b11 + x1);
  const float c00 = lerpf(c000, c100, fx), c10 = lerpf(c010, c110, fx);
  const float c01 = lerpf(c001, c101, fx), c11 = lerpf(c011, c111, fx);
  return lerpf(lerpf(c00, c10, fy), lerpf(c01, c11, fy), fz);
}
// tex3D on any slot: unbound reads 0, a 1x1x1 volume is its single value (filtering a constant
// is exact with a + f*(b-a) lerps), anything else is fetched.
template <int LAYOUT, typename E>
VR_HD float tex3d_any(const TexDesc &t, float x, float y, float z, int frac_bits) {
  if (t.kind == TEX_VOLUME) return tex3d_volume<LAYOUT, E>(t, x, y, z, frac_bits);
  return t.kind == TEX_CONST ? t.cval : 0.0f;
}

// ---- ray state ---------------------------------------------------------------------------
struct Ray {
  F3 pos, step;
  float t, tfar;
  float sx, sy, sz, sw;  // float4 sum
  uint32_t n_marched, n_shaded;
};

// intersectBox (kernel.cu:155-199): Williams et al. slab test with sign-indexed bounds; relies
// on IEEE inf/NaN semantics for axis-parallel rays exactly like the reference.
VR_HD bool intersect_box(F3 o, F3 d, F3 bmin, F3 bmax, float &tnear, float &tfar) {
  const F3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  const bool sx = inv.x < 0, sy = inv.y < 0, sz = inv.z < 0;
  float tmin = ((sx ? bmax.x : bmin.x) - o.x) * inv.x;
  float tmax = ((sx ? bmin.x : bmax.x) - o.x) * inv.x;
  const float tymin = ((sy ? bmax.y : bmin.y) - o.y) * inv.y;
  const float tymax = ((sy ? bmin.y : bmax.y) - o.y) * inv.y;
  if ((tmin > tymax) || (tymin > tmax)) return false;
  if (tymin > tmin) tmin = tymin;
  if (tymax < tmax) tmax = tymax;
  const float tzmin = ((sz ? bmax.z : bmin.z) - o.z) * inv.z;
  const float tzmax = ((sz ? bmin.z : bmax.z) - o.z) * inv.z;
  if ((tmin > tzmax) || (tzmin > tmax)) return false;
  if (tzmin > tmin) tmin = tzmin;
  if (tzmax < tmax) tmax = tzmax;
  tnear = tmin;
  tfar = tmax;
  return true;
}

VR_HD F3 ray_origin(const KernelParams &P) {
  // kernel.cu:399-410: origin = camXOffset * xVec + (-1 * objDist * zVec)
  const F3 xV = ld3(P.m[0]), zV = ld3(P.m[2]);
  return (P.m[3][0] * xV) + (-1.0f * P.m[3][2] * zV);
}

// Ray generation + box clip (kernel.cu:388-431).  Returns false when the ray misses the box
// (the pixel is then never written and keeps the memset 0).
VR_HD bool ray_init(const KernelParams &P, uint32_t x, uint32_t y, F3 origin, Ray &r) {
  const float u = (x / (float)P.W) * 2.0f - 1.0f;
  const float ratio = P.H / (float)P.W;
  const float v = (y / (float)P.H) * 2.0f * ratio - 1.0f * ratio;
  const F3 xV = ld3(P.m[0]), yV = ld3(P.m[1]), zV = ld3(P.m[2]);
  const F3 dir = normalize3(u * normalize3(xV) + v * yV + P.m[3][1] * zV);
  float tnear = 0.0f, tfar = 0.0f;
  if (!intersect_box(origin, dir, ld3(P.boxmin), ld3(P.boxmax), tnear, tfar)) return false;
  if (tnear < 0.0f) tnear = 0.0f;
  r.t = tnear;
  r.tfar = tfar;
  r.pos = origin + dir * tnear;
  r.step = dir * P.tstep;
  r.sx = r.sy = r.sz = r.sw = 0.0f;
  r.n_marched = r.n_shaded = 0;
  return true;
}

// angle (kernel.cu:284-287)
VR_HD float angle3(F3 a, F3 b) { return acosf(dot3(a, b) / (length3(a) * length3(b))); }

// One full iteration of the loop kernel.cu:435-493 up to and including the "under" update.
// GRAD: 0 = computeGradient (kernel.cu:212-253), 1 = lookupGradient (kernel.cu:266-276).
template <int GRAD, int LAYOUT, typename E>
VR_HD void shade_sample(const KernelParams &P, F3 origin, Ray &r) {
  const F3 bmin = ld3(P.boxmin), bmax = ld3(P.boxmax);
  const F3 ext = bmax - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);  // kernel.cu:396
  const F3 ps = (r.pos - bmin) * bscale;                            // kernel.cu:437
  const int fb = P.frac_bits;
  const float em_raw = tex3d_any<LAYOUT, E>(P.em, ps.x, ps.y, ps.z, fb);
  const float ab_raw = P.ab_is_em ? em_raw : tex3d_any<LAYOUT, E>(P.ab, ps.x, ps.y, ps.z, fb);
  const float emission = P.fEm * em_raw;  // kernel.cu:444-447
  const float absorption = P.fAb * ab_raw;
  const float alpha = 1 - VR_EXPF(-absorption * P.tstep);  // kernel.cu:455-456
  r.n_marched++;
  // alpha == 0 makes the premultiplied sample exactly (0,0,0,0) for finite inputs, so the
  // illumination (the expensive part) contributes nothing and is not evaluated.
  if (alpha == 0.0f) return;
  r.n_shaded++;
  const F3 color = ld3(P.color);
  const F3 colored = mk3(emission, emission, emission) * P.tstep * color;  // kernel.cu:459-460

  // shade (kernel.cu:308-353)
  F3 illum = mk3(0.0f, 0.0f, 0.0f);
  if (P.n_lights > 0) {
    F3 g;
    if (GRAD == 0) {
      const F3 gs = ld3(P.gstep);
      F3 s1, s2;
      s1 = (r.pos + mk3(gs.x, 0, 0) - bmin) * bscale;
      s2 = (r.pos - mk3(gs.x, 0, 0) - bmin) * bscale;
      g.x = tex3d_any<LAYOUT, E>(P.em_direct, s1.x, s1.y, s1.z, fb) - tex3d_any<LAYOUT, E>(P.em_direct, s2.x, s2.y, s2.z, fb);
      s1 = (r.pos + mk3(0, gs.y, 0) - bmin) * bscale;
      s2 = (r.pos - mk3(0, gs.y, 0) - bmin) * bscale;
      g.y = tex3d_any<LAYOUT, E>(P.em_direct, s1.x, s1.y, s1.z, fb) - tex3d_any<LAYOUT, E>(P.em_direct, s2.x, s2.y, s2.z, fb);
      s1 = (r.pos + mk3(0, 0, gs.z) - bmin) * bscale;
      s2 = (r.pos - mk3(0, 0, gs.z) - bmin) * bscale;
      g.z = tex3d_any<LAYOUT, E>(P.em_direct, s1.x, s1.y, s1.z, fb) - tex3d_any<LAYOUT, E>(P.em_direct, s2.x, s2.y, s2.z, fb);
      g = g * mk3(0.5f, 0.5f, 0.5f);
    } else {
      g = mk3(tex3d_any<LAYOUT, E>(P.gx, ps.x, ps.y, ps.z, fb), tex3d_any<LAYOUT, E>(P.gy, ps.x, ps.y, ps.z, fb),
              tex3d_any<LAYOUT, E>(P.gz, ps.x, ps.y, ps.z, fb));
    }
    const F3 N = -1.0f * normalize3(g);
    // the reflection fetch sits inside the reference's light loop (kernel.cu:342-344) but does
    // not depend on the light: fetched once
    const float re_raw = P.re_is_em ? em_raw : tex3d_any<LAYOUT, E>(P.re, ps.x, ps.y, ps.z, fb);
    const float reflection = P.fRe * re_raw;
    const F3 lightIn = origin - r.pos;
    for (int i = 0; i < P.n_lights; ++i) {
      const float *L = i < MAX_PARAM_LIGHTS ? P.lights[i] : P.lights_more + 6 * (i - MAX_PARAM_LIGHTS);
      const F3 lpos = mk3(L[0], L[1], L[2]), lcol = mk3(L[3], L[4], L[5]);
      const F3 lightOut = lpos - r.pos;
      const float a = angle3(N, lightIn) / VR_PI_F;
      const float b = angle3(N, lightOut) / VR_PI_F;
      const F3 lightOutProj = lightOut - (dot3(lightOut, N) * N);
      const F3 lightInProj = lightIn - (dot3(lightIn, N) * N);
      const float c = angle3(lightInProj, lightOutProj) / VR_PI_F;
      // the LUT is always fp32 in the linear padded layout
      const float light = tex3d_any<LAYOUT_LINEAR, ElemF32>(P.lut, a, b, c, fb);
      illum = illum + reflection * light * lcol * color;  // kernel.cu:349
    }
  }
  const F3 lit = colored + illum;
  const float shx = lit.x * alpha, shy = lit.y * alpha, shz = lit.z * alpha;  // kernel.cu:477-479
  const float k = 1 - r.sw;                                                    // kernel.cu:482
  r.sx = k * shx + r.sx;
  r.sy = k * shy + r.sy;
  r.sz = k * shz + r.sz;
  r.sw = k * alpha + r.sw;
}

// Loop tail (kernel.cu:485-492).  Returns false when the ray is finished.
VR_HD bool ray_advance(const KernelParams &P, Ray &r) {
  if (r.sw > P.thr) return false;
  r.t += P.tstep;
  if (r.t > r.tfar) return false;
  r.pos = r.pos + r.step;
  return true;
}

// ---- empty-space skipping ------------------------------------------------------------------
// A sample contributes exactly nothing when the absorption-role fetch is 0 (alpha = 1 -
// __expf(-0) = 0 makes the premultiplied sample 0 and `sum` unchanged).  The occupancy grid has
// one bit per 8^3 macro cell of the absorption-role texture, set when any voxel of the cell
// DILATED BY ONE VOXEL is non-zero.  A sample whose low tap index i = floor(x*N-0.5) lies in an
// empty cell reads only zeros.  For such a sample the lane computes how many steps keep it
// inside the cell and runs only the reference's serial recurrences `t += tstep; pos += step`
// (kernel.cu:488-492) for them, so positions of later samples stay bit-identical to a march
// without skipping.  The one-voxel dilation absorbs any rounding in the step-count estimate.
// Returns false when the ray ends inside the skipped span.
VR_HD bool skip_empty(const KernelParams &P, Ray &r, F3 bmin, F3 bscale, F3 dq, F3 inv_dq) {
  for (;;) {
    const F3 ps = (r.pos - bmin) * bscale;
    const float qx = fminf(fmaxf(ps.x, 0.0f), 1.0f) * P.ab.fw - 0.5f;
    const float qy = fminf(fmaxf(ps.y, 0.0f), 1.0f) * P.ab.fh - 0.5f;
    const float qz = fminf(fmaxf(ps.z, 0.0f), 1.0f) * P.ab.fd - 0.5f;
    int ix = (int)floorf(qx), iy = (int)floorf(qy), iz = (int)floorf(qz);
    ix = ix < 0 ? 0 : ix, iy = iy < 0 ? 0 : iy, iz = iz < 0 ? 0 : iz;
    const int cx = ix >> CELL_SHIFT, cy = iy >> CELL_SHIFT, cz = iz >> CELL_SHIFT;
    const uint32_t cell = (uint32_t)((cz * P.ocy + cy) * P.ocx + cx);
    const uint32_t word = VR_LDG(P.occ + (cell >> 5));
    if ((word >> (cell & 31)) & 1u) return true;  // something to shade here
    // steps until the low-tap index leaves the cell along each axis
    const float bx = (float)((dq.x >= 0.0f ? cx + 1 : cx) << CELL_SHIFT);
    const float by = (float)((dq.y >= 0.0f ? cy + 1 : cy) << CELL_SHIFT);
    const float bz = (float)((dq.z >= 0.0f ? cz + 1 : cz) << CELL_SHIFT);
    float n_exit = fminf(fminf((bx - qx) * inv_dq.x, (by - qy) * inv_dq.y), (bz - qz) * inv_dq.z);
    n_exit = fminf(n_exit, 1048576.0f);
    int n = (int)ceilf(n_exit);
    n = n < 1 ? 1 : n;
    for (int k = 0; k < n; ++k) {
      r.n_marched++;
      r.t += P.tstep;
      if (r.t > r.tfar) return false;
      r.pos = r.pos + r.step;
    }
  }
}

// per-ray constants of the skip phase: the ray's step in the absorption texture's voxel space
VR_HD void skip_setup(const KernelParams &P, const Ray &r, F3 bscale, F3 &dq, F3 &inv_dq) {
  dq = mk3(r.step.x * bscale.x * P.ab.fw, r.step.y * bscale.y * P.ab.fh, r.step.z * bscale.z * P.ab.fd);
  const float big = 3.0e38f;
  inv_dq = mk3(dq.x != 0.0f ? 1.0f / dq.x : big, dq.y != 0.0f ? 1.0f / dq.y : big, dq.z != 0.0f ? 1.0f / dq.z : big);
}

// The whole march of one pixel, lane-serial: what the kernel does per lane, without the warp
// scheduling.  Used by the host emulation (tests/host_emul) and by the single-lane debug kernel.
template <int GRAD, int LAYOUT, typename E>
VR_HD bool march_pixel(const KernelParams &P, uint32_t x, uint32_t y, float rgb[3], uint32_t &marched, uint32_t &shaded) {
  const F3 origin = ray_origin(P);
  Ray r;
  if (!ray_init(P, x, y, origin, r)) return false;
  const F3 bmin = ld3(P.boxmin), ext = ld3(P.boxmax) - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);
  F3 dq, inv_dq;
  skip_setup(P, r, bscale, dq, inv_dq);
  bool active = true;
  while (active) {
    if (P.skip) active = skip_empty(P, r, bmin, bscale, dq, inv_dq);
    if (!active) break;
    shade_sample<GRAD, LAYOUT, E>(P, origin, r);
    active = ray_advance(P, r);
  }
  rgb[0] = r.sx, rgb[1] = r.sy, rgb[2] = r.sz;
  marched = r.n_marched, shaded = r.n_shaded;
  return true;
}

}  // namespace vrb
#endif
