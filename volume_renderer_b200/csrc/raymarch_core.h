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
#include <string.h>

#include "vr_types.h"

// Arithmetic flavour of the march (one set of kernels per flavour, selected per context by VR_OPT_ARITHMETIC):
//   0  "exact": every operation individually rounded (the file is compiled with -fmad=false), IEEE sqrt and divide - bit-equal
//      to the CPU checker, which is what makes the bit-level tests possible;
//   1  "reference": what nvcc makes of the reference's own source - FMA contraction on, normalize() by rsqrtf
//      (helper_math.h:1325-1328), IEEE sqrtf / divide in angle() (kernel.cu:284-287);
//   2  "fast": flavour 1 with the cosine of angle() as dot * rsqrt(|a|^2 |b|^2) (one MUFU instead of two IEEE square roots and
//      an IEEE divide) and x / pi as a multiplication.  The angles only address a 64^3 table through 8-bit filter weights.
#ifndef VR_ARITH
#define VR_ARITH 0
#endif
#if defined(__CUDA_ARCH__)
#define VR_LDG(ptr) __ldg(ptr)
#define VR_EXPF(x) __expf(x) /* the reference's intrinsic, kernel.cu:456 */
#if VR_ARITH == 2
#define VR_RSQRT(x) vrb::rsqrt_ftz(x) /* MUFU.RSQ alone: no scaling of subnormal arguments (a squared length below 1e-38 reads as 0) */
#elif VR_ARITH == 1
#define VR_RSQRT(x) rsqrtf(x) /* helper_math.h:1325-1328 */
#else
#define VR_RSQRT(x) (1.0f / sqrtf(x)) /* IEEE divide + sqrt: the normal must not depend on MUFU.RSQ's last bits */
#endif
#else
#define VR_LDG(ptr) (*(ptr))
#define VR_EXPF(x) exp2f((x) * 1.4426950408889634f)
#define VR_RSQRT(x) (1.0f / sqrtf(x))
#endif

#ifndef VR_MAGIC_FLOOR
#define VR_MAGIC_FLOOR 1
#endif
#ifndef VR_LUT_ADJACENT
#define VR_LUT_ADJACENT 1
#endif
#ifndef VR_EARLY_RE
#define VR_EARLY_RE 0
#endif
#ifndef VR_SPLIT_LIGHT_LOOP
#define VR_SPLIT_LIGHT_LOOP 1
#endif
#ifndef VR_TEX_NO_VS
#define VR_TEX_NO_VS 1
#endif
#ifndef VR_TEX_FAST
#define VR_TEX_FAST 1
#endif
#ifndef VR_FAST_CELL
#define VR_FAST_CELL 1
#endif
// texture-unit kernels, lookupGradient: the reflection and the three gradient fetches are issued WITH the emission fetch,
// before the alpha test - one texture round trip less in the dependent chain of a sample, ~10 % of those fetches wasted
#ifndef VR_TEX_SPECULATE
#define VR_TEX_SPECULATE 0
#endif
// texture-unit kernels: the lights are shaded two at a time - both sets of angles first, then both LUT fetches back to back -
// so that the second light's fetch does not wait for the first light's result.  Measured on B200 (configs[1], ms per frame):
// 13.17 -> 12.53; with the speculative fetches above as well 13.04, those alone 13.60 (hence off).  Same operations per light in
// the same order of accumulation: pixels are bit-identical.
#ifndef VR_LIGHT_PAIRS
#define VR_LIGHT_PAIRS 1
#endif

namespace vrb {

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float rsqrt_ftz(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
#endif

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
  // voxel `off` of a 4-element interleaved storage: one 16-byte load
  static VR_HD void load3(const void *p, uint32_t off, float &a, float &b, float &c) {
#if defined(__CUDA_ARCH__)
    const float4 v = __ldg(static_cast<const float4 *>(p) + off);
    a = v.x, b = v.y, c = v.z;
#else
    const float *q = static_cast<const float *>(p) + 4 * (size_t)off;
    a = q[0], b = q[1], c = q[2];
#endif
  }
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
  static VR_HD void load3(const void *p, uint32_t off, float &a, float &b, float &c) {  // one 8-byte load
#if defined(__CUDA_ARCH__)
    const uint2 v = __ldg(static_cast<const uint2 *>(p) + off);
    a = __uint_as_float(v.x << 16), b = __uint_as_float(v.x & 0xffff0000u), c = __uint_as_float(v.y << 16);
#else
    a = load(p, 4 * off), b = load(p, 4 * off + 1), c = load(p, 4 * off + 2);
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
  static VR_HD size_t z(int Z, uint32_t sz) { return (size_t)(uint32_t)Z * sz; }  // the z term alone may pass 2^32 elements
};
template <>
struct Addr<LAYOUT_TEXHW> : Addr<LAYOUT_LINEAR> {};  // never addressed (texture-unit launches), only instantiated
template <>
struct Addr<LAYOUT_BRICK> {
  static VR_HD uint32_t x(int X) { return (((uint32_t)X >> BRICK_SHIFT) << 9) + ((uint32_t)X & BRICK_MASK); }
  static VR_HD uint32_t y(int Y, uint32_t sy) { return ((uint32_t)Y >> BRICK_SHIFT) * sy + (((uint32_t)Y & BRICK_MASK) << 3); }
  static VR_HD size_t z(int Z, uint32_t sz) { return (size_t)((uint32_t)Z >> BRICK_SHIFT) * sz + (((uint32_t)Z & BRICK_MASK) << 6); }
};

// One axis of a normalized-coordinate linear-filter fetch (CUDA Programming Guide, "Linear
// Filtering"; configured at kernel.cu:544-548): xB = x*N - 0.5, i = floor(xB), a = frac(xB).
// x is clamped to [0,1] in float first: that IS clamp-to-edge addressing (any x<0 filters to
// the edge voxel, as does x=0), it maps a NaN coordinate to 0 (fmaxf returns the non-NaN
// operand - the texture unit's rule), and it bounds i to [-1, N-1] so that with the padded
// storage no integer clamps are needed.  X0 = i+1 is the padded index of the low tap.
// floor(xb) as a float and as an int for xb in [-0.5, 2^22).  On the device: one add rounded towards
// -inf onto 1.5 * 2^23 leaves floor(xb) in the low mantissa bits (ulp there is 1), which replaces
// FRND + F2I - both quarter-rate conversion-pipe instructions - by two FADDs and an integer
// subtract; the values are those of floorf / (int).
struct AxisCoord {
  float xb;  // x*N - 0.5 of the clamped coordinate
  float t;   // device: floor(xb) + 1.5 * 2^23; host: floor(xb)
  VR_HD float fl() const {  // floor(xb)
#if defined(__CUDA_ARCH__) && VR_MAGIC_FLOOR
    return t - 12582912.0f;
#else
    return t;
#endif
  }
  VR_HD int i() const {  // (int)floor(xb), in [-1, N-1]
#if defined(__CUDA_ARCH__) && VR_MAGIC_FLOOR
    return __float_as_int(t) - 0x4B400000;
#else
    return (int)t;
#endif
  }
};
VR_HD AxisCoord axis_coord(float x, float fN) {
  AxisCoord a;
  x = fminf(fmaxf(x, 0.0f), 1.0f);
  a.xb = x * fN - 0.5f;
#if defined(__CUDA_ARCH__) && VR_MAGIC_FLOOR
  a.t = __fadd_rd(a.xb, 12582912.0f);
#else
  a.t = floorf(a.xb);
#endif
  return a;
}
// the macro cell of a sample (see cell_at)
struct CellAt {
  AxisCoord ax, ay, az;
  int cx, cy, cz;
  uint32_t cell;
};
VR_HD void axis_finish(const AxisCoord &a, int frac_bits, int &X0, float &f) {
  f = a.xb - a.fl();
  if (frac_bits) f = floorf(f * 256.0f + 0.5f) * (1.0f / 256.0f);  // 1.8 fixed-point weights
  X0 = a.i() + 1;
}
VR_HD void axis_setup(float x, float fN, int frac_bits, int &X0, float &f) { axis_finish(axis_coord(x, fN), frac_bits, X0, f); }
// one fused multiply-add per lerp, written explicitly so that the checker (oracle.cpp: lerp) and
// this file round identically whatever the contraction flags are
VR_HD float lerpf(float a, float b, float f) { return fmaf(f, b - a, a); }

// tex3D on a resident volume (kind == TEX_VOLUME).
template <int LAYOUT, typename E>
VR_HD float tex3d_volume(const TexDesc &t, float x, float y, float z, int frac_bits) {
  int X, Y, Z;
  float fx, fy, fz;
  axis_setup(x, t.fw, frac_bits, X, fx);
  axis_setup(y, t.fh, frac_bits, Y, fy);
  axis_setup(z, t.fd, frac_bits, Z, fz);
  X -= t.ox, Y -= t.oy, Z -= t.oz;  // global padded index -> local storage (0 for a whole volume)
  typedef Addr<LAYOUT> A;
  typedef typename E::T T;
  const uint32_t x0 = A::x(X), x1 = A::x(X + 1);
  const uint32_t y0 = A::y(Y, t.sy), y1 = A::y(Y + 1, t.sy);
  // offsets inside one z layer fit 32 bits (checked at upload); the layer itself is a 64-bit term
  const size_t z0 = A::z(Z, t.sz), z1 = A::z(Z + 1, t.sz);
  const uint32_t o[4] = {y0 + x0, y0 + x1, y1 + x0, y1 + x1};
  float c[8];
  if (t.estride > 1) {  // a component of the interleaved gradient storage
    const T *q = static_cast<const T *>(t.p) + t.ecomp;
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = E::load(q + ((i < 4 ? z0 : z1) + o[i & 3]) * (size_t)t.estride, 0u);
  } else {
    const T *q0 = static_cast<const T *>(t.p) + z0, *q1 = static_cast<const T *>(t.p) + z1;
    if (LAYOUT == LAYOUT_LINEAR && VR_LUT_ADJACENT) {
      // the high x tap is the next element: four addresses, the +1 rides in the load's immediate offset
      // (measured: -2.4 % in the lookupGradient kernels, +2.5 % in the computeGradient ones - see raymarch_kernel.cu)
      c[0] = E::load(q0 + o[0], 0u), c[1] = E::load(q0 + o[0], 1u), c[2] = E::load(q0 + o[2], 0u), c[3] = E::load(q0 + o[2], 1u);
      c[4] = E::load(q1 + o[0], 0u), c[5] = E::load(q1 + o[0], 1u), c[6] = E::load(q1 + o[2], 0u), c[7] = E::load(q1 + o[2], 1u);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) c[i] = E::load(i < 4 ? q0 : q1, o[i & 3]);
    }
  }
  const float c000 = c[0], c100 = c[1], c010 = c[2], c110 = c[3], c001 = c[4], c101 = c[5], c011 = c[6], c111 = c[7];
  const float c00 = lerpf(c000, c100, fx), c10 = lerpf(c010, c110, fx);
  const float c01 = lerpf(c001, c101, fx), c11 = lerpf(c011, c111, fx);
  return lerpf(lerpf(c00, c10, fy), lerpf(c01, c11, fy), fz);
}
// tex3D on any slot: unbound reads 0, a 1x1x1 volume is its single value (filtering a constant
// is exact with a + f*(b-a) lerps), anything else is fetched.
template <int LAYOUT, typename E>
VR_HD float tex3d_any(const TexDesc &t, float x, float y, float z, int frac_bits) {
  if (LAYOUT == LAYOUT_TEXHW) {
#if defined(__CUDA_ARCH__)
    // the reference's own fetch (kernel.cu:439-447 on textures configured at :544-548): normalized coordinates,
    // clamp-to-edge, hardware linear filter with 1.8 fixed-point weights; a NaN coordinate reads as 0 in hardware
    if (t.kind == TEX_VOLUME) return tex3D<float>((cudaTextureObject_t)t.tex, x, y, z);
#endif
    return t.kind == TEX_CONST ? t.cval : 0.0f;
  } else {
    if (t.kind == TEX_VOLUME) return tex3d_volume<LAYOUT == LAYOUT_TEXHW ? LAYOUT_LINEAR : LAYOUT, E>(t, x, y, z, frac_bits);
    return t.kind == TEX_CONST ? t.cval : 0.0f;
  }
}

// the texture-unit fetch itself, for slots the host has verified to be volumes (KernelParams::tex_fast)
VR_HD float tex_raw(const TexDesc &t, float x, float y, float z) {
#if defined(__CUDA_ARCH__)
  return tex3D<float>((cudaTextureObject_t)t.tex, x, y, z);
#else
  (void)t, (void)x, (void)y, (void)z;
  return 0.0f;
#endif
}

// ---- shared tap sets ---------------------------------------------------------------------
// When every volume texture of a launch has the same geometry (the examples' case: emission,
// reflection and the gradient volumes are all w x h x d), the eight tap offsets and the three
// weights of a sample position are computed ONCE and reused for every volume fetched there:
// the arithmetic per fetch is exactly tex3d_volume's, only the common subexpressions are shared.
struct AxisTap {
  uint32_t o0, o1;  // storage offsets contributed by the two taps of this axis (x, y: inside one z layer)
  float f;          // weight of the high tap
};
// The z term is the one that can pass 2^32 elements (a resident volume beyond 1625^3 voxels): ZT is
// uint32_t in the regular kernels and size_t in the WIDE ones, which cost two more registers per lane.
template <typename ZT>
struct AxisTapZ {
  ZT o0, o1;
  float f;
};
template <int LAYOUT>
VR_HD AxisTap axis_tap_x(float x, float fN, int off, int fb) {
  int X;
  AxisTap a;
  axis_setup(x, fN, fb, X, a.f);
  X -= off;
  a.o0 = Addr<LAYOUT>::x(X), a.o1 = Addr<LAYOUT>::x(X + 1);
  return a;
}
template <int LAYOUT>
VR_HD AxisTap axis_tap_y(float y, float fN, uint32_t sy, int off, int fb) {
  int Y;
  AxisTap a;
  axis_setup(y, fN, fb, Y, a.f);
  Y -= off;
  a.o0 = Addr<LAYOUT>::y(Y, sy), a.o1 = Addr<LAYOUT>::y(Y + 1, sy);
  return a;
}
// UNI (template parameter of the kernels): 0 = per-texture addressing, 1 = shared tap sets,
// 2 = shared tap sets with 64-bit z terms (a resident volume of more than 2^32 elements)
template <int UNI>
struct ZSel {
  typedef uint32_t type;
};
template <>
struct ZSel<2> {
  typedef size_t type;
};
template <int LAYOUT, typename ZT>
VR_HD AxisTapZ<ZT> axis_tap_z(float z, float fN, uint32_t sz, int off, int fb) {
  int Z;
  AxisTapZ<ZT> a;
  axis_setup(z, fN, fb, Z, a.f);
  Z -= off;
  a.o0 = (ZT)Addr<LAYOUT>::z(Z, sz), a.o1 = (ZT)Addr<LAYOUT>::z(Z + 1, sz);
  return a;
}
template <int LAYOUT, typename ZT>
VR_HD void axis_taps_from(const TexDesc &g, const AxisCoord &cx, const AxisCoord &cy, const AxisCoord &cz, int fb, AxisTap &ax, AxisTap &ay,
                          AxisTapZ<ZT> &az) {
  int X, Y, Z;
  axis_finish(cx, fb, X, ax.f), axis_finish(cy, fb, Y, ay.f), axis_finish(cz, fb, Z, az.f);
  X -= g.ox, Y -= g.oy, Z -= g.oz;
  ax.o0 = Addr<LAYOUT>::x(X), ax.o1 = Addr<LAYOUT>::x(X + 1);
  ay.o0 = Addr<LAYOUT>::y(Y, g.sy), ay.o1 = Addr<LAYOUT>::y(Y + 1, g.sy);
  az.o0 = (ZT)Addr<LAYOUT>::z(Z, g.sz), az.o1 = (ZT)Addr<LAYOUT>::z(Z + 1, g.sz);
}
template <typename E, typename ZT>
VR_HD float taps_fetch(const void *p, const AxisTap &ax, const AxisTap &ay, const AxisTapZ<ZT> &az) {
  typedef typename E::T T;
  const T *p0 = static_cast<const T *>(p) + az.o0, *p1 = static_cast<const T *>(p) + az.o1;
  const uint32_t o00 = ax.o0 + ay.o0, o10 = ax.o1 + ay.o0, o01 = ax.o0 + ay.o1, o11 = ax.o1 + ay.o1;  // shared by every fetch of a tap set
  const float c000 = E::load(p0, o00), c100 = E::load(p0, o10);
  const float c010 = E::load(p0, o01), c110 = E::load(p0, o11);
  const float c001 = E::load(p1, o00), c101 = E::load(p1, o10);
  const float c011 = E::load(p1, o01), c111 = E::load(p1, o11);
  const float c00 = lerpf(c000, c100, ax.f), c10 = lerpf(c010, c110, ax.f);
  const float c01 = lerpf(c001, c101, ax.f), c11 = lerpf(c011, c111, ax.f);
  return lerpf(lerpf(c00, c10, ay.f), lerpf(c01, c11, ay.f), az.f);
}
// Gx, Gy, Gz of the interleaved gradient storage through one tap set: eight 16-byte loads, then the
// seven lerps of taps_fetch per channel (same operations on the same values as three taps_fetch).
template <typename E, typename ZT>
VR_HD F3 taps_fetch3(const void *p, const AxisTap &ax, const AxisTap &ay, const AxisTapZ<ZT> &az) {
  typedef typename E::T T;
  const T *p0 = static_cast<const T *>(p) + 4 * (size_t)az.o0, *p1 = static_cast<const T *>(p) + 4 * (size_t)az.o1;
  const uint32_t o00 = ax.o0 + ay.o0, o10 = ax.o1 + ay.o0, o01 = ax.o0 + ay.o1, o11 = ax.o1 + ay.o1;
  // row by row (two 16-byte loads, three x-lerps) to keep few values live: r[row] = x-interpolated (Gx, Gy, Gz)
  F3 r[4];
#pragma unroll
  for (int row = 0; row < 4; ++row) {
    const T *q = row < 2 ? p0 : p1;
    const uint32_t lo = (row & 1) ? o01 : o00, hi = (row & 1) ? o11 : o10;
    float a0, b0, c0, a1, b1, c1;
    E::load3(q, lo, a0, b0, c0), E::load3(q, hi, a1, b1, c1);
    r[row] = mk3(lerpf(a0, a1, ax.f), lerpf(b0, b1, ax.f), lerpf(c0, c1, ax.f));
  }
  F3 g;
  g.x = lerpf(lerpf(r[0].x, r[1].x, ay.f), lerpf(r[2].x, r[3].x, ay.f), az.f);
  g.y = lerpf(lerpf(r[0].y, r[1].y, ay.f), lerpf(r[2].y, r[3].y, ay.f), az.f);
  g.z = lerpf(lerpf(r[0].z, r[1].z, ay.f), lerpf(r[2].z, r[3].z, ay.f), az.f);
  return g;
}
// any slot through a shared tap set (the slot is a volume of the common geometry, a constant,
// or unbound)
template <typename E, typename ZT>
VR_HD float taps_any(const TexDesc &t, const AxisTap &ax, const AxisTap &ay, const AxisTapZ<ZT> &az) {
  if (t.kind == TEX_VOLUME) return taps_fetch<E>(t.p, ax, ay, az);
  return t.kind == TEX_CONST ? t.cval : 0.0f;
}

// ---- ray state ---------------------------------------------------------------------------
struct Ray {
  F3 pos, step;
  float t, tfar;
  float sx, sy, sz, sw;  // float4 sum
  uint32_t n_marched, n_shaded;
#ifdef VR_TRACE_PIXEL
  uint32_t trace_slot, trace_n;  // developer build: 0 = this lane is not traced
#endif
};
// Developer build (-DVR_TRACE_PIXEL -DVR_TRACE_X0=.. -DVR_TRACE_Y0=..): the lanes of an 8x8 pixel box keep their latest state in
// host-mapped memory (KernelParams::warp_times), so that it survives a fault of the kernel.
#if defined(VR_TRACE_PIXEL) && defined(__CUDA_ARCH__)
#define VR_TRACE(P, r, stage, a, b)                                                                                              \
  do {                                                                                                                           \
    if ((r).trace_slot) {                                                                                                        \
      volatile unsigned long long *e_ = (P).warp_times + 8ull * (r).trace_slot;                                                  \
      e_[1] = __float_as_uint((r).t) | ((unsigned long long)__float_as_uint((r).tfar) << 32);                                    \
      e_[2] = __float_as_uint((r).pos.x) | ((unsigned long long)__float_as_uint((r).pos.y) << 32);                               \
      e_[3] = __float_as_uint((r).pos.z) | ((unsigned long long)(r).n_marched << 32);                                            \
      e_[4] = (unsigned long long)(a), e_[5] = (unsigned long long)(b);                                                          \
      e_[0] = (unsigned long long)(stage) | ((unsigned long long)(++const_cast<Ray &>(r).trace_n) << 8);                                         \
      __threadfence_system();                                                                                                    \
    }                                                                                                                            \
  } while (0)
#else
#define VR_TRACE(P, r, stage, a, b) \
  do {                              \
  } while (0)
#endif

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

VR_HD F3 ray_origin(const Camera &C) {
  // kernel.cu:399-410: origin = camXOffset * xVec + (-1 * objDist * zVec)
  const F3 xV = ld3(C.m[0]), zV = ld3(C.m[2]);
  return (C.m[3][0] * xV) + (-1.0f * C.m[3][2] * zV);
}

// Ray generation + box clip (kernel.cu:388-431).  Returns false when the ray misses the box
// (the pixel is then never written and keeps the memset 0).
VR_HD bool ray_init(const KernelParams &P, const Camera &C, uint32_t x, uint32_t y, F3 origin, Ray &r) {
  const float u = (x / (float)P.W) * 2.0f - 1.0f;
  const float ratio = P.H / (float)P.W;
  const float v = (y / (float)P.H) * 2.0f * ratio - 1.0f * ratio;
  const F3 xV = ld3(C.m[0]), yV = ld3(C.m[1]), zV = ld3(C.m[2]);
  const F3 dir = normalize3(u * normalize3(xV) + v * yV + C.m[3][1] * zV);
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

// acos for the illumination angles.  The device's own acosf and glibc's already differ in the
// last place, so the angle is the one quantity of the march that is compared by tolerance, not
// bit for bit; what must agree exactly is WHEN it is NaN (argument outside [-1,1] or NaN), and
// that is preserved: sqrt of a negative number is NaN.  Abramowitz & Stegun 4.4.46,
// acos(x) = sqrt(1-x) * P7(x) on [0,1], |error| <= 2e-8, mirrored for x < 0: 13 instructions
// instead of ~30.  The host emulation keeps acosf (bit-equal to the checker).
VR_HD float acos_poly(float x) {
  const float ax = fabsf(x);
  float s;
#if defined(__CUDA_ARCH__)
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(1.0f - ax));
#else
  s = sqrtf(1.0f - ax);
#endif
  float p = -0.0012624911f;
  p = fmaf(p, ax, 0.0066700901f);
  p = fmaf(p, ax, -0.0170881256f);
  p = fmaf(p, ax, 0.0308918810f);
  p = fmaf(p, ax, -0.0501743046f);
  p = fmaf(p, ax, 0.0889789874f);
  p = fmaf(p, ax, -0.2145988016f);
  p = fmaf(p, ax, 1.5707963050f);
  const float r = s * p;
  return x < 0.0f ? VR_PI_F - r : r;
}
VR_HD float vr_acosf(float x) {
#if defined(__CUDA_ARCH__)
  return acos_poly(x);
#else
  return acosf(x);
#endif
}
// x / pi, correctly rounded, as reciprocal multiply + one exact-residual correction (3 instructions
// instead of an IEEE divide).  Bit-equal to `x / VR_PI_F` for every float in (1e-30, 4], for 0 and
// for NaN - checked exhaustively (tests/test_host_logic.py::test_div_pi_is_exact); the angles fed
// to it are 0, NaN or >= 3e-4.
VR_HD float div_pi(float x) {
#if VR_ARITH == 2 && defined(__CUDA_ARCH__)
  return x * (1.0f / VR_PI_F);
#endif
  const float rcp = 1.0f / VR_PI_F;
  const float q0 = x * rcp;
  return fmaf(fmaf(-q0, VR_PI_F, x), rcp, q0);
}

// angle(a, b) / pi (kernel.cu:284-287, :331-339) from dot(a, b) and the two SQUARED lengths.  Flavours 0 and 1: the reference's
// own operations, dot / (sqrt(|a|^2) * sqrt(|b|^2)) with IEEE square roots and divide (the square roots of a vector that
// several angles share are common subexpressions).  Flavour 2: dot * rsqrt(|a|^2 |b|^2).  NaN whenever the reference's is
// (a zero vector, an argument beyond [-1, 1]): acos_poly takes the square root of a negative number.
// (A/B on B200, configs[1], ms per frame: no guard 12.53 | guard inline at 1e-5 13.01, at 2e-6 13.02 | out of line 12.92, at 2e-6
// 12.90: the rare path is not what costs, keeping its operands live is)
#ifndef VR_COS_GUARD
#define VR_COS_GUARD 0.999998f
#endif
#ifndef VR_COS_GUARD_NOINLINE
#define VR_COS_GUARD_NOINLINE 1
#endif
#if defined(__CUDA_ARCH__)
#if VR_COS_GUARD_NOINLINE
__device__ __noinline__
#else
__device__ __forceinline__
#endif
float cos_reference(float d, float la2, float lb2) { return d / (sqrtf(la2) * sqrtf(lb2)); }  // kernel.cu:286, IEEE sqrt and divide
#endif
VR_HD float angle_over_pi(float d, float la2, float lb2) {
#if VR_ARITH == 2 && defined(__CUDA_ARCH__)
  float cs = d * rsqrt_ftz(la2 * lb2);
  // Within 2e-6 of +-1 the last bit decides between an angle and NaN (an argument beyond [-1, 1]; the texture unit reads a NaN
  // coordinate as 0, so at -1 the LUT coordinate jumps from 1 to 0): there - a few samples in a million - the cosine is the
  // reference's own expression, so that the decision falls as it does in flavour 1.  Measured at 1024^3 / 2048^2 against the
  // reference kernel's frame: 47 -> 11 pixels beyond 1e-3 of 4.19 M, PSNR 102.4 -> 111.5 dB - flavour 1's figures.
  if (fabsf(cs) > VR_COS_GUARD) cs = cos_reference(d, la2, lb2);
  return div_pi(vr_acosf(cs));
#else
  return div_pi(vr_acosf(d / (sqrtf(la2) * sqrtf(lb2))));
#endif
}

// shade (kernel.cu:308-353) given the gradient and the reflection sample: N = -normalize(g),
// then per light the three angles, the LUT fetch and the accumulation.  The reference recomputes
// angle(N, lightIn), the projection of lightIn and the reflection fetch for every light although
// none of them depends on the light (kernel.cu:329-344); they are evaluated once here - the same
// operations on the same operands, so the values are those of the reference's loop.
template <int LAYOUT>
VR_HD F3 shade_lights(const KernelParams &P, F3 g, float reflection, F3 pos, F3 origin, F3 color, int fb) {
  const F3 N = -1.0f * normalize3(g);
  const F3 lightIn = origin - pos;
  const float nn = dot3(N, N), ii = dot3(lightIn, lightIn);
  const float a = angle_over_pi(dot3(N, lightIn), nn, ii);  // angle(N, lightIn)/pi
  const F3 lightInProj = lightIn - (dot3(lightIn, N) * N);
  const float pp = dot3(lightInProj, lightInProj);
  F3 illum = mk3(0.0f, 0.0f, 0.0f);
  // (measured on B200, all slower under the 72-register budget: the loop unrolled by two +3.5 %; the angles of two lights
  // first and their sixteen LUT loads together +2.2 %; see also the notes in shade_sample)
  for (int i = 0; i < P.n_lights; ++i) {
    F3 lpos, lcol;
    if (i < MAX_PARAM_LIGHTS) {  // kernel-parameter space: register-indexed constant loads (a pointer select would go generic)
      lpos = mk3(P.lights[i][0], P.lights[i][1], P.lights[i][2]), lcol = mk3(P.lights[i][3], P.lights[i][4], P.lights[i][5]);
    } else {
      const float *L = P.lights_more + 6 * (i - MAX_PARAM_LIGHTS);
      lpos = mk3(VR_LDG(L), VR_LDG(L + 1), VR_LDG(L + 2)), lcol = mk3(VR_LDG(L + 3), VR_LDG(L + 4), VR_LDG(L + 5));
    }
    const F3 lightOut = lpos - pos;
    const float b = angle_over_pi(dot3(N, lightOut), nn, dot3(lightOut, lightOut));
    const F3 lightOutProj = lightOut - (dot3(lightOut, N) * N);
    const float c = angle_over_pi(dot3(lightInProj, lightOutProj), pp, dot3(lightOutProj, lightOutProj));
    // the LUT is always fp32 in the linear padded layout
    float light;
    if (VR_TEX_FAST && LAYOUT == LAYOUT_TEXHW && (P.tex_fast & 16)) light = tex_raw(P.lut, a, b, c);
    else light = tex3d_any<LAYOUT == LAYOUT_TEXHW ? LAYOUT_TEXHW : LAYOUT_LINEAR, ElemF32>(P.lut, a, b, c, fb);
    illum = illum + reflection * light * lcol * color;  // kernel.cu:349
  }
  return illum;
}

// The same loop for the texture-unit kernels, as TWO loops instead of a select per light: in one loop the compiler predicates
// both ways of reading a light and issues the six (switched-off) global loads of the overflow list for every light that sits
// in the kernel parameters.  Measured: texture-unit kernels 15.25 -> 15.15 ms; the load kernels lose 1.5 % with it
// (23.76 -> 24.12), so they keep the loop above.
VR_HD F3 shade_lights_split(const KernelParams &P, F3 g, float reflection, F3 pos, F3 origin, F3 color, int fb) {
  const F3 N = -1.0f * normalize3(g);
  const F3 lightIn = origin - pos;
  const float nn = dot3(N, N), ii = dot3(lightIn, lightIn);
  const float a = angle_over_pi(dot3(N, lightIn), nn, ii);  // angle(N, lightIn)/pi
  const F3 lightInProj = lightIn - (dot3(lightIn, N) * N);
  const float pp = dot3(lightInProj, lightInProj);
  F3 illum = mk3(0.0f, 0.0f, 0.0f);
  auto one_light = [&](const F3 lpos, const F3 lcol) {
    const F3 lightOut = lpos - pos;
    const float b = angle_over_pi(dot3(N, lightOut), nn, dot3(lightOut, lightOut));
    const F3 lightOutProj = lightOut - (dot3(lightOut, N) * N);
    const float c = angle_over_pi(dot3(lightInProj, lightOutProj), pp, dot3(lightOutProj, lightOutProj));
    const float light = (P.tex_fast & 16) ? tex_raw(P.lut, a, b, c) : tex3d_any<LAYOUT_TEXHW, ElemF32>(P.lut, a, b, c, fb);
    illum = illum + reflection * light * lcol * color;  // kernel.cu:349
  };
  const int n_param = P.n_lights < MAX_PARAM_LIGHTS ? P.n_lights : MAX_PARAM_LIGHTS;
  int i0 = 0;
#if VR_LIGHT_PAIRS
  if (P.tex_fast & 16) {
    auto angles = [&](const F3 lpos, float &b, float &c) {
      const F3 lightOut = lpos - pos;
      b = angle_over_pi(dot3(N, lightOut), nn, dot3(lightOut, lightOut));
      const F3 lightOutProj = lightOut - (dot3(lightOut, N) * N);
      c = angle_over_pi(dot3(lightInProj, lightOutProj), pp, dot3(lightOutProj, lightOutProj));
    };
    for (; i0 + 1 < n_param; i0 += 2) {  // the same operations per light, in the same order of accumulation
      float b0, c0, b1, c1;
      angles(mk3(P.lights[i0][0], P.lights[i0][1], P.lights[i0][2]), b0, c0);
      angles(mk3(P.lights[i0 + 1][0], P.lights[i0 + 1][1], P.lights[i0 + 1][2]), b1, c1);
      const float l0 = tex_raw(P.lut, a, b0, c0), l1 = tex_raw(P.lut, a, b1, c1);
      illum = illum + reflection * l0 * mk3(P.lights[i0][3], P.lights[i0][4], P.lights[i0][5]) * color;
      illum = illum + reflection * l1 * mk3(P.lights[i0 + 1][3], P.lights[i0 + 1][4], P.lights[i0 + 1][5]) * color;
    }
  }
#endif
  for (int i = i0; i < n_param; ++i)  // kernel-parameter space: register-indexed constant loads
    one_light(mk3(P.lights[i][0], P.lights[i][1], P.lights[i][2]), mk3(P.lights[i][3], P.lights[i][4], P.lights[i][5]));
  for (int i = MAX_PARAM_LIGHTS; i < P.n_lights; ++i) {
    const float *L = P.lights_more + 6 * (i - MAX_PARAM_LIGHTS);
    one_light(mk3(VR_LDG(L), VR_LDG(L + 1), VR_LDG(L + 2)), mk3(VR_LDG(L + 3), VR_LDG(L + 4), VR_LDG(L + 5)));
  }
  return illum;
}
template <int LAYOUT>
struct Lights {
  static VR_HD F3 shade(const KernelParams &P, F3 g, float reflection, F3 pos, F3 origin, F3 color, int fb) {
    return shade_lights<LAYOUT>(P, g, reflection, pos, origin, color, fb);
  }
};
#if VR_SPLIT_LIGHT_LOOP
template <>
struct Lights<LAYOUT_TEXHW> {
  static VR_HD F3 shade(const KernelParams &P, F3 g, float reflection, F3 pos, F3 origin, F3 color, int fb) {
    return shade_lights_split(P, g, reflection, pos, origin, color, fb);
  }
};
#endif

// One full iteration of the loop kernel.cu:435-493 up to and including the "under" update.
// GRAD: 0 = computeGradient (kernel.cu:212-253), 1 = lookupGradient (kernel.cu:266-276).
// UNI : every volume texture of the launch shares one geometry -> shared tap sets.
// Fixed-point weights (frac_bits, an A/B option) are only compiled into the UNI == 0 kernels; the
// host selects those when the option is set.
template <int GRAD, int LAYOUT, typename E, int UNI>
VR_HD void shade_sample(const KernelParams &P, F3 origin, Ray &r) {
  const F3 bmin = ld3(P.boxmin), bmax = ld3(P.boxmax);
  const F3 ext = bmax - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);  // kernel.cu:396
  const F3 ps = (r.pos - bmin) * bscale;                            // kernel.cu:437
  const int fb = UNI ? 0 : P.frac_bits;
  // UNI means "shared tap sets" for the load kernels; for the texture-unit kernels it marks the instantiation for contexts
  // that own a region (volume-space passes): region test compiled in, coordinates remapped into the region's local arrays
  constexpr bool TAPS = UNI != 0 && LAYOUT != LAYOUT_TEXHW;
  constexpr bool TEXVS = UNI != 0 && LAYOUT == LAYOUT_TEXHW;
  // plain texture-unit launches are never volume-space passes (render_common): their kernels carry no region test
  const int vs_mode = (VR_TEX_NO_VS && LAYOUT == LAYOUT_TEXHW && !TEXVS) ? 0 : P.vs_mode;
  if (vs_mode && !(ps.x >= P.reg_lo[0] && ps.x < P.reg_hi[0] && ps.y >= P.reg_lo[1] && ps.y < P.reg_hi[1] && ps.z >= P.reg_lo[2] &&
                     ps.z < P.reg_hi[2])) {
    r.n_marched++;  // a sample of the global lattice that another brick owns
    return;
  }
  typedef typename ZSel<UNI>::type ZT;
  AxisTap ax, ay;
  AxisTapZ<ZT> az;
  float em_raw, ab_raw;
#if VR_TEX_SPECULATE
  bool spec = false;
  F3 spec_g = mk3(0.0f, 0.0f, 0.0f);
  float spec_re = 0.0f;
#endif
  // where the volume textures are fetched: the normalized volume coordinate, remapped for a region's local arrays
  const F3 tcs = TEXVS ? ld3(P.tex_cs) : mk3(1.0f, 1.0f, 1.0f), tco = TEXVS ? ld3(P.tex_co) : mk3(0.0f, 0.0f, 0.0f);
  const F3 pt = TEXVS ? mk3(fmaf(ps.x, tcs.x, tco.x), fmaf(ps.y, tcs.y, tco.y), fmaf(ps.z, tcs.z, tco.z)) : ps;
  if (TAPS) {
    // (handing the skip phase's x*N - 0.5 and floors of this position over instead of computing
    // them again costs more in live registers than it saves: measured +2.5 %)
    axis_taps_from<LAYOUT>(P.geo, axis_coord(ps.x, P.geo.fw), axis_coord(ps.y, P.geo.fh), axis_coord(ps.z, P.geo.fd), fb, ax, ay, az);
    em_raw = taps_any<E>(P.em, ax, ay, az);
    ab_raw = P.ab_is_em ? em_raw : taps_any<E>(P.ab, ax, ay, az);
  } else if (VR_TEX_FAST && LAYOUT == LAYOUT_TEXHW && (P.tex_fast & 8)) {
    em_raw = tex_raw(P.em, pt.x, pt.y, pt.z);
#if VR_TEX_SPECULATE
    if (GRAD == 1 && (P.tex_fast & 3) == 3 && P.n_lights > 0) {
      spec = true;
      spec_g = mk3(tex_raw(P.gx, pt.x, pt.y, pt.z), tex_raw(P.gy, pt.x, pt.y, pt.z), tex_raw(P.gz, pt.x, pt.y, pt.z));
      spec_re = P.re_is_em ? em_raw : tex_raw(P.re, pt.x, pt.y, pt.z);
    }
#endif
    ab_raw = P.ab_is_em ? em_raw : tex3d_any<LAYOUT, E>(P.ab, pt.x, pt.y, pt.z, fb);
  } else {
    em_raw = tex3d_any<LAYOUT, E>(P.em, pt.x, pt.y, pt.z, fb);
    ab_raw = P.ab_is_em ? em_raw : tex3d_any<LAYOUT, E>(P.ab, pt.x, pt.y, pt.z, fb);
  }
#if VR_EARLY_RE
  // The reflection taps are issued together with the emission taps, before the alpha test: nine of ten samples that
  // reach this point are shaded (the grid check has already removed the empty cells), so one exposed load latency
  // per sample is traded for a few wasted loads.  Measured: lookupGradient 23.9 -> 23.5 ms, computeGradient 27.5 -> 27.9
  // (hence per family, raymarch_kernel.cu).  Same operations on the same values.
  float re_early = 0.0f;
  if (TAPS && P.n_lights > 0 && !P.re_is_em) re_early = taps_any<E>(P.re, ax, ay, az);
#endif
  // (the gradient taps issued here as well: 23.5 -> 25.1 ms, eight 16-byte loads in flight spill; the emission and
  // reflection loads all issued before the first lerp: +1.1 %.  Texture-unit launches: reflection + gradient fetches issued
  // here 15.9 -> 16.3 ms, the LUT fetches of two lights back to back 16.1, both 17.1 - the kernel is issue bound, the
  // speculative work of the unlit samples costs more than the latency it hides.  The emission fetch of pos + step issued one
  // sample ahead, behind this sample's shading: 15.13 -> 15.26 ms, one more live register and 16 more bytes of spills)
  const float emission = P.fEm * em_raw;  // kernel.cu:444-447
  const float absorption = P.fAb * ab_raw;
  const float alpha = 1 - VR_EXPF(-absorption * P.tstep);  // kernel.cu:455-456
  r.n_marched++;
  if (vs_mode == 1) {  // transmittance pass: only sum.w of kernel.cu:482, no colour
    r.sw = (1 - r.sw) * alpha + r.sw;
    return;
  }
  // alpha == 0 makes the premultiplied sample exactly (0,0,0,0) for finite inputs, so the
  // illumination (the expensive part) contributes nothing and is not evaluated.
  if (alpha == 0.0f) return;
  r.n_shaded++;
  const F3 color = ld3(P.color);
  const F3 colored = mk3(emission, emission, emission) * P.tstep * color;  // kernel.cu:459-460

  F3 illum = mk3(0.0f, 0.0f, 0.0f);
  if (P.n_lights > 0) {
    F3 g;
    float re_raw;
    if (TAPS) {
      if (GRAD == 0) {
        // computeGradient: the six taps differ from the sample position in ONE coordinate each
        // ((pos.y + 0 - bmin.y) * bscale.y is ps.y bit for bit), so only that axis is set up anew.
        if (P.em_direct.kind == TEX_VOLUME) {
          const void *p = P.em_direct.p;
          const F3 gs = ld3(P.gstep);
          const AxisTap xp = axis_tap_x<LAYOUT>((r.pos.x + gs.x - bmin.x) * bscale.x, P.geo.fw, P.geo.ox, fb);
          const AxisTap xm = axis_tap_x<LAYOUT>((r.pos.x - gs.x - bmin.x) * bscale.x, P.geo.fw, P.geo.ox, fb);
          g.x = taps_fetch<E>(p, xp, ay, az) - taps_fetch<E>(p, xm, ay, az);
          const AxisTap yp = axis_tap_y<LAYOUT>((r.pos.y + gs.y - bmin.y) * bscale.y, P.geo.fh, P.geo.sy, P.geo.oy, fb);
          const AxisTap ym = axis_tap_y<LAYOUT>((r.pos.y - gs.y - bmin.y) * bscale.y, P.geo.fh, P.geo.sy, P.geo.oy, fb);
          g.y = taps_fetch<E>(p, ax, yp, az) - taps_fetch<E>(p, ax, ym, az);
          const AxisTapZ<ZT> zp = axis_tap_z<LAYOUT, ZT>((r.pos.z + gs.z - bmin.z) * bscale.z, P.geo.fd, P.geo.sz, P.geo.oz, fb);
          const AxisTapZ<ZT> zm = axis_tap_z<LAYOUT, ZT>((r.pos.z - gs.z - bmin.z) * bscale.z, P.geo.fd, P.geo.sz, P.geo.oz, fb);
          g.z = taps_fetch<E>(p, ax, ay, zp) - taps_fetch<E>(p, ax, ay, zm);
          g = g * mk3(0.5f, 0.5f, 0.5f);
        } else {
          g = mk3(0.0f, 0.0f, 0.0f);  // a constant or unbound tex_emission: equal fetches, (c - c) * 0.5 = 0
        }
      } else {
        if (P.grad4) g = taps_fetch3<E>(P.gx.p, ax, ay, az);
        else g = mk3(taps_any<E>(P.gx, ax, ay, az), taps_any<E>(P.gy, ax, ay, az), taps_any<E>(P.gz, ax, ay, az));
      }
#if VR_EARLY_RE
      re_raw = P.re_is_em ? em_raw : re_early;
#else
      re_raw = P.re_is_em ? em_raw : taps_any<E>(P.re, ax, ay, az);
#endif
    } else {
#if VR_TEX_SPECULATE
      if (spec) {
        g = spec_g;
      } else
#endif
      if (VR_TEX_FAST && LAYOUT == LAYOUT_TEXHW && GRAD == 1 && (P.tex_fast & 1)) {
        // straight-line: the three fetches (and the reflection fetch below) share the coordinate registers of ps
        g = mk3(tex_raw(P.gx, pt.x, pt.y, pt.z), tex_raw(P.gy, pt.x, pt.y, pt.z), tex_raw(P.gz, pt.x, pt.y, pt.z));
      } else if (VR_TEX_FAST && LAYOUT == LAYOUT_TEXHW && GRAD == 0 && (P.tex_fast & 4)) {
        const F3 gs = ld3(P.gstep);
        // the offset taps differ from ps in ONE coordinate each ((pos.y + 0 - bmin.y) * bscale.y is ps.y bit for bit)
        const float xp = (r.pos.x + gs.x - bmin.x) * bscale.x, xm = (r.pos.x - gs.x - bmin.x) * bscale.x;
        const float yp = (r.pos.y + gs.y - bmin.y) * bscale.y, ym = (r.pos.y - gs.y - bmin.y) * bscale.y;
        const float zp = (r.pos.z + gs.z - bmin.z) * bscale.z, zm = (r.pos.z - gs.z - bmin.z) * bscale.z;
        if (TEXVS) {
          g.x = tex_raw(P.em_direct, fmaf(xp, tcs.x, tco.x), pt.y, pt.z) - tex_raw(P.em_direct, fmaf(xm, tcs.x, tco.x), pt.y, pt.z);
          g.y = tex_raw(P.em_direct, pt.x, fmaf(yp, tcs.y, tco.y), pt.z) - tex_raw(P.em_direct, pt.x, fmaf(ym, tcs.y, tco.y), pt.z);
          g.z = tex_raw(P.em_direct, pt.x, pt.y, fmaf(zp, tcs.z, tco.z)) - tex_raw(P.em_direct, pt.x, pt.y, fmaf(zm, tcs.z, tco.z));
        } else {
        g.x = tex_raw(P.em_direct, xp, ps.y, ps.z) - tex_raw(P.em_direct, xm, ps.y, ps.z);
        g.y = tex_raw(P.em_direct, ps.x, yp, ps.z) - tex_raw(P.em_direct, ps.x, ym, ps.z);
        g.z = tex_raw(P.em_direct, ps.x, ps.y, zp) - tex_raw(P.em_direct, ps.x, ps.y, zm);
        }
        g = g * mk3(0.5f, 0.5f, 0.5f);
      } else if (GRAD == 0) {
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
        g = mk3(tex3d_any<LAYOUT, E>(P.gx, pt.x, pt.y, pt.z, fb), tex3d_any<LAYOUT, E>(P.gy, pt.x, pt.y, pt.z, fb),
                tex3d_any<LAYOUT, E>(P.gz, pt.x, pt.y, pt.z, fb));
      }
      // the reflection fetch sits inside the reference's light loop (kernel.cu:342-344) but does
      // not depend on the light: fetched once
#if VR_TEX_SPECULATE
      if (spec) re_raw = spec_re;
      else
#endif
      if (VR_TEX_FAST && LAYOUT == LAYOUT_TEXHW && (P.tex_fast & 2)) re_raw = P.re_is_em ? em_raw : tex_raw(P.re, pt.x, pt.y, pt.z);
      else re_raw = P.re_is_em ? em_raw : tex3d_any<LAYOUT, E>(P.re, pt.x, pt.y, pt.z, fb);
    }
    illum = Lights<LAYOUT>::shade(P, g, P.fRe * re_raw, r.pos, origin, color, fb);
  }
  const F3 lit = colored + illum;
  const float shx = lit.x * alpha, shy = lit.y * alpha, shz = lit.z * alpha;  // kernel.cu:477-479
  const float k = 1 - r.sw;                                                    // kernel.cu:482
  r.sx = k * shx + r.sx;
  r.sy = k * shy + r.sy;
  r.sz = k * shz + r.sz;
  r.sw = k * alpha + r.sw;
}

// Volume-space shading pass: the opacity the ray has accumulated in the bricks it crosses BEFORE
// this one, 1 - prod(1 - a_b'), from the gathered per-brick opacities.  Bricks are disjoint convex
// boxes, so their intervals along a ray are ordered by entry distance.  Returns false when the ray
// never enters this brick's box (no owned samples).
VR_HD bool vs_opacity_in(const KernelParams &P, F3 origin, const Ray &r, size_t pixel, float &a_in) {
  const float *mine = P.vs_boxes + 6 * P.vs_index;
  float tn = 0.f, tf = 0.f;
  // a little slack: ownership is decided per sample by the region test, the boxes only order bricks
  const F3 e = mk3(1e-4f, 1e-4f, 1e-4f);
  if (!intersect_box(origin, r.step, ld3(mine) - e, ld3(mine + 3) + e, tn, tf)) return false;
  const size_t plane = (size_t)P.W * P.H;
  float T = 1.0f;
  for (int b = 0; b < P.vs_count; ++b) {
    if (b == P.vs_index) continue;
    const float *box = P.vs_boxes + 6 * b;
    float bn = 0.f, bf = 0.f;
    if (!intersect_box(origin, r.step, ld3(box), ld3(box + 3), bn, bf)) continue;
    if (bn < tn || (bn == tn && b < P.vs_index)) T = T * (1.0f - VR_LDG(P.vs_opacity_all + (size_t)b * plane + pixel));
  }
  a_in = 1.0f - T;
  return true;
}

// Depth slices of the volume-space passes.  The ray's stretch inside this context's region box (from reg_lo / reg_hi, the
// volume's faces where the region ends there) is cut into S equal parts of the ray parameter: slice s covers [bound(s), bound(s+1))
// with bound(0) = -inf and bound(S) = +inf, so samples the box test puts a hair outside still have an owner.  Every kernel
// evaluates the same expressions on the same values (the fma is spelled out: no contraction choices), and r.t itself is the
// bit-exact serial recurrence, so each lattice sample falls into exactly one slice, in every pass alike.  A ray that never
// enters the box has one slice, the first (ownership is decided per sample by the region test regardless).
struct SliceCut {
  float a, w;
  bool inside;
};
VR_HD SliceCut vs_slice_cut(const KernelParams &P, F3 origin, const Ray &r, F3 bmin, F3 ext, int S) {
  const F3 lo = mk3(P.reg_lo[0] > 0.0f ? bmin.x + ext.x * P.reg_lo[0] : bmin.x, P.reg_lo[1] > 0.0f ? bmin.y + ext.y * P.reg_lo[1] : bmin.y,
                    P.reg_lo[2] > 0.0f ? bmin.z + ext.z * P.reg_lo[2] : bmin.z);
  const F3 hi = mk3(P.reg_hi[0] < 1.0f ? bmin.x + ext.x * P.reg_hi[0] : bmin.x + ext.x, P.reg_hi[1] < 1.0f ? bmin.y + ext.y * P.reg_hi[1] : bmin.y + ext.y,
                    P.reg_hi[2] < 1.0f ? bmin.z + ext.z * P.reg_hi[2] : bmin.z + ext.z);
  float un = 0.0f, uf = 0.0f;  // in steps: pos = origin + step * u
  SliceCut c;
  c.inside = intersect_box(origin, r.step, lo, hi, un, uf);
  c.a = un * P.tstep;
  c.w = (uf * P.tstep - c.a) / (float)S;
  // (a ray that lies IN a face of the box gives 0 * inf = NaN in the slab test, as in the reference's intersectBox: one slice then)
  c.inside = c.inside && c.w >= 0.0f && c.a == c.a;
  return c;
}
VR_HD float vs_slice_bound(const SliceCut &c, int s, int S) {
  if (s <= 0) return -3.0e38f;
  if (s >= S || !c.inside) return 3.0e38f;
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(c.w, (float)s, c.a);
#else
  return fmaf(c.w, (float)s, c.a);
#endif
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
// one entry per 8^3 macro cell of the absorption-role texture; a cell is EMPTY when every voxel
// of the cell DILATED BY TWO VOXELS is zero.  A sample whose low tap index i = floor(x*N-0.5)
// lies in an empty cell reads only zeros, and so does any sample less than one voxel further
// along (taps i, i+1).  Sample positions of later samples must stay bit-identical to a march
// without skipping, i.e. to the reference's serial recurrences `t += tstep; pos += step`
// (kernel.cu:488-492).  Two implementations:
//
//  skip == 1 (serial): the lane computes how many steps keep it inside the cell and runs the
//    recurrences that many times.
//  skip == 2 (closed form, default): the grid stores the Chebyshev distance (in cells) to the
//    nearest non-empty cell, so a lane knows a whole cube of empty cells around it, and the
//    recurrences are fast-forwarded EXACTLY: while a float stays inside one binade (same sign
//    and exponent) repeated addition of a constant rounds the same way every time, so after
//    two real steps v0 -> v1 -> v2 inside a binade every later value is v2 + k*(v2 - v1)
//    exactly (the product and the sum are representable, see DESIGN.md "exact fast-forward").
//    The lane jumps k = min(steps to leave the empty cube, steps that keep t, x, y, z inside
//    their binades) samples at once and falls back to real steps at binade borders.
// Both return false when the ray ends inside the skipped span.

// per-ray constants of the skip phase: the ray's step in the absorption texture's voxel space
VR_HD void skip_setup(const KernelParams &P, const Ray &r, F3 bscale, F3 &dq, F3 &inv_dq) {
  dq = mk3(r.step.x * bscale.x * P.ab.fw, r.step.y * bscale.y * P.ab.fh, r.step.z * bscale.z * P.ab.fd);
  const float big = 3.0e38f;
  inv_dq = mk3(dq.x != 0.0f ? 1.0f / dq.x : big, dq.y != 0.0f ? 1.0f / dq.y : big, dq.z != 0.0f ? 1.0f / dq.z : big);
}

VR_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
VR_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}
// the largest float below x (x finite): t > float_below(x)  <=>  t >= x
VR_HD float float_below(float x) {
  const uint32_t u = f2u(x);
  if (x > 0.0f) return u2f(u - 1u);
  if (x < 0.0f) return u2f(u + 1u);
  return u2f(0x80000001u);  // below +-0: the smallest negative denormal
}
// v0, v1, v2 are normal numbers of one binade (same sign, same exponent), or all equal
VR_HD bool one_binade(float v0, float v1, float v2) {
  const uint32_t e0 = f2u(v0) >> 23, e1 = f2u(v1) >> 23, e2 = f2u(v2) >> 23;
  const uint32_t ex = e0 & 0xffu;
  // a coordinate that does not move at all (axis-parallel ray) trivially keeps its increment of 0
  return (e0 == e1 && e1 == e2 && ex != 0u && ex != 0xffu) || (v0 == v1 && v1 == v2);
}
// How many more additions of d keep v inside its binade (conservative; v normal, d a multiple of ulp(v)).
// Jumping straight to the border matters: around 0 the binades are short and every ray through the
// middle of the volume crosses dozens of them (halving the jump until it fits costs 15 %).
VR_HD float binade_room(float v, float d) {
  if (d == 0.0f) return 1.0e9f;
  const float a = fabsf(v), lo = u2f(f2u(a) & 0x7f800000u);
  const bool away = (d > 0.0f) == (v > 0.0f);
  const float span = away ? (lo + lo) - a : a - lo;  // exact: both operands in/at the binade
  return floorf(span / fabsf(d)) - 2.0f;
}

// the macro cell of the sample at r.pos and the lane's low-tap coordinates there
VR_HD CellAt cell_at(const KernelParams &P, const Ray &r, F3 bmin, F3 bscale) {
  CellAt c;
  const F3 ps = (r.pos - bmin) * bscale;
  c.ax = axis_coord(ps.x, P.ab.fw), c.ay = axis_coord(ps.y, P.ab.fh), c.az = axis_coord(ps.z, P.ab.fd);
  int ix = c.ax.i(), iy = c.ay.i(), iz = c.az.i();
  ix = ix < 0 ? 0 : ix, iy = iy < 0 ? 0 : iy, iz = iz < 0 ? 0 : iz;
  c.cx = ix >> CELL_SHIFT, c.cy = iy >> CELL_SHIFT, c.cz = iz >> CELL_SHIFT;
  c.cell = (uint32_t)((c.cz * P.ocy + c.cy) * P.ocx + c.cx);
  return c;
}
// steps until the low-tap index leaves the cube of (2*rad+1)^3 cells around the cell, for a
// per-step move of dq voxels
VR_HD float steps_to_exit(const CellAt &c, int rad, F3 dq, F3 inv_dq) {
  const float bx = (float)((dq.x >= 0.0f ? c.cx + 1 + rad : c.cx - rad) << CELL_SHIFT);
  const float by = (float)((dq.y >= 0.0f ? c.cy + 1 + rad : c.cy - rad) << CELL_SHIFT);
  const float bz = (float)((dq.z >= 0.0f ? c.cz + 1 + rad : c.cz - rad) << CELL_SHIFT);
  return fminf(fminf((bx - c.ax.xb) * inv_dq.x, (by - c.ay.xb) * inv_dq.y), (bz - c.az.xb) * inv_dq.z);
}
// one iteration of the reference's loop tail for an empty sample (kernel.cu:488-492)
VR_HD bool step_empty(const KernelParams &P, Ray &r) {
  r.n_marched++;
  r.t += P.tstep;
  if (r.t > r.tfar) return false;
  r.pos = r.pos + r.step;
  return true;
}

VR_HD bool skip_empty_serial(const KernelParams &P, Ray &r, F3 bmin, F3 bscale, F3 dq, F3 inv_dq) {
  for (;;) {
    const CellAt c = cell_at(P, r, bmin, bscale);
    const uint32_t word = VR_LDG(P.occ + (c.cell >> 5));
    if ((word >> (c.cell & 31)) & 1u) return true;  // something to shade here
    float n_exit = fminf(steps_to_exit(c, 0, dq, inv_dq), 1048576.0f);
    int n = (int)ceilf(n_exit);
    n = n < 1 ? 1 : n;
    for (int k = 0; k < n; ++k)
      if (!step_empty(P, r)) return false;
  }
}

// the sample at r.pos lies in a cell of the distance grid that is empty (P.skip == 2)
// On the device the cell comes from the world position with one multiply-add, two clamps and a magic-number floor per axis
// (19 instructions with the load instead of 35): the result may differ from cell_at's by one cell when the low tap sits
// within ~1e-4 voxels of a cell border, and either answer is safe there - a cell counts as empty only when every voxel of
// the cell DILATED BY TWO VOXELS is zero, while the taps of such a sample reach at most one voxel beyond the cell.  A lane
// that is told "empty" where the exact cell is not simply finds dist == 0 in skip_empty_jump and shades its (zero) sample;
// pixels and sample counts are those of the exact test.
VR_HD bool on_empty_cell(const KernelParams &P, const Ray &r, F3 bmin, F3 bscale) {
#if defined(__CUDA_ARCH__) && VR_FAST_CELL
  (void)bmin, (void)bscale;
  const float vx = fminf(fmaxf(fmaf(r.pos.x, P.cellA[0], P.cellB[0]), 0.0f), P.cellMax[0]);
  const float vy = fminf(fmaxf(fmaf(r.pos.y, P.cellA[1], P.cellB[1]), 0.0f), P.cellMax[1]);
  const float vz = fminf(fmaxf(fmaf(r.pos.z, P.cellA[2], P.cellB[2]), 0.0f), P.cellMax[2]);
  const int cx = __float_as_int(__fadd_rd(vx, 12582912.0f)) - 0x4B400000;
  const int cy = __float_as_int(__fadd_rd(vy, 12582912.0f)) - 0x4B400000;
  const int cz = __float_as_int(__fadd_rd(vz, 12582912.0f)) - 0x4B400000;
  VR_TRACE(P, r, 2, (uint32_t)((cz * P.ocy + cy) * P.ocx + cx), 0);
  return VR_LDG(P.dist + (uint32_t)((cz * P.ocy + cy) * P.ocx + cx)) != 0;
#else
  const CellAt c = cell_at(P, r, bmin, bscale);
  return VR_LDG(P.dist + c.cell) != 0;
#endif
}

// (The same test through the texture unit - the distance grid as a point-sampled 3D texture of bytes, one fetch at
// (x*N - 0.5)/8 per axis instead of the cell arithmetic and the load, 35 -> 10 instructions - was built and measured on the
// texture-unit kernels: 15.13 -> 15.28 ms, computeGradient 14.44 -> 14.96.  The byte is needed before the sample can be
// fetched, and a texture fetch takes several times the latency of a load that hits L1.)
// (Letting a lane shade the rest of a non-empty cell without coming back to the grid - a "fuel"
// counter - was tried: it saves the 38-instruction check per sample but lanes then reach the next
// empty cell out of phase and run the jump below one by one: 27 -> 42 ms.  Checking every sample
// keeps neighbouring rays in step.)
// (Depth slices need nothing here: a slice's end is the ray's tfar - see the kernel - and the jump stops there as it does at the
// end of a whole ray, counting every lattice sample up to it once.)
// STOP = 1 keeps an earlier formulation - an explicit end t_stop of the lane's stretch, 3e38 = none - alive for the WHOLE-RAY
// kernels only, which pass 3e38: with it they compile to the very machine code that ran every test, benchmark and the 360-frame
// turns of this round (checked by comparing the SASS); the sliced kernels use STOP = 0.  See DESIGN 4.5 for why that matters here.
template <int STOP>
VR_HD bool skip_empty_jump_t(const KernelParams &P, Ray &r, F3 bmin, F3 bscale, float t_stop) {
  for (;;) {
    if (STOP && !(r.t < t_stop)) return true;
    CellAt c = cell_at(P, r, bmin, bscale);
#if defined(VR_DEBUG_BOUNDS) && defined(__CUDA_ARCH__)
    if (c.cell >= (uint32_t)(P.ocx * P.ocy * P.ocz) && P.warp_times) {
      const unsigned long long slot = atomicAdd(P.warp_times + 2 * 8192, 1ull);
      if (slot < 1000) {
        unsigned long long *e = P.warp_times + 2 * 8192 + 8 + 6 * slot;
        e[0] = c.cell, e[1] = __float_as_uint(r.t) | ((unsigned long long)__float_as_uint(r.tfar) << 32);
        e[2] = __float_as_uint(r.pos.x) | ((unsigned long long)__float_as_uint(r.pos.y) << 32), e[3] = __float_as_uint(r.pos.z);
        e[4] = (unsigned long long)c.cx | ((unsigned long long)c.cy << 20) | ((unsigned long long)c.cz << 40), e[5] = r.n_marched;
      }
      c.cell = 0;
    }
#endif
    VR_TRACE(P, r, 3, c.cell, 0);
    const int dist = (int)VR_LDG(P.dist + c.cell);
    VR_TRACE(P, r, 4, c.cell, dist);
    if (dist == 0) return true;  // something to shade here
    // two real steps: both samples are still less than one voxel past an empty cell
    const float t0 = r.t;
    const F3 p0 = r.pos;
    if (!step_empty(P, r)) return false;
    const float t1 = r.t;
    const F3 p1 = r.pos;
    if (!step_empty(P, r)) return false;
    if (!(one_binade(t0, t1, r.t) && one_binade(p0.x, p1.x, r.pos.x) && one_binade(p0.y, p1.y, r.pos.y) &&
          one_binade(p0.z, p1.z, r.pos.z)))
      continue;  // at a binade border: keep stepping for real
    // the increments every further addition will produce (exact differences)
    const float dt = r.t - t1;
    const F3 d = r.pos - p1;
    // steps that keep the lane inside the empty cube, measured from where it stands now with the
    // increments it will really take
    const F3 ps = (r.pos - bmin) * bscale;
    CellAt now = c;
    now.ax.xb = fminf(fmaxf(ps.x, 0.0f), 1.0f) * P.ab.fw - 0.5f;
    now.ay.xb = fminf(fmaxf(ps.y, 0.0f), 1.0f) * P.ab.fh - 0.5f;
    now.az.xb = fminf(fmaxf(ps.z, 0.0f), 1.0f) * P.ab.fd - 0.5f;
    const F3 dq = mk3(d.x * bscale.x * P.ab.fw, d.y * bscale.y * P.ab.fh, d.z * bscale.z * P.ab.fd);
    const float big = 3.0e38f;
    const F3 inv_dq = mk3(dq.x != 0.0f ? 1.0f / dq.x : big, dq.y != 0.0f ? 1.0f / dq.y : big, dq.z != 0.0f ? 1.0f / dq.z : big);
    float kf = ceilf(steps_to_exit(now, dist - 1, dq, inv_dq));
    if (STOP && t_stop < 3.0e38f) kf = fminf(kf, fmaxf(ceilf((t_stop - r.t) / dt), 1.0f));
    kf = fminf(kf, fminf(fminf(binade_room(r.t, dt), binade_room(r.pos.x, d.x)), fminf(binade_room(r.pos.y, d.y), binade_room(r.pos.z, d.z))));
    kf = fminf(kf, 4194304.0f);
    if (!(kf >= 1.0f)) continue;
    // t after kf additions; does the ray end inside the span?
    const float t_end = r.t + kf * dt;
    if (t_end > r.tfar) {
      // smallest j with t + j*dt > tfar, found exactly: every t + j*dt, j <= kf, is representable
      float j = floorf((r.tfar - r.t) / dt) + 1.0f;
      j = fminf(fmaxf(j, 1.0f), kf);
      while (j > 1.0f && r.t + (j - 1.0f) * dt > r.tfar) j -= 1.0f;
      while (!(r.t + j * dt > r.tfar)) j += 1.0f;
      r.n_marched += (uint32_t)j;
      r.t = r.t + j * dt;
      return false;
    }
    const F3 p_end = r.pos + kf * d;
    if (!(one_binade(r.t, r.t, t_end) && one_binade(r.pos.x, r.pos.x, p_end.x) && one_binade(r.pos.y, r.pos.y, p_end.y) &&
          one_binade(r.pos.z, r.pos.z, p_end.z)))
      continue;  // cannot happen with the conservative room; stay exact regardless
    r.n_marched += (uint32_t)kf;
    r.t = t_end;
    r.pos = p_end;
    VR_TRACE(P, r, 5, (uint32_t)kf, dist);
  }
}

// Depth slices: run the reference's recurrences `t += tstep; pos += step` (kernel.cu:488-492) forward, EXACTLY, to the first
// lattice sample with t >= t_until - whatever lies on the way (it is another worker's to shade).  The same binade argument as in
// skip_empty_jump; no grid, no counting.  Returns false when the ray ends first.
VR_HD bool fast_forward_to(const KernelParams &P, Ray &r, float t_until) {
  for (;;) {
    if (!(r.t < t_until)) return true;
    if (!(r.t + 3.0f * P.tstep < t_until)) {  // within three samples: real steps, one at a time
      r.t += P.tstep;
      if (r.t > r.tfar) return false;
      r.pos = r.pos + r.step;
      continue;
    }
    const float t0 = r.t;
    const F3 p0 = r.pos;
    r.t += P.tstep;
    if (r.t > r.tfar) return false;
    r.pos = r.pos + r.step;
    const float t1 = r.t;
    const F3 p1 = r.pos;
    r.t += P.tstep;
    if (r.t > r.tfar) return false;
    r.pos = r.pos + r.step;
    if (!(one_binade(t0, t1, r.t) && one_binade(p0.x, p1.x, r.pos.x) && one_binade(p0.y, p1.y, r.pos.y) &&
          one_binade(p0.z, p1.z, r.pos.z)))
      continue;
    const float dt = r.t - t1;
    const F3 d = r.pos - p1;
    float kf = floorf((t_until - r.t) / dt) - 1.0f;  // up to, not into, the slice: the last steps are real ones
    kf = fminf(kf, fminf(fminf(binade_room(r.t, dt), binade_room(r.pos.x, d.x)), fminf(binade_room(r.pos.y, d.y), binade_room(r.pos.z, d.z))));
    kf = fminf(kf, 4194304.0f);
    if (!(kf >= 1.0f)) continue;
    const float t_end = r.t + kf * dt;
    if (t_end > r.tfar) return false;  // t_end < t_until: the ray ends before the slice starts
    const F3 p_end = r.pos + kf * d;
    if (!(one_binade(r.t, r.t, t_end) && one_binade(r.pos.x, r.pos.x, p_end.x) && one_binade(r.pos.y, r.pos.y, p_end.y) &&
          one_binade(r.pos.z, r.pos.z, p_end.z)))
      continue;
    r.t = t_end;
    r.pos = p_end;
  }
}

VR_HD bool skip_empty_jump(const KernelParams &P, Ray &r, F3 bmin, F3 bscale, float t_stop = 3.0e38f) {
  return skip_empty_jump_t<1>(P, r, bmin, bscale, t_stop);
}

VR_HD bool skip_empty(const KernelParams &P, Ray &r, F3 bmin, F3 bscale, F3 dq, F3 inv_dq) {
  return P.skip == 2 ? skip_empty_jump(P, r, bmin, bscale) : skip_empty_serial(P, r, bmin, bscale, dq, inv_dq);
}

// The whole march of one pixel, lane-serial: what the kernel does per lane, without the warp
// scheduling.  Used by the host emulation (tests/host_emul) and by the single-lane debug kernel.
template <int GRAD, int LAYOUT, typename E, int UNI>
VR_HD bool march_pixel(const KernelParams &P, uint32_t x, uint32_t y, float rgba[4], uint32_t &marched, uint32_t &shaded) {
  const F3 origin = ray_origin(P.cam);
  Ray r;
  if (!ray_init(P, P.cam, x, y, origin, r)) return false;
  const F3 bmin = ld3(P.boxmin), ext = ld3(P.boxmax) - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);
  F3 dq, inv_dq;
  skip_setup(P, r, bscale, dq, inv_dq);
  bool active = true;
  if (P.vs_mode == 2) {
    float a_in = 0.0f;
    active = vs_opacity_in(P, origin, r, (size_t)x * P.H + y, a_in);
    r.sw = a_in;
    if (a_in > P.thr) active = false;
  }
  while (active) {
    if (P.skip) active = skip_empty(P, r, bmin, bscale, dq, inv_dq);
    if (!active) break;
    shade_sample<GRAD, LAYOUT, E, UNI>(P, origin, r);
    active = ray_advance(P, r);
  }
  rgba[0] = r.sx, rgba[1] = r.sy, rgba[2] = r.sz, rgba[3] = r.sw;
  marched = r.n_marched, shaded = r.n_shaded;
  return true;
}

}  // namespace vrb
#endif
