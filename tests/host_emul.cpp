/* host_emul.cpp - TEST INFRASTRUCTURE: runs the product's per-lane device code
 * (volume_renderer_b200/csrc/raymarch_core.h, every function is __host__ __device__) on the CPU,
 * lane-serially, over padded (bricked) storage built here with the same geometry rules as
 * context.cu / resident_kernels.cu.  It lets the `-m "not gpu"` suite compare the very
 * arithmetic the kernel executes with the oracle before any GPU time is spent.  It is not a
 * product path: nothing under volume_renderer_b200/ builds or loads it.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../include/vr_abi.h"
#include "../volume_renderer_b200/csrc/raymarch_core.h"
#include "../volume_renderer_b200/csrc/render_setup.h"

using namespace vrb;

namespace {

struct HostTex {
  std::vector<float> store;
  std::vector<uint32_t> occ;
  std::vector<uint8_t> dist;
  TexDesc t{};
  int ocx = 0, ocy = 0, ocz = 0;
};

template <int LAYOUT>
void build(HostTex &T, const float *src, int w, int h, int d, bool with_occ) {
  memset(&T.t, 0, sizeof T.t);
  if (!src || w == 0) {
    T.t.kind = TEX_UNBOUND;
    return;
  }
  if (w == 1 && h == 1 && d == 1) {
    T.t.kind = TEX_CONST;
    T.t.cval = src[0];
    T.t.w = T.t.h = T.t.d = 1;
    return;
  }
  const uint64_t pw = w + 2, ph = h + 2, pd = d + 2;
  uint64_t elems;
  if (LAYOUT == LAYOUT_BRICK) {
    const uint64_t nbx = (pw + 7) / 8, nby = (ph + 7) / 8, nbz = (pd + 7) / 8;
    T.t.sy = (uint32_t)(nbx * 512), T.t.sz = (uint32_t)(nbx * nby * 512);
    elems = nbx * nby * nbz * 512;
  } else {
    T.t.sy = (uint32_t)pw, T.t.sz = (uint32_t)(pw * ph);
    elems = pw * ph * pd;
  }
  T.store.assign(elems, 0.0f);
  typedef Addr<LAYOUT> A;
  for (int Z = 0; Z < (int)pd; ++Z)
    for (int Y = 0; Y < (int)ph; ++Y)
      for (int X = 0; X < (int)pw; ++X) {
        const int sx = X - 1 < 0 ? 0 : (X - 1 > w - 1 ? w - 1 : X - 1);
        const int sy = Y - 1 < 0 ? 0 : (Y - 1 > h - 1 ? h - 1 : Y - 1);
        const int sz = Z - 1 < 0 ? 0 : (Z - 1 > d - 1 ? d - 1 : Z - 1);
        T.store[A::x(X) + A::y(Y, T.t.sy) + A::z(Z, T.t.sz)] = src[((size_t)sz * h + sy) * w + sx];
      }
  T.t.p = T.store.data();
  T.t.w = w, T.t.h = h, T.t.d = d;
  T.t.fw = (float)w, T.t.fh = (float)h, T.t.fd = (float)d;
  T.t.kind = TEX_VOLUME;
  if (with_occ) {
    const int dil = 2;
    T.ocx = (w + 7) / 8, T.ocy = (h + 7) / 8, T.ocz = (d + 7) / 8;
    T.occ.assign(((size_t)T.ocx * T.ocy * T.ocz + 31) / 32, 0u);
    for (int cz = 0; cz < T.ocz; ++cz)
      for (int cy = 0; cy < T.ocy; ++cy)
        for (int cx = 0; cx < T.ocx; ++cx) {
          bool any = false;
          for (int z = cz * 8 - dil; z <= cz * 8 + 7 + dil && !any; ++z)
            for (int y = cy * 8 - dil; y <= cy * 8 + 7 + dil && !any; ++y)
              for (int x = cx * 8 - dil; x <= cx * 8 + 7 + dil && !any; ++x) {
                const int sx = x < 0 ? 0 : (x > w - 1 ? w - 1 : x), sy = y < 0 ? 0 : (y > h - 1 ? h - 1 : y),
                          sz = z < 0 ? 0 : (z > d - 1 ? d - 1 : z);
                any = !(src[((size_t)sz * h + sy) * w + sx] == 0.0f);
              }
          const size_t cell = ((size_t)cz * T.ocy + cy) * T.ocx + cx;
          if (any) T.occ[cell >> 5] |= 1u << (cell & 31);
        }
    // distance grid (resident_kernels.cu: k_dist_pass), brute force here
    const size_t ncell = (size_t)T.ocx * T.ocy * T.ocz;
    T.dist.assign(ncell, 255);
    std::vector<int> occl;
    for (size_t cell = 0; cell < ncell; ++cell)
      if ((T.occ[cell >> 5] >> (cell & 31)) & 1u) occl.push_back((int)cell);
    for (size_t cell = 0; cell < ncell; ++cell) {
      const int cx = (int)(cell % T.ocx), cy = (int)((cell / T.ocx) % T.ocy), cz = (int)(cell / ((size_t)T.ocx * T.ocy));
      int best = 255;
      for (int o : occl) {
        const int ox = o % T.ocx, oy = (o / T.ocx) % T.ocy, oz = o / (T.ocx * T.ocy);
        const int dd = std::max(std::abs(cx - ox), std::max(std::abs(cy - oy), std::abs(cz - oz)));
        best = std::min(best, dd);
      }
      T.dist[cell] = (uint8_t)best;
    }
  }
}

// Depth slices of the volume-space passes, lane-serial: what the workers of one pixel's S slices do in
// raymarch_kernel.cu - slice starts by the exact fast-forward (k_slice_states), the slice's end as the ray's tfar, pass 1 one
// opacity per slice folded as 1 - prod(1 - a_s), pass 2 every slice started from the bricks in front and the slices in front and
// the shares summed in slice order - with the product's own helpers.
int g_slices = 1;
template <int GRAD, int LAYOUT, int UNI>
bool march_pixel_sliced(const KernelParams &P, uint32_t x, uint32_t y, float rgba[4], uint32_t &marched, uint32_t &shaded) {
  const int S = P.vs_slices;
  const F3 origin = ray_origin(P.cam);
  Ray r0;
  if (!ray_init(P, P.cam, x, y, origin, r0)) return false;
  const F3 bmin = ld3(P.boxmin), ext = ld3(P.boxmax) - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);
  const size_t plane = (size_t)P.W * P.H, pixel = (size_t)x * P.H + y;
  float a_front = 0.0f;
  const bool in_brick = P.vs_mode == 2 ? vs_opacity_in(P, origin, r0, pixel, a_front) : true;
  const SliceCut cut = vs_slice_cut(P, origin, r0, bmin, ext, S);
  Ray walker = r0;
  bool walker_alive = cut.inside;
  float T1 = 1.0f, sum[3] = {0.0f, 0.0f, 0.0f};
  marched = shaded = 0;
  for (int s = 0; s < S; ++s) {
    Ray r = r0;
    bool active = true;
    if (s > 0) {
      if (walker_alive) walker_alive = fast_forward_to(P, walker, vs_slice_bound(cut, s, S));
      if (walker_alive) r.t = walker.t, r.pos = walker.pos;
      else active = false;
    }
    const float ts1 = vs_slice_bound(cut, s + 1, S);
    if (ts1 < 3.0e38f) r.tfar = fminf(r.tfar, float_below(ts1));
    if (r.t > r.tfar) active = false;
    if (P.vs_mode == 2) {
      float T = 1.0f - a_front;
      for (int q = 0; q < s; ++q) T = T * (1.0f - P.vs_slice_opacity[(size_t)q * plane + pixel]);
      r.sw = 1.0f - T;
      if (!in_brick || r.sw > P.thr) active = false;
    }
    F3 dq, inv_dq;
    skip_setup(P, r, bscale, dq, inv_dq);
    while (active) {
      if (P.skip) active = skip_empty(P, r, bmin, bscale, dq, inv_dq);
      if (!active) break;
      shade_sample<GRAD, LAYOUT, ElemF32, UNI>(P, origin, r);
      active = ray_advance(P, r);
    }
    if (P.vs_mode == 1) {
      P.vs_slice_opacity[(size_t)s * plane + pixel] = r.sw;
      T1 = T1 * (1.0f - r.sw);
    } else if (s == 0) {
      sum[0] = r.sx, sum[1] = r.sy, sum[2] = r.sz;
    } else {
      sum[0] += r.sx, sum[1] += r.sy, sum[2] += r.sz;
    }
    marched += r.n_marched, shaded += r.n_shaded;
  }
  rgba[0] = sum[0], rgba[1] = sum[1], rgba[2] = sum[2], rgba[3] = 1.0f - T1;
  return true;
}

template <int GRAD, int LAYOUT, int UNI>
void run(const KernelParams &P, float *out, uint64_t *stats) {
  const size_t size = (size_t)P.W * P.H;
  if (P.vs_mode == 1)
    memset(P.vs_opacity_out, 0, size * sizeof(float));
  else
    memset(out, 0, size * 3 * sizeof(float));
  uint64_t marched = 0, shaded = 0, hits = 0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : marched, shaded, hits)
  for (int64_t x = 0; x < P.W; ++x)
    for (int y = 0; y < P.H; ++y) {
      float rgb[4];
      uint32_t m = 0, s = 0;
      const bool hit = (P.vs_mode && P.vs_slices > 1) ? march_pixel_sliced<GRAD, LAYOUT, UNI>(P, (uint32_t)x, (uint32_t)y, rgb, m, s)
                                                        : march_pixel<GRAD, LAYOUT, ElemF32, UNI>(P, (uint32_t)x, (uint32_t)y, rgb, m, s);
      if (!hit) continue;
      const size_t k = (size_t)x * P.H + y;
      if (P.vs_mode == 1) {
        P.vs_opacity_out[k] = rgb[3];
        continue;
      }
      out[k] = rgb[0], out[k + size] = rgb[1], out[k + 2 * size] = rgb[2];
      marched += m, shaded += s, hits += 1;
    }
  if (stats) stats[0] = marched, stats[1] = shaded, stats[2] = hits;
}

template <int LAYOUT>
int emul(const vr_volume *em, const vr_volume *re, const vr_volume *gx, const vr_volume *gy, const vr_volume *gz,
         const vr_volume *lut, const vr_light *lights, size_t n_lights, const float factors[3], const float es[3],
         const uint64_t res[2], const float R[9], const float props[3], float thr, const float color[3], int skip,
         int frac_bits, int variant, int n_bricks, const uint64_t *boxes_vox, float *out, uint64_t *stats) {
  HostTex Em, Re, Gx, Gy, Gz, Lut;
  auto ptr = [](const vr_volume *v) { return v ? static_cast<const float *>(v->data) : nullptr; };
  auto dim = [](const vr_volume *v, int i) { return v ? (int)v->dims[i] : 0; };
  build<LAYOUT>(Em, ptr(em), dim(em, 0), dim(em, 1), dim(em, 2), true);
  build<LAYOUT>(Re, ptr(re), dim(re, 0), dim(re, 1), dim(re, 2), false);
  build<LAYOUT>(Gx, ptr(gx), dim(gx, 0), dim(gx, 1), dim(gx, 2), false);
  build<LAYOUT>(Gy, ptr(gy), dim(gy, 0), dim(gy, 1), dim(gy, 2), false);
  build<LAYOUT>(Gz, ptr(gz), dim(gz, 0), dim(gz, 1), dim(gz, 2), false);
  build<LAYOUT_LINEAR>(Lut, ptr(lut), dim(lut, 0), dim(lut, 1), dim(lut, 2), false);
  KernelParams P;
  memset(&P, 0, sizeof P);
  setup_frame(P, factors, es, res[1], res[0], R, props, thr, color, em->dims[0], em->dims[1], em->dims[2]);
  P.em = P.ab = P.em_direct = Em.t;  // Em = Ab aliased, the examples' case
  P.ab_is_em = 1;
  P.re = Re.t;
  P.gx = Gx.t, P.gy = Gy.t, P.gz = Gz.t;
  P.lut = Lut.t;
  P.gradient_method = gx ? 1 : 0;
  P.frac_bits = frac_bits;
  P.skip = Em.t.kind == TEX_VOLUME ? skip : 0;
  P.occ = Em.occ.data(), P.dist = Em.dist.data(), P.ocx = Em.ocx, P.ocy = Em.ocy, P.ocz = Em.ocz;
  // shared tap sets exactly as context.cu decides them
  P.uniform = 1;
  bool have_geo = false;
  const TexDesc *vt[] = {&P.em, &P.ab, &P.re, &P.em_direct, &P.gx, &P.gy, &P.gz};
  for (const TexDesc *t : vt) {
    if (t->kind != TEX_VOLUME) continue;
    if (!have_geo) P.geo = *t, P.geo.p = nullptr, have_geo = true;
    if (t->w != P.geo.w || t->h != P.geo.h || t->d != P.geo.d || t->sy != P.geo.sy || t->sz != P.geo.sz) P.uniform = 0;
  }
  if (variant == 1) P.uniform = 0;
  if (P.frac_bits) P.uniform = 0;
  P.count = 1;
  P.n_lights = lut ? (int)n_lights : 0;
  if (P.n_lights > MAX_PARAM_LIGHTS) return 2;
  for (int i = 0; i < P.n_lights; ++i) {
    P.lights[i][0] = lights[i].position[2], P.lights[i][1] = lights[i].position[1], P.lights[i][2] = lights[i].position[0];
    memcpy(P.lights[i] + 3, lights[i].color, 12);
  }
  auto go = [&](float *dst, uint64_t *st) {
    if (P.gradient_method) {
      if (P.uniform) run<1, LAYOUT, 1>(P, dst, st); else run<1, LAYOUT, 0>(P, dst, st);
    } else {
      if (P.uniform) run<0, LAYOUT, 1>(P, dst, st); else run<0, LAYOUT, 0>(P, dst, st);
    }
  };
  if (n_bricks <= 0) {
    go(out, stats);
    return 0;
  }
  // volume-space mode on virtual shards (whole storage, region test only): opacity pass per brick,
  // shading pass per brick, sum of the shares in brick order - what multigpu.VolumeSpace does on GPUs
  const size_t plane = (size_t)P.W * P.H;
  const int64_t dims[3] = {(int64_t)em->dims[0], (int64_t)em->dims[1], (int64_t)em->dims[2]};
  std::vector<float> opac((size_t)n_bricks * plane, 0.f), boxes((size_t)n_bricks * 6), part(plane * 3);
  // depth slices (emul_set_slices): one set of per-slice opacity planes per brick, written by pass 1, read by pass 2
  const int S = g_slices > 1 && P.skip == 2 ? g_slices : 1;
  std::vector<float> slice_opac(S > 1 ? (size_t)n_bricks * S * plane : 0, 0.f);
  P.vs_slices = S;
  for (int b = 0; b < n_bricks; ++b)
    for (int i = 0; i < 3; ++i) {
      const float ext = P.boxmax[i] - P.boxmin[i];
      boxes[6 * b + i] = P.boxmin[i] + ext * ((float)boxes_vox[6 * b + i] / (float)dims[i]);
      boxes[6 * b + 3 + i] = P.boxmin[i] + ext * ((float)boxes_vox[6 * b + 3 + i] / (float)dims[i]);
    }
  auto region = [&](int b) {
    for (int i = 0; i < 3; ++i) {
      const int64_t lo = (int64_t)boxes_vox[6 * b + i], hi = (int64_t)boxes_vox[6 * b + 3 + i];
      P.reg_lo[i] = lo <= 0 ? -INFINITY : (float)lo / (float)dims[i];
      P.reg_hi[i] = hi >= dims[i] ? INFINITY : (float)hi / (float)dims[i];
    }
    P.vs_index = b, P.vs_count = n_bricks;
  };
  for (int b = 0; b < n_bricks; ++b) {
    region(b);
    P.vs_mode = 1;
    P.vs_opacity_out = opac.data() + (size_t)b * plane;
    P.vs_slice_opacity = S > 1 ? slice_opac.data() + (size_t)b * S * plane : nullptr;
    go(nullptr, nullptr);
  }
  memset(out, 0, plane * 3 * sizeof(float));
  if (stats) stats[0] = stats[1] = stats[2] = 0;
  P.vs_opacity_all = opac.data(), P.vs_boxes = boxes.data();
  for (int b = 0; b < n_bricks; ++b) {
    region(b);
    P.vs_mode = 2;
    P.vs_slice_opacity = S > 1 ? slice_opac.data() + (size_t)b * S * plane : nullptr;
    uint64_t st[3] = {0, 0, 0};
    go(part.data(), st);
    for (size_t i = 0; i < plane * 3; ++i) out[i] += part[i];
    if (stats) stats[0] += st[0], stats[1] += st[1], stats[2] += st[2];  // the shading passes of all bricks
  }
  return 0;
}

}  // namespace

extern "C" int emul_render(const vr_volume *em, const vr_volume *re, const vr_volume *gx, const vr_volume *gy,
                           const vr_volume *gz, const vr_volume *lut, const vr_light *lights, size_t n_lights,
                           const float factors[3], const float es[3], const uint64_t res[2], const float R[9],
                           const float props[3], float thr, const float color[3], int layout, int skip, int frac_bits,
                           int variant, int n_bricks, const uint64_t *boxes_vox, float *out, uint64_t *stats) {
  if (!em) return 1;
  if (layout == LAYOUT_BRICK)
    return emul<LAYOUT_BRICK>(em, re, gx, gy, gz, lut, lights, n_lights, factors, es, res, R, props, thr, color, skip, frac_bits, variant, n_bricks, boxes_vox, out, stats);
  return emul<LAYOUT_LINEAR>(em, re, gx, gy, gz, lut, lights, n_lights, factors, es, res, R, props, thr, color, skip, frac_bits, variant, n_bricks, boxes_vox, out, stats);
}

extern "C" void emul_set_slices(int slices) { g_slices = slices; }

// unit-test hooks for the small numeric helpers of raymarch_core.h
extern "C" float emul_acos_poly(float x) { return acos_poly(x); }
extern "C" uint64_t emul_div_pi_mismatches(uint32_t first_bits, uint32_t last_bits, uint32_t stride) {
  uint64_t bad = 0;
  for (uint64_t u = first_bits; u <= last_bits; u += stride) {
    float x;
    const uint32_t b = (uint32_t)u;
    memcpy(&x, &b, 4);
    const float q = div_pi(x), ref = x / VR_PI_F;
    if (!(q == ref)) ++bad;
  }
  return bad;
}

// Depth slices: fast_forward_to must land on the very (t, pos) the reference's serial recurrences
// `t += tstep; pos += step` (volumeRender_kernel.cu:488-492) reach at the first sample with t >= t_until - bit for bit.
// Random rays through the box [-1,1]^3 (directions with tiny and zero components, starts on a face so that coordinates cross
// zero and many binades); returns the number of rays on which the two disagree (state or "the ray ended first").
static uint32_t lcg(uint32_t &s) { return s = s * 1664525u + 1013904223u; }
static float urand(uint32_t &s) { return (float)(lcg(s) >> 8) * (1.0f / 16777216.0f); }
extern "C" uint64_t emul_fast_forward_mismatches(uint32_t seed, uint32_t n_rays) {
  uint64_t bad = 0;
  uint32_t s = seed;
  for (uint32_t i = 0; i < n_rays; ++i) {
    KernelParams P;
    memset(&P, 0, sizeof P);
    const int n = 64 << (lcg(s) % 7u);  // 64 .. 4096 voxels across
    P.tstep = 1.0f / (2.2f * (float)n * 1.41421356f);
    float d[3] = {urand(s) * 2.0f - 1.0f, urand(s) * 2.0f - 1.0f, urand(s) * 2.0f - 1.0f};
    const uint32_t kind = lcg(s) % 8u;
    if (kind == 0) d[lcg(s) % 3u] = 0.0f;          // axis-parallel in one coordinate
    if (kind == 1) d[lcg(s) % 3u] *= 1.0e-6f;      // ... nearly
    if (kind == 2) d[lcg(s) % 3u] *= 1.0e-3f;
    const float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (!(len > 1.0e-3f)) continue;
    Ray r;
    memset(&r, 0, sizeof r);
    r.step = mk3(d[0] / len * P.tstep, d[1] / len * P.tstep, d[2] / len * P.tstep);
    r.pos = mk3(urand(s) * 2.0f - 1.0f, urand(s) * 2.0f - 1.0f, urand(s) * 2.0f - 1.0f);
    const uint32_t face = lcg(s) % 3u;  // start on a face of the box
    (face == 0 ? r.pos.x : face == 1 ? r.pos.y : r.pos.z) = (lcg(s) & 1u) ? 1.0f : -1.0f;
    r.t = 3.0f + urand(s) * 3.0f;  // the examples' distances: t in the binades [2,4) and [4,8)
    const uint32_t steps_total = 1u + lcg(s) % (uint32_t)(4 * n);
    r.tfar = r.t + (float)steps_total * P.tstep;
    const float t_until = r.t + urand(s) * 1.2f * (r.tfar - r.t);  // sometimes beyond the ray's end
    Ray a = r, b = r;
    bool alive_a = true;
    while (a.t < t_until) {  // the reference's loop tail, one sample at a time
      a.t += P.tstep;
      if (a.t > a.tfar) {
        alive_a = false;
        break;
      }
      a.pos = a.pos + a.step;
    }
    const bool alive_b = fast_forward_to(P, b, t_until);
    if (alive_a != alive_b) {
      ++bad;
      continue;
    }
    if (alive_a && (f2u(a.t) != f2u(b.t) || f2u(a.pos.x) != f2u(b.pos.x) || f2u(a.pos.y) != f2u(b.pos.y) || f2u(a.pos.z) != f2u(b.pos.z))) ++bad;
  }
  return bad;
}
extern "C" uint64_t emul_float_below_mismatches(uint32_t seed, uint32_t n) {
  uint64_t bad = 0;
  uint32_t s = seed;
  for (uint32_t i = 0; i < n; ++i) {
    float x = u2f(lcg(s));
    if (!(x == x) || std::isinf(x)) continue;
    if (float_below(x) != nextafterf(x, -INFINITY)) ++bad;
  }
  const float special[] = {0.0f, -0.0f, 1.0f, 1.17549435e-38f, 1.4e-45f, -1.4e-45f, 3.0e38f};
  for (float x : special)
    if (float_below(x) != nextafterf(x, -INFINITY)) ++bad;
  return bad;
}
