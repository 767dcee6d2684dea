// resident_kernels.cu - kernels that build what stays resident in HBM between renders:
//   * repack: MATLAB-layout slices -> padded (bricked) storage (replaces cudaMalloc3DArray +
//     cudaMemcpy3D into an opaque cudaArray, src/C/vr/volumeRender_kernel.cu:522-554)
//   * occupancy: the empty-space bit grid of a resident volume (no reference counterpart)
//   * hg_lut: the Henyey-Greenstein table (src/C/mex/HenyeyGreenstein.cc:47-91) on the GPU
//   * synth: seeded synthetic ViBE-Z-shaped volumes generated straight into HBM
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <mutex>

#include "raymarch_core.h"
#include "raymarch_launch.h"

namespace vrb {

// ---- repack ---------------------------------------------------------------------------------
__device__ __forceinline__ float load_src(const void *p, size_t i, int dtype) {
  if (dtype == 0) return __ldg(static_cast<const float *>(p) + i);
  return __uint_as_float((uint32_t)__ldg(static_cast<const uint16_t *>(p) + i) << 16);
}
__device__ __forceinline__ void store_dst(void *p, size_t i, int dtype, float v) {
  if (dtype == 0) {
    static_cast<float *>(p)[i] = v;
  } else {
    static_cast<__nv_bfloat16 *>(p)[i] = __float2bfloat16_rn(v);
  }
}

// One thread per destination element, destination-offset order: a warp writes 128 contiguous
// bytes (fp32) and reads 4 source rows of 8 voxels (bricked) or one row of 32 (linear).
template <int LAYOUT>
__global__ void __launch_bounds__(256) k_repack(const RepackArgs a) {
  const uint64_t n = a.o1 - a.o0;
  const int pw = a.lw, ph = a.lh, pd = a.ld;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t o = a.o0 + i;
    int X, Y, Z;
    if (LAYOUT == LAYOUT_BRICK) {
      const uint64_t bz = o / a.sz;
      const uint32_t rz = (uint32_t)(o - bz * a.sz);
      const uint32_t by = rz / a.sy, ry = rz - by * a.sy;
      const uint32_t bx = ry >> 9, in = ry & 511;
      X = (int)(bx << 3) + (int)(in & 7);
      Y = (int)(by << 3) + (int)((in >> 3) & 7);
      Z = (int)(bz << 3) + (int)(in >> 6);
    } else {
      const uint64_t z = o / a.sz;
      const uint32_t rz = (uint32_t)(o - z * a.sz);
      Y = (int)(rz / a.sy);
      X = (int)(rz - (uint32_t)Y * a.sy);
      Z = (int)z;
    }
    float v = 0.0f;
    if (X < pw && Y < ph && Z < pd) {
      const int sx = min(max(X + a.gx0, 0), a.w - 1);
      const int sy = min(max(Y + a.gy0, 0), a.h - 1);
      const int sz = min(max(Z + a.gz0, 0), a.d - 1);
      if (sx < a.box[0] || sx >= a.box[3] || sy < a.box[1] || sy >= a.box[4] || sz < a.box[2] || sz >= a.box[5]) continue;  // not part of this update
      if (a.src_surface) v = surf3Dread<float>((cudaSurfaceObject_t)a.src_surface, (sx - a.sx0) * (int)sizeof(float), sy - a.sy0, sz - a.src_z0);
      else v = load_src(a.src, ((size_t)(sz - a.src_z0) * a.sh + (sy - a.sy0)) * a.sw + (sx - a.sx0), a.src_dtype);
    } else if (a.box[0] > 0 || a.box[1] > 0 || a.box[2] > 0 || a.box[3] < a.w || a.box[4] < a.h || a.box[5] < a.d) {
      continue;  // brick filler beyond the padded extent: written once by the full upload
    }
    store_dst(a.dst, o * (uint64_t)a.estride + (uint64_t)a.ecomp, a.dst_dtype, v);
  }
}

cudaError_t launch_repack(const RepackArgs &a, cudaStream_t stream) {
  const uint64_t n = a.o1 - a.o0;
  if (n == 0) return cudaSuccess;
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  if (a.layout == LAYOUT_BRICK)
    k_repack<LAYOUT_BRICK><<<(unsigned)blocks, 256, 0, stream>>>(a);
  else
    k_repack<LAYOUT_LINEAR><<<(unsigned)blocks, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---- texture-unit mode: padded storage -> cudaArray ---------------------------------------------
// One thread per voxel, x fastest: a warp reads one or two brick rows and writes 128 contiguous bytes of the array's
// row (the array's own tiling is the driver's business).  HBM bound: 4 bytes read + 4 written per voxel.
// The array holds the voxel box [a0, a0 + an) of the volume (the whole volume, or a region context's resident box clipped to
// the volume); bf16 storage is widened exactly.
struct Box6 {
  int a0[3], an[3];
};
template <int LAYOUT, int DTYPE>
__global__ void __launch_bounds__(256) k_fill_array(const TexDesc t, cudaSurfaceObject_t surf, const Box6 b) {
  const uint64_t n = (uint64_t)b.an[0] * b.an[1] * b.an[2];
  const int es = t.estride > 1 ? t.estride : 1;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % (uint64_t)b.an[0]), y = (int)((i / (uint64_t)b.an[0]) % (uint64_t)b.an[1]), z = (int)(i / ((uint64_t)b.an[0] * b.an[1]));
    typedef Addr<LAYOUT> A;
    // voxel v sits at padded index v + 1 (minus the local storage's origin)
    const size_t off = A::z(z + b.a0[2] + 1 - t.oz, t.sz) + A::y(y + b.a0[1] + 1 - t.oy, t.sy) + A::x(x + b.a0[0] + 1 - t.ox);
    const float v = load_src(t.p, off * (size_t)es + (size_t)t.ecomp, DTYPE);
    surf3Dwrite(v, surf, x * (int)sizeof(float), y, z);
  }
}
cudaError_t launch_fill_array(const TexDesc &t, int layout, int dtype, const int a0[3], const int an[3], unsigned long long surface, cudaStream_t stream) {
  Box6 b;
  for (int i = 0; i < 3; ++i) b.a0[i] = a0[i], b.an[i] = an[i];
  const uint64_t n = (uint64_t)an[0] * an[1] * an[2];
  if (n == 0) return cudaSuccess;
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  const cudaSurfaceObject_t s_ = (cudaSurfaceObject_t)surface;
  if (layout == LAYOUT_BRICK) {
    if (dtype == 0) k_fill_array<LAYOUT_BRICK, 0><<<(unsigned)blocks, 256, 0, stream>>>(t, s_, b);
    else k_fill_array<LAYOUT_BRICK, 1><<<(unsigned)blocks, 256, 0, stream>>>(t, s_, b);
  } else {
    if (dtype == 0) k_fill_array<LAYOUT_LINEAR, 0><<<(unsigned)blocks, 256, 0, stream>>>(t, s_, b);
    else k_fill_array<LAYOUT_LINEAR, 1><<<(unsigned)blocks, 256, 0, stream>>>(t, s_, b);
  }
  return cudaGetLastError();
}

// ---- occupancy ------------------------------------------------------------------------------
// One warp per macro cell: scans padded voxels [8c+1-dilate, 8c+8+dilate] per axis (voxel index
// v sits at padded index v+1; the replicated border only repeats edge voxels, so reading it is
// harmless) and sets the cell's bit when any value is not exactly +-0.
struct Reg6 {
  int v[6];
};
template <int LAYOUT, typename E>
__global__ void __launch_bounds__(256) k_occupancy(const TexDesc t, uint32_t *occ, int ocx, int ocy, int ocz, int dilate, const Reg6 reg) {
  const int lane = threadIdx.x & 31;
  const uint64_t ncell = (uint64_t)ocx * ocy * ocz;
  const int span = 8 + 2 * dilate;
  const int pw = t.lw, ph = t.lh, pd = t.ld;
  for (uint64_t cell = (uint64_t)blockIdx.x * 8 + (threadIdx.x >> 5); cell < ncell; cell += (uint64_t)gridDim.x * 8) {
    const int cx = (int)(cell % ocx), cy = (int)((cell / ocx) % ocy), cz = (int)(cell / ((uint64_t)ocx * ocy));
    // an owned sample has its low tap in [lo-1, hi-1]: a cell outside that range stays empty
    if (cx * 8 + 7 < reg.v[0] - 1 || cx * 8 > reg.v[3] - 1 || cy * 8 + 7 < reg.v[1] - 1 || cy * 8 > reg.v[4] - 1 || cz * 8 + 7 < reg.v[2] - 1 ||
        cz * 8 > reg.v[5] - 1)
      continue;
    // global padded index of the first scanned voxel, moved into the local storage
    const int x0 = cx * 8 + 1 - dilate - t.ox, y0 = cy * 8 + 1 - dilate - t.oy, z0 = cz * 8 + 1 - dilate - t.oz;
    bool any = false;
    const int total = span * span * span;
    for (int i = lane; i < total && !any; i += 32) {
      const int dx = i % span, dy = (i / span) % span, dz = i / (span * span);
      const int X = min(max(x0 + dx, 0), pw - 1), Y = min(max(y0 + dy, 0), ph - 1), Z = min(max(z0 + dz, 0), pd - 1);
      typedef Addr<LAYOUT> A;
      const float v = E::load(static_cast<const typename E::T *>(t.p) + A::z(Z, t.sz), A::x(X) + A::y(Y, t.sy));
      any = !(v == 0.0f);  // NaN counts as occupied
    }
    if (__any_sync(0xffffffffu, any) && lane == 0) atomicOr(occ + (cell >> 5), 1u << (cell & 31));
  }
}

// Bricked storage, two streaming passes instead of one gather per cell:
//  A. one warp per 8^3 storage brick (2 KiB fp32, read once, coalesced): a 64-bit mask with one bit
//     per 2^3 sub-cell of the brick, set when any of its voxels is not +-0;
//  B. one thread per macro cell: OR of the sub-cell bits that cover the cell dilated by `dilate`
//     voxels (rounded outwards to sub-cells, i.e. at most one voxel more conservative).
template <typename E>
__global__ void __launch_bounds__(256) k_subcell_masks(const void *storage, uint64_t nbricks, unsigned long long *masks) {
  const int lane = threadIdx.x & 31;
  for (uint64_t b = (uint64_t)blockIdx.x * 8 + (threadIdx.x >> 5); b < nbricks; b += (uint64_t)gridDim.x * 8) {
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint32_t e = (uint32_t)lane + 32u * j;  // element of the brick: i = e&7, j = (e>>3)&7, k = e>>6
      const float v = E::load(static_cast<const typename E::T *>(storage) + b * 512, e);
      if (!(v == 0.0f)) {
        const uint32_t bit = ((e & 7) >> 1) + 4u * (((e >> 3) & 7) >> 1) + 16u * (e >> 7);
        if (bit < 32) lo |= 1u << bit; else hi |= 1u << (bit - 32);
      }
    }
    lo = __reduce_or_sync(0xffffffffu, lo), hi = __reduce_or_sync(0xffffffffu, hi);
    if (lane == 0) masks[b] = ((unsigned long long)hi << 32) | lo;
  }
}
__global__ void __launch_bounds__(256) k_occupancy_from_masks(const TexDesc t, const unsigned long long *masks, uint32_t *occ, int ocx, int ocy,
                                                             int ocz, int dilate, const Reg6 reg) {
  const uint64_t ncell = (uint64_t)ocx * ocy * ocz;
  const uint32_t nbx = t.sy >> 9, nby = t.sz / t.sy;  // bricks per row / per layer (TexDesc strides of the bricked layout)
  for (uint64_t cell = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; cell < ncell; cell += (uint64_t)gridDim.x * blockDim.x) {
    const int c[3] = {(int)(cell % ocx), (int)((cell / ocx) % ocy), (int)(cell / ((uint64_t)ocx * ocy))};
    if (c[0] * 8 + 7 < reg.v[0] - 1 || c[0] * 8 > reg.v[3] - 1 || c[1] * 8 + 7 < reg.v[1] - 1 || c[1] * 8 > reg.v[4] - 1 || c[2] * 8 + 7 < reg.v[2] - 1 ||
        c[2] * 8 > reg.v[5] - 1)
      continue;
    // sub-cell range of the dilated cell in local storage coordinates (voxel v sits at padded index v + 1 - o)
    const int off[3] = {t.ox, t.oy, t.oz}, ldim[3] = {t.lw, t.lh, t.ld};
    int s0[3], s1[3];
    for (int a = 0; a < 3; ++a) {
      const int x0 = min(max(c[a] * 8 + 1 - dilate - off[a], 0), ldim[a] - 1), x1 = min(max(c[a] * 8 + 8 + dilate - off[a], 0), ldim[a] - 1);
      s0[a] = x0 >> 1, s1[a] = x1 >> 1;
    }
    bool any = false;
    for (int bz = s0[2] >> 2; bz <= s1[2] >> 2 && !any; ++bz)
      for (int by = s0[1] >> 2; by <= s1[1] >> 2 && !any; ++by)
        for (int bx = s0[0] >> 2; bx <= s1[0] >> 2 && !any; ++bx) {
          // sub-cells of this brick inside the range, per axis, as 4-bit masks
          uint32_t m[3];
          const int bb[3] = {bx, by, bz};
          for (int a = 0; a < 3; ++a) {
            const int lo = max(s0[a] - 4 * bb[a], 0), hi = min(s1[a] - 4 * bb[a], 3);
            m[a] = ((1u << (hi + 1)) - 1u) & ~((1u << lo) - 1u);
          }
          unsigned long long want = 0;
          for (int kz = 0; kz < 4; ++kz)
            if (m[2] >> kz & 1u)
              for (int ky = 0; ky < 4; ++ky)
                if (m[1] >> ky & 1u) want |= (unsigned long long)m[0] << (4 * ky + 16 * kz);
          any = (__ldg(masks + ((uint64_t)bz * nby + by) * nbx + bx) & want) != 0ull;
        }
    if (any) atomicOr(occ + (cell >> 5), 1u << (cell & 31));
  }
}

cudaError_t launch_occupancy(const TexDesc &t, int layout, int dtype, uint32_t *occ, int ocx, int ocy, int ocz,
                             int dilate, const int reg6[6], cudaStream_t stream) {
  Reg6 reg;
  for (int i = 0; i < 6; ++i) reg.v[i] = reg6[i];
  const uint64_t ncell = (uint64_t)ocx * ocy * ocz;
  uint64_t blocks = (ncell + 7) / 8;
  if (blocks > 148ull * 32) blocks = 148ull * 32;
  if (blocks < 1) blocks = 1;
  const unsigned g = (unsigned)blocks;
  if (layout == LAYOUT_BRICK) {
    const uint64_t nbricks = (uint64_t)((t.lw + 7) / 8) * ((t.lh + 7) / 8) * ((t.ld + 7) / 8);
    unsigned long long *masks = nullptr;
    cudaError_t e = cudaMallocAsync(&masks, nbricks * 8, stream);
    if (e != cudaSuccess) return e;
    uint64_t ba = (nbricks + 7) / 8;
    if (ba > 148ull * 64) ba = 148ull * 64;
    if (dtype == 0)
      k_subcell_masks<ElemF32><<<(unsigned)ba, 256, 0, stream>>>(t.p, nbricks, masks);
    else
      k_subcell_masks<ElemBF16><<<(unsigned)ba, 256, 0, stream>>>(t.p, nbricks, masks);
    uint64_t bb = (ncell + 255) / 256;
    if (bb > 148ull * 32) bb = 148ull * 32;
    k_occupancy_from_masks<<<(unsigned)bb, 256, 0, stream>>>(t, masks, occ, ocx, ocy, ocz, dilate, reg);
    e = cudaGetLastError();
    cudaFreeAsync(masks, stream);
    return e;
  } else {
    if (dtype == 0)
      k_occupancy<LAYOUT_LINEAR, ElemF32><<<g, 256, 0, stream>>>(t, occ, ocx, ocy, ocz, dilate, reg);
    else
      k_occupancy<LAYOUT_LINEAR, ElemBF16><<<g, 256, 0, stream>>>(t, occ, ocx, ocy, ocz, dilate, reg);
  }
  return cudaGetLastError();
}

// ---- distance grid ------------------------------------------------------------------------------
// Chebyshev distance (in macro cells, capped at 255) from every cell to the nearest non-empty
// cell: D(c) = min over occupied o of max(|cx-ox|, |cy-oy|, |cz-oz|).  The max-norm separates
// exactly into three 1-D passes: Dx = distance along the row, Dxy(c) = min_k max(|k|, Dx(c + k*ey)),
// D(c) = min_k max(|k|, Dxy(c + k*ez)).  One thread per cell; the scan stops at k = current best.
__global__ void __launch_bounds__(256) k_dist_pass(const uint32_t *occ, const uint8_t *in, uint8_t *out, int ocx, int ocy, int ocz, int axis) {
  const uint64_t ncell = (uint64_t)ocx * ocy * ocz;
  for (uint64_t cell = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; cell < ncell; cell += (uint64_t)gridDim.x * blockDim.x) {
    const int cx = (int)(cell % ocx), cy = (int)((cell / ocx) % ocy), cz = (int)(cell / ((uint64_t)ocx * ocy));
    const int pos = axis == 0 ? cx : (axis == 1 ? cy : cz);
    const int n = axis == 0 ? ocx : (axis == 1 ? ocy : ocz);
    const int64_t stride = axis == 0 ? 1 : (axis == 1 ? ocx : (int64_t)ocx * ocy);
    auto f = [&](int64_t c) -> int {
      if (axis == 0) return ((__ldg(occ + (c >> 5)) >> (c & 31)) & 1u) ? 0 : 255;
      return (int)__ldg(in + c);
    };
    int best = f((int64_t)cell);
    for (int k = 1; k < best; ++k) {
      if (pos - k >= 0) best = min(best, max(k, f((int64_t)cell - k * stride)));
      if (pos + k < n) best = min(best, max(k, f((int64_t)cell + k * stride)));
    }
    out[cell] = (uint8_t)best;
  }
}
cudaError_t launch_distance(const uint32_t *occ, uint8_t *dist, uint8_t *tmp, int ocx, int ocy, int ocz, cudaStream_t stream) {
  const uint64_t ncell = (uint64_t)ocx * ocy * ocz;
  uint64_t blocks = (ncell + 255) / 256;
  if (blocks > 148ull * 32) blocks = 148ull * 32;
  if (blocks < 1) blocks = 1;
  const unsigned g = (unsigned)blocks;
  k_dist_pass<<<g, 256, 0, stream>>>(occ, nullptr, dist, ocx, ocy, ocz, 0);
  k_dist_pass<<<g, 256, 0, stream>>>(occ, dist, tmp, ocx, ocy, ocz, 1);
  k_dist_pass<<<g, 256, 0, stream>>>(occ, tmp, dist, ocx, ocy, ocz, 2);
  return cudaGetLastError();
}

// ---- Henyey-Greenstein LUT --------------------------------------------------------------------
// HenyeyGreenstein.cc:47-91: index k = c*n*n + a*n + b with gamma = c*pi/n, alpha = a*pi/n,
// beta = b*pi/n; cos(theta) = dot(Rx(gamma) * (sin a, 0, cos a), (sin b, 0, cos b))
//                           = sin a * sin b + cos g * cos a * cos b          (float3.h:74-104)
// value = 1/(4 pi) * (1 - g^2) / sqrt((1 + g^2 - 2 g cos(theta))^3).  fp32 like the reference;
// the explicit _rn intrinsics keep the reference's operation order (no FMA contraction).
__global__ void k_hg_lut(float *out, uint32_t n, float g) {
  const float PI = 3.14159265358979323846f;
  const uint64_t total = (uint64_t)n * n * n;
  const float frac = PI / (float)n;
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t b = (uint32_t)(k % n), a = (uint32_t)((k / n) % n), c = (uint32_t)(k / ((uint64_t)n * n));
    const float gamma = __fmul_rn((float)c, frac), alpha = __fmul_rn((float)a, frac), beta = __fmul_rn((float)b, frac);
    const float sa = sinf(alpha), ca = cosf(alpha), sb = sinf(beta), cb = cosf(beta), cg = cosf(gamma);
    // rotated lightOut = (sa, sin(g)*ca, cg*ca); the y component meets lightIn.y = 0
    const float rz = __fmul_rn(cg, ca);
    const float cosTheta = __fadd_rn(__fadd_rn(__fmul_rn(sa, sb), 0.0f), __fmul_rn(rz, cb));
    const float g2 = __fmul_rn(g, g);
    const float num = __fsub_rn(1.0f, g2);
    const float base = __fsub_rn(__fadd_rn(1.0f, g2), __fmul_rn(__fmul_rn(2.0f, g), cosTheta));
    const float den = sqrtf(__fmul_rn(__fmul_rn(base, base), base));
    out[k] = __fmul_rn(__fdiv_rn(1.0f, __fmul_rn(4.0f, PI)), __fdiv_rn(num, den));
  }
}
cudaError_t launch_hg_lut(float *d_out, uint32_t n, float g, cudaStream_t stream) {
  const uint64_t total = (uint64_t)n * n * n;
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  k_hg_lut<<<(unsigned)blocks, 256, 0, stream>>>(d_out, n, g);
  return cudaGetLastError();
}

// ---- the same LUT without any allocation (vr_henyey_greenstein_static) ----------------------------
// Scratch for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it): the n^3 device buffer of vr_henyey_greenstein lives in this file-scope array instead, 64^3 floats = 1 MiB, held by
// every GPU that loads the library until the process exits.  In HOST code the name `g_hg_scratch` is only a stand-in for the
// symbol, not a device address: a kernel handed it would fault (or, where the GPU can reach host memory, silently use the
// wrong memory).  The address comes from cudaGetSymbolAddress, once per device with that device current.
// Every call shares the array and it keeps the last call's contents; k_hg_lut writes every one of the n^3 elements the call
// then reads, so nothing has to be cleared.  Calls that overlap on the GPU (two streams, two host threads) race on it.
static __device__ __align__(256) float g_hg_scratch[HG_STATIC_MAX_N * HG_STATIC_MAX_N * HG_STATIC_MAX_N];

namespace {
enum { kMaxDevices = 64 };
std::mutex g_hg_scratch_mutex;
float *g_hg_scratch_addr[kMaxDevices];  // per device: the array's DEVICE address, resolved on first use
}  // namespace

cudaError_t launch_hg_lut_static(uint32_t n, float g, cudaStream_t stream, float **d_lut) {
  if (!d_lut || n == 0 || n > HG_STATIC_MAX_N) return cudaErrorInvalidValue;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  float *ptr = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_hg_scratch_mutex);
    if (!g_hg_scratch_addr[dev]) {  // set-up, once per device, with that device current
      void *p = nullptr;
      e = cudaGetSymbolAddress(&p, g_hg_scratch);
      if (e != cudaSuccess) return e;
      g_hg_scratch_addr[dev] = static_cast<float *>(p);
    }
    ptr = g_hg_scratch_addr[dev];
  }
  e = launch_hg_lut(ptr, n, g, stream);  // the kernel gets the device address, never the symbol's host-side name
  if (e != cudaSuccess) return e;
  *d_lut = ptr;
  return cudaSuccess;
}

// ---- synthetic volumes ------------------------------------------------------------------------
// Counter-based generator: every voxel is a pure function of (x, y, z, dims, seed), written with
// explicit round-to-nearest intrinsics and no transcendental functions so that the numpy twin in
// tests/synth.py reproduces it bit for bit.
__host__ __device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t hash3(uint32_t i, uint32_t j, uint32_t k, uint32_t seed) {
  return hash_u32(i * 0x9e3779b1u ^ hash_u32(j * 0x85ebca77u ^ hash_u32(k * 0xc2b2ae3du ^ seed)));
}
__device__ __forceinline__ float u01(uint32_t h) { return __fmul_rn((float)(h >> 8), 1.0f / 16777216.0f); }
__device__ __forceinline__ float lerp_rn(float a, float b, float f) { return __fadd_rn(a, __fmul_rn(f, __fsub_rn(b, a))); }
__device__ __forceinline__ float smooth_rn(float f) {  // f*f*(3-2f)
  return __fmul_rn(__fmul_rn(f, f), __fsub_rn(3.0f, __fmul_rn(2.0f, f)));
}
__device__ float value_noise(float px, float py, float pz, uint32_t seed) {
  const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
  const uint32_t i = (uint32_t)(int)fx, j = (uint32_t)(int)fy, k = (uint32_t)(int)fz;
  const float wx = smooth_rn(__fsub_rn(px, fx)), wy = smooth_rn(__fsub_rn(py, fy)), wz = smooth_rn(__fsub_rn(pz, fz));
  const float c000 = u01(hash3(i, j, k, seed)), c100 = u01(hash3(i + 1, j, k, seed));
  const float c010 = u01(hash3(i, j + 1, k, seed)), c110 = u01(hash3(i + 1, j + 1, k, seed));
  const float c001 = u01(hash3(i, j, k + 1, seed)), c101 = u01(hash3(i + 1, j, k + 1, seed));
  const float c011 = u01(hash3(i, j + 1, k + 1, seed)), c111 = u01(hash3(i + 1, j + 1, k + 1, seed));
  const float c00 = lerp_rn(c000, c100, wx), c10 = lerp_rn(c010, c110, wx);
  const float c01 = lerp_rn(c001, c101, wx), c11 = lerp_rn(c011, c111, wx);
  return lerp_rn(lerp_rn(c00, c10, wy), lerp_rn(c01, c11, wy), wz);
}
// voxel-centre coordinate in [-1,1]: ((i + 0.5) / n) * 2 - 1
__device__ __forceinline__ float ncoord(int i, int n) {
  return __fsub_rn(__fmul_rn(__fdiv_rn(__fadd_rn((float)i, 0.5f), (float)n), 2.0f), 1.0f);
}

// kind 0: "brain" - soft-edged ellipsoid (semi-axes 0.8, 0.55, 0.4 of the box) times
// (0.6 + 0.4 * 3-octave value noise), clamped to [0,1], exactly 0 outside the ellipsoid.
__device__ float synth_brain(int x, int y, int z, int w, int h, int d, uint32_t seed) {
  const float u = ncoord(x, w), v = ncoord(y, h), s = ncoord(z, d);
  const float a = __fdiv_rn(u, 0.8f), b = __fdiv_rn(v, 0.55f), c = __fdiv_rn(s, 0.4f);
  const float r2 = __fadd_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)), __fmul_rn(c, c));
  if (!(r2 < 1.0f)) return 0.0f;
  const float m = smooth_rn(fminf(__fmul_rn(__fsub_rn(1.0f, r2), 4.0f), 1.0f));
  const float qx = __fmul_rn(__fadd_rn(u, 1.0f), 0.5f), qy = __fmul_rn(__fadd_rn(v, 1.0f), 0.5f), qz = __fmul_rn(__fadd_rn(s, 1.0f), 0.5f);
  const float n0 = value_noise(__fmul_rn(qx, 8.0f), __fmul_rn(qy, 8.0f), __fmul_rn(qz, 8.0f), seed);
  const float n1 = value_noise(__fmul_rn(qx, 16.0f), __fmul_rn(qy, 16.0f), __fmul_rn(qz, 16.0f), seed + 1u);
  const float n2 = value_noise(__fmul_rn(qx, 32.0f), __fmul_rn(qy, 32.0f), __fmul_rn(qz, 32.0f), seed + 2u);
  const float nz = __fadd_rn(__fadd_rn(__fmul_rn(0.5f, n0), __fmul_rn(0.3f, n1)), __fmul_rn(0.2f, n2));
  const float val = __fmul_rn(m, __fadd_rn(0.6f, __fmul_rn(0.4f, nz)));
  return fminf(fmaxf(val, 0.0f), 1.0f);
}

// kind 1: "tubes" - union of 200 seeded capsules in [-1,1]^3, radius 3..6 voxels (of dim w),
// soft edge of 1.5 voxels; exactly 0 away from every capsule.
enum { kTubes = 200 };
__device__ float synth_tubes(int x, int y, int z, int w, int h, int d, uint32_t seed) {
  const float px = ncoord(x, w), py = ncoord(y, h), pz = ncoord(z, d);
  const float vox = __fdiv_rn(2.0f, (float)w);
  float best = 0.0f;
  for (uint32_t t = 0; t < kTubes; ++t) {
    const uint32_t s = seed ^ (t * 0x632be5abu);
    const float ax = __fsub_rn(__fmul_rn(u01(hash_u32(s + 1u)), 1.6f), 0.8f);
    const float ay = __fsub_rn(__fmul_rn(u01(hash_u32(s + 2u)), 1.6f), 0.8f);
    const float az = __fsub_rn(__fmul_rn(u01(hash_u32(s + 3u)), 1.6f), 0.8f);
    const float dx = __fsub_rn(__fmul_rn(u01(hash_u32(s + 4u)), 0.8f), 0.4f);
    const float dy = __fsub_rn(__fmul_rn(u01(hash_u32(s + 5u)), 0.8f), 0.4f);
    const float dz = __fsub_rn(__fmul_rn(u01(hash_u32(s + 6u)), 0.8f), 0.4f);
    const float rad = __fmul_rn(__fadd_rn(3.0f, __fmul_rn(u01(hash_u32(s + 7u)), 3.0f)), vox);
    const float ex = __fsub_rn(px, ax), ey = __fsub_rn(py, ay), ez = __fsub_rn(pz, az);
    const float dd = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    const float ed = __fadd_rn(__fadd_rn(__fmul_rn(ex, dx), __fmul_rn(ey, dy)), __fmul_rn(ez, dz));
    const float tt = fminf(fmaxf(__fdiv_rn(ed, fmaxf(dd, 1e-12f)), 0.0f), 1.0f);
    const float cx = __fsub_rn(ex, __fmul_rn(tt, dx)), cy = __fsub_rn(ey, __fmul_rn(tt, dy)), cz = __fsub_rn(ez, __fmul_rn(tt, dz));
    const float dist = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz)));
    const float edge = __fmul_rn(1.5f, vox);
    const float vv = fminf(fmaxf(__fdiv_rn(__fsub_rn(rad, dist), edge), 0.0f), 1.0f);
    best = fmaxf(best, vv);
  }
  return best;
}

__global__ void __launch_bounds__(256) k_synth(void *out, int kind, uint32_t seed, int w, int h, int d, int dtype) {
  const uint64_t total = (uint64_t)w * h * d;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % w), y = (int)((i / w) % h), z = (int)(i / ((uint64_t)w * h));
    const float v = kind == 0 ? synth_brain(x, y, z, w, h, d, seed) : synth_tubes(x, y, z, w, h, d, seed);
    store_dst(out, i, dtype, v);
  }
}
cudaError_t launch_synth(void *d_out, int kind, uint32_t seed, int w, int h, int d, int dtype, cudaStream_t stream) {
  const uint64_t total = (uint64_t)w * h * d;
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  if (blocks < 1) blocks = 1;
  k_synth<<<(unsigned)blocks, 256, 0, stream>>>(d_out, kind, seed, w, h, d, dtype);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_synth_region(void *out, int kind, uint32_t seed, int w, int h, int d, int x0, int y0, int z0, int bw, int bh,
                                                      int bd, int dtype) {
  const uint64_t total = (uint64_t)bw * bh * bd;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const int x = x0 + (int)(i % bw), y = y0 + (int)((i / bw) % bh), z = z0 + (int)(i / ((uint64_t)bw * bh));
    const float v = kind == 0 ? synth_brain(x, y, z, w, h, d, seed) : synth_tubes(x, y, z, w, h, d, seed);
    store_dst(out, i, dtype, v);
  }
}
cudaError_t launch_synth_region(void *d_out, int kind, uint32_t seed, const int dims[3], const int lo[3], const int hi[3], int dtype,
                                cudaStream_t stream) {
  const uint64_t total = (uint64_t)(hi[0] - lo[0]) * (hi[1] - lo[1]) * (hi[2] - lo[2]);
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  if (blocks < 1) blocks = 1;
  k_synth_region<<<(unsigned)blocks, 256, 0, stream>>>(d_out, kind, seed, dims[0], dims[1], dims[2], lo[0], lo[1], lo[2], hi[0] - lo[0], hi[1] - lo[1],
                                                       hi[2] - lo[2], dtype);
  return cudaGetLastError();
}

// ---- sort-last compositing ------------------------------------------------------------------------
struct PartialPtrs {
  const float *p[64];
};
__global__ void __launch_bounds__(256) k_compose(const PartialPtrs in, int n, uint64_t n4, uint64_t n_floats, float *out) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (uint64_t)gridDim.x * blockDim.x) {
    float4 acc = reinterpret_cast<const float4 *>(in.p[0])[i];  // 128-bit loads; peers' partials come over NVLink
    for (int b = 1; b < n; ++b) {
      const float4 v = reinterpret_cast<const float4 *>(in.p[b])[i];
      acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
    }
    reinterpret_cast<float4 *>(out)[i] = acc;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (uint64_t i = n4 * 4; i < n_floats; ++i) {
      float a = in.p[0][i];
      for (int b = 1; b < n; ++b) a += in.p[b][i];
      out[i] = a;
    }
}
cudaError_t launch_compose(const float *const *d_partials, int n, uint64_t n_floats, float *d_out, cudaStream_t stream) {
  PartialPtrs pp;
  for (int i = 0; i < 64; ++i) pp.p[i] = i < n ? d_partials[i] : nullptr;
  const uint64_t n4 = n_floats / 4;
  uint64_t blocks = (n4 + 255) / 256;
  if (blocks > 148ull * 8) blocks = 148ull * 8;
  if (blocks < 1) blocks = 1;
  k_compose<<<(unsigned)blocks, 256, 0, stream>>>(pp, n, n4, n_floats, d_out);
  return cudaGetLastError();
}

// Depth slices of one brick (VR_OPT_VS_SLICES): S planes of n floats each -> one.  product = 1: opacities, 1 - prod(1 - a_s),
// front to back (what a march through the whole brick accumulates: sum.w += (1 - sum.w) * alpha, kernel.cu:470-483);
// product = 0: the additive colour shares, summed in slice order.
__global__ void __launch_bounds__(256) k_fold_slices(const float *in, int S, uint64_t n, float *out, int product, const FoldOwner own) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    if (own.nranks > 1) {  // an image-space partition: only the pixels of this context's tiles hold anything (k = x * H + y)
      const uint64_t k = i % own.plane;
      const uint32_t x = (uint32_t)(k / own.H), y = (uint32_t)(k - (uint64_t)x * own.H);
      const uint32_t tx = x / own.tile_w, ty = y / own.tile_h;
      const uint32_t key = own.columns ? tx : ty * own.ntx + tx;
      if (key % own.nranks != own.rank) continue;
    }
    if (product) {
      float T = 1.0f;
      for (int s = 0; s < S; ++s) T = T * (1.0f - __ldg(in + (uint64_t)s * n + i));
      out[i] = 1.0f - T;
    } else {
      float a = __ldg(in + i);
      for (int s = 1; s < S; ++s) a += __ldg(in + (uint64_t)s * n + i);
      out[i] = a;
    }
  }
}
cudaError_t launch_fold_slices(const float *d_slices, int S, uint64_t n, float *d_out, int product, const FoldOwner &own, cudaStream_t stream) {
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148ull * 8) blocks = 148ull * 8;
  if (blocks < 1) blocks = 1;
  k_fold_slices<<<(unsigned)blocks, 256, 0, stream>>>(d_slices, S, n, d_out, product, own);
  return cudaGetLastError();
}

// ---- image epilogue (what the reference does in MATLAB after the mex returns) ---------------------------
// Stereo compose (VolumeRender.m:286-306): both eyes were rendered delta columns wider; the left image
// keeps columns delta.., the right one columns ..Wp-delta; RedCyan takes R from the left eye and G,B from
// the right, LeftRightHorizontal puts the crops side by side.  Images are H x W x 3 column-major planar.
__global__ void __launch_bounds__(256) k_stereo(const float *left, const float *right, int H, int Wp, int delta, int mode, float *out) {
  const int Wc = Wp - delta, Wo = mode == 0 ? Wc : 2 * Wc;
  const size_t plane_in = (size_t)H * Wp, plane_out = (size_t)H * Wo, total = plane_out * 3;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int ch = (int)(i / plane_out);
    const size_t k = i - (size_t)ch * plane_out;
    const int x = (int)(k / H), y = (int)(k - (size_t)x * H);
    float v;
    if (mode == 0) {
      v = ch == 0 ? left[(size_t)ch * plane_in + (size_t)(x + delta) * H + y] : right[(size_t)ch * plane_in + (size_t)x * H + y];
    } else {
      v = x < Wc ? left[(size_t)ch * plane_in + (size_t)(x + delta) * H + y] : right[(size_t)ch * plane_in + (size_t)(x - Wc) * H + y];
    }
    out[i] = v;
  }
}
cudaError_t launch_stereo(const float *left, const float *right, int H, int Wp, int delta, int mode, float *out, cudaStream_t stream) {
  const size_t total = (size_t)H * (mode == 0 ? Wp - delta : 2 * (Wp - delta)) * 3;
  size_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  k_stereo<<<(unsigned)blocks, 256, 0, stream>>>(left, right, H, Wp, delta, mode, out);
  return cudaGetLastError();
}

// normalizeSequence / normalizeImage (VolumeRender.m:587-701): global min and max over all frames, then
// (like the reference, including its `+ min` shift for negative minima) every value divided by the max.
__device__ __forceinline__ unsigned ord(float f) {  // order-preserving float -> uint
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unord(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }
__global__ void __launch_bounds__(256) k_minmax(const float *img, size_t n, unsigned *mm) {
  unsigned lo = 0xffffffffu, hi = 0u;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned o = ord(img[i]);
    lo = min(lo, o), hi = max(hi, o);
  }
  lo = __reduce_min_sync(0xffffffffu, lo), hi = __reduce_max_sync(0xffffffffu, hi);
  if ((threadIdx.x & 31) == 0) atomicMin(mm, lo), atomicMax(mm + 1, hi);
}
__global__ void __launch_bounds__(256) k_normalize(float *img, size_t n, const unsigned *mm) {
  float mn = unord(mm[0]), mx = unord(mm[1]);
  const float shift = mn < 0.0f ? mn : 0.0f;  // VolumeRender.m:676-683: r = r + minValue, max = max + abs(min)
  if (mn < 0.0f) mx = mx + fabsf(mn);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) img[i] = (img[i] + shift) / mx;
}
cudaError_t launch_normalize(float *const *imgs, int n_imgs, size_t n_each, unsigned *d_mm, cudaStream_t stream) {
  const unsigned init[2] = {0xffffffffu, 0u};
  cudaError_t e = cudaMemcpyAsync(d_mm, init, sizeof init, cudaMemcpyHostToDevice, stream);
  if (e != cudaSuccess) return e;
  size_t blocks = (n_each + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  for (int i = 0; i < n_imgs; ++i) k_minmax<<<(unsigned)blocks, 256, 0, stream>>>(imgs[i], n_each, d_mm);
  for (int i = 0; i < n_imgs; ++i) k_normalize<<<(unsigned)blocks, 256, 0, stream>>>(imgs[i], n_each, d_mm);
  return cudaGetLastError();
}

// ---- MATLAB gradient() ---------------------------------------------------------------------------
// Volume.grad (src/matlab/VolumeRender/Volume.m:181-205) calls MATLAB's gradient(Data): unit
// spacing, central differences inside, one-sided at the ends, zero along a singleton axis; its
// first output is the derivative along dim 2, the second along dim 1, the third along dim 3.
__device__ __forceinline__ float grad1(const void *p, int dtype, size_t k, int64_t i, int64_t n, int64_t stride) {
  if (n < 2) return 0.0f;
  if (i == 0) return __fsub_rn(load_src(p, k + stride, dtype), load_src(p, k, dtype));
  if (i == n - 1) return __fsub_rn(load_src(p, k, dtype), load_src(p, k - stride, dtype));
  return __fdiv_rn(__fsub_rn(load_src(p, k + stride, dtype), load_src(p, k - stride, dtype)), 2.0f);
}
__global__ void __launch_bounds__(256) k_gradient(const void *in, int w, int h, int d, int dtype, void *gx, void *gy, void *gz) {
  const uint64_t total = (uint64_t)w * h * d;
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (uint64_t)gridDim.x * blockDim.x) {
    const int x = (int)(k % w), y = (int)((k / w) % h), z = (int)(k / ((uint64_t)w * h));
    store_dst(gy, k, dtype, grad1(in, dtype, k, x, w, 1));
    store_dst(gx, k, dtype, grad1(in, dtype, k, y, h, w));
    store_dst(gz, k, dtype, grad1(in, dtype, k, z, d, (int64_t)w * h));
  }
}
cudaError_t launch_gradient(const void *d_in, int w, int h, int d, int dtype, void *gx, void *gy, void *gz, cudaStream_t stream) {
  const uint64_t total = (uint64_t)w * h * d;
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  if (blocks < 1) blocks = 1;
  k_gradient<<<(unsigned)blocks, 256, 0, stream>>>(d_in, w, h, d, dtype, gx, gy, gz);
  return cudaGetLastError();
}

}  // namespace vrb
