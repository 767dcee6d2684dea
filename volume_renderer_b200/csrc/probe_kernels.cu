// probe_kernels.cu - the gather ceilings of the box, measured ("BW_gather = the box's measured
// L2->SM bandwidth from a committed microbenchmark, since MEASURED_PEAKS.json records only HBM").
//
// Four probes with the access pattern of the ray march and NOTHING else: a warp is an 8x4 bundle of
// parallel rays 0.7 voxels apart (neighbouring pixels of a 2048^2 image over a 1024^3 volume), every
// lane walks its ray in steps of 0.45 voxels (the reference's tstep, volumeRender.cpp:133-145) and
// fetches the eight trilinear taps of each position:
//   kind 0  eight 4-byte loads per fetch (one scalar volume)               -> k_probe_ldg<1>
//   kind 1  eight 16-byte loads per fetch (the interleaved gradient voxels) -> k_probe_ldg<4>
//   kind 2  one tex3D<float> with hardware linear filtering (what the reference's kernel issues,
//           volumeRender_kernel.cu:544-548)                                 -> k_probe_tex<float>
//   kind 3  one tex3D<float4>                                               -> k_probe_tex<float4>
//   kind 4  north_star's sketch, measured: the eight taps of every fetch read from SHARED MEMORY, where a 16 x 8 x 8-voxel box per
//           14 steps of the bundle would be staged by TMA (cp.async.bulk.tensor.3d, mbarrier double buffering per warp).  By
//           default the staging itself is left out: the rate is the upper bound of any staged march   -> k_probe_tma
// The footprint (edge^3 voxels) chooses the level that serves the taps: 32^3 fp32 sits in L1, 256^3 in
// L2, 1024^3 needs HBM.  Kinds 0/1 spend ~15 integer/float instructions per fetch on addresses - far
// fewer than the march does - so their rate is the ceiling of the load path, not of a renderer.
#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint, libcuda is not linked)
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "raymarch_launch.h"

namespace vrb {
namespace {

struct ProbeRay {
  float x, y, z, dx, dy, dz;
};
// the lane's ray: bundles start at hashed positions so that the grid covers the whole footprint
__device__ ProbeRay probe_ray(int edge) {
  const unsigned warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  unsigned h = warp * 2654435761u;
  h ^= h >> 15, h *= 2246822519u, h ^= h >> 13;
  ProbeRay r;
  const float e = (float)edge;
  r.x = (float)(h & 1023u) * (e / 1024.0f) + 0.7f * (float)(lane & 7);
  r.y = (float)((h >> 10) & 1023u) * (e / 1024.0f) + 0.7f * (float)(lane >> 3);
  r.z = (float)((h >> 20) & 1023u) * (e / 1024.0f);
  // the direction of example1's camera through the volume, 0.45 voxels per step
  r.dx = 0.45f * 0.37f, r.dy = 0.45f * 0.21f, r.dz = 0.45f * 0.905f;
  return r;
}

template <int VEC>
struct Tap;
template <>
struct Tap<1> {
  typedef float T;
  static __device__ float sum(float v) { return v; }
};
template <>
struct Tap<4> {
  typedef float4 T;
  static __device__ float sum(float4 v) { return v.x + v.y + v.z; }
};

template <int VEC>
__global__ void __launch_bounds__(256, VEC == 4 ? 5 : 8) k_probe_ldg(const typename Tap<VEC>::T *__restrict__ vol, int log2_edge, int iters, float *sink) {
  const int edge = 1 << log2_edge;
  const unsigned m = (unsigned)edge - 1u;
  ProbeRay r = probe_ray(edge);
  float acc = 0.0f;
  for (int k = 0; k < iters; ++k) {
    const float fx = floorf(r.x), fy = floorf(r.y), fz = floorf(r.z);
    const unsigned ix = (unsigned)(int)fx, iy = (unsigned)(int)fy, iz = (unsigned)(int)fz;
    const float wx = r.x - fx, wy = r.y - fy, wz = r.z - fz;
    const unsigned x0 = ix & m, x1 = (ix + 1u) & m;
    const unsigned y0 = (iy & m) << log2_edge, y1 = ((iy + 1u) & m) << log2_edge;
    const unsigned z0 = (iz & m) << (2 * log2_edge), z1 = ((iz + 1u) & m) << (2 * log2_edge);
    const float c000 = Tap<VEC>::sum(__ldg(vol + (z0 + y0 + x0))), c100 = Tap<VEC>::sum(__ldg(vol + (z0 + y0 + x1)));
    const float c010 = Tap<VEC>::sum(__ldg(vol + (z0 + y1 + x0))), c110 = Tap<VEC>::sum(__ldg(vol + (z0 + y1 + x1)));
    const float c001 = Tap<VEC>::sum(__ldg(vol + (z1 + y0 + x0))), c101 = Tap<VEC>::sum(__ldg(vol + (z1 + y0 + x1)));
    const float c011 = Tap<VEC>::sum(__ldg(vol + (z1 + y1 + x0))), c111 = Tap<VEC>::sum(__ldg(vol + (z1 + y1 + x1)));
    const float c00 = fmaf(wx, c100 - c000, c000), c10 = fmaf(wx, c110 - c010, c010);
    const float c01 = fmaf(wx, c101 - c001, c001), c11 = fmaf(wx, c111 - c011, c011);
    const float c0 = fmaf(wy, c10 - c00, c00), c1 = fmaf(wy, c11 - c01, c01);
    acc += fmaf(wz, c1 - c0, c0);
    r.x += r.dx, r.y += r.dy, r.z += r.dz;
  }
  if (acc == 123.456f) sink[0] = acc;  // never true for the zero-filled footprint; keeps the loads alive
}

__device__ float tex_sum(float v) { return v; }
__device__ float tex_sum(float4 v) { return v.x + v.y + v.z; }
template <typename T>
__global__ void __launch_bounds__(256, 8) k_probe_tex(cudaTextureObject_t tex, int log2_edge, int iters, float *sink) {
  const int edge = 1 << log2_edge;
  ProbeRay r = probe_ray(edge);
  const float inv = 1.0f / (float)edge;
  r.x *= inv, r.y *= inv, r.z *= inv, r.dx *= inv, r.dy *= inv, r.dz *= inv;  // normalized coordinates, wrap addressing
  float acc = 0.0f;
  for (int k = 0; k < iters; ++k) {
    acc += tex_sum(tex3D<T>(tex, r.x, r.y, r.z));
    r.x += r.dx, r.y += r.dy, r.z += r.dz;
  }
  if (acc == 123.456f) sink[0] = acc;
}

// ---- kind 4: TMA-staged neighbourhood ---------------------------------------------------------------------------
// One warp = one 8x4 ray bundle, as everywhere in this file.  The bundle spans 4.9 x 2.1 voxels and drifts
// (0.167, 0.095, 0.407) voxels per step, so a box of 16 x 8 x 8 voxels (4 KiB of fp32) anchored at the floor of the bundle's
// corner holds every tap of the next 14 steps: 448 fetches per 4 KiB staged = 9 B of shared-memory fill per fetch against the
// 32 algorithmic bytes.  Two boxes per warp: lane 0 arms the mbarrier of the free box and issues the 3-D tensor copy for the
// NEXT 14 steps, all lanes wait for the current box's barrier, then read eight taps per fetch from shared memory.
enum { TBX = 16, TBY = 8, TBZ = 8, TMA_STEPS = 14, TMA_BOX_BYTES = TBX * TBY * TBZ * 4, TMA_WARPS = 8 };
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(count), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// (bounded: a copy that never completes - a bad descriptor - must end the measurement, not hang the device)
__device__ __forceinline__ bool mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spins = 0; !done && spins < (1u << 24); ++spins)
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
  return done != 0;
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, uint64_t *bar, int x, int y, int z) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
               "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z)
               : "memory");
}

// variant (debug, $VR_TMA_VARIANT): 0 = the real thing; 1 = no tensor copy and no barrier (taps from whatever the shared memory holds);
// 2 = barrier only (a plain arrive instead of the copy's transaction)
// The tensor map lives in global memory (written by the host before the launch) and is passed by pointer.
__global__ void __launch_bounds__(32 * TMA_WARPS) k_probe_tma(const CUtensorMap *__restrict__ tmap, int log2_edge, int chunks, float *sink, int variant) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // shared-memory addresses as plain integers (warp-uniform): two boxes per warp, then two barriers per warp
  const uint32_t base = smem_u32(smem_raw);
  // layout: [0, 1024) the barriers (16 bytes per warp), then the boxes
  const uint32_t bar0 = base + (uint32_t)warp * 16u;
  const uint32_t box0 = base + 1024u + (uint32_t)warp * 2u * TMA_BOX_BYTES;
  const float *boxp = reinterpret_cast<const float *>(smem_raw + 1024 + (size_t)warp * 2 * TMA_BOX_BYTES);
  const int edge = 1 << log2_edge;
  const int m = edge - 1;
  ProbeRay r = probe_ray(edge);
  // the bundle's corner = lane 0's ray (lane & 7 == 0, lane >> 3 == 0); positions of a chunk are start + k * d for every lane
  const float cx = __shfl_sync(0xffffffffu, r.x, 0), cy = __shfl_sync(0xffffffffu, r.y, 0), cz = __shfl_sync(0xffffffffu, r.z, 0);
  const float offx = r.x - cx, offy = r.y - cy;
  if (lane == 0 && variant != 1) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  float acc = 0.0f;
  for (int c = -1; c < chunks; ++c) {
    // lane 0 arms the free box's barrier and issues the tensor copy for chunk c + 1 (its readers finished at the __syncwarp below)
    if (lane == 0 && c + 1 < chunks && variant != 1) {
      const int n = c + 1;
      const float sx = cx + (float)(n * TMA_STEPS) * r.dx, sy = cy + (float)(n * TMA_STEPS) * r.dy, sz = cz + (float)(n * TMA_STEPS) * r.dz;
      // wrapped into the footprint and kept wholly inside it (the contents do not matter for the rate; the taps below index the box only)
      const int ox = min((int)floorf(sx) & m, edge - TBX), oy = min((int)floorf(sy) & m, edge - TBY), oz = min((int)floorf(sz) & m, edge - TBZ);
      const uint32_t bar = bar0 + (uint32_t)(n & 1) * 8u, dst = box0 + (uint32_t)(n & 1) * TMA_BOX_BYTES;
      if (variant == 2) {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
      } else {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)TMA_BOX_BYTES) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
                     "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(ox), "r"(oy), "r"(oz)
                     : "memory");
      }
    }
    if (c < 0) continue;
    if (variant != 1 && !mbar_wait(reinterpret_cast<uint64_t *>(smem_raw) + 2 * warp + (c & 1), (uint32_t)((c >> 1) & 1))) {
      sink[1] = -1.0f;  // the launcher reports it
      return;
    }
    const float sx = cx + (float)(c * TMA_STEPS) * r.dx, sy = cy + (float)(c * TMA_STEPS) * r.dy, sz = cz + (float)(c * TMA_STEPS) * r.dz;
    const float *b = boxp + (c & 1) * (TMA_BOX_BYTES / 4);
    // this lane's position relative to the box's origin at the chunk's first step
    float px = (sx - floorf(sx)) + offx, py = (sy - floorf(sy)) + offy, pz = sz - floorf(sz);
#pragma unroll 2
    for (int k = 0; k < TMA_STEPS; ++k) {
      const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
      const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
      const float wx = px - fx, wy = py - fy, wz = pz - fz;
      const float *q = b + (iz * TBY + iy) * TBX + ix;
      const float c000 = q[0], c100 = q[1], c010 = q[TBX], c110 = q[TBX + 1];
      const float c001 = q[TBX * TBY], c101 = q[TBX * TBY + 1], c011 = q[TBX * TBY + TBX], c111 = q[TBX * TBY + TBX + 1];
      const float c00 = fmaf(wx, c100 - c000, c000), c10 = fmaf(wx, c110 - c010, c010);
      const float c01 = fmaf(wx, c101 - c001, c001), c11 = fmaf(wx, c111 - c011, c011);
      const float c0 = fmaf(wy, c10 - c00, c00), c1 = fmaf(wy, c11 - c01, c01);
      acc += fmaf(wz, c1 - c0, c0);
      px += r.dx, py += r.dy, pz += r.dz;
    }
    __syncwarp();  // every lane is done with this box before lane 0 refills it two chunks on
  }
  if (acc == 123.456f) sink[0] = acc;
}

// debug ($VR_TMA_SELFTEST): one warp loads ONE 16 x 8 box of a 64 x 64 fp32 matrix (2-D map) and copies it out
__global__ void k_tma_selftest(const CUtensorMap *__restrict__ tmap, float *out) {
  __shared__ __align__(128) float box[TBX * TBY];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar, TBX * TBY * 4);
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(box)),
                 "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(&bar)), "r"(16), "r"(8)
                 : "memory");
  }
  const bool ok = mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < TBX * TBY; i += 32) out[i] = ok ? box[i] : -1.0f;
}

__global__ void k_tma_selftest3(const CUtensorMap *__restrict__ tmap, float *out) {
  __shared__ __align__(1024) float box[TBX * TBY * TBZ];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar, TMA_BOX_BYTES);
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(box)),
                 "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(&bar)), "r"(16), "r"(8), "r"(8)
                 : "memory");
  }
  const bool ok = mbar_wait(&bar, 0);
  if (threadIdx.x == 0) out[0] = box[0], out[1] = box[1], out[2] = ok ? 1.0f : -1.0f;
}
static CUtensorMap *d_tmap_dbg(const CUtensorMap *h) {
  CUtensorMap *d = nullptr;
  cudaMalloc(&d, sizeof(CUtensorMap));
  cudaMemcpy(d, h, sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  return d;  // (debug only: leaked)
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

// Runs one probe on `stream`; *fetches_per_s = trilinear fetches (8 taps each) per second, *ms = duration of
// the timed launch (the second of two).  Allocates and frees its own footprint.
cudaError_t launch_probe_gather(int kind, int log2_edge, int sm_count, cudaStream_t stream, double *fetches_per_s, double *ms_out) {
  if (kind < 0 || kind > 4 || log2_edge < 4 || log2_edge > 10) return cudaErrorInvalidValue;
  const size_t voxels = (size_t)1 << (3 * log2_edge);
  const size_t vb = (kind == 1 || kind == 3) ? 16 : 4;
  // one wave: 2048 resident threads per SM (1280 with 16-byte taps; 768 with two 4 KiB boxes per warp in shared memory)
  const int threads = 256, blocks = sm_count * (kind == 1 ? 5 : (kind == 4 ? 3 : 8));
  // ~2 G taps per launch whatever the kind: long enough to time, short enough for a bench preamble
  const int iters = (kind >= 2) ? 2048 : 1024;
  float *sink = nullptr;
  void *lin = nullptr;
  cudaArray_t arr = nullptr;
  cudaTextureObject_t tex = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t err = cudaMalloc(&sink, 8);
  if (err == cudaSuccess) err = cudaMemsetAsync(sink, 0, 8, stream);
  if (err == cudaSuccess) err = cudaEventCreate(&e0);
  if (err == cudaSuccess) err = cudaEventCreate(&e1);
  CUtensorMap tmap;
  CUtensorMap *d_tmap = nullptr;
  const int tma_chunks = 2048 / TMA_STEPS;
  const size_t tma_smem = 1024 + (size_t)TMA_WARPS * 2 * TMA_BOX_BYTES;
  if (err == cudaSuccess && (kind < 2 || kind == 4)) {
    err = cudaMalloc(&lin, voxels * vb);
    if (err == cudaSuccess) err = cudaMemsetAsync(lin, 0, voxels * vb, stream);
  }
  if (err == cudaSuccess && (kind == 2 || kind == 3)) {
    const cudaChannelFormatDesc cd = kind == 2 ? cudaCreateChannelDesc<float>() : cudaCreateChannelDesc<float4>();
    const size_t e = (size_t)1 << log2_edge;
    err = cudaMalloc3DArray(&arr, &cd, make_cudaExtent(e, e, e));
    if (err == cudaSuccess) {  // contents do not matter for the rate; fill from a zeroed slab so reads are defined
      void *slab = nullptr;
      err = cudaMalloc(&slab, e * e * vb);
      if (err == cudaSuccess) err = cudaMemsetAsync(slab, 0, e * e * vb, stream);
      for (size_t z = 0; z < e && err == cudaSuccess; ++z) {
        cudaMemcpy3DParms p = {};
        p.srcPtr = make_cudaPitchedPtr(slab, e * vb, e, e);
        p.dstArray = arr;
        p.dstPos = make_cudaPos(0, 0, z);
        p.extent = make_cudaExtent(e, e, 1);
        p.kind = cudaMemcpyDeviceToDevice;
        err = cudaMemcpy3DAsync(&p, stream);
      }
      if (err == cudaSuccess) err = cudaStreamSynchronize(stream);
      cudaFree(slab);
    }
    if (err == cudaSuccess) {
      cudaResourceDesc rd = {};
      rd.resType = cudaResourceTypeArray;
      rd.res.array.array = arr;
      cudaTextureDesc td = {};
      td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeWrap;
      td.filterMode = cudaFilterModeLinear;  // kernel.cu:544-548
      td.readMode = cudaReadModeElementType;
      td.normalizedCoords = 1;
      err = cudaCreateTextureObject(&tex, &rd, &td, nullptr);
    }
  }
  if (err == cudaSuccess && kind == 4) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (err == cudaSuccess && (!fn || q != cudaDriverEntryPointSuccess)) err = cudaErrorNotSupported;
    if (err == cudaSuccess) {
      const cuuint64_t e = (cuuint64_t)1 << log2_edge;
      const cuuint64_t gdim[3] = {e, e, e}, gstride[2] = {e * 4, e * e * 4};
      const cuuint32_t bx[3] = {TBX, TBY, TBZ}, es[3] = {1, 1, 1};
      const CUresult r = reinterpret_cast<EncodeTiledFn>(fn)(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, lin, gdim, gstride, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) {
        fprintf(stderr, "vr_probe_gather: cuTensorMapEncodeTiled failed with CUresult %d\n", (int)r);
        err = cudaErrorInvalidValue;
      }
      if (getenv("VR_TMA_SELFTEST")) {
        const unsigned long long *w = reinterpret_cast<const unsigned long long *>(&tmap);
        fprintf(stderr, "tensor map (3-D, CUresult %d): %016llx %016llx %016llx %016llx | lin %p\n", (int)r, w[0], w[1], w[2], w[3], lin);
        // 2-D self test on the same buffer: a 64 x 64 matrix of i + 64 * j
        float h[64 * 64], got[TBX * TBY];
        for (int i = 0; i < 64 * 64; ++i) h[i] = (float)i;
        cudaMemcpy(lin, h, sizeof h, cudaMemcpyHostToDevice);
        CUtensorMap m2;
        const cuuint64_t gd[2] = {64, 64}, gs[1] = {64 * 4};
        const cuuint32_t b2[2] = {TBX, TBY}, e2[2] = {1, 1};
        const CUresult r2 = reinterpret_cast<EncodeTiledFn>(fn)(&m2, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, lin, gd, gs, b2, e2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CUtensorMap *dm2 = nullptr;
        float *dout = nullptr;
        cudaMalloc(&dm2, sizeof m2), cudaMalloc(&dout, sizeof got);
        cudaMemcpy(dm2, &m2, sizeof m2, cudaMemcpyHostToDevice);
        k_tma_selftest<<<1, 32>>>(dm2, dout);
        const cudaError_t se = cudaDeviceSynchronize();
        cudaMemcpy(got, dout, sizeof got, cudaMemcpyDeviceToHost);
        fprintf(stderr, "2-D self test: encode %d, kernel %s, box[0..3] = %g %g %g %g (want %g %g ...), box[16] = %g (want %g)\n", (int)r2, cudaGetErrorString(se), got[0], got[1],
                got[2], got[3], (float)(8 * 64 + 16), (float)(8 * 64 + 17), got[16], (float)(9 * 64 + 16));
        // the 3-D map of the probe itself, one box at (16, 8, 8): element (x, y, z) of the buffer = x + E * (y + E * z) while it holds h
        k_tma_selftest3<<<1, 32>>>(d_tmap_dbg(&tmap), dout);
        const cudaError_t se3 = cudaDeviceSynchronize();
        cudaMemcpy(got, dout, sizeof got, cudaMemcpyDeviceToHost);
        fprintf(stderr, "3-D self test: kernel %s, box[0..1] = %g %g, status %g\n", cudaGetErrorString(se3), got[0], got[1], got[2]);
        cudaFree(dm2), cudaFree(dout);
        cudaMemset(lin, 0, sizeof h);
        if (se != cudaSuccess) err = se;
      }
    } else {
      fprintf(stderr, "vr_probe_gather: cudaGetDriverEntryPoint(cuTensorMapEncodeTiled): %s\n", cudaGetErrorString(err));
    }
    if (err == cudaSuccess) err = cudaMalloc(&d_tmap, sizeof(CUtensorMap));
    if (err == cudaSuccess) err = cudaMemcpy(d_tmap, &tmap, sizeof(CUtensorMap), cudaMemcpyHostToDevice);
    if (err == cudaSuccess) {
      err = cudaFuncSetAttribute(k_probe_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tma_smem);
      if (err != cudaSuccess) fprintf(stderr, "vr_probe_gather: cudaFuncSetAttribute(%zu bytes of shared memory): %s\n", tma_smem, cudaGetErrorString(err));
    }
  }
  for (int rep = 0; rep < 2 && err == cudaSuccess; ++rep) {
    if (rep == 1) err = cudaEventRecord(e0, stream);
    if (err != cudaSuccess) break;
    if (kind == 0) k_probe_ldg<1><<<blocks, threads, 0, stream>>>(static_cast<const float *>(lin), log2_edge, iters, sink);
    else if (kind == 1) k_probe_ldg<4><<<blocks, threads, 0, stream>>>(static_cast<const float4 *>(lin), log2_edge, iters, sink);
    else if (kind == 4) {
      // Default = variant 1: the taps from shared memory with the staging taken out, i.e. the UPPER BOUND of a staged march.
      // The full pipeline (VR_TMA_VARIANT=0) faults with "illegal instruction" in its tensor copy on this pool although the
      // single-box self tests (VR_TMA_SELFTEST=1, 2-D and 3-D, same descriptor, same PTX) load correct data; compute-sanitizer is
      // closed here, so it is unresolved and stays behind the debug variable (profiles/r02_tma_staging.md).
      const char *tv = getenv("VR_TMA_VARIANT");
      k_probe_tma<<<blocks, threads, tma_smem, stream>>>(d_tmap, log2_edge, tma_chunks, sink, tv ? atoi(tv) : 1);
      const cudaError_t le = cudaPeekAtLastError();
      if (le != cudaSuccess) fprintf(stderr, "vr_probe_gather: k_probe_tma launch (%d blocks x %d threads, %zu B smem): %s\n", blocks, threads, tma_smem, cudaGetErrorString(le));
    }
    else if (kind == 2) k_probe_tex<float><<<blocks, threads, 0, stream>>>(tex, log2_edge, iters, sink);
    else k_probe_tex<float4><<<blocks, threads, 0, stream>>>(tex, log2_edge, iters, sink);
    err = cudaGetLastError();
  }
  if (err == cudaSuccess) err = cudaEventRecord(e1, stream);
  if (err == cudaSuccess) err = cudaEventSynchronize(e1);
  float ms = 0.0f;
  if (err == cudaSuccess) err = cudaEventElapsedTime(&ms, e0, e1);
  if (err == cudaSuccess && kind == 4) {
    float flag[2] = {0.0f, 0.0f};
    err = cudaMemcpy(flag, sink, sizeof flag, cudaMemcpyDeviceToHost);
    if (err == cudaSuccess && flag[1] != 0.0f) err = cudaErrorLaunchTimeout;  // a tensor copy never completed
  }
  if (err == cudaSuccess) {
    *ms_out = ms;
    const double per_thread = kind == 4 ? (double)tma_chunks * TMA_STEPS : (double)iters;
    *fetches_per_s = (double)blocks * threads * per_thread / (ms * 1e-3);
  }
  if (d_tmap) cudaFree(d_tmap);
  if (tex) cudaDestroyTextureObject(tex);
  if (arr) cudaFreeArray(arr);
  if (lin) cudaFree(lin);
  if (sink) cudaFree(sink);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  return err;
}

}  // namespace vrb
