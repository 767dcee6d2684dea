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
// The footprint (edge^3 voxels) chooses the level that serves the taps: 32^3 fp32 sits in L1, 256^3 in
// L2, 1024^3 needs HBM.  Kinds 0/1 spend ~15 integer/float instructions per fetch on addresses - far
// fewer than the march does - so their rate is the ceiling of the load path, not of a renderer.
#include <cuda_runtime.h>
#include <stdint.h>

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

}  // namespace

// Runs one probe on `stream`; *fetches_per_s = trilinear fetches (8 taps each) per second, *ms = duration of
// the timed launch (the second of two).  Allocates and frees its own footprint.
cudaError_t launch_probe_gather(int kind, int log2_edge, int sm_count, cudaStream_t stream, double *fetches_per_s, double *ms_out) {
  if (kind < 0 || kind > 3 || log2_edge < 4 || log2_edge > 10) return cudaErrorInvalidValue;
  const size_t voxels = (size_t)1 << (3 * log2_edge);
  const size_t vb = (kind == 1 || kind == 3) ? 16 : 4;
  const int threads = 256, blocks = sm_count * (kind == 1 ? 5 : 8);  // one wave: 2048 resident threads per SM (1280 with 16-byte taps)
  // ~2 G taps per launch whatever the kind: long enough to time, short enough for a bench preamble
  const int iters = (kind >= 2) ? 2048 : 1024;
  float *sink = nullptr;
  void *lin = nullptr;
  cudaArray_t arr = nullptr;
  cudaTextureObject_t tex = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t err = cudaMalloc(&sink, 4);
  if (err == cudaSuccess) err = cudaEventCreate(&e0);
  if (err == cudaSuccess) err = cudaEventCreate(&e1);
  if (err == cudaSuccess && kind < 2) {
    err = cudaMalloc(&lin, voxels * vb);
    if (err == cudaSuccess) err = cudaMemsetAsync(lin, 0, voxels * vb, stream);
  }
  if (err == cudaSuccess && kind >= 2) {
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
  for (int rep = 0; rep < 2 && err == cudaSuccess; ++rep) {
    if (rep == 1) err = cudaEventRecord(e0, stream);
    if (err != cudaSuccess) break;
    if (kind == 0) k_probe_ldg<1><<<blocks, threads, 0, stream>>>(static_cast<const float *>(lin), log2_edge, iters, sink);
    else if (kind == 1) k_probe_ldg<4><<<blocks, threads, 0, stream>>>(static_cast<const float4 *>(lin), log2_edge, iters, sink);
    else if (kind == 2) k_probe_tex<float><<<blocks, threads, 0, stream>>>(tex, log2_edge, iters, sink);
    else k_probe_tex<float4><<<blocks, threads, 0, stream>>>(tex, log2_edge, iters, sink);
    err = cudaGetLastError();
  }
  if (err == cudaSuccess) err = cudaEventRecord(e1, stream);
  if (err == cudaSuccess) err = cudaEventSynchronize(e1);
  float ms = 0.0f;
  if (err == cudaSuccess) err = cudaEventElapsedTime(&ms, e0, e1);
  if (err == cudaSuccess) {
    *ms_out = ms;
    *fetches_per_s = (double)blocks * threads * iters / (ms * 1e-3);
  }
  if (tex) cudaDestroyTextureObject(tex);
  if (arr) cudaFreeArray(arr);
  if (lin) cudaFree(lin);
  if (sink) cudaFree(sink);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  return err;
}

}  // namespace vrb
