// hwfilter_probe.cu - developer probe: what exactly does the texture unit compute for a linear-filtered tex3D<float> with
// normalized coordinates (the reference's texture mode, volumeRender_kernel.cu:544-548)?  Dumps raw samples so that the
// oracle's emulation of the filter (oracle.cpp: tex3D, 1.8 fixed-point weights) can be fitted to the hardware.
//   build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/hwfilter_probe tools/hwfilter_probe.cu
//   run  : tools/hwfilter_probe gpurun_out/hwfilter   (writes <prefix>_{coords,ramp,rand,vol}.f32)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

static cudaTextureObject_t make_tex(const std::vector<float> &v, int nx, int ny, int nz, cudaArray_t *arr) {
  cudaChannelFormatDesc cd = cudaCreateChannelDesc<float>();
  CK(cudaMalloc3DArray(arr, &cd, make_cudaExtent(nx, ny, nz)));
  cudaMemcpy3DParms p = {};
  p.srcPtr = make_cudaPitchedPtr((void *)v.data(), nx * sizeof(float), nx, ny);
  p.dstArray = *arr;
  p.extent = make_cudaExtent(nx, ny, nz);
  p.kind = cudaMemcpyHostToDevice;
  CK(cudaMemcpy3D(&p));
  cudaResourceDesc rd = {};
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = *arr;
  cudaTextureDesc td = {};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear;
  td.readMode = cudaReadModeElementType;
  td.normalizedCoords = 1;
  cudaTextureObject_t t;
  CK(cudaCreateTextureObject(&t, &rd, &td, nullptr));
  return t;
}

__global__ void k_sample(cudaTextureObject_t t, const float *c, int m, float *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) out[i] = tex3D<float>(t, c[3 * i], c[3 * i + 1], c[3 * i + 2]);
}

static void dump(const char *prefix, const char *name, const float *p, size_t n) {
  char path[512];
  snprintf(path, sizeof path, "%s_%s.f32", prefix, name);
  FILE *f = fopen(path, "wb");
  fwrite(p, 4, n, f);
  fclose(f);
}

int main(int argc, char **argv) {
  const char *prefix = argc > 1 ? argv[1] : "hwfilter";
  const int NX = 16, NY = 12, NZ = 20, M = 1 << 20;
  unsigned s = 12345u;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (float)(s >> 8) * (1.0f / 16777216.0f); };
  std::vector<float> rampx(NX * NY * NZ), rampy(NX * NY * NZ), rampz(NX * NY * NZ), rv(NX * NY * NZ);
  for (int z = 0; z < NZ; ++z)
    for (int y = 0; y < NY; ++y)
      for (int x = 0; x < NX; ++x) {
        const int i = (z * NY + y) * NX + x;
        rampx[i] = (float)x, rampy[i] = (float)y, rampz[i] = (float)z, rv[i] = rnd();
      }
  std::vector<float> c(3 * (size_t)M);
  for (int i = 0; i < M; ++i) {
    if (i < M / 4) {  // fine sweep along x at fixed y, z: the weight as a function of the coordinate
      c[3 * i] = -0.05f + 1.1f * (float)i / (float)(M / 4), c[3 * i + 1] = 0.4f, c[3 * i + 2] = 0.6f;
    } else if (i < M / 2) {  // fine sweep inside one voxel
      c[3 * i] = (5.0f + (float)(i - M / 4) / (float)(M / 4) * 1.5f) / NX, c[3 * i + 1] = 0.37f, c[3 * i + 2] = 0.53f;
    } else {
      c[3 * i] = -0.1f + 1.2f * rnd(), c[3 * i + 1] = -0.1f + 1.2f * rnd(), c[3 * i + 2] = -0.1f + 1.2f * rnd();
    }
  }
  float *dc, *dout;
  CK(cudaMalloc(&dc, c.size() * 4));
  CK(cudaMalloc(&dout, (size_t)M * 4));
  CK(cudaMemcpy(dc, c.data(), c.size() * 4, cudaMemcpyHostToDevice));
  std::vector<float> out(M);
  dump(prefix, "coords", c.data(), c.size());
  dump(prefix, "vol", rv.data(), rv.size());
  const std::vector<float> *vols[4] = {&rampx, &rampy, &rampz, &rv};
  const char *names[4] = {"rampx", "rampy", "rampz", "rand"};
  for (int k = 0; k < 4; ++k) {
    cudaArray_t arr;
    cudaTextureObject_t t = make_tex(*vols[k], NX, NY, NZ, &arr);
    k_sample<<<(M + 255) / 256, 256>>>(t, dc, M, dout);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(out.data(), dout, (size_t)M * 4, cudaMemcpyDeviceToHost));
    dump(prefix, names[k], out.data(), M);
    cudaDestroyTextureObject(t);
    cudaFreeArray(arr);
  }
  printf("wrote %s_*.f32 (NX %d NY %d NZ %d, M %d)\n", prefix, NX, NY, NZ, M);
  return 0;
}
