/* texref_shim.h - TEST INFRASTRUCTURE (oracle/_ref).  The reference's kernel file
 * (src/C/vr/volumeRender_kernel.cu) uses CUDA texture REFERENCES, which CUDA 12 removed.  This
 * header, force-included in front of the reference source, gives the removed names back on top
 * of texture OBJECTS so that the reference's own d_render and its own upload / binding code run
 * on this box:
 *   texture<T, dim, mode>            a struct holding a cudaTextureObject_t + the fields the
 *                                    reference sets (normalized, filterMode, addressMode[])
 *   tex3D(texture, x, y, z)          -> tex3D<T>(object, x, y, z)  (the hardware fetch)
 *   cudaBindTextureToArray / cudaUnbindTexture
 * The only edit made to the reference source (by sed, at build time, into oracle/_ref/) is the
 * storage class of its seven namespace-scope `texture<...> tex_*;` variables, which become
 * `__managed__` so that both host and device code can name them.  Nothing else is touched; the
 * build is labelled "shimmed reference" wherever its numbers appear. */
#ifndef TEXREF_SHIM_H_
#define TEXREF_SHIM_H_
#include <cuda_runtime.h>
#include <string.h>

template <class T, int Dim = cudaTextureType3D, enum cudaTextureReadMode Mode = cudaReadModeElementType>
struct texture {
  cudaTextureObject_t obj;
  int normalized;
  enum cudaTextureFilterMode filterMode;
  enum cudaTextureAddressMode addressMode[3];
};

template <class T, int Dim, enum cudaTextureReadMode Mode>
__device__ __forceinline__ T tex3D(texture<T, Dim, Mode> t, float x, float y, float z) {
  return tex3D<T>(t.obj, x, y, z);
}

template <class T, int Dim, enum cudaTextureReadMode Mode>
inline cudaError_t cudaUnbindTexture(texture<T, Dim, Mode> &t) {
  cudaError_t e = cudaSuccess;
  if (t.obj) e = cudaDestroyTextureObject(t.obj);
  t.obj = 0;
  return e;
}

template <class T, int Dim, enum cudaTextureReadMode Mode>
inline cudaError_t cudaBindTextureToArray(texture<T, Dim, Mode> &t, cudaArray *array, const cudaChannelFormatDesc &) {
  cudaDeviceSynchronize(); /* the texture variables are __managed__: no kernel may be in flight while the host writes them */
  if (t.obj) cudaDestroyTextureObject(t.obj);
  t.obj = 0;
  cudaResourceDesc res;
  memset(&res, 0, sizeof res);
  res.resType = cudaResourceTypeArray;
  res.res.array.array = array;
  cudaTextureDesc td;
  memset(&td, 0, sizeof td);
  td.normalizedCoords = t.normalized;
  td.filterMode = t.filterMode;
  td.addressMode[0] = t.addressMode[0], td.addressMode[1] = t.addressMode[1], td.addressMode[2] = t.addressMode[2];
  td.readMode = Mode;
  cudaTextureObject_t o = 0;
  cudaError_t e = cudaCreateTextureObject(&o, &res, &td, NULL);
  t.obj = o;
  return e;
}
#endif
