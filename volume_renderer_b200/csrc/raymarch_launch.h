// raymarch_launch.h - host-callable launchers of the device code (kernels live in the .cu files).
#ifndef RAYMARCH_LAUNCH_H_
#define RAYMARCH_LAUNCH_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include "vr_types.h"

namespace vrb {

// raymarch_kernel.cu: the ray march.  layout / dtype describe the volume textures of the launch
// (all volume slots of one launch share them; the LUT is always fp32 linear).
// One module per gradient method (volumeRender.h:141-144).
// ... and per arithmetic flavour (raymarch_core.h: VR_ARITH; 0 exact, 1 reference, 2 fast).
#define VR_DECLARE_LAUNCH(a)                                                                                                  \
  cudaError_t launch_raymarch_lookup_a##a(const KernelParams &P, int layout, int dtype, cudaStream_t stream, int sm_count); \
  cudaError_t launch_raymarch_compute_a##a(const KernelParams &P, int layout, int dtype, cudaStream_t stream, int sm_count);
VR_DECLARE_LAUNCH(0)
VR_DECLARE_LAUNCH(1)
VR_DECLARE_LAUNCH(2)
#undef VR_DECLARE_LAUNCH
inline cudaError_t launch_raymarch(const KernelParams &P, int layout, int dtype, cudaStream_t stream, int sm_count, int arith) {
  typedef cudaError_t (*Fn)(const KernelParams &, int, int, cudaStream_t, int);
  static const Fn lookup[3] = {launch_raymarch_lookup_a0, launch_raymarch_lookup_a1, launch_raymarch_lookup_a2};
  static const Fn compute[3] = {launch_raymarch_compute_a0, launch_raymarch_compute_a1, launch_raymarch_compute_a2};
  if (arith < 0 || arith > 2) return cudaErrorInvalidValue;
  return (P.gradient_method ? lookup : compute)[arith](P, layout, dtype, stream, sm_count);
}

// depth slices of the volume-space passes: the first lattice sample of every slice (raymarch_kernel.cu: k_slice_states)
cudaError_t launch_slice_states_a0(const KernelParams &P, cudaStream_t stream, int sm_count);
cudaError_t launch_slice_states_a1(const KernelParams &P, cudaStream_t stream, int sm_count);
cudaError_t launch_slice_states_a2(const KernelParams &P, cudaStream_t stream, int sm_count);
inline cudaError_t launch_slice_states(const KernelParams &P, cudaStream_t stream, int sm_count, int arith) {
  typedef cudaError_t (*Fn)(const KernelParams &, cudaStream_t, int);
  static const Fn fn[3] = {launch_slice_states_a0, launch_slice_states_a1, launch_slice_states_a2};
  if (arith < 0 || arith > 2) return cudaErrorInvalidValue;
  return fn[arith](P, stream, sm_count);
}

// resident_kernels.cu: everything that builds the resident representation of a volume.
struct RepackArgs {
  void *dst;           // padded (bricked) storage of the whole volume
  const void *src;     // staged source slices, MATLAB layout (dims[0] fastest)
  int w, h, d;         // GLOBAL voxel dims
  int lw, lh, ld;      // local padded dims of dst (w+2, h+2, d+2 for a whole volume)
  int gx0, gy0, gz0;   // dst padded index (X,Y,Z) holds global voxel clamp(X+gx0, 0, w-1)... (-1 for a whole volume)
  int sx0, sy0;        // src covers global voxels [sx0, sx0+sw) x [sy0, sy0+sh) per slice
  int sw, sh;
  int src_z0;          // global z index of the first staged slice
  uint64_t o0, o1;     // destination element range [o0, o1) to produce
  uint32_t sy, sz;     // destination strides (see TexDesc)
  int estride, ecomp;  // destination element o is written to dst[o * estride + ecomp] (1, 0 unless interleaved)
  int box[6];          // only elements whose (clamped) source voxel lies in [box[0..2], box[3..5]) are written;
                       // {0,0,0,w,h,d} = all (partial update of a resident volume, vr_update_box)
  int layout;          // LAYOUT_*
  int src_dtype;       // vr_dtype of src
  int dst_dtype;       // vr_dtype of dst
  unsigned long long src_surface;  // != 0: the source is the volume's own w x h x d fp32 cudaArray (a surface object) instead of
                                   // `src` - rebuilds a padded storage that was dropped in texture-unit mode (context.cu)
};
cudaError_t launch_repack(const RepackArgs &a, cudaStream_t stream);

// occupancy bit grid: bit set when any voxel of the 8^3 macro cell dilated by `dilate` voxels is
// non-zero (read from the padded storage).
// reg = {lo x,y,z, hi x,y,z} in voxels: cells that cannot hold an owned sample stay empty.
cudaError_t launch_occupancy(const TexDesc &t, int layout, int dtype, uint32_t *occ, int ocx, int ocy, int ocz,
                             int dilate, const int reg[6], cudaStream_t stream);

// distance grid of an occupancy bit grid: dist[cell] = Chebyshev distance in cells to the nearest
// set bit (0 on a set bit, capped at 255).  tmp: scratch of the same size.
cudaError_t launch_distance(const uint32_t *occ, uint8_t *dist, uint8_t *tmp, int ocx, int ocy, int ocz, cudaStream_t stream);

// Henyey-Greenstein LUT (HenyeyGreenstein.cc:47-91): n^3 floats on the device.
cudaError_t launch_hg_lut(float *d_out, uint32_t n, float g, cudaStream_t stream);
// The same LUT into the library's file-scope device scratch (no allocation; n <= HG_STATIC_MAX_N): enqueues the kernel on
// `stream` for the CURRENT device and returns the scratch's device address in *d_lut (valid until the next call).
enum { HG_STATIC_MAX_N = 64 };
cudaError_t launch_hg_lut_static(uint32_t n, float g, cudaStream_t stream, float **d_lut);

// synthetic ViBE-Z-shaped volumes; kind 0 = brain, 1 = tubes
cudaError_t launch_synth(void *d_out, int kind, uint32_t seed, int w, int h, int d, int dtype, cudaStream_t stream);

// the voxel box [lo, hi) of the same synthetic volume, stored densely (lo[0] fastest)
cudaError_t launch_synth_region(void *d_out, int kind, uint32_t seed, const int dims[3], const int lo[3], const int hi[3], int dtype,
                                cudaStream_t stream);
// sort-last compositing of the volume-space mode: out[i] = sum over bricks (in index order) of partial_b[i].
// Every partial is already weighted by the transmittance of the bricks in front of it, so the "under"
// operator over depth-ordered partials reduces to an order-free sum.  Partials may live on peer GPUs.
cudaError_t launch_compose(const float *const *d_partials, int n, uint64_t n_floats, float *d_out, cudaStream_t stream);
// which pixels of a plane a context owns under an image-space partition (nranks <= 1: all of them)
struct FoldOwner {
  uint64_t plane;  // W * H
  uint32_t H, tile_w, tile_h, ntx, rank, nranks, columns;
};
cudaError_t launch_fold_slices(const float *d_slices, int S, uint64_t n, float *d_out, int product, const FoldOwner &own, cudaStream_t stream);

// image epilogue: stereo crop + compose (mode 0 RedCyan, 1 LeftRightHorizontal) and sequence normalisation
cudaError_t launch_stereo(const float *left, const float *right, int H, int Wp, int delta, int mode, float *out, cudaStream_t stream);
cudaError_t launch_normalize(float *const *imgs, int n_imgs, size_t n_each, unsigned *d_mm, cudaStream_t stream);

// MATLAB gradient() of a MATLAB-layout volume (Volume.m:181-205): gx = d/d dim2, gy = d/d dim1, gz = d/d dim3
cudaError_t launch_gradient(const void *d_in, int w, int h, int d, int dtype, void *gx, void *gy, void *gz, cudaStream_t stream);

// texture-unit mode (VR_OPT_TEXTURE_UNIT): copies the voxels of a whole resident fp32 volume (padded storage, any layout,
// interleaved component or not) into a w x h x d cudaArray through a surface object
// ... or of the voxel box [a0, a0 + an) of it (a region context's resident box clipped to the volume); dtype = vr_dtype of the storage
cudaError_t launch_fill_array(const TexDesc &t, int layout, int dtype, const int a0[3], const int an[3], unsigned long long surface, cudaStream_t stream);

// probe_kernels.cu: gather ceilings of the box with the march's access pattern; kind 0 = 8 x 4-byte
// loads per fetch, 1 = 8 x 16-byte loads, 2 = tex3D<float> linear, 3 = tex3D<float4> linear; footprint (2^log2_edge)^3 voxels
cudaError_t launch_probe_gather(int kind, int log2_edge, int sm_count, cudaStream_t stream, double *fetches_per_s, double *ms);

}  // namespace vrb
#endif
