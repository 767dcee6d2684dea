// vr_types.h - POD types shared by the host side (context.cu) and the kernels.
#ifndef VR_TYPES_H_
#define VR_TYPES_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define VR_HD __host__ __device__ __forceinline__
#else
#define VR_HD inline
#endif

namespace vrb {

// Texture "slots" = the reference's seven texture references (volumeRender_kernel.cu:90-120),
// numbered like vr::VolumeType (volumeRender.h:152-160).
enum Slot { S_EM = 0, S_AB = 1, S_RE = 2, S_DX = 3, S_DY = 4, S_DZ = 5, S_LIGHT = 6, S_N = 7 };

enum TexKind {
  TEX_UNBOUND = 0,  // unbound texture reference: reads 0 (undefined in the reference)
  TEX_CONST = 1,    // 1x1x1 volume (the default VolumeReflection, VolumeRender.m:131): a constant
  TEX_VOLUME = 2
};

// LAYOUT_TEXHW: the launch fetches through the texture unit (TexDesc::tex, hardware linear filtering as the reference
// configures it, kernel.cu:544-548); the padded storage then only feeds the empty-space grids
enum Layout { LAYOUT_LINEAR = 0, LAYOUT_BRICK = 1, LAYOUT_TEXHW = 2 };
enum { BRICK = 8, BRICK_SHIFT = 3, BRICK_MASK = 7, BRICK_ELEMS = 512 };
enum { CELL_SHIFT = 3 };  // empty-space macro cell = 8^3 voxels

// Resident volume as the kernels see it.  Storage is PADDED by one replicated voxel on every
// side (element (X,Y,Z), X in [0,w+1], holds voxel clamp(X-1,0,w-1)...), so that clamp-to-edge
// addressing (kernel.cu:546-548) needs no integer clamps: coordinates are clamped to [0,1] in
// float, which gives i = floor(x*N-0.5) in [-1,N-1], and the taps i,i+1 are always stored.
//   linear : off = X + Y*sy + Z*sz                       (sy = w+2, sz = sy*(h+2))
//   bricked: off = (X>>3)*512 + (X&7) + (Y>>3)*sy + (Y&7)*8 + (Z>>3)*sz + (Z&7)*64
//            (sy = nbx*512, sz = nbx*nby*512; a brick is 8x8x8 voxels = 2 KiB of fp32)
// A context that owns only a sub-box of the volume (volume-space mode) stores just that region plus
// a halo: w, h, d stay the GLOBAL dims (they define the coordinate math), ox/oy/oz are subtracted
// from the global padded tap index to get the index into the local storage (0 for a whole volume),
// lw/lh/ld are the local padded dims.
struct TexDesc {
  const void *p;
  int w, h, d;
  float fw, fh, fd;
  uint32_t sy, sz;
  int kind;
  float cval;
  int ox, oy, oz;
  int lw, lh, ld;
  // element (X,Y,Z) sits at p[off * estride + ecomp].  estride is 1 (0 reads as 1) except for the
  // three gradient volumes, which share ONE storage of 4-element voxels (Gx, Gy, Gz, unused) so
  // that lookupGradient costs one 16-byte load per tap instead of three 4-byte ones.
  int estride, ecomp;
  unsigned long long tex;  // cudaTextureObject_t of the volume's cudaArray (VR_OPT_TEXTURE_UNIT), else 0
};

enum { MAX_PARAM_LIGHTS = 8 };

// One camera of a launch: what differs between the frames of a batch (a stereo pair, a rotation movie).
struct Camera {
  float m[4][3];  // rows 0..2 camera axes, row 3 = (camXOffset, focal, objDist)  (volumeRender.h:37-44)
  float *out;     // W*H*3 floats, column-major planar: k = x*H + y (kernel.cu:496-506)
};

// Everything one ray-march launch needs; passed by value as a __grid_constant__ parameter.
struct KernelParams {
  Camera cam;           // the frame of a single-frame launch
  const Camera *cams;   // n_frames > 1: the cameras of the batch (device memory); work items of all
  int n_frames;         // frames share ONE queue, so the long rays of one frame fill the tail of another
  uint32_t tiles_total; // ntx * nty
  int W, H;
  // image-space partition (multi-GPU): the image is cut into tile_w x tile_h tiles, tile j is
  // rendered by the context with j % part_nranks == part_rank.
  int part_rank, part_nranks;
  int tile_w, tile_h, ntx, nty;
  int wt_x, wt_y;          // warp tiles per partition tile
  int wshift;              // a warp covers (1 << wshift) x (32 >> wshift) pixels (default 3: 8 x 4)
  uint32_t num_owned;      // partition tiles owned by this rank
  uint32_t num_work;       // warp work items = num_owned * wt_x * wt_y
  unsigned int *work_counter;
  // owned tiles in the order they are handed out (heaviest first, see context.cu: tile_order);
  // entry j is a tile index ty * ntx + tx.  NULL = part_rank + j * part_nranks.
  const uint32_t *tile_order;
  unsigned int *tile_cost;         // per tile: shaded samples of the launch (feeds the next launch's hand-out order); NULL = off
  unsigned long long *warp_times;  // debug: per resident warp [start, end] in globaltimer ns (NULL = off)
  unsigned long long *stats;  // [0] samples marched, [1] samples shaded, [2] rays hit
  // vr::RenderOptions (volumeRender.h:81-109)
  float fEm, fAb, fRe;
  float boxmin[3], boxmax[3];
  float thr, tstep;
  float color[3];
  float gstep[3];  // (1/w, 1/h, 1/d) of the emission volume (volumeRender.cpp:273-275)
  // role -> texture, already resolved through d_idx* on the host (kernel.cu:75-85,127-144)
  TexDesc em, ab, re;
  TexDesc em_direct;  // tex_emission itself, as computeGradient reads it (kernel.cu:225-248)
  TexDesc gx, gy, gz, lut;
  TexDesc geo;   // the geometry shared by every TEX_VOLUME slot above but the LUT (valid when `uniform`)
  int uniform;   // all volume textures of the launch share one geometry -> shared tap sets
  int ab_is_em;  // absorption role resolves to the same texture as the emission role
  int re_is_em;  // reflection role resolves to the same texture as the emission role
  int grad4;     // gx, gy, gz are components 0..2 of one interleaved storage (gx.p)
  int tex_fast;  // texture-unit launches: bit 0 = gx, gy, gz are all volumes, bit 1 = re is a volume, bit 2 = em_direct is a
                 // volume, bit 3 = em, bit 4 = the LUT -> straight-line fetches that share one set of coordinate registers (no per-slot kind tests)
  int gradient_method;  // 0 compute, 1 lookup (volumeRender.h:141-144)
  int frac_bits;        // 0 = fp32 weights, 8 = 1.8 fixed point
  int skip;             // empty-space skipping: 0 off, 1 serial per cell, 2 closed-form jumps (default)
  int count;            // maintain stats
  // volume-space mode (sort-last): this launch marches only the samples whose normalized position
  // lies in [reg_lo, reg_hi) - the global sample lattice is kept, other samples count as empty.
  //   vs_mode 0: off.  1: transmittance pass - per pixel the opacity a_b this brick alone would
  //   accumulate (no shading), written to vs_opacity_out[pixel].  2: shading pass - sum.w starts at
  //   the opacity of the bricks in front, 1 - prod(1 - a_b'), read from vs_opacity_all
  //   [brick][pixel]; the output is this brick's additive share of the final image.
  int vs_mode;
  int vs_index, vs_count;          // this brick, number of bricks
  float reg_lo[3], reg_hi[3];      // owned region in normalized texture coordinates
  // Depth slices (vs_slices > 1): the stretch of a ray inside this brick is cut into vs_slices equal parts handled by DIFFERENT
  // work items (item = (tile, warp footprint, slice)), so that the longest chain of a launch is a slice, not a whole ray.  Pass 1
  // writes the opacity of every slice to vs_slice_opacity[slice][pixel] (the brick's total, 1 - prod(1 - a_s), is derived from
  // them and is what the bricks exchange); pass 2 starts slice s from the bricks in front AND the slices s' < s of this brick and
  // writes its additive share to cam.out + slice * 3 * W * H (summed in slice order afterwards).
  // Where slice s >= 1 of a pixel starts - (t, pos) of its first lattice sample, t = 3e38 when the ray ends before - is found ONCE
  // per frame by k_slice_states (the exact fast-forward of the reference's recurrences) and read by both passes:
  // entry (s - 1) * W * H + pixel of vs_slice_state.
  int vs_slices;
  int part_columns;  // 1: the image-space partition deals tile COLUMNS (VR_OPT_PART_MODE); only k_slice_states needs to know
  int measure_tail;  // whole-ray shading pass: stats[3..5] = ~(first warp start), ~(first warp end), last warp end, ns (the automatic choice of slices)
  float *vs_slice_opacity;
  float *vs_slice_state;  // 4 floats per entry (16-byte aligned)
  float *vs_opacity_out;           // mode 1
  const float *vs_opacity_all;     // mode 2: vs_count planes of W*H floats (pixel k = x*H + y)
  const float *vs_boxes;           // mode 2: vs_count x 6 floats, world-space brick boxes (lo xyz, hi xyz)
  // lights (volumeRender.h:50-55), position already axis-reversed
  int n_lights;
  float lights[MAX_PARAM_LIGHTS][6];
  const float *lights_more;  // lights beyond MAX_PARAM_LIGHTS, 6 floats each (device memory)
  // occupancy bit grid of the absorption-role texture: bit (cz*ocy+cy)*ocx+cx set = not empty
  const uint32_t *occ;
  // the same grid as bytes: Chebyshev distance in cells to the nearest non-empty cell (0 = the
  // cell itself is not empty, 255 = at least 255)
  const uint8_t *dist;
  int ocx, ocy, ocz;
  // the macro cell of a sample's low tap straight from the world position: floor(clamp(pos * cellA + cellB, 0, cellMax))
  // per axis (= floor((x*N - 0.5) / 8) up to ~1e-4 voxels, far inside the grid's two-voxel dilation); see on_empty_cell
  float cellA[3], cellB[3], cellMax[3];
  // texture-unit launches of a context that owns a REGION (k_raymarch<.., LAYOUT_TEXHW, .., UNI = 1>): the arrays hold the
  // region's voxel box [a0, a0 + aw) only, so a normalized volume coordinate x becomes x * tex_cs + tex_co
  // (= (x * N - a0) / aw: the same texel position x * N - 0.5 - a0 in the local array)
  float tex_cs[3], tex_co[3];
};

}  // namespace vrb
#endif
