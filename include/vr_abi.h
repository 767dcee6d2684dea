/* vr_abi.h - C ABI of the B200 ray-march engine (libvr_b200.so).
 *
 * This is the drop-in boundary for the ONE hot path of raphiniert-com/volume_renderer: the
 * per-pixel ray march.  Every entry point replaces one command of the reference's mex gateway
 * `volumeRender(cmd, handle, ...)` (reference: src/C/mex/render.cpp:50-277) or one of its two
 * helper mex files.  The gateway (mex/volumeRender_mex.cpp in this repo) only unpacks mxArrays
 * into the plain pointers below; ALL argument re-ordering the reference does while unpacking
 * (axis reversal of positions, element sizes and the rotation matrix) happens behind this ABI,
 * so the arrays are passed exactly as MATLAB hands them over.
 *
 * Conventions
 *  - plain C types only, no exceptions / longjmp across the boundary; every call returns a
 *    vr_status, the message is available from vr_last_error().
 *  - all state is context-local (the reference keeps process-wide globals,
 *    volumeRender_kernel.cu:55-120); a context is bound to one CUDA device.
 *  - host volume memory is borrowed for the duration of the call only (render.cpp:96-98 borrows
 *    MATLAB's buffer through mxGetPropertyShared, volumeRender.cpp:310).
 *  - MATLAB layout everywhere: column-major, dims[0] = size(Data,1) is the fastest axis and
 *    becomes texture x (volumeRender.cpp:338-339).
 */
#ifndef VR_ABI_H_
#define VR_ABI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define VR_API __declspec(dllexport)
#else
#define VR_API __attribute__((visibility("default")))
#endif

#define VR_ABI_VERSION 1

/* Opaque persistent context = the reference's mm::MManager behind a class_handle
 * (src/C/vr/mm/mmanager.hxx:25-105, src/C/vr/mm/class_handle.hpp:42-62). */
typedef struct vr_ctx vr_ctx;

typedef enum vr_status {
  VR_OK = 0,
  VR_ERR_INVALID_ARGUMENT = 1,  /* mexErrMsgTxt("insufficient parameter!") etc., render.cpp:51-56,136-140 */
  VR_ERR_INVALID_HANDLE = 2,    /* "Handle not valid.", class_handle.hpp:69-70 */
  VR_ERR_INSUFFICIENT_VRAM = 3, /* "insufficient free VRAM!", mmanager.hxx:161-172 */
  VR_ERR_CUDA = 4,              /* CUDA runtime error (the reference drops these in release, common.h:46) */
  VR_ERR_NO_DEVICE = 5,         /* no CUDA device / wrong architecture: there is NO CPU fallback */
  VR_ERR_UNSUPPORTED = 6
} vr_status;

typedef enum vr_dtype {
  VR_DTYPE_F32 = 0,  /* the reference's only type (volumeRender.h:26) */
  VR_DTYPE_BF16 = 1  /* storage only, arithmetic stays fp32 */
} vr_dtype;

typedef enum vr_memspace {
  VR_MEM_HOST = 0,   /* pageable or pinned host memory (the MATLAB case) */
  VR_MEM_DEVICE = 1, /* already resident on this context's device (C harness / pipelines) */
  VR_MEM_DEVICE_REGION = 2 /* device buffer holding ONLY this context's region plus halo, clipped to the volume
                              (vr_set_region); dims still describe the whole volume */
} vr_memspace;

/* One MATLAB `Volume` object as mxMake_volume reads it (volumeRender.cpp:307-342):
 * Data pointer + dims, TimeLastUpdate.  memory_size is derived (volumeRender.cpp:71-72). */
typedef struct vr_volume {
  const void *data;
  uint64_t dims[3];     /* size(Data,1), size(Data,2), size(Data,3); 1 for a 2-D array */
  uint64_t last_update; /* Volume.TimeLastUpdate (ms timestamp, Volume.m:225-238) */
  int32_t dtype;        /* vr_dtype */
  int32_t memspace;     /* vr_memspace */
} vr_volume;

/* One MATLAB `LightSource` (LightSource.m:31-37): single 1x3 Position and Color exactly as
 * stored in MATLAB.  Position is axis-reversed internally like render.cpp:162-168. */
typedef struct vr_light {
  float position[3];
  float color[3];
} vr_light;

/* Counters of the last vr_render*/
typedef struct vr_stats {
  double kernel_ms;          /* device time of the ray-march launch(es), CUDA events */
  double upload_ms;          /* device time of uploads done by the last vr_sync_volumes */
  uint64_t samples_marched;  /* iterations of the loop volumeRender_kernel.cu:435-493 (skipped ones count) */
  uint64_t samples_shaded;   /* iterations that actually sampled/shaded (not elided by empty-space skipping) */
  uint64_t rays_hit;         /* pixels whose ray hit the box */
  uint64_t bytes_resident;   /* device bytes owned by the context */
  uint64_t bytes_uploaded;   /* host->device bytes moved by the last vr_sync_volumes / vr_render */
  uint32_t kernel_launches;  /* kernels launched by the last vr_render */
  uint32_t uploads;          /* volume uploads done by the last vr_sync_volumes */
  uint32_t fetch_mode;       /* how the last render fetched: 0 = loads + fp32 lerps, 1 = texture unit (VR_OPT_TEXTURE_UNIT) */
  uint32_t arithmetic;       /* arithmetic flavour of the last render's kernel (VR_OPT_ARITHMETIC) */
  uint32_t vs_slices;        /* depth slices the last volume-space pass marched with (VR_OPT_VS_SLICES resolved); 1 otherwise */
  uint32_t reserved_;
} vr_stats;

/* ---- 'new'  (render.cpp:58-65) -------------------------------------------------------- */
/* Creates a context on the current CUDA device (or $VR_DEVICE).  The reference picks the
 * "max GFLOPS" device (volumeRender.cpp:77-87,163-220); one process per GPU makes that moot. */
VR_API vr_status vr_create(vr_ctx **out_ctx);
VR_API vr_status vr_create_on_device(vr_ctx **out_ctx, int device);

/* ---- 'delete'  (render.cpp:72-79, class_handle.hpp:79-83) ------------------------------ */
/* Frees everything the context owns.  Unlike ~MManager (mmanager.hxx:103-105) it does NOT call
 * cudaDeviceReset().  Tolerates a stale / already deleted handle (VolumeRender.m memClear then
 * delete double-deletes, VolumeRender.m:184-186,230-237): returns VR_ERR_INVALID_HANDLE. */
VR_API vr_status vr_destroy(vr_ctx *ctx);

/* ---- 'mem_info'  (render.cpp:85-87, mmanager.hxx:218-284) ------------------------------ */
/* Writes the NUL-terminated report into buf (truncated to cap); *needed gets the full length. */
VR_API vr_status vr_mem_info(vr_ctx *ctx, char *buf, size_t cap, size_t *needed);

/* ---- 'sync_volumes'  (render.cpp:93-129, mmanager.hxx:178-201, kernel.cu:739-867) ------- */
/* Argument order is the mex's: emission, REFLECTION, absorption.  gx/gy/gz all non-NULL = the
 * nrhs==9 form (gradient lookup); all NULL = the nrhs==6 form (resetGradients, on-the-fly). */
VR_API vr_status vr_sync_volumes(vr_ctx *ctx, uint64_t time_last_mem_sync,
                                 const vr_volume *emission, const vr_volume *reflection,
                                 const vr_volume *absorption, const vr_volume *gradient_x,
                                 const vr_volume *gradient_y, const vr_volume *gradient_z);

/* ---- 'render'  (render.cpp:134-277) --------------------------------------------------- */
/* lights / illumination: NULL mirrors the `logical` placeholders (render.cpp:145-146): the
 *   light set and LUT of the previous render stay in force (none on a fresh context).
 * factors        single[3] = [FactorEmission, FactorReflection, FactorAbsorption]  (VolumeRender.m:558)
 * element_size_um single[3] = ElementSizeUm as in MATLAB (reversed inside, render.cpp:195)
 * resolution     uint64[2] = [H, W]                                                (render.cpp:196,237-243)
 * rotation       single[9] = flip(RotationMatrix), column-major                    (render.cpp:211-219)
 * props          single[3] = [CameraXOffset, FocalLength, DistanceToObject]        (render.cpp:220-221)
 * color          single[3]                                                         (render.cpp:240)
 * out_image      H*W*3 floats, MATLAB H x W x 3 single (column-major planar, kernel.cu:496-506) */
VR_API vr_status vr_render(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                           const vr_volume *illumination, const float factors[3],
                           const float element_size_um[3], const uint64_t resolution[2],
                           const float rotation[9], const float props[3],
                           float opacity_threshold, const float color[3], float *out_image);

/* Same call, but the image stays on the device: d_out_image is a device pointer to H*W*3 floats
 * and the work is enqueued on `cuda_stream` (a cudaStream_t, NULL = the context's stream) without
 * a host synchronisation.  This is the entry for pipelines that post-process on the GPU. */
VR_API vr_status vr_render_device(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                  const vr_volume *illumination, const float factors[3],
                                  const float element_size_um[3], const uint64_t resolution[2],
                                  const float rotation[9], const float props[3],
                                  float opacity_threshold, const float color[3],
                                  float *d_out_image, void *cuda_stream);

/* A batch of frames that differ only in camera (a stereo pair = the two p_render passes of
 * VolumeRender.m:284-285; a rotation movie = the loop of examples/example2.m:59-66) in ONE launch:
 * rotations = n_frames x single[9], props = n_frames x single[3], d_out_images = n_frames device
 * pointers.  All frames share one work queue, so one frame's long rays overlap another's tail. */
VR_API vr_status vr_render_batch_device(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                        const vr_volume *illumination, const float factors[3],
                                        const float element_size_um[3], const uint64_t resolution[2],
                                        size_t n_frames, const float *rotations, const float *props,
                                        float opacity_threshold, const float color[3],
                                        float *const *d_out_images, void *cuda_stream);

/* ---- several GPUs behind the same boundary (no reference counterpart; the reference renders on ONE device,
 * volumeRender.cpp:77-87,153) ---------------------------------------------------------------------------------
 * The context becomes the leader of a group with one member context per listed device (device_ids[0] must be the
 * context's own device; n = 0 or 1 dissolves the group).  vr_sync_volumes then makes the volumes resident on every
 * device (one host thread per device), and every render is cut into tile columns dealt round-robin to the devices:
 * host-image renders (vr_render, vr_render_batch*) have each device copy its own columns into the caller's image over
 * its own PCIe link, device-image renders have the kernels of all devices store into the caller's image over NVLink.
 * Pixels are those of one device, bit for bit.  A MATLAB user adds one call after 'new':
 * volumeRender('set_devices', handle, int32([0 1 2 3])).  Call it before the first vr_sync_volumes. */
VR_API vr_status vr_set_devices(vr_ctx *ctx, uint32_t n, const int *device_ids);

/* Pipelined movie batches (examples/example2.m:59-66 renders 360 frames of one scene): _begin enqueues the launch of a
 * batch and the device->host copies of its frames (of the pixels this context owns when it is partitioned,
 * VR_OPT_PART_MODE) and returns; the next _begin renders into a second set of device images while those copies are on the
 * wire.  out_images must stay valid - and should be pinned, a pageable destination makes the copy synchronous - until
 * vr_render_batch_end, which waits for everything begun; vr_get_stats then holds the counters summed over the batches
 * (kernel_ms = sum of the launches).  Any other call on the context ends the pipeline first. */
VR_API vr_status vr_render_batch_begin(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                       const vr_volume *illumination, const float factors[3],
                                       const float element_size_um[3], const uint64_t resolution[2],
                                       size_t n_frames, const float *rotations, const float *props,
                                       float opacity_threshold, const float color[3], float *const *out_images);
VR_API vr_status vr_render_batch_end(vr_ctx *ctx);

/* The same batch with host images (n_frames pointers to H*W*3 floats each), synchronous like vr_render. */
VR_API vr_status vr_render_batch(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                 const vr_volume *illumination, const float factors[3],
                                 const float element_size_um[3], const uint64_t resolution[2],
                                 size_t n_frames, const float *rotations, const float *props,
                                 float opacity_threshold, const float color[3], float *const *out_images);

/* ---- HenyeyGreenstein(n[,g]) mex  (src/C/mex/HenyeyGreenstein.cc:29-96) ---------------- */
/* n^3 floats, index k = c*n*n + a*n + b (HenyeyGreenstein.cc:86); evaluated on the GPU. */
VR_API vr_status vr_henyey_greenstein(uint32_t n, float g, float *out_lut);

/* The same LUT WITHOUT ANY ALLOCATION.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * vr_henyey_greenstein allocates and frees an n^3 device scratch per call; here that scratch is a file-scope
 * `static __device__` array of the library, VR_HG_STATIC_MAX_N^3 floats = 1 MiB, which every GPU that loads the library
 * holds until the process exits.  n > VR_HG_STATIC_MAX_N is rejected with VR_ERR_INVALID_ARGUMENT (64 is the size every
 * example of the reference uses, example1.m:35).  The call runs on the CURRENT device, enqueues the kernel and the copy
 * on `cuda_stream` (NULL = the legacy default stream), waits for them, and calls no cudaMalloc / cudaFree or any
 * allocating library routine.  out_lut = n^3 floats in host OR device memory (same layout as above).  Every call shares
 * the one array per device, which keeps the last call's contents (each call overwrites all it reads): calls that
 * overlap on the GPU - two streams, two host threads - race; serialise them. */
#define VR_HG_STATIC_MAX_N 64
VR_API vr_status vr_henyey_greenstein_static(uint32_t n, float g, float *out_lut, void *cuda_stream);

/* ---- timestamp() mex  (src/C/mex/timestamp.cpp:13-33) ---------------------------------- */
/* Milliseconds since the epoch.  The reference stores them through an int* (timestamp.cpp:25,33),
 * i.e. only the low 32 bits reach MATLAB; truncate_like_reference != 0 reproduces that. */
VR_API uint64_t vr_timestamp_ms(int truncate_like_reference);

/* ---- errors --------------------------------------------------------------------------- */
VR_API const char *vr_last_error(const vr_ctx *ctx); /* ctx may be NULL: last error of this thread */
VR_API const char *vr_status_string(vr_status s);
VR_API int vr_abi_version(void);

/* ---- extensions with no reference counterpart ------------------------------------------ */
/* Options (ints).  Defaults reproduce the reference as written. */
typedef enum vr_option {
  VR_OPT_FILTER_FRAC_BITS = 0,  /* 0 = fp32 trilinear weights (default); 8 = the texture unit's 1.8 fixed point */
  VR_OPT_EMPTY_SPACE_SKIP = 1,  /* exact empty-space skipping: 2 (default) = closed-form jumps over the distance grid,
                                   1 = serial recurrences cell by cell, 0 = march every sample; pixels are bit-identical.
                                   PRECONDITION of the bit-identity: voxels and factors are finite.  A sample whose
                                   absorption fetch is 0 is not shaded (alpha = 0 makes its premultiplied contribution
                                   exactly 0 for finite inputs); the reference would carry a NaN/Inf emission or reflection
                                   voxel at such a sample into the pixel.  A negative OpacityThreshold (the reference then
                                   stops after its first sample) switches skipping off for that render. */
  VR_OPT_KERNEL_VARIANT = 2,    /* 0 (default) = shared tap sets when all volumes share a geometry; 1 = per-texture addressing;
                                   3 = force the kernels with 64-bit z offsets (chosen automatically for a resident
                                   volume of more than 2^32 elements).  Pixels are bit-identical */
  VR_OPT_ABSORPTION_ROLE_FIX = 3, /* 0 (default) = reference as written: absorption samples the emission texture
                                     (kernel.cu:80); 1 = sample the uploaded absorption volume */
  VR_OPT_PART_RANK = 4,         /* image-space partition: this context renders tiles t with t % nranks == rank */
  VR_OPT_PART_NRANKS = 5,
  VR_OPT_TILE_W = 6,            /* partition tile size in pixels */
  VR_OPT_TILE_H = 7,
  VR_OPT_BRICK_LAYOUT = 8,      /* 1 = bricked resident layout (default), 0 = linear */
  VR_OPT_COUNT_SAMPLES = 9,     /* 1 (default) = maintain vr_stats sample counters */
  VR_OPT_CLEAR_IMAGE = 10,      /* 1 (default) = zero the whole output image first (volumeRender.cpp:271); 0 = only the
                                   pixels of owned tiles are written (every owned pixel is, misses as 0): required when
                                   several ranks render into ONE shared image */
  VR_OPT_TILE_ORDER = 11,       /* hand-out order of tiles: 1 (default) = the tiles with the longest rays first (measured on
                                   the previous launch, else centre-first), one list over all frames of a batch; 2 = the
                                   same, frame after frame; 3 = by the tile's total of shaded samples (A/B); 0 = raster order */
  VR_OPT_WARP_SHAPE = 12,       /* s in [0,5]: one warp marches a (1<<s) x (32>>s) pixel footprint; default 3 = 8 x 4 */
  VR_OPT_GRADIENT_INTERLEAVE = 13, /* 1 (default): the three gradient volumes of lookupGradient are stored interleaved
                                     (16-byte voxels Gx,Gy,Gz,-; one load per tap); 0: three separate volumes.  Takes
                                     effect at the next upload of the gradient volumes; pixels are bit-identical */
  VR_OPT_TEXTURE_UNIT = 14,     /* how a trilinear fetch is done.  2 (default): through the texture unit on fp32 cudaArray copies of
                                   the volumes, configured as the reference configures its textures (volumeRender_kernel.cu:
                                   544-548) - the reference's own hardware filtering - where this context can (fp32 or bf16 volumes,
                                   whole or a vr_set_region shard, whose arrays fit), else as 0.  1: texture unit or
                                   VR_ERR_UNSUPPORTED.  0: eight loads + fp32 lerps, the documented filter formula with fp32 weights
                                   (or VR_OPT_FILTER_FRAC_BITS) */
  VR_OPT_ARITHMETIC = 15,       /* arithmetic flavour of the ray-march kernels.  0 "exact": every fp32 operation individually rounded,
                                   IEEE square roots and divides - bit-equal to the CPU checker of the test suite.  1 "reference": what nvcc
                                   makes of the reference's own source - FMA contraction, normalize() by rsqrtf (helper_math.h:
                                   1325-1328), IEEE sqrtf / divide in angle() (volumeRender_kernel.cu:284-287).  2 "fast": flavour 1
                                   with the cosine in angle() as dot * rsqrt(|a|^2 |b|^2) and x / pi as a multiplication.  All three are
                                   held to the 1e-3 bar against the checker and to the measured envelope against the reference
                                   kernel's own images (tests/test_golden_refcuda.py); $VR_ARITHMETIC sets the default of new contexts */
  VR_OPT_PART_MODE = 16,        /* image-space partition (VR_OPT_PART_RANK / _NRANKS): 0 (default) = tiles dealt round-robin, 1 = tile
                                   COLUMNS dealt round-robin (tile tx belongs to rank tx % nranks).  With columns the pixels a rank
                                   owns are, per image plane, runs of tile_w * H contiguous floats: a rank can copy exactly its own
                                   pixels to the host (vr_render*, vr_render_batch_begin write only owned pixels of the host image
                                   when nranks > 1), so the read-back of a frame is spread over the PCIe links of all GPUs */
  VR_OPT_TEXTURE_SINGLE_COPY = 17, /* 1 (default): in texture-unit mode a volume is resident ONCE - when its cudaArray and the empty-space
                                   grids stand, the padded brick storage they were built from is freed (4 instead of 8 GiB per 1024^3 fp32
                                   volume) and rebuilt from the array only when something needs it: vr_update_box (such a volume keeps
                                   both copies from then on), a switch to the load path, new skip grids.  0: keep both copies */
  VR_OPT_VS_SLICES = 18,        /* volume-space passes (vr_render_opacity / vr_render_partial): cut every ray's stretch inside the brick
                                   into this many equal depth slices marched by DIFFERENT warps; 1 = whole rays; 0 (default) =
                                   automatic: frames march whole rays and their shading pass measures its own tail - the part of the
                                   launch during which warps had already run out of work; 30 % or more -> 8 slices from then on, less ->
                                   whole rays (needs VR_OPT_COUNT_SAMPLES, else whole rays); explicit values up to 16; needs the closed-form
                                   empty-space skip and kernels with one common volume geometry, else whole rays.  A launch then ends with its
                                   longest SLICE instead of its longest ray - what matters when a brick's rays are thousands of samples
                                   long (a lone warp shades ~0.3 samples per microsecond).  Where each slice of each pixel starts - the
                                   bit-exact state of the reference's recurrences there - is found once per frame by a small kernel of
                                   the opacity pass and reused by the shading pass, which must follow with the same camera and value.
                                   Every lattice sample is still shaded once, at its bit-exact position; a brick's opacity and colour
                                   are folded from the slices' (1 - prod(1 - a_s); sum), so images agree with whole rays to float
                                   rounding, not bit for bit.  Costs (S-1) * 16 + S * 16 bytes per pixel of scratch */
  VR_OPT_RAY_SLICES = 19,       /* latency mode of single frames (vr_render, vr_render_device, one-frame batches; also on device groups):
                                   1 (default) = the plain march; 2 .. 16 = the frame is rendered as the two passes of the volume-space
                                   mode on the whole volume - an opacity pass, then the shading pass - both cut into that many depth
                                   slices, so that the launch no longer waits for its longest rays (what limits ONE frame spread over
                                   several GPUs).  Costs the opacity pass (about a third of a frame on one GPU: not a throughput option),
                                   agrees with the plain march to float rounding (the colour shares of the slices are summed afterwards).
                                   Scenes the passes do not take - volumes of different sizes, no closed-form empty-space skip,
                                   per-texture addressing - are rendered plainly */
  VR_OPT_COUNT_
} vr_option;
/* Default: 2.  Measured on a B200 against the reference kernel's OWN images (tests/golden/refcuda_*.npy, seven scenes): flavours 1
 * and 2 agree with them to PSNR 112-125 dB, maximum 7.8e-5 ... 3.2e-4 of the image maximum, bit for bit on the scene without
 * shading; flavour 0 - like the CPU checker - to 98-106 dB, maximum 3.6e-4 ... 1.5e-3 and one pixel at 2.5e-2
 * (profiles/r02_arith_ab.md).  Flavour 2 renders the headline workload 12 % faster than flavour 0, 7 % faster than 1. */
#define VR_DEFAULT_ARITHMETIC 2
VR_API vr_status vr_set_option(vr_ctx *ctx, int option, int64_t value);
VR_API vr_status vr_get_option(vr_ctx *ctx, int option, int64_t *value);
/* vr_get_stats is also a synchronisation point: it returns when everything the context has enqueued is complete. */
VR_API vr_status vr_get_stats(vr_ctx *ctx, vr_stats *out);
/* Work of later calls is enqueued on this cudaStream_t (NULL = legacy default stream). */
VR_API vr_status vr_set_stream(vr_ctx *ctx, void *cuda_stream);
VR_API vr_status vr_get_device(vr_ctx *ctx, int *device);

/* Device-side synthetic "ViBE-Z-shaped" volume generator; fills d_out (device
 * pointer, dims[0] fastest) so that 1024^3+ test volumes never cross PCIe.  kind: 0 = brain
 * ellipsoid with value noise, 1 = sparse tubes.  The host twin lives in the test suite. */
VR_API vr_status vr_synth_volume(vr_ctx *ctx, int kind, uint32_t seed, const uint64_t dims[3],
                                 int32_t dtype, void *d_out);
/* Volume.grad (src/matlab/VolumeRender/Volume.m:181-205), i.e. MATLAB's [gx,gy,gz] = gradient(Data):
 * unit spacing, central differences inside, one-sided at the ends; gx = derivative along dim 2,
 * gy along dim 1, gz along dim 3.  `in` may be host or device memory; the three outputs (same
 * dims and dtype as `in`) live in `out_memspace`.  SURVEY 8(f) row 3: saves 3 volume uploads. */
VR_API vr_status vr_volume_gradient(vr_ctx *ctx, const vr_volume *in, int32_t out_memspace, void *gx, void *gy,
                                    void *gz);
/* ---- volume-space mode (sort-last), no reference counterpart -------------------------------
 * For volumes that are sharded over GPUs (README.md:5 "huge data"; the reference renders volumes
 * that do not fit together in separate passes and adds the images in MATLAB, example1.m:85).
 * Each context owns the voxel box [lo, hi) of every volume it syncs after vr_set_region (plus
 * `halo` >= 2 voxels of neighbours; 0 = 4) and marches, on the GLOBAL sample lattice, only the
 * samples inside its box.  One frame = two launches per context with one exchange between them:
 *   1. vr_render_opacity_device: per pixel the opacity this brick alone accumulates (H*W floats);
 *      all-gather these planes (n_bricks x H*W floats, brick-major);
 *   2. vr_render_partial_device: shades the owned samples starting from the opacity of the bricks
 *      in front (so the reference's `sum.w > threshold` termination, volumeRender_kernel.cu:485,
 *      acts exactly where it would in one pass) and writes this brick's ADDITIVE share of the image;
 *   3. vr_compose_partials_device: sums the shares in brick order (pointers may be peer memory).
 * brick_boxes = n_bricks x {lo x,y,z, hi x,y,z} in voxels.  vr_set_region(ctx, NULL, NULL, 0) ends
 * the mode.  Volumes must share the emission volume's dims. */
VR_API vr_status vr_set_region(vr_ctx *ctx, const uint64_t lo[3], const uint64_t hi[3], uint32_t halo);

/* Partial update (extension): rewrites the voxel box [lo, hi) of the RESIDENT volume that `role`
 * reads (0 emission, 1 absorption, 2 reflection, 3..5 gradient x, y, z) from `volume`, which must have the dims and
 * dtype of the resident one and hold the new contents at least inside the box (the rest of it is not read).  Only the
 * z slices of the box cross PCIe.  examples/example3.m:135-161 rewrites a masked half of the volume every frame; the
 * reference re-uploads all of it (volumeRender_kernel.cu:659-672).  The slot takes `volume`'s identity (pointer,
 * timestamp), so a following vr_sync_volumes with a time_last_mem_sync at or after that timestamp moves nothing. */
VR_API vr_status vr_update_box(vr_ctx *ctx, int role, const vr_volume *volume, const uint64_t lo[3], const uint64_t hi[3]);
VR_API vr_status vr_render_opacity_device(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                          const vr_volume *illumination, const float factors[3],
                                          const float element_size_um[3], const uint64_t resolution[2],
                                          const float rotation[9], const float props[3],
                                          float opacity_threshold, const float color[3],
                                          float *d_opacity, void *cuda_stream);
VR_API vr_status vr_render_partial_device(vr_ctx *ctx, const vr_light *lights, size_t n_lights,
                                          const vr_volume *illumination, const float factors[3],
                                          const float element_size_um[3], const uint64_t resolution[2],
                                          const float rotation[9], const float props[3],
                                          float opacity_threshold, const float color[3],
                                          uint32_t n_bricks, uint32_t brick_index,
                                          const uint64_t *brick_boxes, const float *d_opacity_all,
                                          float *d_out_image, void *cuda_stream);
VR_API vr_status vr_compose_partials_device(vr_ctx *ctx, uint32_t n, const float *const *d_partials,
                                            uint64_t n_floats, float *d_out, void *cuda_stream);
/* The voxel box [lo, hi) of vr_synth_volume's volume, dense, lo[0] fastest (a VR_MEM_DEVICE_REGION
 * source when the box is the region plus halo clipped to the volume). */
VR_API vr_status vr_synth_volume_region(vr_ctx *ctx, int kind, uint32_t seed, const uint64_t dims[3],
                                        const uint64_t lo[3], const uint64_t hi[3], int32_t dtype, void *d_out);

/* Sharing one device image between the per-GPU processes of the image-space mode (CUDA IPC over
 * NVLink peer access): the root exports the image it allocated with vr_device_alloc, every other
 * rank opens it and passes the mapped pointer as d_out_image of vr_render_device, so the ray-march
 * kernel stores its tiles straight into the root's HBM. handle = 64 bytes (cudaIpcMemHandle_t). */
#define VR_IPC_HANDLE_BYTES 64
VR_API vr_status vr_ipc_export(vr_ctx *ctx, void *d_ptr, unsigned char handle[VR_IPC_HANDLE_BYTES]);
VR_API vr_status vr_ipc_open(vr_ctx *ctx, const unsigned char handle[VR_IPC_HANDLE_BYTES], void **d_ptr);
VR_API vr_status vr_ipc_close(vr_ctx *ctx, void *d_ptr);
/* ---- image epilogue on the device (done in MATLAB by the reference) -----------
 * Stereo compose of VolumeRender.m:286-306: d_left / d_right are the two eyes rendered at
 * res_wide = [H, W + delta]; mode 0 = RedCyan -> H x W x 3, mode 1 = LeftRightHorizontal -> H x 2W x 3.
 * Sequence normalisation of VolumeRender.m:587-701 in place over n_images frames (global min / max,
 * including the reference's `+ min` shift for negative minima).  Multi-channel combine
 * (example3.m:239, plain addition) is vr_compose_partials_device. */
VR_API vr_status vr_compose_stereo_device(vr_ctx *ctx, const float *d_left, const float *d_right,
                                          const uint64_t res_wide[2], uint64_t delta, int mode,
                                          float *d_out, void *cuda_stream);
VR_API vr_status vr_normalize_sequence_device(vr_ctx *ctx, float *const *d_images, uint32_t n_images,
                                              uint64_t n_floats_each, void *cuda_stream);
VR_API vr_status vr_device_upload(vr_ctx *ctx, void *d_dst, const void *host_src, size_t bytes);
/* Debug: [start,end] globaltimer ns of every resident warp of the last ray-march launch
 * (2 x max_warps uint64, at most 8192 warps); only when $VR_WARP_TIMES was set at vr_create. */
VR_API vr_status vr_debug_warp_times(vr_ctx *ctx, uint64_t *host_pairs, size_t max_warps);
/* Device memory helpers for callers without a CUDA runtime binding (C harness, ctypes). */
VR_API vr_status vr_device_alloc(vr_ctx *ctx, size_t bytes, void **d_ptr);
VR_API vr_status vr_device_free(vr_ctx *ctx, void *d_ptr);
VR_API vr_status vr_device_download(vr_ctx *ctx, void *host_dst, const void *d_src, size_t bytes);
/* Measurement only: the gather ceilings of this device with the march's access pattern (8x4 bundles of
 * parallel rays, one trilinear fetch = 8 taps per step; SURVEY 8(d) "BW_gather from a committed
 * microbenchmark").  kind 0 = eight 4-byte loads per fetch, 1 = eight 16-byte loads, 2 = one tex3D<float>
 * with hardware linear filtering (the reference's fetch, volumeRender_kernel.cu:544-548), 3 = tex3D<float4>,
 * 4 = the taps read from shared memory, where a TMA-staged march would keep the voxel neighbourhood; the staging itself is left out, so the
 * rate is the upper bound of that design.
 * Footprint (2^log2_edge)^3 voxels: 5 -> L1, 8 -> L2, 10 -> HBM.  Returns trilinear fetches per second. */
VR_API vr_status vr_probe_gather(vr_ctx *ctx, int kind, int log2_edge, double *fetches_per_s, double *ms);

#ifdef __cplusplus
}
#endif
#endif /* VR_ABI_H_ */
