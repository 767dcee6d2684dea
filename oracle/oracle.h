/* oracle.h - C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a CPU restatement of the reference's ray-march path
 * (raphiniert-com/volume_renderer, src/C/vr/volumeRender_kernel.cu + volumeRender.cpp +
 * mex/render.cpp), used as the parity checker by tests/, by __graft_entry__.smoke() and as the
 * timed CPU baseline of bench.py.  Nothing in the product (volume_renderer_b200/, mex/,
 * harness/) may include, link or call it.
 *
 * Parity pin: the reference ships no tests, golden images or CPU ray march.  The pins are
 * (i) bit-equality of orc_henyey_greenstein / orc_init_render with the reference's own
 * HenyeyGreenstein.cc / volumeRender.cpp compiled into oracle/_ref (tests/test_oracle_ref.py), and
 * (ii) golden images produced on a B200 by the reference's OWN d_render kernel and upload code,
 * compiled through a texture-object shim (oracle/ref_cuda -> oracle/_ref/libref_cuda.so; fixtures
 * tests/golden/refcuda_*.npy, generator tests/golden/make_golden_refcuda.py, test
 * tests/test_golden_refcuda.py).  Against that hardware path - 1.8 fixed-point texture filtering,
 * nvcc's own FMA contraction, approximate rsqrtf/__expf, and a discontinuity of the illumination at
 * NaN angles - this restatement agrees to a median 1e-4 of the image maximum, 98-99.9 % of the pixels
 * within 1e-3, PSNR 74-85 dB, and to 6e-6 where no discontinuity is in play.
 *
 * The call surface mirrors include/vr_abi.h one to one so the same test drives both.
 */
#ifndef ORACLE_H_
#define ORACLE_H_

#include "../include/vr_abi.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

orc_ctx *orc_create(void);
void orc_destroy(orc_ctx *ctx);

/* options (same numbering as vr_option where they overlap) */
enum {
  ORC_OPT_FILTER_FRAC_BITS = 0,    /* 0 = fp32 weights, 8 = documented 1.8 fixed-point weights, 9 = the filter as measured on the B200 texture unit */
  ORC_OPT_ABSORPTION_ROLE_FIX = 3, /* 0 = reference as written */
  ORC_OPT_BORROW_HOST_DATA = 100,  /* 1 = "upload" keeps a pointer instead of a copy (big baselines) */
  ORC_OPT_THREADS = 101,           /* OpenMP threads, 0 = all */
  ORC_OPT_FAST_EXP = 102           /* 0 = expf, 1 = exp2f(x*log2e) like __expf's decomposition (default) */
};
int orc_set_option(orc_ctx *ctx, int option, int64_t value);

int orc_sync_volumes(orc_ctx *ctx, uint64_t time_last_mem_sync, const vr_volume *emission,
                     const vr_volume *reflection, const vr_volume *absorption,
                     const vr_volume *gradient_x, const vr_volume *gradient_y,
                     const vr_volume *gradient_z);

/* window = {x0, x1, y0, y1} (pixels, half-open) restricts the rendered pixels (the rest of the
 * image stays 0), NULL = full frame. */
int orc_render(orc_ctx *ctx, const vr_light *lights, size_t n_lights,
               const vr_volume *illumination, const float factors[3],
               const float element_size_um[3], const uint64_t resolution[2],
               const float rotation[9], const float props[3], float opacity_threshold,
               const float color[3], const uint64_t *window, float *out_image);

/* introspection of the upload state machine (kernel.cu:739-867) */
typedef struct orc_state {
  int32_t idx_emission, idx_absorption, idx_reflection; /* d_idx* (kernel.cu:75-85) */
  int32_t gradient_method;                              /* dc_activeGradientMethod */
  int32_t bound[7];                                     /* texture bound to an array? */
  uint64_t uploads[7];                                  /* cumulative uploads per texture slot */
  uint64_t num_lights;
} orc_state;
void orc_get_state(const orc_ctx *ctx, orc_state *out);

/* counters of the last orc_render */
typedef struct orc_stats {
  uint64_t samples_marched;
  uint64_t rays_hit;
  double seconds;  /* wall time of the pixel loop */
  int32_t threads;
} orc_stats;
void orc_get_stats(const orc_ctx *ctx, orc_stats *out);

/* vr::initRender (volumeRender.cpp:112-156): boxmin/boxmax/tstep for a volume extent. */
void orc_init_render(const uint64_t dims[3], const float element_size_um_xyz[3], float boxmin[3],
                     float boxmax[3], float *tstep);

/* HenyeyGreenstein(n, g) (HenyeyGreenstein.cc:29-96) */
void orc_henyey_greenstein(uint32_t n, float g, float *out);

/* software tex3D (normalized, linear, clamp; kernel.cu:544-548), for unit tests */
float orc_tex3d(const float *data, uint64_t w, uint64_t h, uint64_t d, float x, float y, float z,
                int frac_bits);

/* CPU twins of vr_synth_volume / vr_volume_gradient (bit-identical; MATLAB layout) */
void orc_synth_volume(int kind, uint32_t seed, const uint64_t dims[3], float *out);
void orc_volume_gradient(const float *in, const uint64_t dims[3], float *gx, float *gy, float *gz);

const char *orc_last_error(const orc_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif
