/* vrbench.c - plain-C driver of the C ABI (include/vr_abi.h): renders example1-shaped frames of a
 * synthetic ViBE-Z-like volume without MATLAB.  It performs the same call sequence the MATLAB
 * class does per frame (VolumeRender.m:497-583): 'sync_volumes' then 'render'.
 *
 *   vrbench [--size N] [--image W H] [--frames K] [--lights L] [--grad] [--bf16] [--stereo] [--gpus G] [--ray-slices S] [--out file.pfm]
 *
 * --gpus G: the context leads a group of devices 0..G-1 (vr_set_devices): volumes resident on each, tile columns dealt
 * round-robin, the same H x W x 3 image comes back - several GPUs behind the one 'render' call, no torch, no MPI.
 *
 * --ray-slices S: VR_OPT_RAY_SLICES, the latency mode of single frames (every frame here is one 'render' call).
 *
 * Links only against libvr_b200.so (no CPU fallback, no checker code).
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/vr_abi.h"

#define CHECK(call)                                                                      \
  do {                                                                                   \
    vr_status s_ = (call);                                                               \
    if (s_ != VR_OK) {                                                                   \
      fprintf(stderr, "%s: %s: %s\n", #call, vr_status_string(s_), vr_last_error(ctx)); \
      return 1;                                                                          \
    }                                                                                    \
  } while (0)

/* VolumeRender.rotate (VolumeRender.m:239-262): R = R * Rx * Ry * Rz, degrees */
static void mat_mul(double a[9], const double b[9]) {
  double r[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
  memcpy(a, r, sizeof r);
}
static void rotate(double R[9], double al, double be, double ga) {
  const double k = 3.14159265358979323846 / 180.0;
  const double ca = cos(al * k), sa = sin(al * k), cb = cos(be * k), sb = sin(be * k), cg = cos(ga * k), sg = sin(ga * k);
  const double Rx[9] = {1, 0, 0, 0, ca, -sa, 0, sa, ca}, Ry[9] = {cb, 0, sb, 0, 1, 0, -sb, 0, cb}, Rz[9] = {cg, -sg, 0, sg, cg, 0, 0, 0, 1};
  mat_mul(R, Rx), mat_mul(R, Ry), mat_mul(R, Rz);
}
/* flip(RotationMatrix) as MATLAB stores it: column-major single 3x3 (VolumeRender.m:560) */
static void flipped_colmajor(const double R[9], float out[9]) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) out[c * 3 + r] = (float)R[3 * (2 - r) + c];
}

static int write_pfm(const char *path, const float *img, uint64_t H, uint64_t W) {
  FILE *f = fopen(path, "wb");
  if (!f) return 1;
  fprintf(f, "PF\n%llu %llu\n-1.0\n", (unsigned long long)W, (unsigned long long)H);
  for (uint64_t y = H; y-- > 0;) /* PFM rows go bottom-up */
    for (uint64_t x = 0; x < W; ++x)
      for (int c = 0; c < 3; ++c) fwrite(&img[(uint64_t)c * W * H + x * H + y], 4, 1, f); /* H x W x 3 column-major */
  fclose(f);
  return 0;
}

int main(int argc, char **argv) {
  uint64_t n = 256, W = 512, H = 512;
  int frames = 3, nl = 1, grad = 0, bf16 = 0, stereo = 0, gpus = 1, ray_slices = 1;
  const char *out = NULL;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--size") && i + 1 < argc) n = strtoull(argv[++i], NULL, 10);
    else if (!strcmp(argv[i], "--image") && i + 2 < argc) W = strtoull(argv[++i], NULL, 10), H = strtoull(argv[++i], NULL, 10);
    else if (!strcmp(argv[i], "--frames") && i + 1 < argc) frames = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--lights") && i + 1 < argc) nl = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--grad")) grad = 1;
    else if (!strcmp(argv[i], "--bf16")) bf16 = 1;
    else if (!strcmp(argv[i], "--stereo")) stereo = 1;
    else if (!strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--ray-slices") && i + 1 < argc) ray_slices = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--out") && i + 1 < argc) out = argv[++i];
    else { fprintf(stderr, "usage: vrbench [--size N] [--image W H] [--frames K] [--lights L] [--grad] [--bf16] [--stereo] [--gpus G] [--out f.pfm]\n"); return 2; }
  }
  if (nl > 2) nl = 2;
  vr_ctx *ctx = NULL;
  CHECK(vr_create_on_device(&ctx, 0));
  if (gpus > 1) { /* before the first sync: every device of the group receives the volumes */
    int ids[64];
    if (gpus > 64) gpus = 64;
    for (int i = 0; i < gpus; ++i) ids[i] = i;
    CHECK(vr_set_devices(ctx, (uint32_t)gpus, ids));
  }
  if (ray_slices > 1) CHECK(vr_set_option(ctx, VR_OPT_RAY_SLICES, ray_slices));
  const int dt = bf16 ? VR_DTYPE_BF16 : VR_DTYPE_F32;
  const size_t vbytes = (size_t)n * n * n * (bf16 ? 2 : 4);
  const uint64_t dims[3] = {n, n, n};
  void *dE = NULL, *dG[3] = {NULL, NULL, NULL};
  CHECK(vr_device_alloc(ctx, vbytes, &dE));
  CHECK(vr_synth_volume(ctx, 0, 0x5EED01BEu, dims, dt, dE)); /* generated straight into HBM */
  vr_volume E = {dE, {n, n, n}, 1, dt, VR_MEM_DEVICE}, G[3];
  float one = 1.0f;
  vr_volume Re = {&one, {1, 1, 1}, 2, VR_DTYPE_F32, VR_MEM_HOST}; /* default VolumeReflection = Volume(1), VolumeRender.m:131 */
  if (grad) {
    for (int i = 0; i < 3; ++i) CHECK(vr_device_alloc(ctx, vbytes, &dG[i]));
    CHECK(vr_volume_gradient(ctx, &E, VR_MEM_DEVICE, dG[0], dG[1], dG[2]));
    for (int i = 0; i < 3; ++i) { vr_volume g = {dG[i], {n, n, n}, (uint64_t)(3 + i), dt, VR_MEM_DEVICE}; G[i] = g; }
  }
  float *lut = (float *)malloc(64 * 64 * 64 * sizeof(float));
  CHECK(vr_henyey_greenstein(64, 0.8f, lut));
  vr_volume L = {lut, {64, 64, 64}, 9, VR_DTYPE_F32, VR_MEM_HOST};
  const vr_light lights[2] = {{{500, 1000, 550}, {0, 1, 1}}, {{550, 1000, 500}, {1, 0, 1}}}; /* example1.m:36 */
  const float factors[3] = {1.0f, 0.4f, 0.6f}, es[3] = {1, 1, 1}, color[3] = {1, 1, 0};       /* [fEm, fRe, fAb], example1.m:55-58 */
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  rotate(R, 125, 25, 0); /* example1.m:39-42 */
  const float base = stereo ? 0.03f : 0.0f; /* CameraXOffset 0.06 / 2, example3.m:105 */
  const uint64_t delta = stereo ? (uint64_t)floor(base * (double)W / 2.0 + 0.5) : 0; /* VolumeRender.m:277-282 */
  const uint64_t res[2] = {H, W + delta};
  float *img = (float *)malloc((size_t)res[0] * res[1] * 3 * sizeof(float));
  uint64_t stamp = 0; /* TimeLastMemSync: 0 on the first frame -> everything uploads once */
  double tot_ms = 0, tot_samples = 0;
  for (int f = 0; f < frames; ++f) {
    for (int eye = 0; eye < (stereo ? 2 : 1); ++eye) {
      CHECK(vr_sync_volumes(ctx, stamp, &E, &Re, &E, grad ? &G[0] : NULL, grad ? &G[1] : NULL, grad ? &G[2] : NULL));
      stamp = 100; /* later syncs: nothing is newer than the last sync, nothing moves (example2.m) */
      float rot[9];
      flipped_colmajor(R, rot);
      const float props[3] = {stereo ? (eye ? -base : base) : 0.0f, 3.0f, 6.0f};
      CHECK(vr_render(ctx, nl ? lights : NULL, (size_t)nl, nl ? &L : NULL, factors, es, res, rot, props, 0.9f, color, img));
      vr_stats st;
      CHECK(vr_get_stats(ctx, &st));
      printf("frame %d eye %d: kernel %.3f ms, %llu samples marched (%llu shaded), %.2f Gsamples/s, %llu rays hit\n", f, eye, st.kernel_ms,
             (unsigned long long)st.samples_marched, (unsigned long long)st.samples_shaded, st.samples_marched / (st.kernel_ms * 1e6),
             (unsigned long long)st.rays_hit);
      if (f > 0 || frames == 1) tot_ms += st.kernel_ms, tot_samples += (double)st.samples_marched;
    }
    rotate(R, 0, 1, 0); /* example2.m: rotation movie */
  }
  if (tot_ms > 0) printf("total: %.2f Gsamples/s, %.2f frames/s (kernel time)\n", tot_samples / (tot_ms * 1e6), (frames > 1 ? frames - 1 : 1) * 1e3 / tot_ms);
  if (out && write_pfm(out, img, res[0], res[1])) fprintf(stderr, "cannot write %s\n", out);
  char info[4096];
  size_t need = 0;
  CHECK(vr_mem_info(ctx, info, sizeof info, &need));
  fputs(info, stdout);
  for (int i = 0; i < 3; ++i) if (dG[i]) CHECK(vr_device_free(ctx, dG[i]));
  CHECK(vr_device_free(ctx, dE));
  free(img), free(lut);
  CHECK(vr_destroy(ctx));
  return 0;
}
