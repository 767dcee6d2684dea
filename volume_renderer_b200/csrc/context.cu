// context.cu - the C ABI of include/vr_abi.h: persistent context, volume residency, render.
//
// Replaces, behind the same command protocol, the reference's host runtime for the ray-march
// path: mm::MManager (src/C/vr/mm/mmanager.hxx), syncWithDevice and friends
// (src/C/vr/volumeRender_kernel.cu:522-867), vr::initRender / vr::render
// (src/C/vr/volumeRender.cpp:112-300) and the argument marshalling of the mex gateway
// (src/C/mex/render.cpp:93-277).  Differences by design: context-local state instead of process
// globals, volumes resident as padded bricks in HBM instead of cudaArrays, one persistent image
// buffer / light block instead of a per-frame cudaMalloc, every CUDA error checked and returned.
// There is no CPU path in here: without a B200 every call fails with VR_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/vr_abi.h"
#include "raymarch_launch.h"
#include "render_setup.h"
#include "vr_types.h"

using namespace vrb;

namespace {

thread_local std::string tl_error;
std::mutex g_registry_mutex;
std::set<const vr_ctx *> g_registry;  // live contexts: lets a stale handle be recognised

// vr::Volume (volumeRender.h:62-71) and its operator== (volumeRender.cpp:24-26)
struct HostVol {
  const void *data = nullptr;
  uint64_t w = 0, h = 0, d = 0;
  uint64_t bytes = 0;
  uint64_t last_update = 0;
  int dtype = VR_DTYPE_F32;
  int memspace = VR_MEM_HOST;
};
inline bool same(const HostVol &a, const HostVol &b) {
  return a.data == b.data && a.last_update == b.last_update && a.bytes == b.bytes;
}
inline size_t elem_size(int dtype) { return dtype == VR_DTYPE_BF16 ? 2 : 4; }
HostVol make_volume(const vr_volume *v) {
  HostVol r;
  if (!v) return r;
  r.data = v->data;
  r.w = v->dims[0], r.h = v->dims[1], r.d = v->dims[2];
  r.dtype = v->dtype;
  r.memspace = v->memspace;
  r.bytes = r.w * r.h * r.d * elem_size(r.dtype);
  r.last_update = v->last_update;
  return r;
}

// One texture slot = what a bound texture reference + its cudaArray was (kernel.cu:90-120).
enum { kStageBufs = 3 };

struct Resident {
  void *storage = nullptr;  // padded (bricked) voxels
  size_t storage_bytes = 0;
  uint32_t *occ = nullptr;  // occupancy bit grid
  size_t occ_bytes = 0;
  int reg[6] = {0, 0, 0, 0, 0, 0};  // owned voxel box (the whole volume unless vr_set_region)
  uint8_t *dist = nullptr;  // distance grid (cells to the nearest non-empty cell)
  size_t dist_bytes = 0;
  int ocx = 0, ocy = 0, ocz = 0;
  int w = 0, h = 0, d = 0;        // global voxel dims
  int ox = 0, oy = 0, oz = 0;     // global padded index -> local padded index
  int lw = 0, lh = 0, ld = 0;     // local padded dims
  int dtype = VR_DTYPE_F32;
  int layout = LAYOUT_LINEAR;
  uint32_t sy = 0, sz = 0;
  int kind = TEX_UNBOUND;
  float cval = 0.0f;
  HostVol src;  // what was uploaded (identity: pointer, timestamp, bytes)
  // a gradient volume that lives as component `ecomp` of the context's interleaved gradient storage
  // (`storage` then points at vr_ctx::g4 and is not owned)
  bool shared = false;
  int estride = 1, ecomp = 0;
  // texture-unit mode (VR_OPT_TEXTURE_UNIT): the same voxels as a w x h x d cudaArray behind a texture object configured
  // like the reference's (kernel.cu:544-548); built lazily from `storage`, dropped whenever `storage` is rewritten
  cudaArray_t array = nullptr;
  cudaTextureObject_t texobj = 0;
  size_t array_bytes = 0;
  bool array_valid = false;
  int a0[3] = {0, 0, 0}, an[3] = {0, 0, 0};  // the voxel box [a0, a0 + an) the array holds: the resident box clipped to the volume
  // texture-unit mode keeps ONE copy: once the array and the empty-space grids exist the padded storage is freed
  // (drop_storage) and rebuilt from the array only when something needs it (restore_storage); a volume that receives
  // partial updates keeps both copies
  bool keep_storage = false;
};

// geometry of the interleaved gradient storage: a gradient volume joins it when it matches
struct G4Geo {
  int w = 0, h = 0, d = 0, lw = 0, lh = 0, ld = 0, ox = 0, oy = 0, oz = 0, dtype = 0, layout = 0;
  uint32_t sy = 0, sz = 0;
  bool operator==(const G4Geo &o) const {
    return w == o.w && h == o.h && d == o.d && lw == o.lw && lh == o.lh && ld == o.ld && ox == o.ox && oy == o.oy && oz == o.oz &&
           dtype == o.dtype && layout == o.layout && sy == o.sy && sz == o.sz;
  }
};

// Everything one launch owns while it is in flight.  The synchronous entry points use slot 0; the pipelined batches
// (vr_render_batch_begin / _end) alternate between the two, so that the frames of batch k leave for the host while the
// kernel of batch k+1 runs.
struct LaunchSlot {
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the ray-march launch (timing)
  cudaEvent_t ev_cost = nullptr;             // the launch's tile costs have reached h_tile_cost
  cudaEvent_t ev_rendered = nullptr;         // launch + its counters complete (the copy stream waits on it)
  cudaEvent_t ev_copied = nullptr;           // the frames of this slot have reached host memory
  unsigned long long *d_stats = nullptr;     // 3 counters
  unsigned long long *h_stats = nullptr;     // pinned
  unsigned int *h_tile_cost = nullptr;       // pinned: measured cost per tile of this slot's last launch
  size_t cost_tiles = 0;                     // capacity in tiles
  uint64_t cost_sig = 0;                     // geometry the host copy belongs to (W, H, tile size)
  bool cost_valid = false;
  float *d_image = nullptr;                  // persistent output image(s) of the host-image entry points
  size_t image_cap = 0;
  bool stats_pending = false, timing_pending = false;
  uint32_t launches = 1;  // kernels of the batch in flight (a latency-mode frame is several)
  bool in_flight = false;                    // a pipelined batch of this slot has not been harvested yet
};

}  // namespace

#ifdef VR_TRACE_PIXEL
static void *g_trace_host = nullptr;
#endif
// what the slice states of a region context (d_slice_state) were computed for
struct SliceStateKey {
  float cam[4][3], box[3], tstep;
  uint64_t W, H;
  int slices, arith;
};

struct vr_ctx {
  uint32_t magic = 0xB200F0A5u;
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cudaEvent_t evu0 = nullptr, evu1 = nullptr;
  LaunchSlot slot[2];
  int cur = 0;                       // slot of the launch being set up
  cudaStream_t copy_stream = nullptr;  // device->host copies of pipelined batches
  bool pipe_open = false;            // between vr_render_batch_begin and vr_render_batch_end
  uint64_t pipe_batches = 0;
  vr_stats pipe_total{};             // counters summed over the batches of the open pipeline
  // a group of contexts driven through this one (vr_set_devices): members[0] is this context itself
  std::vector<vr_ctx *> members;
  vr_ctx *leader = nullptr;          // set in the other members
  bool ever_synced = false;
  cudaEvent_t ev_group = nullptr;    // orders this member's stream against the leader's (device-image renders of a group)
  bool group_peer = false;           // every pair of the group's devices has peer access
  cudaEvent_t ev_order = nullptr;  // orders a caller's stream behind work enqueued on the context's stream
  // mm::MManager state (mmanager.hxx:25-100)
  uint64_t timeLastMemSync = 0;
  HostVol host[S_N];
  bool arr[S_N] = {false, false, false, false, false, false, false};  // ptr_d_volume* != 0
  // device globals of the reference (kernel.cu:55-120), context-local here
  Resident tex[S_N];
  int idxEm = S_EM, idxAb = S_EM /* sic: kernel.cu:80 initialises d_idxAbsorption to emission */, idxRe = S_RE;
  int gradientMethod = 0;
  // Gx, Gy, Gz interleaved as 4-element voxels (vr_types.h: TexDesc::estride): allocated by the first
  // gradient volume uploaded, freed when the last one using it is released
  void *g4 = nullptr;
  size_t g4_bytes = 0;
  int g4_users = 0;
  G4Geo g4geo;
  std::vector<vr_light> lights;  // positions already axis-reversed (render.cpp:167)
  float *d_lights_more = nullptr;
  size_t lights_more_cap = 0;
  // persistent buffers
  void *d_stage[kStageBufs] = {nullptr, nullptr, nullptr};  // device side of the upload staging
  void *h_stage[kStageBufs] = {nullptr, nullptr, nullptr};  // pinned host side
  cudaEvent_t ev_stage[kStageBufs] = {nullptr, nullptr, nullptr};
  size_t stage_cap = 0;
  unsigned int *d_counter = nullptr;      // work counter + padding
  uint32_t *d_tile_order = nullptr;       // owned tiles, heaviest first
  Camera *d_cams = nullptr;               // cameras of a batch
  // measured cost per tile of the previous launch (device counters, pinned host copy, completion event)
  unsigned int *d_tile_cost = nullptr;
  size_t tile_cost_tiles = 0;     // capacity in tiles
  size_t cams_cap = 0;
  unsigned long long *d_warp_times = nullptr;  // debug buffer ($VR_WARP_TIMES=1): 2 x 8192 entries
  size_t tile_order_cap = 0;
  std::vector<uint32_t> h_tile_order;
  // volume-space mode: the voxel sub-box this context owns (region_on == false: everything)
  bool texhw_failed = false;  // VR_OPT_TEXTURE_UNIT = 2: the array copies did not fit; retried after the next upload
  bool region_on = false;
  int64_t rlo[3] = {0, 0, 0}, rhi[3] = {0, 0, 0};
  int halo = 2;
  float *d_boxes = nullptr;  // world-space brick boxes of the last partial render
  // depth slices of the volume-space passes (VR_OPT_VS_SLICES): per-slice opacity planes (written by the opacity pass, read by the
  // shading pass of the same frame) and per-slice colour shares
  float *d_slice_opac = nullptr, *d_slice_img = nullptr;
  float *d_slice_state = nullptr;   // where the slices of every pixel start: written once per frame, by the opacity pass
  size_t slice_opac_cap = 0, slice_img_cap = 0, slice_state_cap = 0;
  SliceStateKey key_of_states{};  // the frame d_slice_state belongs to
  int slices_of_opacity_pass = 0;  // what the planes in d_slice_opac hold (the shading pass must use the same cut)
  float *d_self_opac = nullptr;    // VR_OPT_RAY_SLICES: the opacity plane of a plain frame rendered in two sliced passes
  size_t self_opac_cap = 0;
  int auto_slices = 0;             // VR_OPT_VS_SLICES = 0: the number decided from the first frames' measurements (0 = not yet)
  double last_tail = -1.0;         // tail fraction of the last launch that measured it (-1: none)
  uint64_t auto_slices_sig = 0;    // image and tiling that decision belongs to
  uint64_t pixels_of_opacity_pass = 0;
  size_t boxes_cap = 0;
  // options
  int64_t opt[32];
  // bookkeeping
  vr_stats stats{};
  std::string err;
};

namespace {

vr_status fail(vr_ctx *c, vr_status s, const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  tl_error = buf;
  if (c) c->err = buf;
  return s;
}
#define VR_CUDA(c, call)                                                                                      \
  do {                                                                                                        \
    cudaError_t e_ = (call);                                                                                  \
    if (e_ != cudaSuccess)                                                                                    \
      return fail(c, e_ == cudaErrorMemoryAllocation ? VR_ERR_INSUFFICIENT_VRAM : VR_ERR_CUDA, "%s: %s (%s:%d)", #call, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                                \
  } while (0)
#define VR_TRY(expr)                 \
  do {                               \
    vr_status s_ = (expr);           \
    if (s_ != VR_OK) return s_;      \
  } while (0)

bool live(const vr_ctx *c) {
  if (!c) return false;
  std::lock_guard<std::mutex> g(g_registry_mutex);
  return g_registry.count(c) != 0 && c->magic == 0xB200F0A5u;
}
#define VR_CHECK_HANDLE(c) \
  if (!live(c)) return fail(nullptr, VR_ERR_INVALID_HANDLE, "Handle not valid.")

void release(vr_ctx *c, Resident &r) {
  if (r.shared) {
    if (--c->g4_users == 0) {
      cudaFree(c->g4);
      c->g4 = nullptr, c->g4_bytes = 0;
    }
  } else if (r.storage) {
    cudaFree(r.storage);
  }
  if (r.occ) cudaFree(r.occ);
  if (r.dist) cudaFree(r.dist);
  if (r.texobj) cudaDestroyTextureObject(r.texobj);
  if (r.array) cudaFreeArray(r.array);
  r = Resident();
}

uint64_t resident_bytes(const vr_ctx *c) {
  uint64_t b = 0;
  for (int i = 0; i < S_N; ++i) b += c->tex[i].storage_bytes + c->tex[i].occ_bytes + c->tex[i].dist_bytes + c->tex[i].array_bytes;
  b += c->slot[0].image_cap + c->slot[1].image_cap + kStageBufs * c->stage_cap + c->lights_more_cap + c->g4_bytes;
  b += c->slice_opac_cap + c->slice_img_cap + c->slice_state_cap + c->self_opac_cap;  // scratch of the depth slices (volume-space passes)
  return b;
}

// storage geometry of a padded volume
void geometry(int layout, uint64_t pw, uint64_t ph, uint64_t pd, uint32_t *sy, uint32_t *sz, uint64_t *elems,
              uint64_t *layers, uint64_t *layer_elems) {
  if (layout == LAYOUT_BRICK) {
    const uint64_t nbx = (pw + 7) / 8, nby = (ph + 7) / 8, nbz = (pd + 7) / 8;
    *sy = (uint32_t)(nbx * 512);
    *sz = (uint32_t)(nbx * nby * 512);
    *layers = nbz;
    *layer_elems = nbx * nby * 512;
  } else {
    *sy = (uint32_t)pw;
    *sz = (uint32_t)(pw * ph);
    *layers = pd;
    *layer_elems = pw * ph;
  }
  *elems = *layers * *layer_elems;
}

TexDesc describe(const Resident &r);
vr_status restore_storage(vr_ctx *c, Resident &r, cudaStream_t stream);


// memcpy with a few threads: one core copies ~10 GB/s, the PCIe 5 link takes 50
void parallel_copy(void *dst, const void *src, size_t bytes) {
  unsigned nt = std::thread::hardware_concurrency();
  nt = nt > 8 ? 8 : (nt < 1 ? 1 : nt);
  if (bytes < ((size_t)8 << 20) || nt == 1) {
    memcpy(dst, src, bytes);
    return;
  }
  const size_t per = ((bytes / nt) + 4095) & ~(size_t)4095;
  std::vector<std::thread> th;
  for (unsigned i = 0; i < nt; ++i) {
    const size_t o = (size_t)i * per;
    if (o >= bytes) break;
    const size_t n = o + per < bytes ? per : bytes - o;
    th.emplace_back([=] { memcpy(static_cast<char *>(dst) + o, static_cast<const char *>(src) + o, n); });
  }
  for (auto &t : th) t.join();
}

// A rendered frame into the caller's host memory (volumeRender.cpp:280-281 + the copy loop of render.cpp:267-269).
// MATLAB's mxArray is pageable; the driver stages such a copy itself at ~16 GB/s (3.1 ms for a 2048^2 frame, measured).
// A pinned mirror filled in eight DMA pieces with four host threads moving piece k while piece k+1 is on the wire was
// tried: 2.1 ms at best, but every few frames the copy threads were descheduled on the bench box and a frame took
// 40-77 ms instead of 29 (gpurun_out r4).  A steady 3 ms beats a fragile 2: the plain copy stays.
// Enqueues on `stream` the copy of the pixels THIS context renders: the whole frame for an unpartitioned context; with the
// column partition (VR_OPT_PART_MODE = 1) the owned tile columns - per image plane `rows` runs of tile_w * H contiguous floats
// at a pitch of nranks * tile_w * H floats, one 2-D copy per plane (plus the clipped last column).
vr_status enqueue_download(vr_ctx *c, void *out, const void *d_src, uint64_t W, uint64_t H, cudaStream_t stream) {
  const int64_t nranks = c->opt[VR_OPT_PART_NRANKS], rank = c->opt[VR_OPT_PART_RANK];
  const size_t plane = (size_t)W * H;
  if (nranks <= 1 || c->opt[VR_OPT_PART_MODE] != 1) {
    // unpartitioned, or tiles dealt round-robin: the whole (cleared) image with this context's tiles in it
    VR_CUDA(c, cudaMemcpyAsync(out, d_src, plane * 3 * sizeof(float), cudaMemcpyDeviceToHost, stream));
    return VR_OK;
  }
  const size_t tw = (size_t)c->opt[VR_OPT_TILE_W];
  const size_t ntx = (W + tw - 1) / tw;
  const size_t run = tw * H * sizeof(float), pitch = (size_t)nranks * run;
  // full columns owned: tx = rank, rank + nranks, ... with (tx + 1) * tw <= W; a clipped last column is copied on its own
  size_t full = 0, clipped_tx = (size_t)-1;
  for (size_t tx = (size_t)rank; tx < ntx; tx += (size_t)nranks) {
    if ((tx + 1) * tw <= W) ++full;
    else clipped_tx = tx;
  }
  for (int p = 0; p < 3; ++p) {
    const size_t base = (size_t)p * plane * sizeof(float);
    const char *src = static_cast<const char *>(d_src) + base;
    char *dst = static_cast<char *>(out) + base;
    if (full) VR_CUDA(c, cudaMemcpy2DAsync(dst + (size_t)rank * run, pitch, src + (size_t)rank * run, pitch, run, full, cudaMemcpyDeviceToHost, stream));
    if (clipped_tx != (size_t)-1) {
      const size_t o = clipped_tx * run, n = (W - clipped_tx * tw) * H * sizeof(float);
      VR_CUDA(c, cudaMemcpyAsync(dst + o, src + o, n, cudaMemcpyDeviceToHost, stream));
    }
  }
  return VR_OK;
}
// the same, and wait for it (the synchronous 'render')
vr_status download_frame(vr_ctx *c, void *out, const void *d_src, size_t bytes) {
  VR_CUDA(c, cudaMemcpyAsync(out, d_src, bytes, cudaMemcpyDeviceToHost, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  return VR_OK;
}

// staging of host sources: kStageBufs buffers in pinned host memory and as many on the device, so
// that the CPU copy of slab k+1 (pageable -> pinned), the DMA of slab k and its repack overlap
vr_status ensure_stage(vr_ctx *c, size_t bytes) {
  if (c->stage_cap >= bytes) return VR_OK;
  for (int i = 0; i < kStageBufs; ++i) {
    if (c->d_stage[i]) cudaFree(c->d_stage[i]);
    if (c->h_stage[i]) cudaFreeHost(c->h_stage[i]);
    c->d_stage[i] = c->h_stage[i] = nullptr;
  }
  c->stage_cap = 0;
  for (int i = 0; i < kStageBufs; ++i) {
    VR_CUDA(c, cudaMalloc(&c->d_stage[i], bytes));
    VR_CUDA(c, cudaMallocHost(&c->h_stage[i], bytes));
  }
  c->stage_cap = bytes;
  return VR_OK;
}

// Writes the voxels of `v` that lie in the voxel box `box` = {lo xyz, hi xyz} into the resident
// storage of `r` (whose geometry is already set): the whole volume at upload time, a sub-box for
// vr_update_box.  Host sources are streamed in slabs of whole destination layers (H2D of the source
// slices a slab needs, then a repack kernel that writes the padded, bricked storage); only the z
// slices inside the box cross PCIe.
vr_status fill_resident(vr_ctx *c, Resident &r, const HostVol &v, const int box[6]) {
  r.array_valid = false;  // the texture-unit copy (if any) is refilled from the storage before the next launch that uses it
  c->texhw_failed = false;
  const int64_t dims[3] = {r.w, r.h, r.d};
  const int64_t g0[3] = {r.ox - 1, r.oy - 1, r.oz - 1};  // global voxel of local padded index 0
  const size_t es = elem_size(v.dtype);
  uint64_t elems, layers, layer_elems;
  uint32_t sy, sz;
  geometry(r.layout, (uint64_t)r.lw, (uint64_t)r.lh, (uint64_t)r.ld, &sy, &sz, &elems, &layers, &layer_elems);
  RepackArgs a;
  a.src_surface = 0;
  a.dst = r.storage;
  a.w = r.w, a.h = r.h, a.d = r.d;
  a.lw = r.lw, a.lh = r.lh, a.ld = r.ld;
  a.gx0 = (int)g0[0], a.gy0 = (int)g0[1], a.gz0 = (int)g0[2];
  a.sx0 = 0, a.sy0 = 0, a.sw = r.w, a.sh = r.h;
  a.sy = sy, a.sz = sz;
  a.estride = r.estride, a.ecomp = r.ecomp;
  a.layout = r.layout;
  a.src_dtype = a.dst_dtype = v.dtype;
  for (int i = 0; i < 6; ++i) a.box[i] = box[i];
  const uint64_t zper = r.layout == LAYOUT_BRICK ? 8 : 1;  // padded Z planes per layer
  // local padded planes that hold a voxel of the box (the planes below / above the volume replicate its faces)
  const int64_t Zlo = box[2] <= 0 ? 0 : std::max<int64_t>(box[2] - g0[2], 0);
  const int64_t Zhi = box[5] >= dims[2] ? (int64_t)r.ld : std::min<int64_t>(box[5] - g0[2], r.ld);
  if (Zhi <= Zlo) return VR_OK;
  const uint64_t lay0 = (uint64_t)Zlo / zper, lay1 = ((uint64_t)Zhi + zper - 1) / zper;
  const size_t slice_bytes = (size_t)v.w * v.h * es;
  if (v.memspace == VR_MEM_DEVICE_REGION) {
    // the caller's device buffer holds exactly the resident box (owned region plus halo, clipped to the volume)
    int64_t b0[3], b1[3];
    const int64_t lext[3] = {r.lw, r.lh, r.ld};
    for (int i = 0; i < 3; ++i) b0[i] = g0[i] < 0 ? 0 : g0[i], b1[i] = g0[i] + lext[i] > dims[i] ? dims[i] : g0[i] + lext[i];
    a.src = v.data;
    a.sx0 = (int)b0[0], a.sy0 = (int)b0[1], a.sw = (int)(b1[0] - b0[0]), a.sh = (int)(b1[1] - b0[1]);
    a.src_z0 = (int)b0[2];
    a.o0 = lay0 * layer_elems, a.o1 = lay1 * layer_elems;
    VR_CUDA(c, launch_repack(a, c->stream));
  } else if (v.memspace == VR_MEM_DEVICE) {
    a.src = v.data;
    a.src_z0 = 0;
    a.o0 = lay0 * layer_elems, a.o1 = lay1 * layer_elems;
    VR_CUDA(c, launch_repack(a, c->stream));
  } else {
    // slabs of ~64 MiB of source slices through kStageBufs staging buffers
    const size_t budget = (size_t)64 << 20;
    uint64_t slab_layers = budget / (slice_bytes * zper);
    if (slab_layers < 1) slab_layers = 1;
    if (slab_layers > lay1 - lay0) slab_layers = lay1 - lay0;
    VR_TRY(ensure_stage(c, (size_t)(slab_layers * zper + 1) * slice_bytes));
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, v.data) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();  // an unregistered host pointer is reported as an error by older drivers
    auto clampz = [&](int64_t z) { return z < box[2] ? (int64_t)box[2] : (z > box[5] - 1 ? (int64_t)box[5] - 1 : z); };
    int k = 0;
    for (uint64_t l0 = lay0; l0 < lay1; l0 += slab_layers, ++k) {
      const uint64_t l1 = l0 + slab_layers < lay1 ? l0 + slab_layers : lay1;
      // padded planes [l0*zper, l1*zper) hold global voxel z = clamp(Z + gz0); only slices inside the box are read
      const int64_t z0 = clampz((int64_t)(l0 * zper) + g0[2]), z1 = clampz((int64_t)(l1 * zper) - 1 + g0[2]);
      const size_t nbytes = (size_t)(z1 - z0 + 1) * slice_bytes;
      const int bi = k % kStageBufs;
      // the DMA that last read this pinned buffer - a slab of this call or of the PREVIOUS upload (k restarts at 0 on
      // every call and nothing else orders the CPU copy behind that DMA) - must have finished before the CPU rewrites
      // it; on a never-recorded event this returns at once
      VR_CUDA(c, cudaEventSynchronize(c->ev_stage[bi]));
      const char *src = static_cast<const char *>(v.data) + (size_t)z0 * slice_bytes;
      if (!pinned) {
        parallel_copy(c->h_stage[bi], src, nbytes);
        src = static_cast<const char *>(c->h_stage[bi]);
      }
      VR_CUDA(c, cudaMemcpyAsync(c->d_stage[bi], src, nbytes, cudaMemcpyHostToDevice, c->stream));
      a.src = c->d_stage[bi];
      a.src_z0 = (int)z0;
      a.o0 = l0 * layer_elems, a.o1 = l1 * layer_elems;
      VR_CUDA(c, launch_repack(a, c->stream));
      VR_CUDA(c, cudaEventRecord(c->ev_stage[bi], c->stream));
      c->stats.bytes_uploaded += nbytes;
    }
    // the caller's memory is borrowed for the duration of the call only (MATLAB owns it): a pinned
    // source is read by the DMA engine itself, so wait for the last copies
    if (pinned) VR_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return VR_OK;
}

// createTextureFromVolume (kernel.cu:522-554): make `v` resident in slot `slot`.  The volume is
// streamed in slabs of whole destination layers: H2D of the source slices a slab needs, then a
// repack kernel that writes the padded (bricked) storage.  A 1x1x1 volume becomes a constant.
vr_status upload(vr_ctx *c, int slot, const HostVol &v) {
  Resident &r = c->tex[slot];
  release(c, r);
  c->arr[slot] = true;
  if (!v.data || v.w == 0 || v.h == 0 || v.d == 0) return VR_OK;  // stays unbound, reads 0
  if (v.dtype != VR_DTYPE_F32 && v.dtype != VR_DTYPE_BF16) return fail(c, VR_ERR_INVALID_ARGUMENT, "unknown volume dtype %d", v.dtype);
  if (v.w > 0x3fffffff || v.h > 0x3fffffff || v.d > 0x3fffffff) return fail(c, VR_ERR_UNSUPPORTED, "volume dimension too large");
  c->stats.uploads++;
  if (v.w == 1 && v.h == 1 && v.d == 1) {
    uint32_t raw = 0;
    if (v.memspace == VR_MEM_DEVICE) {
      VR_CUDA(c, cudaMemcpyAsync(&raw, v.data, elem_size(v.dtype), cudaMemcpyDeviceToHost, c->stream));
      VR_CUDA(c, cudaStreamSynchronize(c->stream));
    } else {
      memcpy(&raw, v.data, elem_size(v.dtype));
    }
    if (v.dtype == VR_DTYPE_BF16) raw <<= 16;
    memcpy(&r.cval, &raw, 4);
    r.kind = TEX_CONST;
    r.w = r.h = r.d = 1;
    r.dtype = v.dtype;
    r.src = v;
    c->stats.bytes_uploaded += elem_size(v.dtype);
    return VR_OK;
  }
  const int layout = slot == S_LIGHT ? LAYOUT_LINEAR : (c->opt[VR_OPT_BRICK_LAYOUT] ? LAYOUT_BRICK : LAYOUT_LINEAR);
  // what becomes resident: the whole volume padded by one replicated voxel, or (volume-space mode)
  // the owned voxel box [rlo, rhi) plus `halo` voxels of real neighbours (replicated at the volume's faces)
  const bool region = c->region_on && slot != S_LIGHT;
  const int64_t dims[3] = {(int64_t)v.w, (int64_t)v.h, (int64_t)v.d};
  int64_t lo[3] = {0, 0, 0}, hi[3] = {dims[0], dims[1], dims[2]};
  int halo = 1;
  if (region) {
    halo = c->halo;
    for (int i = 0; i < 3; ++i) {
      lo[i] = c->rlo[i], hi[i] = c->rhi[i];
      if (lo[i] < 0 || hi[i] > dims[i] || lo[i] >= hi[i]) return fail(c, VR_ERR_INVALID_ARGUMENT, "region [%lld,%lld) does not fit axis %d of a %lld-voxel volume", (long long)lo[i], (long long)hi[i], i, (long long)dims[i]);
    }
  }
  const int64_t g0[3] = {lo[0] - halo, lo[1] - halo, lo[2] - halo};  // global voxel of local padded index 0
  const uint64_t pw = (uint64_t)(hi[0] - lo[0] + 2 * halo), ph = (uint64_t)(hi[1] - lo[1] + 2 * halo), pd = (uint64_t)(hi[2] - lo[2] + 2 * halo);
  uint64_t elems, layers, layer_elems;
  uint32_t sy, sz;
  geometry(layout, pw, ph, pd, &sy, &sz, &elems, &layers, &layer_elems);
  if (layer_elems > 0xffffffffull) return fail(c, VR_ERR_UNSUPPORTED, "volume slice too large for 32-bit in-layer offsets");
  const size_t es = elem_size(v.dtype);
  size_t free_b = 0, total_b = 0;
  VR_CUDA(c, cudaMemGetInfo(&free_b, &total_b));
  if (free_b < elems * es) {  // MManager::checkFreeDeviceMemory (mmanager.hxx:144-173)
    return fail(c, VR_ERR_INSUFFICIENT_VRAM,
                "insufficient free VRAM!\n\tTotal Memory (MB): \t%zu\n\tFree Memory (MB): \t%zu\n\tRequired memory (MB): \t%zu\n",
                total_b >> 20, free_b >> 20, (size_t)(elems * es) >> 20);
  }
  r.w = (int)v.w, r.h = (int)v.h, r.d = (int)v.d;
  r.ox = (int)(1 + g0[0]), r.oy = (int)(1 + g0[1]), r.oz = (int)(1 + g0[2]);
  r.lw = (int)pw, r.lh = (int)ph, r.ld = (int)pd;
  r.dtype = v.dtype;
  r.layout = layout;
  r.sy = sy, r.sz = sz;
  if (slot >= S_DX && slot <= S_DZ && c->opt[VR_OPT_GRADIENT_INTERLEAVE] && elems <= 0xffffffffull) {
    // lookupGradient reads the three gradient volumes at the same taps: they share one storage of
    // 4-element voxels, this volume is component slot - S_DX of it
    G4Geo geo;
    geo.w = r.w, geo.h = r.h, geo.d = r.d, geo.lw = r.lw, geo.lh = r.lh, geo.ld = r.ld, geo.ox = r.ox, geo.oy = r.oy, geo.oz = r.oz;
    geo.dtype = r.dtype, geo.layout = r.layout, geo.sy = sy, geo.sz = sz;
    if (!c->g4 && free_b >= elems * es * 4) {
      VR_CUDA(c, cudaMalloc(&c->g4, elems * es * 4));
      VR_CUDA(c, cudaMemsetAsync(c->g4, 0, elems * es * 4, c->stream));
      c->g4_bytes = elems * es * 4;
      c->g4geo = geo;
    }
    if (c->g4 && c->g4geo == geo) {
      r.storage = c->g4;
      r.shared = true;
      r.estride = 4, r.ecomp = slot - S_DX;
      c->g4_users++;
    }
  }
  if (!r.shared) {
    VR_CUDA(c, cudaMalloc(&r.storage, elems * es));
    r.storage_bytes = elems * es;
  }

  if (v.memspace == VR_MEM_DEVICE_REGION && !region) return fail(c, VR_ERR_INVALID_ARGUMENT, "a region buffer needs vr_set_region first");
  const int full[6] = {0, 0, 0, r.w, r.h, r.d};
  VR_TRY(fill_resident(c, r, v, full));
  r.kind = TEX_VOLUME;
  r.src = v;
  for (int i = 0; i < 3; ++i) r.reg[i] = (int)lo[i], r.reg[3 + i] = (int)hi[i];
  // the empty-space grids are built on first use as the absorption role (ensure_skip_grids)
  return VR_OK;
}

// Occupancy bits + distance bytes of a resident volume, built the first time it serves as the
// absorption role (only that texture decides which samples are empty): one streaming pass over
// the bricks, a pass over the cells, three passes of the distance transform.
vr_status ensure_skip_grids(vr_ctx *c, Resident &r, cudaStream_t stream) {
  if (r.kind != TEX_VOLUME || (r.occ && r.dist)) return VR_OK;
  VR_TRY(restore_storage(c, r, stream));  // the grids are built from the padded storage
  r.ocx = (r.w + 7) / 8, r.ocy = (r.h + 7) / 8, r.ocz = (r.d + 7) / 8;
  const uint64_t ncell = (uint64_t)r.ocx * r.ocy * r.ocz;
  r.occ_bytes = (size_t)((ncell + 31) / 32) * 4;
  VR_CUDA(c, cudaMalloc(&r.occ, r.occ_bytes));
  VR_CUDA(c, cudaMemsetAsync(r.occ, 0, r.occ_bytes, stream));
  const TexDesc t = describe(r);
  VR_CUDA(c, launch_occupancy(t, r.layout, r.dtype, r.occ, r.ocx, r.ocy, r.ocz, 2, r.reg, stream));
  r.dist_bytes = (size_t)ncell;
  VR_CUDA(c, cudaMalloc(&r.dist, r.dist_bytes));
  uint8_t *tmp = nullptr;
  VR_CUDA(c, cudaMallocAsync(&tmp, r.dist_bytes, stream));
  VR_CUDA(c, launch_distance(r.occ, r.dist, tmp, r.ocx, r.ocy, r.ocz, stream));
  VR_CUDA(c, cudaFreeAsync(tmp, stream));
  return VR_OK;
}

// Texture-unit mode: the volume as a cudaArray behind a texture object with the reference's settings
// (createTextureFromVolume, kernel.cu:522-554: normalized coordinates, linear filter, clamp, element type), filled from
// the resident padded storage - one path whatever the source was (host, device, partial update, interleaved component).
// The voxel box a context holds of a volume: everything for a whole volume, the owned region plus halo clipped to the
// volume for a region context (vr_set_region).
void array_box(const Resident &r, int a0[3], int an[3]) {
  const int g0[3] = {r.ox - 1, r.oy - 1, r.oz - 1}, lext[3] = {r.lw, r.lh, r.ld}, dims[3] = {r.w, r.h, r.d};
  for (int i = 0; i < 3; ++i) {
    a0[i] = g0[i] < 0 ? 0 : g0[i];
    const int hi = g0[i] + lext[i] > dims[i] ? dims[i] : g0[i] + lext[i];
    an[i] = hi - a0[i];
  }
}
bool array_eligible(const Resident &r) {
  if (r.kind != TEX_VOLUME) return true;  // constants and unbound slots need no array
  int a0[3], an[3];
  array_box(r, a0, an);
  return (r.dtype == VR_DTYPE_F32 || r.dtype == VR_DTYPE_BF16) && an[0] > 0 && an[1] > 0 && an[2] > 0 && an[0] <= 16384 && an[1] <= 16384 && an[2] <= 16384;
}
void drop_array(Resident &r) {
  if (r.texobj) cudaDestroyTextureObject(r.texobj);
  if (r.array) cudaFreeArray(r.array);
  r.texobj = 0, r.array = nullptr, r.array_bytes = 0, r.array_valid = false;
}
// fp32 arrays whatever the storage type: bf16 voxels widen exactly, and the texture unit has no bf16 element type (a half-float
// array would round them).  A bf16 volume in texture-unit mode therefore holds 4 bytes per voxel once its storage is dropped.
vr_status ensure_array(vr_ctx *c, Resident &r, cudaStream_t stream) {
  if (r.kind != TEX_VOLUME || r.array_valid) return VR_OK;
  if (!array_eligible(r)) return fail(c, VR_ERR_UNSUPPORTED, "texture-unit mode: a 3D array holds at most 16384 voxels per axis");
  if (!r.array) {
    array_box(r, r.a0, r.an);
    const cudaChannelFormatDesc cd = cudaCreateChannelDesc<float>();
    VR_CUDA(c, cudaMalloc3DArray(&r.array, &cd, make_cudaExtent((size_t)r.an[0], (size_t)r.an[1], (size_t)r.an[2]), cudaArraySurfaceLoadStore));
    r.array_bytes = (size_t)r.an[0] * r.an[1] * r.an[2] * sizeof(float);
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = r.array;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;  // kernel.cu:546-548
    td.filterMode = cudaFilterModeLinear;                                                // kernel.cu:545
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 1;                                                             // kernel.cu:544
    VR_CUDA(c, cudaCreateTextureObject(&r.texobj, &rd, &td, nullptr));
  }
  cudaResourceDesc rd = {};
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = r.array;
  cudaSurfaceObject_t surf = 0;
  VR_CUDA(c, cudaCreateSurfaceObject(&surf, &rd));
  const cudaError_t e = launch_fill_array(describe(r), r.layout, r.dtype, r.a0, r.an, (unsigned long long)surf, stream);
  // the kernel has the handle by value; destroying the object after the launch is ordered behind it by the stream sync
  if (e == cudaSuccess) cudaStreamSynchronize(stream);
  cudaDestroySurfaceObject(surf);
  VR_CUDA(c, e);
  r.array_valid = true;
  return VR_OK;
}

// ---- one copy per volume in texture-unit mode ----------------------------------------------------------------
// The padded storage feeds the skip grids, the array fill, partial updates and the load kernels - none of which a
// texture-unit render needs once the array and the grids stand.  Dropping it halves the resident bytes of the default mode
// (8 -> 4 GiB per 1024^3 fp32 volume), so a scene that fits once into the 180 GB stays on the fast path.
void drop_storage(vr_ctx *c, Resident &r) {
  if (r.kind != TEX_VOLUME || !r.storage || !r.array_valid || r.keep_storage) return;
  if (r.shared) {
    // the interleaved gradient storage goes when all of its components have their arrays
    for (int s = S_DX; s <= S_DZ; ++s) {
      const Resident &q = c->tex[s];
      if (q.shared && q.storage && (!q.array_valid || q.keep_storage)) return;
    }
    cudaFree(c->g4);
    c->g4 = nullptr, c->g4_bytes = 0;
    for (int s = S_DX; s <= S_DZ; ++s)
      if (c->tex[s].shared) c->tex[s].storage = nullptr;
    return;
  }
  cudaFree(r.storage);
  r.storage = nullptr, r.storage_bytes = 0;
}
// Rebuilds a dropped storage from the volume's array: the very repack of an upload, its source read through a surface,
// so the storage holds the bits an upload would have written.
vr_status restore_storage(vr_ctx *c, Resident &r, cudaStream_t stream) {
  if (r.kind != TEX_VOLUME || r.storage) return VR_OK;
  if (!r.array || !r.array_valid) return fail(c, VR_ERR_CUDA, "internal: a volume has neither storage nor a valid array");
  uint64_t elems, layers, layer_elems;
  uint32_t sy, sz;
  geometry(r.layout, (uint64_t)r.lw, (uint64_t)r.lh, (uint64_t)r.ld, &sy, &sz, &elems, &layers, &layer_elems);
  const size_t es = elem_size(r.dtype);
  if (r.shared) {
    G4Geo geo;
    geo.w = r.w, geo.h = r.h, geo.d = r.d, geo.lw = r.lw, geo.lh = r.lh, geo.ld = r.ld, geo.ox = r.ox, geo.oy = r.oy, geo.oz = r.oz;
    geo.dtype = r.dtype, geo.layout = r.layout, geo.sy = sy, geo.sz = sz;
    if (!c->g4) {
      VR_CUDA(c, cudaMalloc(&c->g4, elems * es * 4));
      VR_CUDA(c, cudaMemsetAsync(c->g4, 0, elems * es * 4, stream));
      c->g4_bytes = elems * es * 4;
      c->g4geo = geo;
    }
    if (c->g4geo == geo) {
      r.storage = c->g4;
    } else {  // the interleaved storage was re-made for another geometry meanwhile: this component lives on its own from now on
      r.shared = false, r.estride = 1, r.ecomp = 0;
      c->g4_users--;
    }
  }
  if (!r.shared) {
    VR_CUDA(c, cudaMalloc(&r.storage, elems * es));
    r.storage_bytes = elems * es;
  }
  cudaResourceDesc rd = {};
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = r.array;
  cudaSurfaceObject_t surf = 0;
  VR_CUDA(c, cudaCreateSurfaceObject(&surf, &rd));
  RepackArgs a;
  a.dst = r.storage, a.src = nullptr, a.src_surface = (unsigned long long)surf;
  a.w = r.w, a.h = r.h, a.d = r.d, a.lw = r.lw, a.lh = r.lh, a.ld = r.ld;
  a.gx0 = r.ox - 1, a.gy0 = r.oy - 1, a.gz0 = r.oz - 1;
  a.sx0 = r.a0[0], a.sy0 = r.a0[1], a.src_z0 = r.a0[2], a.sw = r.an[0], a.sh = r.an[1];  // the array's origin in the volume
  a.o0 = 0, a.o1 = elems;
  a.sy = sy, a.sz = sz, a.estride = r.estride, a.ecomp = r.ecomp;
  a.box[0] = a.box[1] = a.box[2] = 0, a.box[3] = r.w, a.box[4] = r.h, a.box[5] = r.d;
  a.layout = r.layout, a.src_dtype = a.dst_dtype = r.dtype;
  cudaError_t e = launch_repack(a, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);  // the kernel holds the surface handle by value
  cudaDestroySurfaceObject(surf);
  VR_CUDA(c, e);
  return VR_OK;
}

// syncVolume (kernel.cu:659-672): whatever the flag says, the old array is released and the
// volume is uploaded again.
vr_status syncVolume(vr_ctx *c, int slot, const HostVol &v) { return upload(c, slot, v); }

// The reference re-uploads the LUT on every 'render' and the three gradient volumes on every
// 'sync_volumes' (syncVolume frees and re-creates whatever its flag says, kernel.cu:659-672 +
// :522-541).  The data does not change between those calls unless the volume was stamped, so
// the copy is elided when the slot already holds exactly this (pointer, timestamp, size).
bool slot_is_current(const vr_ctx *c, int slot, const HostVol &v) {
  const Resident &r = c->tex[slot];
  return c->arr[slot] && r.kind != TEX_UNBOUND && same(r.src, v) && r.src.w == v.w && r.src.h == v.h && r.src.d == v.d &&
         r.src.dtype == v.dtype && v.last_update != 0;
}
vr_status syncVolumeCached(vr_ctx *c, int slot, const HostVol &v) {
  if (slot_is_current(c, slot, v)) return VR_OK;
  return upload(c, slot, v);
}

// referenceTexture (kernel.cu:631-648): unbind the slot's texture, drop its array (the reference
// leaks it; freed here) and redirect the role index.
void referenceTexture(vr_ctx *c, int slot, int *idx, int assigned) {
  if (c->arr[slot]) {
    release(c, c->tex[slot]);
    c->arr[slot] = false;
  }
  *idx = assigned;
}

// syncWithDevice (kernel.cu:739-867): per-role dirty test, alias test, order Em -> Ab -> Re.
vr_status syncWithDevice(vr_ctx *c) {
  const HostVol &Em = c->host[S_EM], &Ab = c->host[S_AB], &Re = c->host[S_RE];
  const uint64_t T = c->timeLastMemSync;
  const bool simEmAb = same(Em, Ab), simEmRe = same(Em, Re), simAbRe = same(Ab, Re);
  const bool reqEm = Em.last_update > T || T == 0;
  const bool reqAb = Ab.last_update > T || T == 0;
  const bool reqRe = Re.last_update > T || T == 0;
  bool updEm = false, updAb = false, updRe = false;
  if (reqEm) {
    VR_TRY(syncVolume(c, S_EM, Em));
    updEm = true;
    if (simEmRe && !updRe) referenceTexture(c, S_RE, &c->idxRe, S_EM), updRe = true;
    if (simEmAb && !updAb) referenceTexture(c, S_AB, &c->idxAb, S_EM), updAb = true;
  }
  if (reqAb) {
    if (!updAb) {
      VR_TRY(syncVolume(c, S_AB, Ab));
      updAb = true;
    }
    if (simAbRe && !updRe) referenceTexture(c, S_RE, &c->idxRe, S_AB), updRe = true;
    if (simEmAb && !updEm) {
      // kernel.cu:818-819 hands tex_emission together with the ABSORPTION array pointer: the
      // emission texture is unbound when an absorption array exists, that pointer is dropped,
      // and the emission role is redirected to the absorption texture.
      if (c->arr[S_AB]) release(c, c->tex[S_EM]);
      c->arr[S_AB] = false;
      c->idxEm = S_AB;
      updEm = true;
    }
  }
  if (reqRe) {
    if (!updRe) {
      VR_TRY(syncVolume(c, S_RE, Re));
      updRe = true;
    }
    if (simAbRe && !updAb) referenceTexture(c, S_AB, &c->idxAb, S_RE), updAb = true;
    if (simEmAb && !updEm) referenceTexture(c, S_EM, &c->idxEm, S_RE), updEm = true;
  }
  return VR_OK;
}

// MManager::resetGradients (mmanager.hxx:206-213): only the host descriptors are cleared; the
// device arrays of live gradient volumes stay (inverted null test) - here they are kept too, so
// that the sampled data matches; they are released when the context dies or they are replaced.
void resetGradients(vr_ctx *c) {
  c->host[S_DX] = HostVol();
  c->host[S_DY] = HostVol();
  c->host[S_DZ] = HostVol();
}

// MManager::sync (mmanager.hxx:178-201)
vr_status mm_sync(vr_ctx *c) {
  if ((c->host[S_DX].last_update == 0 && c->arr[S_DX]) || (c->host[S_DY].last_update == 0 && c->arr[S_DY]) ||
      (c->host[S_DZ].last_update == 0 && c->arr[S_DZ]))
    resetGradients(c);
  VR_TRY(syncWithDevice(c));
  if (c->host[S_DX].last_update != 0 && c->host[S_DY].last_update != 0 && c->host[S_DZ].last_update != 0) {
    // setGradientTextures (kernel.cu:703-722)
    if (!slot_is_current(c, S_DX, c->host[S_DX]) && !slot_is_current(c, S_DY, c->host[S_DY]) && !slot_is_current(c, S_DZ, c->host[S_DZ])) {
      // all three are replaced: drop them together so that the interleaved storage is re-made for the new geometry
      for (int s = S_DX; s <= S_DZ; ++s) release(c, c->tex[s]);
    }
    VR_TRY(syncVolumeCached(c, S_DX, c->host[S_DX]));
    VR_TRY(syncVolumeCached(c, S_DY, c->host[S_DY]));
    VR_TRY(syncVolumeCached(c, S_DZ, c->host[S_DZ]));
    c->gradientMethod = 1;
  }
  return VR_OK;
}

// MManager::getRequiredMemory (mmanager.hxx:110-134)
uint64_t required_memory(const vr_ctx *c) {
  const HostVol &Em = c->host[S_EM], &Ab = c->host[S_AB], &Re = c->host[S_RE];
  uint64_t need = Em.bytes;
  if (!same(Em, Ab) && !same(Re, Ab)) need += Ab.bytes;
  if (!same(Em, Re) && !same(Re, Ab)) need += Re.bytes;
  need += c->host[S_DX].bytes + c->host[S_DY].bytes + c->host[S_DZ].bytes;
  return need;
}

vr_status check_free(vr_ctx *c, uint64_t need) {
  size_t free_b = 0, total_b = 0;
  VR_CUDA(c, cudaMemGetInfo(&free_b, &total_b));
  if (free_b < need)
    return fail(c, VR_ERR_INSUFFICIENT_VRAM,
                "insufficient free VRAM!\n\tTotal Memory (MB): \t%zu\n\tFree Memory (MB): \t%zu\n\tRequired memory (MB): \t%zu\n",
                total_b >> 20, free_b >> 20, (size_t)need >> 20);
  return VR_OK;
}

TexDesc describe(const Resident &r) {
  TexDesc t{};
  t.p = r.storage;
  t.w = r.w, t.h = r.h, t.d = r.d;
  t.fw = (float)r.w, t.fh = (float)r.h, t.fd = (float)r.d;
  t.sy = r.sy, t.sz = r.sz;
  t.kind = r.kind;
  t.cval = r.cval;
  t.ox = r.ox, t.oy = r.oy, t.oz = r.oz;
  t.lw = r.lw, t.lh = r.lh, t.ld = r.ld;
  t.estride = r.estride, t.ecomp = r.ecomp;
  t.tex = r.array_valid ? (unsigned long long)r.texobj : 0ull;
  return t;
}

vr_status set_device(vr_ctx *c) {
  VR_CUDA(c, cudaSetDevice(c->device));
  return VR_OK;
}

// Counters and timing of a finished launch of slot `S` (the stream has been synchronised or S.ev_rendered waited for).
void read_slot(vr_ctx *c, LaunchSlot &S, vr_stats &into, bool accumulate) {
  if (S.timing_pending) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, S.ev0, S.ev1) == cudaSuccess) into.kernel_ms = accumulate ? into.kernel_ms + ms : ms;
    S.timing_pending = false;
  }
  if (S.stats_pending) {
    if (accumulate) {
      into.samples_marched += S.h_stats[0], into.samples_shaded += S.h_stats[1], into.rays_hit += S.h_stats[2];
    } else {
      into.samples_marched = S.h_stats[0], into.samples_shaded = S.h_stats[1], into.rays_hit = S.h_stats[2];
    }
    c->last_tail = -1.0;  // the part of the launch during which warps had already run out of work (measured launches only)
    if (S.h_stats[5]) {
      const double t0 = (double)~S.h_stats[3], idle = (double)~S.h_stats[4], t1 = (double)S.h_stats[5];
      if (t1 > t0) c->last_tail = (t1 - idle) / (t1 - t0);
    }
    S.stats_pending = false;
  }
}

// Pipelined batches: wait until the frames of slot `S` are in host memory and add its counters to the pipeline's totals.
vr_status harvest(vr_ctx *c, LaunchSlot &S) {
  if (!S.in_flight) return VR_OK;
  VR_CUDA(c, cudaEventSynchronize(S.ev_copied));
  read_slot(c, S, c->pipe_total, true);
  c->pipe_total.kernel_launches += S.launches;
  S.in_flight = false;
  return VR_OK;
}
vr_status drain_pipeline(vr_ctx *c) {
  if (!c->pipe_open) return VR_OK;
  // oldest first
  VR_TRY(harvest(c, c->slot[c->pipe_batches & 1]));
  VR_TRY(harvest(c, c->slot[(c->pipe_batches + 1) & 1]));
  const uint64_t resident = c->stats.bytes_resident;
  const double upload_ms = c->stats.upload_ms;
  const uint32_t uploads = c->stats.uploads, fetch = c->stats.fetch_mode, arith = c->stats.arithmetic, vss = c->stats.vs_slices;
  const uint64_t uploaded = c->stats.bytes_uploaded;
  c->stats = c->pipe_total;
  c->stats.bytes_resident = resident, c->stats.upload_ms = upload_ms, c->stats.uploads = uploads, c->stats.bytes_uploaded = uploaded;
  c->stats.fetch_mode = fetch, c->stats.arithmetic = arith, c->stats.vs_slices = vss;
  c->pipe_open = false;
  c->cur = 0;
  return VR_OK;
}

vr_status collect(vr_ctx *c) {
  if (c->pipe_open) return drain_pipeline(c);
  LaunchSlot &S = c->slot[0];
  if (!S.stats_pending && !S.timing_pending) return VR_OK;
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  read_slot(c, S, c->stats, false);
  return VR_OK;
}

// volume-space passes of render_common (NULL = plain render)
struct VsArgs {
  int mode;                    // 1 opacity pass, 2 shading pass
  int index, count;            // this brick, number of bricks
  float *opacity_out;          // mode 1
  const float *opacity_all;    // mode 2
  const uint64_t *boxes_vox;   // mode 2: count x {lo x,y,z, hi x,y,z} in voxels
  // the two passes of ONE plain frame cut into depth slices (VR_OPT_RAY_SLICES): the "brick" is the whole volume of a context
  // without a region; the pair is timed as one launch
  bool self;
};

// 'render' minus the read-back: marshal like render.cpp:142-245, set up like vr::initRender
// (volumeRender.cpp:112-156) and vr::render (:264-277), launch on the context's stream.
vr_status render_common(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                        const float factors[3], const float es[3], const uint64_t res[2], size_t n_frames, const float *R,
                        const float *props, float thr, const float color[3], float *const *d_outs, cudaStream_t stream,
                        const VsArgs *vs = nullptr) {
  if (!vs && !c->region_on && n_frames == 1 && c->opt[VR_OPT_RAY_SLICES] > 1 && d_outs && d_outs[0] && res) {
    // Latency mode of a single frame: an opacity pass and a shading pass, both cut into depth slices.  Scenes the
    // passes do not take (volumes of different sizes, no closed-form skip ...) are rendered as ever.
    const size_t need = (size_t)res[0] * res[1] * sizeof(float);
    if (c->self_opac_cap < need) {
      if (c->d_self_opac) cudaFree(c->d_self_opac);
      c->d_self_opac = nullptr, c->self_opac_cap = 0;
      VR_CUDA(c, cudaMalloc(&c->d_self_opac, need));
      c->self_opac_cap = need;
    }
    VsArgs a{};
    a.mode = 1, a.opacity_out = c->d_self_opac, a.self = true;
    float *dummy = c->d_self_opac;
    const int64_t clear = c->opt[VR_OPT_CLEAR_IMAGE];
    c->opt[VR_OPT_CLEAR_IMAGE] = 0;
    VR_CUDA(c, cudaEventRecord(c->slot[c->cur].ev0, stream));  // the frame's time: both passes (the opacity pass leaves the slot alone)
    const vr_status st = render_common(c, lights, n_lights, illumination, factors, es, res, 1, R, props, thr, color, &dummy, stream, &a);
    c->opt[VR_OPT_CLEAR_IMAGE] = clear;
    if (st == VR_OK) {
      const HostVol &E0 = c->host[S_EM];
      const uint64_t box[6] = {0, 0, 0, E0.w, E0.h, E0.d};
      VsArgs b{};
      b.mode = 2, b.index = 0, b.count = 1, b.opacity_all = c->d_self_opac, b.boxes_vox = box, b.self = true;
      return render_common(c, nullptr, 0, nullptr, factors, es, res, 1, R, props, thr, color, d_outs, stream, &b);  // (lights: set by pass 1)
    }
    if (st != VR_ERR_UNSUPPORTED) return st;
    tl_error.clear();  // not a scene for the passes: the plain render below
  }
  if (c->region_on != (vs != nullptr && !vs->self))
    return fail(c, VR_ERR_INVALID_ARGUMENT, c->region_on ? "this context owns a region (vr_set_region): use vr_render_opacity_device / vr_render_partial_device"
                                                         : "volume-space passes need vr_set_region first");
  if (!factors || !es || !res || !R || !props || !color || !d_outs || n_frames == 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "insufficient parameter!");
  if (n_frames > 4096) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: at most 4096 frames per batch");
  for (size_t f = 0; f < n_frames; ++f)
    if (!d_outs[f]) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: null output");
  c->stats.bytes_uploaded = 0;
  if (lights && illumination) {  // render.cpp:145-191; a logical placeholder keeps the previous set
    c->host[S_LIGHT] = make_volume(illumination);
    c->lights.resize(n_lights);
    for (size_t l = 0; l < n_lights; ++l) {
      vr_light &o = c->lights[l];
      // make_float3Inv on Position (render.cpp:167), plain order on Color
      o.position[0] = lights[l].position[2], o.position[1] = lights[l].position[1], o.position[2] = lights[l].position[0];
      memcpy(o.color, lights[l].color, sizeof o.color);
    }
    // setIlluminationTexture (kernel.cu:682-689) tests `last_update > timeLastMemSync` but
    // syncVolume re-uploads regardless of its flag; the LUT is tiny, the data is what matters.
    const uint32_t uploads_before = c->stats.uploads;
    VR_TRY(syncVolumeCached(c, S_LIGHT, c->host[S_LIGHT]));
    if (c->stats.uploads != uploads_before && stream != c->stream) {
      // the upload ran on the context's stream; a render on a caller's stream must not start before it has landed
      VR_CUDA(c, cudaEventRecord(c->ev_order, c->stream));
      VR_CUDA(c, cudaStreamWaitEvent(stream, c->ev_order, 0));
    }
  }
  const HostVol &E = c->host[S_EM];
  if (E.w == 0 || E.h == 0 || E.d == 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: no emission volume synced");
  const uint64_t H = res[0], W = res[1];  // render.cpp:196,243
  if (H == 0 || W == 0 || H > 0x7fffffff || W > 0x7fffffff || W * H > 0xffffffffull)
    return fail(c, VR_ERR_INVALID_ARGUMENT, "render: bad resolution");

  KernelParams P;
  memset(&P, 0, sizeof P);
  P.cam.out = d_outs[0];
  setup_frame(P, factors, es, W, H, R, props, thr, color, E.w, E.h, E.d);

  // role -> texture through the redirect indices (kernel.cu:127-144)
  int sEm = c->idxEm, sAb = c->idxAb, sRe = c->idxRe;
  if (c->opt[VR_OPT_ABSORPTION_ROLE_FIX] && c->tex[S_AB].kind != TEX_UNBOUND) sAb = S_AB;
  // VR_OPT_TEXTURE_UNIT: 0 = loads + fp32 lerps, 1 = texture unit or an error, 2 (default) = texture unit where this
  // context can (fp32 whole volumes that fit twice), else loads
  const int64_t tu = c->opt[VR_OPT_TEXTURE_UNIT];
  bool texhw = false;
  if (tu && !(tu == 2 && c->texhw_failed)) {
    const int need[] = {sEm, sAb, sRe, S_EM, S_DX, S_DY, S_DZ, S_LIGHT};
    bool eligible = true;
    for (int sl : need) eligible = eligible && array_eligible(c->tex[sl]);
    if (eligible || tu == 1) {  // tu == 1: ensure_array reports what is in the way
      texhw = true;
      for (int sl : need) {
        const vr_status st = ensure_array(c, c->tex[sl], stream);
        if (st == VR_OK) continue;
        if (tu == 1) return st;
        cudaGetLastError();  // auto: no room for the second copy -> loads, until the volumes change
        c->texhw_failed = true, texhw = false;
        for (int s2 : need)
          if (c->tex[s2].storage || c->tex[s2].kind != TEX_VOLUME) drop_array(c->tex[s2]);  // (an array that is the ONLY copy stays)
        break;
      }
      if (texhw && vs) {
        // a region context's arrays hold its resident box only: one remap for all volume textures, so they must agree on it
        const Resident *first = nullptr;
        for (int sl : need) {
          const Resident &q = c->tex[sl];
          if (q.kind != TEX_VOLUME || sl == S_LIGHT) continue;
          if (!first) first = &q;
          if (memcmp(q.a0, first->a0, sizeof q.a0) || memcmp(q.an, first->an, sizeof q.an) || q.w != first->w || q.h != first->h || q.d != first->d) {
            if (tu == 1) return fail(c, VR_ERR_UNSUPPORTED, "texture-unit mode: the volumes of a region context must share one resident box");
            texhw = false;
          }
        }
        if (texhw && first) {
          const int dims[3] = {first->w, first->h, first->d};
          for (int i = 0; i < 3; ++i) {
            P.tex_cs[i] = (float)((double)dims[i] / (double)first->an[i]);
            P.tex_co[i] = (float)(-(double)first->a0[i] / (double)first->an[i]);
          }
        }
      }
    }
  }
  // the load kernels read the padded storage: bring back what an earlier texture-unit render dropped
  if (!texhw) {
    const int need[] = {sEm, sAb, sRe, S_EM, S_DX, S_DY, S_DZ, S_LIGHT};
    for (int sl : need) VR_TRY(restore_storage(c, c->tex[sl], stream));
  }
  if (c->opt[VR_OPT_EMPTY_SPACE_SKIP]) VR_TRY(ensure_skip_grids(c, c->tex[sAb], stream));  // (restores that storage if it must)
  P.em = describe(c->tex[sEm]);
  P.ab = describe(c->tex[sAb]);
  P.re = describe(c->tex[sRe]);
  P.em_direct = describe(c->tex[S_EM]);
  P.gx = describe(c->tex[S_DX]);
  P.gy = describe(c->tex[S_DY]);
  P.gz = describe(c->tex[S_DZ]);
  P.lut = describe(c->tex[S_LIGHT]);
  P.ab_is_em = sAb == sEm;
  P.re_is_em = sRe == sEm;
  P.gradient_method = c->gradientMethod;
  P.frac_bits = (int)c->opt[VR_OPT_FILTER_FRAC_BITS];
  P.count = (vs && vs->self && vs->mode == 1) ? 0 : (int)c->opt[VR_OPT_COUNT_SAMPLES];
  // all volume textures of one launch must share element type and layout
  int dtype = -1, layout = -1;
  const int used[] = {sEm, sAb, sRe, S_EM, S_DX, S_DY, S_DZ};
  for (int s : used) {
    const Resident &r = c->tex[s];
    if (r.kind != TEX_VOLUME) continue;
    if (dtype < 0) dtype = r.dtype, layout = r.layout;
    if (dtype != r.dtype || layout != r.layout)
      return fail(c, VR_ERR_UNSUPPORTED, "volumes of one render must share dtype and layout (slot %d differs)", s);
  }
  if (dtype < 0) dtype = VR_DTYPE_F32, layout = LAYOUT_LINEAR;
  Resident &rab = c->tex[sAb];
  P.skip = (rab.kind == TEX_VOLUME && rab.occ && rab.dist) ? (int)c->opt[VR_OPT_EMPTY_SPACE_SKIP] : 0;
  P.occ = rab.occ, P.dist = rab.dist, P.ocx = rab.ocx, P.ocy = rab.ocy, P.ocz = rab.ocz;
  if (P.skip) {
    // both skipping schemes rely on a sample moving less than half a voxel of the absorption-role
    // texture per step (true whenever that texture is not much finer than the emission volume
    // that defines tstep); otherwise every sample is marched
    float mv = 0.f;
    const float dim[3] = {P.ab.fw, P.ab.fh, P.ab.fd};
    for (int i = 0; i < 3; ++i) mv = fmaxf(mv, P.tstep * dim[i] / (P.boxmax[i] - P.boxmin[i]));
    if (!(mv <= 0.45f)) P.skip = 0;
    // Skipped samples never run the `sum.w > threshold` test (kernel.cu:485): with a NEGATIVE threshold the reference
    // stops after its first sample wherever that is, so every sample is marched then.  (Non-finite voxels: see vr_abi.h.)
    if (!(P.thr >= 0.0f)) P.skip = 0;
    for (int i = 0; i < 3; ++i) {  // on_empty_cell: cell of the low tap = floor(((pos - bmin) * bscale * N - 0.5) / 8)
      const double bscale = 1.0 / ((double)P.boxmax[i] - (double)P.boxmin[i]);
      const int oc = i == 0 ? P.ocx : (i == 1 ? P.ocy : P.ocz);
      P.cellA[i] = (float)(bscale * dim[i] / 8.0);
      P.cellB[i] = (float)((-(double)P.boxmin[i] * bscale * dim[i] - 0.5) / 8.0);
      P.cellMax[i] = (float)oc - 0.5f;
    }
  }
  // shared tap sets: every volume texture of the launch has the geometry of the first one
  P.uniform = 1;
  bool have_geo = false;
  const TexDesc *vt[] = {&P.em, &P.ab, &P.re, &P.em_direct, &P.gx, &P.gy, &P.gz};
  for (const TexDesc *t : vt) {
    if (t->kind != TEX_VOLUME) continue;
    if (!have_geo) P.geo = *t, P.geo.p = nullptr, have_geo = true;
    if (t->w != P.geo.w || t->h != P.geo.h || t->d != P.geo.d || t->sy != P.geo.sy || t->sz != P.geo.sz || t->ox != P.geo.ox ||
        t->oy != P.geo.oy || t->oz != P.geo.oz)
      P.uniform = 0;
  }
  // the gradient trio is either three components of the interleaved storage or three plain volumes;
  // a mix (geometries differed at upload time) goes through the per-texture path
  const Resident &rx = c->tex[S_DX], &ry = c->tex[S_DY], &rz = c->tex[S_DZ];
  P.grad4 = rx.kind == TEX_VOLUME && ry.kind == TEX_VOLUME && rz.kind == TEX_VOLUME && rx.shared && ry.shared && rz.shared;
  if (!P.grad4 && (rx.shared || ry.shared || rz.shared)) P.uniform = 0;
  if (c->opt[VR_OPT_KERNEL_VARIANT] == 1) P.uniform = 0;  // variant 1: per-texture addressing everywhere (A/B measurements)
  // a resident volume of more than 2^32 elements needs the kernels with 64-bit z terms
  bool wide = false;
  for (int s : used) {
    const Resident &r = c->tex[s];
    if (r.kind == TEX_VOLUME && (uint64_t)r.sz * (uint64_t)(r.layout == LAYOUT_BRICK ? (r.ld + 7) / 8 : r.ld) > 0xffffffffull) wide = true;
  }
  if (wide) {
    if (!P.uniform || layout != LAYOUT_BRICK || P.frac_bits)
      return fail(c, VR_ERR_UNSUPPORTED, "volumes of more than 2^32 resident elements need the bricked layout, one common geometry and fp32 weights");
    P.uniform = 2;
  }
  if (c->opt[VR_OPT_KERNEL_VARIANT] == 3 && P.uniform == 1 && layout == LAYOUT_BRICK) P.uniform = 2;  // variant 3: force the 64-bit kernels (tests)
  if (P.frac_bits) P.uniform = 0;  // fixed-point weights (A/B option) are compiled into the per-texture kernels only
  if (texhw) {
    P.uniform = vs ? 1 : 0, P.frac_bits = 0;  // the hardware filters (1.8 fixed-point weights by construction); 1 = the region kernels
    layout = LAYOUT_TEXHW;
    P.tex_fast = (P.gx.kind == TEX_VOLUME && P.gy.kind == TEX_VOLUME && P.gz.kind == TEX_VOLUME ? 1 : 0) | (P.re.kind == TEX_VOLUME ? 2 : 0) |
                 (P.em_direct.kind == TEX_VOLUME ? 4 : 0) | (P.em.kind == TEX_VOLUME ? 8 : 0) | (P.lut.kind == TEX_VOLUME ? 16 : 0);
  }

  // volume-space mode
  P.vs_mode = 0;
  int slices = 1;
  bool measure_tail = false;
  if (vs) {
    if (n_frames != 1) return fail(c, VR_ERR_UNSUPPORTED, "volume-space passes render one frame per launch");
    const int64_t dims[3] = {(int64_t)E.w, (int64_t)E.h, (int64_t)E.d};
    for (const TexDesc *t : vt)
      if (t->kind == TEX_VOLUME && (t->w != dims[0] || t->h != dims[1] || t->d != dims[2]))
        return fail(c, VR_ERR_UNSUPPORTED, "volume-space mode: every volume must have the emission volume's dims");
    // taps of an owned sample reach 1.5 voxels (+ the on-the-fly gradient offset 1/(2*boxmax_i) voxels) beyond the region
    float reach = 1.5f;
    if (!P.gradient_method && !c->lights.empty())
      for (int i = 0; i < 3; ++i) reach = fmaxf(reach, 1.5f + 1.0f / (2.0f * P.boxmax[i]));
    if (!vs->self && (float)c->halo < reach)
      return fail(c, VR_ERR_INVALID_ARGUMENT, "region halo %d too small: this scene needs %d voxels", c->halo, (int)ceilf(reach));
    P.vs_mode = vs->mode, P.vs_index = vs->index, P.vs_count = vs->count;
    for (int i = 0; i < 3; ++i) {
      P.reg_lo[i] = (vs->self || c->rlo[i] <= 0) ? -INFINITY : (float)c->rlo[i] / (float)dims[i];
      P.reg_hi[i] = (vs->self || c->rhi[i] >= dims[i]) ? INFINITY : (float)c->rhi[i] / (float)dims[i];
    }
    P.vs_opacity_out = vs->opacity_out, P.vs_opacity_all = vs->opacity_all;
    if (vs->mode == 2) {
      if (vs->count < 1 || vs->index < 0 || vs->index >= vs->count || !vs->boxes_vox || !vs->opacity_all) return fail(c, VR_ERR_INVALID_ARGUMENT, "partial render: bad brick arguments");
      std::vector<float> boxes((size_t)vs->count * 6);
      for (int b = 0; b < vs->count; ++b)
        for (int i = 0; i < 3; ++i) {  // voxel planes -> world planes of the box [boxmin, boxmax]
          const float ext = P.boxmax[i] - P.boxmin[i];
          boxes[6 * b + i] = P.boxmin[i] + ext * ((float)vs->boxes_vox[6 * b + i] / (float)dims[i]);
          boxes[6 * b + 3 + i] = P.boxmin[i] + ext * ((float)vs->boxes_vox[6 * b + 3 + i] / (float)dims[i]);
        }
      const size_t need = boxes.size() * sizeof(float);
      if (c->boxes_cap < need) {
        if (c->d_boxes) cudaFree(c->d_boxes);
        c->d_boxes = nullptr, c->boxes_cap = 0;
        VR_CUDA(c, cudaMalloc(&c->d_boxes, need));
        c->boxes_cap = need;
      }
      VR_CUDA(c, cudaMemcpyAsync(c->d_boxes, boxes.data(), need, cudaMemcpyHostToDevice, stream));
      P.vs_boxes = c->d_boxes;
    } else if (!vs->opacity_out) {
      return fail(c, VR_ERR_INVALID_ARGUMENT, "opacity pass: null output");
    }
    // depth slices: the workers of a slice find its start with the exact fast-forward of the empty-space jump
    slices = (int)c->opt[VR_OPT_VS_SLICES];
    if (slices == 0) {
      // Automatic.  A lone warp needs microseconds per shaded sample (the loop is one dependent chain), so a ray of 6000 samples
      // through dense matter lasts 18 ms whatever the other 4735 warps do; slices pay when that chain is what the launch waits
      // for, and cost about a millisecond per frame (slice starts, folds, S times the items) when it is not.  So the context
      // MEASURES: frames march whole rays, and their shading pass records when its first warp ran out of work and when its last
      // one finished.  A tail of 30 % of the launch or more -> 8 slices from then on (measured on 2048^3 bricks of a 4096^3
      // volume: tails of 70 - 90 %, 35.6 -> 19.7 ms per frame); shorter tails -> whole rays (two bricks of 1024^3 / 2048^3: tails
      // of 12 - 19 %, slices change the frame by -2 ... +7 %).  Starts over when the image or the tiling changes; needs
      // VR_OPT_COUNT_SAMPLES (the read-back of the counters).
      const uint64_t tw = (uint64_t)c->opt[VR_OPT_TILE_W], th = (uint64_t)c->opt[VR_OPT_TILE_H];
      const uint64_t asig = (W << 40) ^ (H << 20) ^ (tw << 8) ^ th;
      if (c->auto_slices_sig != asig) c->auto_slices = 0, c->auto_slices_sig = asig, c->last_tail = -1.0;
      // (whole rays keep measuring, a cut is kept: a scene that turns a long tail towards the camera later is still cut then)
      if (vs->mode == 1 && c->auto_slices <= 1 && c->stats.vs_slices == 1 && c->last_tail >= 0.0) c->auto_slices = c->last_tail >= 0.3 ? 8 : 1;
      measure_tail = vs->mode == 2 && c->auto_slices <= 1 && P.count;
      slices = vs->mode == 1 ? (c->auto_slices ? c->auto_slices : 1) : (c->slices_of_opacity_pass > 1 ? c->slices_of_opacity_pass : 1);
    }
    if (vs->self) slices = (int)std::min<int64_t>(c->opt[VR_OPT_RAY_SLICES], 16);  // a plain frame in latency mode: explicit only
    // (the kernels with per-texture addressing, P.uniform == 0, are built for whole rays only)
    if (slices < 1 || P.skip != 2 || P.uniform == 0) slices = 1;
    if (vs->self && vs->mode == 1 && slices == 1) return fail(c, VR_ERR_UNSUPPORTED, "this scene cannot be cut into depth slices");
    if (vs->mode == 2 && slices != (c->slices_of_opacity_pass > 1 ? c->slices_of_opacity_pass : 1))
      return fail(c, VR_ERR_INVALID_ARGUMENT, "partial render with %d depth slices after an opacity pass with %d", slices, c->slices_of_opacity_pass);
    if (slices > 1) {
      const size_t plane = (size_t)W * H * sizeof(float);
      if (vs->mode == 2 && c->pixels_of_opacity_pass != W * H) return fail(c, VR_ERR_INVALID_ARGUMENT, "partial render: resolution differs from the opacity pass");
      float **buf = vs->mode == 1 ? &c->d_slice_opac : &c->d_slice_img;
      size_t &cap = vs->mode == 1 ? c->slice_opac_cap : c->slice_img_cap;
      const size_t need = plane * slices * (vs->mode == 1 ? 1 : 3);
      if (cap < need) {
        if (*buf) cudaFree(*buf);
        *buf = nullptr, cap = 0;
        VR_CUDA(c, cudaMalloc(buf, need));
        cap = need;
      }
      const size_t need_st = (size_t)W * H * sizeof(float4) * (slices - 1);
      if (c->slice_state_cap < need_st) {
        if (c->d_slice_state) cudaFree(c->d_slice_state);
        c->d_slice_state = nullptr, c->slice_state_cap = 0;
        VR_CUDA(c, cudaMalloc(&c->d_slice_state, need_st));
        c->slice_state_cap = need_st;
        memset(&c->key_of_states, 0, sizeof c->key_of_states);
      }
      P.vs_slices = slices, P.vs_slice_opacity = c->d_slice_opac, P.vs_slice_state = c->d_slice_state;
      if (vs->mode == 2) P.cam.out = c->d_slice_img;
    }
    if (vs->mode == 1) c->slices_of_opacity_pass = slices, c->pixels_of_opacity_pass = W * H;
    P.measure_tail = measure_tail && slices == 1;
  }

  // lights
  P.n_lights = (int)c->lights.size();
  for (int i = 0; i < P.n_lights && i < MAX_PARAM_LIGHTS; ++i) {
    memcpy(P.lights[i], c->lights[i].position, 12);
    memcpy(P.lights[i] + 3, c->lights[i].color, 12);
  }
  if (P.n_lights > MAX_PARAM_LIGHTS) {
    const size_t need = (size_t)(P.n_lights - MAX_PARAM_LIGHTS) * 24;
    if (c->lights_more_cap < need) {
      if (c->d_lights_more) cudaFree(c->d_lights_more);
      c->d_lights_more = nullptr, c->lights_more_cap = 0;
      VR_CUDA(c, cudaMalloc(&c->d_lights_more, need));
      c->lights_more_cap = need;
    }
    VR_CUDA(c, cudaMemcpyAsync(c->d_lights_more, c->lights.data() + MAX_PARAM_LIGHTS, need, cudaMemcpyHostToDevice, stream));
    P.lights_more = c->d_lights_more;
  }
  if (P.n_lights > 0 && c->tex[S_LIGHT].kind == TEX_UNBOUND) {
    // lights without a LUT cannot happen through render.cpp (both or neither); keep it defined
    P.n_lights = 0;
  }

  // image-space partition
  P.part_rank = (int)c->opt[VR_OPT_PART_RANK];
  P.part_nranks = (int)c->opt[VR_OPT_PART_NRANKS];
  if (P.part_nranks < 1 || P.part_rank < 0 || P.part_rank >= P.part_nranks) return fail(c, VR_ERR_INVALID_ARGUMENT, "bad partition %d/%d", P.part_rank, P.part_nranks);
  P.tile_w = (int)c->opt[VR_OPT_TILE_W], P.tile_h = (int)c->opt[VR_OPT_TILE_H];
  P.wshift = (int)c->opt[VR_OPT_WARP_SHAPE];
  if (P.wshift < 0 || P.wshift > 5) return fail(c, VR_ERR_INVALID_ARGUMENT, "warp shape must be in [0, 5]");
  const int ww = 1 << P.wshift, wh = 32 >> P.wshift;
  if (P.tile_w < ww || P.tile_h < wh || P.tile_w % ww || P.tile_h % wh) return fail(c, VR_ERR_INVALID_ARGUMENT, "tile must be a multiple of the %dx%d warp footprint", ww, wh);
  P.ntx = (P.W + P.tile_w - 1) / P.tile_w, P.nty = (P.H + P.tile_h - 1) / P.tile_h;
  P.wt_x = P.tile_w / ww, P.wt_y = P.tile_h / wh;
  const uint64_t ntiles = (uint64_t)P.ntx * P.nty;
  // which tiles this context renders: tiles dealt round-robin (mode 0), or whole tile COLUMNS dealt round-robin (mode 1: the
  // owned pixels of every image plane are then a few contiguous runs of tile_w * H floats - what a device->host copy wants)
  const bool columns = c->opt[VR_OPT_PART_MODE] == 1 && P.part_nranks > 1;
  P.part_columns = columns ? 1 : 0;
  const uint64_t own_cols = (uint64_t)P.ntx > (uint64_t)P.part_rank ? ((uint64_t)P.ntx - P.part_rank + P.part_nranks - 1) / P.part_nranks : 0;
  const uint64_t owned = columns ? own_cols * P.nty
                                 : (ntiles > (uint64_t)P.part_rank ? (ntiles - P.part_rank + P.part_nranks - 1) / P.part_nranks : 0);
  auto owned_tile = [&](uint64_t j) -> uint32_t {
    if (!columns) return (uint32_t)(P.part_rank + j * P.part_nranks);
    return (uint32_t)((j / own_cols) * P.ntx + P.part_rank + (j % own_cols) * P.part_nranks);
  };
  const uint64_t entries = owned * n_frames;  // (frame, owned tile) pairs of the launch
  P.num_owned = (uint32_t)entries;
  const uint64_t work = entries * P.wt_x * P.wt_y * (uint64_t)slices;  // (slices: volume-space passes only)
  if (work > 0xfffffff0ull || ntiles * n_frames > 0xfffffff0ull) return fail(c, VR_ERR_UNSUPPORTED, "too many work items");
  P.num_work = (uint32_t)work;
  P.work_counter = c->d_counter;
  P.warp_times = c->d_warp_times;
  if (c->d_warp_times) VR_CUDA(c, cudaMemsetAsync(c->d_warp_times, 0, 8 * 8192 * sizeof(unsigned long long), stream));
  P.tiles_total = (uint32_t)ntiles;
  P.n_frames = (int)n_frames;
  // cameras of a batch
  std::vector<Camera> cams(n_frames);
  for (size_t f = 0; f < n_frames; ++f) {
    fill_camera(cams[f], R + 9 * f, props + 3 * f);
    cams[f].out = d_outs[f];
  }
  P.cams = nullptr;
  if (n_frames > 1) {
    const size_t need = n_frames * sizeof(Camera);
    if (c->cams_cap < need) {
      if (c->d_cams) cudaFree(c->d_cams);
      c->d_cams = nullptr, c->cams_cap = 0;
      VR_CUDA(c, cudaMalloc(&c->d_cams, need));
      c->cams_cap = need;
    }
    VR_CUDA(c, cudaMemcpyAsync(c->d_cams, cams.data(), need, cudaMemcpyHostToDevice, stream));  // pageable: staged before return
    P.cams = c->d_cams;
  }
  // Hand the tiles out heaviest first.  One warp marches 32 rays from start to end, so a tile of
  // long rays picked up late would run alone at the end of the kernel; rays through the middle
  // of the volume are the long ones, so owned tiles are ordered by the distance of their centre
  // from the projection of the box centre (longest-processing-time-first, by a geometric guess).
  // In a batch all frames share the one queue: the light tiles of every frame fill the tail.
  P.tile_order = nullptr;
  const bool ordered = c->opt[VR_OPT_TILE_ORDER] != 0;
  const uint64_t sig = ((uint64_t)P.W << 40) ^ ((uint64_t)P.H << 20) ^ ((uint64_t)P.tile_w << 8) ^ (uint64_t)P.tile_h;
  // measured tile costs: those of the most recent launch that has completed (this slot's, else the other slot's - a pipelined
  // batch is set up while its predecessor still runs)
  LaunchSlot &S = c->slot[c->cur];
  const unsigned int *h_cost = nullptr;
  for (int k = 0; k < 2 && !h_cost; ++k) {
    LaunchSlot &Q = c->slot[(c->cur + 1 + k) & 1];  // the other slot holds the more recent launch
    if (Q.cost_valid && Q.cost_sig == sig && Q.cost_tiles >= ntiles && cudaEventQuery(Q.ev_cost) == cudaSuccess) h_cost = Q.h_tile_cost;
    cudaGetLastError();  // cudaErrorNotReady of the query is not an error
  }
  const bool use_cost = h_cost != nullptr;
  if ((ordered && entries > 1) || n_frames > 1 || (columns && entries > 0)) {  // (the kernel's built-in hand-out is the tile round-robin)
    std::vector<std::pair<float, uint32_t>> key((size_t)entries);
    for (size_t f = 0; f < n_frames; ++f) {
      const Camera &C = cams[f];
      const float *xv = C.m[0], *yv = C.m[1], *zv = C.m[2];
      const float xl = sqrtf(xv[0] * xv[0] + xv[1] * xv[1] + xv[2] * xv[2]);
      float o[3];  // camera origin (kernel.cu:399-410)
      for (int i = 0; i < 3; ++i) o[i] = C.m[3][0] * xv[i] - C.m[3][2] * zv[i];
      // the point the long rays pass through: the box centre (the world origin) - or, for a context that owns a region, the
      // centre of ITS box: only the rays through that brick have work here
      float ctr[3] = {0.f, 0.f, 0.f};
      if (vs && !vs->self)
        for (int i = 0; i < 3; ++i) {
          const float lo_n = (float)c->rlo[i] / (float)(i == 0 ? E.w : i == 1 ? E.h : E.d), hi_n = (float)c->rhi[i] / (float)(i == 0 ? E.w : i == 1 ? E.h : E.d);
          ctr[i] = P.boxmin[i] + (P.boxmax[i] - P.boxmin[i]) * 0.5f * (lo_n + hi_n);
        }
      const float d[3] = {ctr[0] - o[0], ctr[1] - o[1], ctr[2] - o[2]};
      const float a = (d[0] * xv[0] + d[1] * xv[1] + d[2] * xv[2]) / (xl > 0 ? xl : 1.f);
      const float b = (d[0] * yv[0] + d[1] * yv[1] + d[2] * yv[2]);
      const float cz = (d[0] * zv[0] + d[1] * zv[1] + d[2] * zv[2]);
      const float ratio = (float)P.H / (float)P.W;
      float px = 0.5f * P.W, py = 0.5f * P.H;
      if (cz > 1e-6f) {
        px = (a / cz * C.m[3][1] + 1.f) * 0.5f * P.W;
        py = (b / cz * C.m[3][1] + ratio) / (2.f * ratio) * P.H;
      }
      for (uint64_t j = 0; j < owned; ++j) {
        const uint32_t tile = owned_tile(j);
        const float tx = ((tile % P.ntx) + 0.5f) * P.tile_w - px, ty = ((tile / P.ntx) + 0.5f) * P.tile_h - py;
        const uint32_t entry = (uint32_t)(f * ntiles + tile);
        float k = ordered ? tx * tx + ty * ty : (float)(f * owned + j);
        // measured beats guessed: the previous launch's shaded samples per tile (a movie turns slowly), most expensive first
        // key: the longest ray of the tile (= the duration of its longest item), then the tile's total
        if (ordered && use_cost) k = -((float)h_cost[ntiles + tile] * 1024.0f + fminf((float)h_cost[tile] * (1.0f / 4096.0f), 1023.0f));
        // order 1: one heaviest-first list over all frames of the batch (measured best: 30.3 vs 31.3 ms per frame);
        // order 2: frame after frame, each heaviest-first
        if (ordered && use_cost && c->opt[VR_OPT_TILE_ORDER] == 3) k = -(float)h_cost[tile];  // A/B: by the tile's total
        if (c->opt[VR_OPT_TILE_ORDER] == 2) k = (float)f * 4.0e9f + fminf(fmaxf(k, -1.9e9f), 1.9e9f);
        key[f * owned + j] = std::make_pair(k, entry);
      }
    }
    std::sort(key.begin(), key.end());
    c->h_tile_order.resize((size_t)entries);
    for (uint64_t j = 0; j < entries; ++j) c->h_tile_order[j] = key[j].second;
    if (c->tile_order_cap < entries * 4) {
      if (c->d_tile_order) cudaFree(c->d_tile_order);
      c->d_tile_order = nullptr, c->tile_order_cap = 0;
      VR_CUDA(c, cudaMalloc(&c->d_tile_order, entries * 4));
      c->tile_order_cap = entries * 4;
    }
    // pageable source: the runtime stages it before returning, so the vector may be reused next frame
    VR_CUDA(c, cudaMemcpyAsync(c->d_tile_order, c->h_tile_order.data(), entries * 4, cudaMemcpyHostToDevice, stream));
    P.tile_order = c->d_tile_order;
  }
  P.stats = S.d_stats;
  P.tile_cost = nullptr;
  // (volume-space: the SHADING pass measures; the opacity pass of the next frame is ordered by the same figures - the same rays
  // through the same brick - and records nothing: its samples are not "shaded")
  if (ordered && (!vs || vs->mode == 2)) {
    if (c->tile_cost_tiles < ntiles) {
      if (c->d_tile_cost) cudaFree(c->d_tile_cost);
      c->d_tile_cost = nullptr, c->tile_cost_tiles = 0;
      VR_CUDA(c, cudaMalloc(&c->d_tile_cost, 2 * ntiles * sizeof(unsigned int)));  // [sum of shaded samples | longest ray] per tile
      c->tile_cost_tiles = ntiles;
    }
    if (S.cost_tiles < ntiles) {
      if (S.h_tile_cost) cudaFreeHost(S.h_tile_cost);
      S.h_tile_cost = nullptr, S.cost_tiles = 0, S.cost_valid = false;
      VR_CUDA(c, cudaMallocHost(&S.h_tile_cost, 2 * ntiles * sizeof(unsigned int)));
      S.cost_tiles = ntiles;
    }
    // the host copy of the previous launch is final before it is overwritten: same stream, and use_cost saw its event
    VR_CUDA(c, cudaMemsetAsync(c->d_tile_cost, 0, 2 * ntiles * sizeof(unsigned int), stream));
    P.tile_cost = c->d_tile_cost;
  }

  // cudaMemset of the image (volumeRender.cpp:271): pixels whose ray misses the box stay 0
  if (c->opt[VR_OPT_CLEAR_IMAGE])
    for (size_t f = 0; f < n_frames; ++f) VR_CUDA(c, cudaMemsetAsync(d_outs[f], 0, (size_t)W * H * 3 * sizeof(float), stream));
  VR_CUDA(c, cudaMemsetAsync(c->d_counter, 0, 64, stream));
  VR_CUDA(c, cudaMemsetAsync(S.d_stats, 0, 6 * sizeof(unsigned long long), stream));  // [marched, shaded, hits | tail measurement]
  const bool self1 = vs && vs->self && vs->mode == 1;  // the opacity pass of a latency-mode frame: no counters, no timing of its own
  if (!(vs && vs->self)) VR_CUDA(c, cudaEventRecord(S.ev0, stream));  // (latency mode: recorded before the opacity pass)
  if (texhw && c->opt[VR_OPT_TEXTURE_SINGLE_COPY]) {
    // texture-unit launches read arrays and grids only: the padded storages go (cudaFree waits for the grid builds that
    // still read them; this happens on the first frame after an upload only)
    const int need[] = {sEm, sAb, sRe, S_EM, S_DX, S_DY, S_DZ, S_LIGHT};
    for (int sl : need) drop_storage(c, c->tex[sl]);
  }
  const int arith = (int)c->opt[VR_OPT_ARITHMETIC];
  int slice_states_ran = 0;
  if (slices > 1) {
    // the slices' first samples: found by the opacity pass, reused by the shading pass of the same frame (same camera, same cut)
    SliceStateKey key;
    memset(&key, 0, sizeof key);
    memcpy(key.cam, P.cam.m, sizeof key.cam), memcpy(key.box, P.boxmax, sizeof key.box);
    key.W = W, key.H = H, key.tstep = P.tstep, key.slices = slices, key.arith = arith;
    if (vs->mode == 1 || memcmp(&key, &c->key_of_states, sizeof key) != 0) {
      VR_CUDA(c, launch_slice_states(P, stream, c->sm_count, arith));
      c->key_of_states = key;
      slice_states_ran = 1;
    }
  }
  if (P.num_work > 0) VR_CUDA(c, launch_raymarch(P, layout, dtype, stream, c->sm_count, arith));
#ifdef VR_DEBUG_BOUNDS  // developer build: grid cells outside the grid that the jump asked for (logged and clamped by the kernel)
  if (c->d_warp_times) {
    std::vector<unsigned long long> log(8 + 6 * 16);
    cudaError_t de = cudaStreamSynchronize(stream);
    if (de == cudaSuccess) de = cudaMemcpy(log.data(), c->d_warp_times + 2 * 8192, log.size() * 8, cudaMemcpyDeviceToHost);
    fprintf(stderr, "[bounds] device %d vs_mode %d: %s, %llu cells out of range\n", c->device, P.vs_mode, cudaGetErrorString(de), log[0]);
    for (unsigned long long i = 0; i < log[0] && i < 16; ++i) {
      const unsigned long long *e = log.data() + 8 + 6 * i;
      auto f = [](unsigned long long u) { float v; unsigned w = (unsigned)u; memcpy(&v, &w, 4); return v; };
      fprintf(stderr, "[bounds]   cell %llu t %.7g tfar %.7g pos %.7g %.7g %.7g c %llu %llu %llu marched %llu\n", e[0], f(e[1]), f(e[1] >> 32), f(e[2]),
              f(e[2] >> 32), f(e[3]), e[4] & 0xfffff, (e[4] >> 20) & 0xfffff, e[4] >> 40, e[5]);
    }
  }
#endif
  c->stats.arithmetic = (uint32_t)arith;
  if (slices > 1) {  // the brick's opacity plane / colour share from those of its slices
    FoldOwner own;
    own.plane = W * H, own.H = (uint32_t)H, own.tile_w = (uint32_t)P.tile_w, own.tile_h = (uint32_t)P.tile_h, own.ntx = (uint32_t)P.ntx;
    own.rank = (uint32_t)P.part_rank, own.nranks = (uint32_t)P.part_nranks, own.columns = columns ? 1u : 0u;
    if (vs->mode == 1) VR_CUDA(c, launch_fold_slices(c->d_slice_opac, slices, W * H, vs->opacity_out, 1, own, stream));
    else VR_CUDA(c, launch_fold_slices(c->d_slice_img, slices, 3 * W * H, d_outs[0], 0, own, stream));
  }
  if (!self1) VR_CUDA(c, cudaEventRecord(S.ev1, stream));
  if (P.tile_cost) {
    VR_CUDA(c, cudaMemcpyAsync(S.h_tile_cost, c->d_tile_cost, 2 * ntiles * sizeof(unsigned int), cudaMemcpyDeviceToHost, stream));
    VR_CUDA(c, cudaEventRecord(S.ev_cost, stream));
    S.cost_valid = true, S.cost_sig = sig;
  }
  if (P.count) {
    VR_CUDA(c, cudaMemcpyAsync(S.h_stats, S.d_stats, 6 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    S.stats_pending = true;
  }
  if (!self1) S.timing_pending = true;
  c->stats.kernel_launches = (P.num_work > 0 ? 1 : 0) + (slices > 1 ? 1 + slice_states_ran : 0);  // march (+ fold, + slice starts)
  if (vs && vs->self && vs->mode == 2) c->stats.kernel_launches += 3;  // the opacity pass of the same frame: slice starts, march, fold
  c->stats.vs_slices = (uint32_t)slices;
  c->stats.fetch_mode = texhw ? 1u : 0u;
  return VR_OK;
}

// ---- device groups (vr_set_devices) ------------------------------------------------------------------------
// the contexts a call on `c` drives: the group it leads, else itself
std::vector<vr_ctx *> group_of(vr_ctx *c) { return c->members.empty() ? std::vector<vr_ctx *>{c} : c->members; }

// counters of the whole group: sums, and the slowest member's times
vr_stats merged_stats(vr_ctx *c) {
  vr_stats t = c->stats;
  for (size_t i = 1; i < c->members.size(); ++i) {
    const vr_stats &m = c->members[i]->stats;
    t.kernel_ms = std::max(t.kernel_ms, m.kernel_ms), t.upload_ms = std::max(t.upload_ms, m.upload_ms);
    t.samples_marched += m.samples_marched, t.samples_shaded += m.samples_shaded, t.rays_hit += m.rays_hit;
    t.bytes_resident += m.bytes_resident, t.bytes_uploaded += m.bytes_uploaded;
    t.kernel_launches += m.kernel_launches, t.uploads += m.uploads;
  }
  return t;
}
#define VR_NOT_A_MEMBER(c) \
  if ((c)->leader) return fail(c, VR_ERR_INVALID_ARGUMENT, "this context is a member of a device group: call its leader")
#define VR_NOT_A_GROUP(c, what) \
  if (!(c)->members.empty()) return fail(c, VR_ERR_UNSUPPORTED, what " is not available on a device group (vr_set_devices)")

vr_status sync_volumes_one(vr_ctx *c, uint64_t T, const vr_volume *em, const vr_volume *re, const vr_volume *ab, const vr_volume *gx,
                           const vr_volume *gy, const vr_volume *gz);

}  // namespace

extern "C" {

vr_status vr_create_on_device(vr_ctx **out, int device) {
  if (!out) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_create: null output");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) return fail(nullptr, VR_ERR_NO_DEVICE, "no CUDA device (%s); there is no CPU fallback", cudaGetErrorString(e));
  if (device < 0) {
    const char *env = getenv("VR_DEVICE");
    if (env) {
      device = atoi(env);
    } else if (cudaGetDevice(&device) != cudaSuccess) {
      device = 0;
    }
  }
  if (device >= n) return fail(nullptr, VR_ERR_NO_DEVICE, "device %d of %d does not exist", device, n);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(nullptr, VR_ERR_NO_DEVICE, "cannot query device %d", device);
  if (prop.major != 10) return fail(nullptr, VR_ERR_NO_DEVICE, "device %d is sm_%d%d; this library holds sm_100a code only", device, prop.major, prop.minor);
  vr_ctx *c = new vr_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  memset(c->opt, 0, sizeof c->opt);
  c->opt[VR_OPT_EMPTY_SPACE_SKIP] = 2;
  c->opt[VR_OPT_PART_NRANKS] = 1;
  c->opt[VR_OPT_GRADIENT_INTERLEAVE] = 1;
  c->opt[VR_OPT_TILE_W] = 64;
  c->opt[VR_OPT_TILE_H] = 64;
  c->opt[VR_OPT_BRICK_LAYOUT] = 1;
  c->opt[VR_OPT_COUNT_SAMPLES] = 1;
  c->opt[VR_OPT_CLEAR_IMAGE] = 1;
  c->opt[VR_OPT_TILE_ORDER] = 1;
  c->opt[VR_OPT_WARP_SHAPE] = 3;
  // opt-in without touching the MATLAB classes: VR_TEXTURE_UNIT=1 in the environment of the MATLAB process
  c->opt[VR_OPT_TEXTURE_UNIT] = 2;
  // without touching the MATLAB classes: VR_TEXTURE_UNIT=0 (loads + fp32 lerps) / 1 / 2 in the environment of the MATLAB process
  if (const char *tu = getenv("VR_TEXTURE_UNIT")) c->opt[VR_OPT_TEXTURE_UNIT] = atoi(tu) < 0 ? 0 : (atoi(tu) > 2 ? 2 : atoi(tu));
  c->opt[VR_OPT_TEXTURE_SINGLE_COPY] = 1;
  c->opt[VR_OPT_VS_SLICES] = 0;
  c->opt[VR_OPT_RAY_SLICES] = 1;
  if (const char *sl = getenv("VR_VS_SLICES")) c->opt[VR_OPT_VS_SLICES] = atoi(sl) < 0 ? 0 : (atoi(sl) > 16 ? 16 : atoi(sl));
  c->opt[VR_OPT_ARITHMETIC] = VR_DEFAULT_ARITHMETIC;
  if (const char *ar = getenv("VR_ARITHMETIC")) c->opt[VR_OPT_ARITHMETIC] = atoi(ar) < 0 ? 0 : (atoi(ar) > 2 ? 2 : atoi(ar));
  cudaError_t ce = cudaSetDevice(device);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  c->own_stream = ce == cudaSuccess;
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (int k = 0; k < 2 && ce == cudaSuccess; ++k) {
    LaunchSlot &S = c->slot[k];
    ce = cudaEventCreate(&S.ev0);
    if (ce == cudaSuccess) ce = cudaEventCreate(&S.ev1);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&S.ev_cost, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&S.ev_rendered, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&S.ev_copied, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaMalloc(&S.d_stats, 64);
    if (ce == cudaSuccess) ce = cudaMallocHost(&S.h_stats, 64);
  }
  if (ce == cudaSuccess) ce = cudaEventCreate(&c->evu0);
  if (ce == cudaSuccess) ce = cudaEventCreate(&c->evu1);
  for (int i = 0; i < kStageBufs && ce == cudaSuccess; ++i) ce = cudaEventCreateWithFlags(&c->ev_stage[i], cudaEventDisableTiming);
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&c->ev_order, cudaEventDisableTiming);
  if (ce == cudaSuccess) ce = cudaMalloc(&c->d_counter, 64);
#ifdef VR_TRACE_PIXEL  // developer build: the buffer lives in host memory the kernels write through a mapping - it survives a fault
  if (ce == cudaSuccess) ce = cudaHostAlloc(&g_trace_host, 8 * 8192 * sizeof(unsigned long long), cudaHostAllocMapped);
  if (ce == cudaSuccess) ce = cudaHostGetDevicePointer((void **)&c->d_warp_times, g_trace_host, 0);
#else
  if (ce == cudaSuccess && getenv("VR_WARP_TIMES")) ce = cudaMalloc(&c->d_warp_times, 8 * 8192 * sizeof(unsigned long long));
#endif  // [start, end] per warp; item records of profile builds behind
  if (ce != cudaSuccess) {
    vr_status s = fail(nullptr, VR_ERR_CUDA, "vr_create: %s", cudaGetErrorString(ce));
    delete c;
    return s;
  }
  {
    std::lock_guard<std::mutex> g(g_registry_mutex);
    g_registry.insert(c);
  }
  *out = c;
  return VR_OK;
}
vr_status vr_create(vr_ctx **out) { return vr_create_on_device(out, -1); }

vr_status vr_destroy(vr_ctx *c) {
  {
    std::lock_guard<std::mutex> g(g_registry_mutex);
    if (!c || !g_registry.count(c)) {
      tl_error = "Handle not valid.";
      return VR_ERR_INVALID_HANDLE;
    }
    g_registry.erase(c);
  }
  if (c->leader) {  // a member destroyed on its own leaves its group
    auto &mm = c->leader->members;
    mm.erase(std::remove(mm.begin(), mm.end(), c), mm.end());
    c->leader = nullptr;
  }
  for (size_t i = 1; i < c->members.size(); ++i) {
    c->members[i]->leader = nullptr;
    vr_destroy(c->members[i]);
  }
  c->members.clear();
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
  if (c->ev_group) cudaEventDestroy(c->ev_group);
  for (int i = 0; i < S_N; ++i) release(c, c->tex[i]);
  for (LaunchSlot &S : c->slot) {
    if (S.d_image) cudaFree(S.d_image);
    if (S.d_stats) cudaFree(S.d_stats);
    if (S.h_stats) cudaFreeHost(S.h_stats);
    if (S.h_tile_cost) cudaFreeHost(S.h_tile_cost);
    cudaEvent_t evs[5] = {S.ev0, S.ev1, S.ev_cost, S.ev_rendered, S.ev_copied};
    for (cudaEvent_t e : evs)
      if (e) cudaEventDestroy(e);
  }
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (int i = 0; i < kStageBufs; ++i) {
    if (c->d_stage[i]) cudaFree(c->d_stage[i]);
    if (c->h_stage[i]) cudaFreeHost(c->h_stage[i]);
    if (c->ev_stage[i]) cudaEventDestroy(c->ev_stage[i]);
  }
  if (c->d_lights_more) cudaFree(c->d_lights_more);
  if (c->d_counter) cudaFree(c->d_counter);
  if (c->d_tile_order) cudaFree(c->d_tile_order);
  if (c->d_cams) cudaFree(c->d_cams);
  if (c->d_tile_cost) cudaFree(c->d_tile_cost);
  if (c->ev_order) cudaEventDestroy(c->ev_order);
  if (c->d_boxes) cudaFree(c->d_boxes);
  if (c->d_slice_opac) cudaFree(c->d_slice_opac);
  if (c->d_slice_img) cudaFree(c->d_slice_img);
  if (c->d_slice_state) cudaFree(c->d_slice_state);
  if (c->d_self_opac) cudaFree(c->d_self_opac);
  if (c->d_warp_times) cudaFree(c->d_warp_times);
  cudaEventDestroy(c->evu0), cudaEventDestroy(c->evu1);
  if (c->own_stream) cudaStreamDestroy(c->stream);
  c->magic = 0;
  delete c;
  return VR_OK;
}

vr_status vr_mem_info(vr_ctx *c, char *buf, size_t cap, size_t *needed) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  size_t free_b = 0, total_b = 0;
  VR_CUDA(c, cudaMemGetInfo(&free_b, &total_b));
  const double MB = 1024.0 * 1024.0;
  std::string s;
  char line[512];
  auto add = [&](const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(line, sizeof line, fmt, ap);
    va_end(ap);
    s += line;
  };
  // same report as MManager::memInfo (mmanager.hxx:218-284), device pointers = resident storage
  add("Memory Information\n------------------------------\nlast sync (timestamp): %llu\n\n", (unsigned long long)c->timeLastMemSync);
  add("\tGPU %d\n\t---\n\t\tTotal Memory (MB): \t%g\n\t\tFree Memory (MB): \t%g\n\t\tUsed Memory (MB): \t%g\n\n", c->device, total_b / MB,
      free_b / MB, (total_b - free_b) / MB);
  add("\t\tVolumes\n\t\t-------\n");
  const char *names[S_N] = {"Emission", "Absorption", "Reflection", "dX", "dY", "dZ", "light"};
  for (int i = 0; i < S_N; ++i)
    add("\t\t%s (MB): %g ptr: %p resident (MB): %g\n", names[i], c->host[i].bytes / MB, c->tex[i].storage, c->tex[i].storage_bytes / MB);
  if (c->g4) add("\t\tdX,dY,dZ interleaved (MB): %g ptr: %p users: %d\n", c->g4_bytes / MB, c->g4, c->g4_users);
  {
    double arrays = 0;
    for (int i = 0; i < S_N; ++i) arrays += (double)c->tex[i].array_bytes;
    add("\t\tfetch: %s; texture-unit copies (MB): %g\n",
        c->opt[VR_OPT_TEXTURE_UNIT] == 0 ? "loads + fp32 lerps" : (c->texhw_failed ? "loads (no room for the texture copies)" : "texture unit where possible"), arrays / MB);
  }
  if (c->members.size() > 1) {
    std::string ids;
    for (vr_ctx *m : c->members) ids += (ids.empty() ? "" : ", ") + std::to_string(m->device);
    add("\t\tdevice group: %zu members on devices [%s], tile columns dealt round-robin, peer access %s\n", c->members.size(), ids.c_str(), c->group_peer ? "yes" : "no");
  }
  add("\t\trole -> texture: Em->%s Ab->%s Re->%s, gradient method: %s\n", names[c->idxEm], names[c->idxAb], names[c->idxRe],
      c->gradientMethod ? "lookup" : "compute");
  add("\n\t\tSimilarity of Volumes\n\t\t---------------------\n\t\t\tEm\tAb\tRe\n\t\tEm\t1\t\n\t\tAb\t%d\t1\n\t\tRe\t%d\t%d\t1\n\n",
      (int)same(c->host[S_EM], c->host[S_AB]), (int)same(c->host[S_EM], c->host[S_RE]), (int)same(c->host[S_AB], c->host[S_RE]));
  if (needed) *needed = s.size();
  if (buf && cap) {
    const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return VR_OK;
}

vr_status vr_sync_volumes(vr_ctx *c, uint64_t T, const vr_volume *em, const vr_volume *re, const vr_volume *ab,
                          const vr_volume *gx, const vr_volume *gy, const vr_volume *gz) {
  VR_CHECK_HANDLE(c);
  VR_NOT_A_MEMBER(c);
  if (c->members.size() < 2) return sync_volumes_one(c, T, em, re, ab, gx, gy, gz);
  // every device of the group holds all volumes (image-space mode): one host thread per device, each over its own PCIe link.
  // A member that has never been synced gets everything (T = 0), whatever the caller's timestamp says.
  std::vector<vr_status> st(c->members.size(), VR_OK);
  std::vector<std::thread> th;
  for (size_t i = 0; i < c->members.size(); ++i)
    th.emplace_back([&, i] {
      vr_ctx *m = c->members[i];
      st[i] = sync_volumes_one(m, m->ever_synced ? T : 0, em, re, ab, gx, gy, gz);
    });
  for (auto &t : th) t.join();
  cudaSetDevice(c->device);
  for (size_t i = 0; i < st.size(); ++i)
    if (st[i] != VR_OK) return i == 0 ? st[i] : fail(c, st[i], "device %d: %s", c->members[i]->device, c->members[i]->err.c_str());
  return VR_OK;
}

extern "C++" {
namespace {
vr_status sync_volumes_one(vr_ctx *c, uint64_t T, const vr_volume *em, const vr_volume *re, const vr_volume *ab, const vr_volume *gx,
                           const vr_volume *gy, const vr_volume *gz) {
  if (!em || !re || !ab) return fail(c, VR_ERR_INVALID_ARGUMENT, "insufficient parameter!");
  const int ng = (gx != nullptr) + (gy != nullptr) + (gz != nullptr);
  if (ng != 0 && ng != 3) return fail(c, VR_ERR_INVALID_ARGUMENT, "gradient volumes: pass all three or none");
  VR_TRY(set_device(c));
  // Inside a pipeline of batches (VolumeRender.m syncs before every render, and nothing is dirty in a movie) the sync must
  // not wait for the frames in flight: uploads, if any, are ordered behind the running launch by the context's stream.
  if (!c->pipe_open) VR_TRY(collect(c));
  c->stats.uploads = 0;
  c->stats.bytes_uploaded = 0;
  c->timeLastMemSync = T;  // render.cpp:94
  c->host[S_EM] = make_volume(em);
  c->host[S_RE] = make_volume(re);
  c->host[S_AB] = make_volume(ab);
  int method = 0;
  if (ng == 3) {  // nrhs == 9 (render.cpp:105-110)
    c->host[S_DX] = make_volume(gx);
    c->host[S_DY] = make_volume(gy);
    c->host[S_DZ] = make_volume(gz);
    method = 1;
  } else {  // nrhs == 6 (render.cpp:111-113)
    resetGradients(c);
  }
  VR_TRY(check_free(c, 0));  // the per-upload check in upload() is the binding one (old copies are freed first)
  (void)required_memory;
  c->gradientMethod = method;  // setGradientMethod (render.cpp:121)
  VR_CUDA(c, cudaEventRecord(c->evu0, c->stream));
  VR_TRY(mm_sync(c));  // render.cpp:122
  VR_CUDA(c, cudaEventRecord(c->evu1, c->stream));
  // host memory is only borrowed for the duration of the call
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, c->evu0, c->evu1);
  c->stats.upload_ms = ms;
  c->stats.bytes_resident = resident_bytes(c);
  c->ever_synced = true;
  return VR_OK;
}
}  // namespace
}  // extern "C++"

extern "C++" {
namespace {
// Device-image render of a context or of a whole device group.  In a group every member launches on its own device and its
// kernel stores the tiles it owns straight into the caller's image on the leader's device (peer stores over NVLink, misses
// as 0, so nobody clears the image): the launch IS the gather.  The caller's stream continues only after every member's
// launch, and no member starts before the work already queued on that stream.
vr_status render_device_group(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination, const float factors[3],
                              const float es[3], const uint64_t res[2], size_t n_frames, const float *rotations, const float *props,
                              float thr, const float color[3], float *const *d_outs, void *cuda_stream) {
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  if (s != c->stream) VR_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->members.size() < 2) return render_common(c, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, d_outs, s);
  if (!c->group_peer) return fail(c, VR_ERR_UNSUPPORTED, "device-image renders of a group need peer access between its devices");
  VR_CUDA(c, cudaEventRecord(c->ev_group, s));
  for (size_t i = 0; i < c->members.size(); ++i) {
    vr_ctx *m = c->members[i];
    VR_TRY(set_device(m));
    if (i) {
      VR_TRY(collect(m));
      VR_CUDA(m, cudaStreamWaitEvent(m->stream, c->ev_group, 0));
    }
    const int64_t clear = m->opt[VR_OPT_CLEAR_IMAGE];
    m->opt[VR_OPT_CLEAR_IMAGE] = 0;  // every owned pixel is written by its owner
    const vr_status st = render_common(m, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, d_outs, i ? m->stream : s);
    m->opt[VR_OPT_CLEAR_IMAGE] = clear;
    if (st != VR_OK) {
      cudaSetDevice(c->device);
      return i == 0 ? st : fail(c, st, "device %d: %s", m->device, m->err.c_str());
    }
    if (i) VR_CUDA(m, cudaEventRecord(m->ev_group, m->stream));
  }
  VR_TRY(set_device(c));
  for (size_t i = 1; i < c->members.size(); ++i) VR_CUDA(c, cudaStreamWaitEvent(s, c->members[i]->ev_group, 0));
  return VR_OK;
}
}  // namespace
}  // extern "C++"

vr_status vr_render_device(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                           const float factors[3], const float es[3], const uint64_t res[2], const float R[9],
                           const float props[3], float thr, const float color[3], float *d_out, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  VR_NOT_A_MEMBER(c);
  if (!d_out) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: null output");
  return render_device_group(c, lights, n_lights, illumination, factors, es, res, 1, R, props, thr, color, &d_out, cuda_stream);
}

vr_status vr_render_batch_device(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                                 const float factors[3], const float es[3], const uint64_t res[2], size_t n_frames,
                                 const float *rotations, const float *props, float thr, const float color[3],
                                 float *const *d_out_images, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  VR_NOT_A_MEMBER(c);
  return render_device_group(c, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, d_out_images, cuda_stream);
}

// One process, several GPUs, behind the reference's own boundary (the reference picks ONE device, volumeRender.cpp:77-87): the
// context becomes the leader of a group with one member context per device.  From then on vr_sync_volumes makes the volumes
// resident on every device, and every render is cut into tile columns dealt round-robin to the devices (VR_OPT_PART_MODE = 1);
// host-image renders have each device copy its own columns into the caller's image over its own PCIe link, device-image
// renders have the kernels of all devices store into the caller's image over NVLink.  Pixels are those of one device, bit
// for bit.  device_ids[0] must be the context's own device; a device may be listed more than once (members then share it).
// n = 0 or 1 dissolves the group.  Call it before the first vr_sync_volumes: a member added later receives the volumes of
// the next sync in full, but not the upload history that decides the role redirects (kernel.cu:739-867).
vr_status vr_set_devices(vr_ctx *c, uint32_t n, const int *device_ids) {
  VR_CHECK_HANDLE(c);
  VR_NOT_A_MEMBER(c);
  if (c->region_on) return fail(c, VR_ERR_UNSUPPORTED, "a context that owns a region (vr_set_region) cannot lead a device group");
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  // dissolve what is there
  for (size_t i = 1; i < c->members.size(); ++i) {
    c->members[i]->leader = nullptr;
    vr_destroy(c->members[i]);
  }
  c->members.clear();
  VR_TRY(set_device(c));
  c->opt[VR_OPT_PART_RANK] = 0, c->opt[VR_OPT_PART_NRANKS] = 1, c->opt[VR_OPT_PART_MODE] = 0;
  if (n <= 1) {
    if (n == 1 && device_ids && device_ids[0] != c->device) return fail(c, VR_ERR_INVALID_ARGUMENT, "set_devices: the first device must be the context's own (%d)", c->device);
    return VR_OK;
  }
  if (!device_ids || n > 64) return fail(c, VR_ERR_INVALID_ARGUMENT, "set_devices: bad arguments");
  if (device_ids[0] != c->device) return fail(c, VR_ERR_INVALID_ARGUMENT, "set_devices: the first device must be the context's own (%d)", c->device);
  int ndev = 0;
  VR_CUDA(c, cudaGetDeviceCount(&ndev));
  for (uint32_t i = 0; i < n; ++i)
    if (device_ids[i] < 0 || device_ids[i] >= ndev) return fail(c, VR_ERR_NO_DEVICE, "set_devices: device %d of %d does not exist", device_ids[i], ndev);
  // peer access between every pair of distinct devices (kernels store into the leader's image, read device-memory sources)
  c->group_peer = true;
  for (uint32_t a = 0; a < n; ++a)
    for (uint32_t b = 0; b < n; ++b) {
      if (device_ids[a] == device_ids[b]) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, device_ids[a], device_ids[b]);
      if (!can) {
        c->group_peer = false;
        continue;
      }
      cudaSetDevice(device_ids[a]);
      const cudaError_t e = cudaDeviceEnablePeerAccess(device_ids[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) c->group_peer = false;
      cudaGetLastError();
    }
  VR_TRY(set_device(c));
  if (!c->ev_group) VR_CUDA(c, cudaEventCreateWithFlags(&c->ev_group, cudaEventDisableTiming));
  c->members.push_back(c);
  for (uint32_t i = 1; i < n; ++i) {
    vr_ctx *m = nullptr;
    const vr_status st = vr_create_on_device(&m, device_ids[i]);
    cudaError_t ce = cudaSuccess;
    if (st == VR_OK) ce = cudaEventCreateWithFlags(&m->ev_group, cudaEventDisableTiming);
    if (st != VR_OK || ce != cudaSuccess) {
      const std::string why = st != VR_OK ? tl_error : std::string(cudaGetErrorString(ce));
      if (m) vr_destroy(m);
      for (size_t k = 1; k < c->members.size(); ++k) {
        c->members[k]->leader = nullptr;
        vr_destroy(c->members[k]);
      }
      c->members.clear();
      cudaSetDevice(c->device);
      return fail(c, st != VR_OK ? st : VR_ERR_CUDA, "set_devices: device %d: %s", device_ids[i], why.c_str());
    }
    m->leader = c;
    memcpy(m->opt, c->opt, sizeof m->opt);
    c->members.push_back(m);
  }
  for (uint32_t i = 0; i < n; ++i) {
    vr_ctx *m = c->members[i];
    m->opt[VR_OPT_PART_RANK] = i, m->opt[VR_OPT_PART_NRANKS] = n, m->opt[VR_OPT_PART_MODE] = 1;
  }
  VR_TRY(set_device(c));
  return VR_OK;
}

vr_status vr_set_region(vr_ctx *c, const uint64_t lo[3], const uint64_t hi[3], uint32_t halo) {
  VR_CHECK_HANDLE(c);
  VR_NOT_A_MEMBER(c);
  VR_NOT_A_GROUP(c, "vr_set_region");
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  if (!lo || !hi || (hi[0] == 0 && hi[1] == 0 && hi[2] == 0)) {
    c->region_on = false;
    return VR_OK;
  }
  for (int i = 0; i < 3; ++i) {
    if (lo[i] >= hi[i] || hi[i] > 0x3fffffff) return fail(c, VR_ERR_INVALID_ARGUMENT, "set_region: empty or oversized box on axis %d", i);
    c->rlo[i] = (int64_t)lo[i], c->rhi[i] = (int64_t)hi[i];
  }
  c->halo = halo ? (int)halo : 4;
  if (c->halo < 2 || c->halo > 64) return fail(c, VR_ERR_INVALID_ARGUMENT, "set_region: halo must be in [2, 64]");
  c->region_on = true;
  return VR_OK;
}

// Partial update of a resident volume (example3.m:135-161 rewrites a masked half of
// the volume every frame and the reference uploads all of it again, kernel.cu:659-672).
vr_status vr_update_box(vr_ctx *c, int role, const vr_volume *vol, const uint64_t lo[3], const uint64_t hi[3]) {
  VR_CHECK_HANDLE(c);
  if (!vol || !lo || !hi) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_update_box: null argument");
  if (role < 0 || role > S_DZ) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_update_box: role %d (0 emission, 1 absorption, 2 reflection, 3..5 gradient x,y,z)", role);
  if (!c->leader && c->members.size() > 1) {  // a group: the box goes to every device
    for (size_t i = 1; i < c->members.size(); ++i) {
      vr_ctx *m = c->members[i];
      m->leader = nullptr;
      const vr_status st = vr_update_box(m, role, vol, lo, hi);
      m->leader = c;
      if (st != VR_OK) return fail(c, st, "device %d: %s", m->device, m->err.c_str());
    }
  }
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  // The slot that HOLDS this host volume: the role's own texture when one was uploaded for it (a distinct absorption
  // volume sits in tex[S_AB] although, as written, d_idxAbsorption keeps pointing at emission, kernel.cu:80), else the
  // texture the role was aliased to by referenceTexture (kernel.cu:631-648; roles that alias it see the update too).
  int slot = role;
  if (role <= S_RE) {
    const int idx = role == S_EM ? c->idxEm : role == S_AB ? c->idxAb : c->idxRe;
    const bool own = c->arr[role] && c->tex[role].kind != TEX_UNBOUND && same(c->tex[role].src, c->host[role]);
    if (!own && idx != role && same(c->host[role], c->host[idx])) slot = idx;
  }
  Resident &r = c->tex[slot];
  const HostVol v = make_volume(vol);
  if (r.kind != TEX_VOLUME) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_update_box: the role has no resident volume (sync it first)");
  if ((int)v.w != r.w || (int)v.h != r.h || (int)v.d != r.d || v.dtype != r.dtype)
    return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_update_box: the volume must have the dims and dtype of the resident one (%d x %d x %d)", r.w, r.h, r.d);
  if (v.memspace == VR_MEM_DEVICE_REGION && !c->region_on) return fail(c, VR_ERR_INVALID_ARGUMENT, "a region buffer needs vr_set_region first");
  int box[6];
  const int dims[3] = {r.w, r.h, r.d};
  for (int i = 0; i < 3; ++i) {
    if (lo[i] >= hi[i] || hi[i] > (uint64_t)dims[i]) return fail(c, VR_ERR_INVALID_ARGUMENT, "vr_update_box: box [%llu,%llu) does not fit axis %d (%d voxels)", (unsigned long long)lo[i], (unsigned long long)hi[i], i, dims[i]);
    box[i] = (int)lo[i], box[3 + i] = (int)hi[i];
  }
  c->stats.bytes_uploaded = 0;
  VR_TRY(restore_storage(c, r, c->stream));  // (dropped by a texture-unit render: rebuilt from the array once)
  r.keep_storage = true;                     // a volume that is updated in place keeps both copies from now on
  VR_CUDA(c, cudaEventRecord(c->evu0, c->stream));
  VR_TRY(fill_resident(c, r, v, box));
  // what the slot now holds: the caller's volume at its new timestamp (the next sync must not upload it again)
  r.src = v;
  const HostVol before = c->host[role];
  for (int s = 0; s <= S_DZ; ++s)
    if (s == role || (s <= S_RE && role <= S_RE && same(c->host[s], before))) c->host[s] = v;  // aliased roles stay aliased
  // the empty-space grids describe the old contents: rebuilt on the next render that needs them
  if (r.occ) cudaFree(r.occ);
  if (r.dist) cudaFree(r.dist);
  r.occ = nullptr, r.dist = nullptr, r.occ_bytes = r.dist_bytes = 0;
  VR_CUDA(c, cudaEventRecord(c->evu1, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, c->evu0, c->evu1);
  c->stats.upload_ms = ms;
  c->stats.uploads = 1;
  return VR_OK;
}

vr_status vr_render_opacity_device(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                                   const float factors[3], const float es[3], const uint64_t res[2], const float R[9],
                                   const float props[3], float thr, const float color[3], float *d_opacity, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  if (!d_opacity) return fail(c, VR_ERR_INVALID_ARGUMENT, "opacity pass: null output");
  VR_NOT_A_MEMBER(c);
  VR_NOT_A_GROUP(c, "the volume-space opacity pass");
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  if (s != c->stream) VR_CUDA(c, cudaStreamSynchronize(c->stream));
  VsArgs vs{};
  vs.mode = 1;
  vs.opacity_out = d_opacity;
  float *dummy = d_opacity;  // no image is written in this pass; any non-null pointer passes the argument check
  const int64_t clear = c->opt[VR_OPT_CLEAR_IMAGE];
  c->opt[VR_OPT_CLEAR_IMAGE] = 0;
  // The opacity pass counts into launch slot 1: the shading pass that follows (slot 0) then finds nothing pending and does not
  // have to wait for this pass and the exchange behind it before it can be set up (the counters of slot 1 are never read).
  c->cur = 1;
  const vr_status st = render_common(c, lights, n_lights, illumination, factors, es, res, 1, R, props, thr, color, &dummy, s, &vs);
  c->cur = 0;
  c->slot[1].stats_pending = c->slot[1].timing_pending = false;
  c->opt[VR_OPT_CLEAR_IMAGE] = clear;
  return st;
}

vr_status vr_render_partial_device(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                                   const float factors[3], const float es[3], const uint64_t res[2], const float R[9],
                                   const float props[3], float thr, const float color[3], uint32_t n_bricks, uint32_t brick_index,
                                   const uint64_t *brick_boxes, const float *d_opacity_all, float *d_out, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  if (!d_out) return fail(c, VR_ERR_INVALID_ARGUMENT, "partial render: null output");
  VR_NOT_A_MEMBER(c);
  VR_NOT_A_GROUP(c, "the volume-space shading pass");
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  if (s != c->stream) VR_CUDA(c, cudaStreamSynchronize(c->stream));
  VsArgs vs{};
  vs.mode = 2, vs.index = (int)brick_index, vs.count = (int)n_bricks;
  vs.opacity_all = d_opacity_all, vs.boxes_vox = brick_boxes;
  return render_common(c, lights, n_lights, illumination, factors, es, res, 1, R, props, thr, color, &d_out, s, &vs);
}

vr_status vr_compose_partials_device(vr_ctx *c, uint32_t n, const float *const *d_partials, uint64_t n_floats, float *d_out, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  if (!d_partials || !d_out || n == 0 || n > 64) return fail(c, VR_ERR_INVALID_ARGUMENT, "compose: bad arguments");
  VR_TRY(set_device(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  VR_CUDA(c, launch_compose(d_partials, (int)n, n_floats, d_out, s));
  return VR_OK;
}

extern "C++" {
namespace {
// the persistent device images of slot `S` for a batch of host-image frames
vr_status slot_images(vr_ctx *c, LaunchSlot &S, size_t bytes, size_t n_frames, std::vector<float *> &d_outs) {
  if (S.image_cap < bytes * n_frames) {
    if (S.d_image) cudaFree(S.d_image);
    S.d_image = nullptr, S.image_cap = 0;
    VR_CUDA(c, cudaMalloc(&S.d_image, bytes * n_frames));
    S.image_cap = bytes * n_frames;
  }
  d_outs.resize(n_frames);
  for (size_t f = 0; f < n_frames; ++f) d_outs[f] = S.d_image + f * (bytes / sizeof(float));
  return VR_OK;
}

// One pipelined batch on ONE context: launch on the context's stream into the images of the next slot, copies of the owned
// pixels on the copy stream behind it.  Returns without waiting.
vr_status batch_begin_one(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination, const float factors[3],
                          const float es[3], const uint64_t res[2], size_t n_frames, const float *rotations, const float *props, float thr,
                          const float color[3], float *const *out_images) {
  if (!out_images || !res || n_frames == 0 || n_frames > 4096) return fail(c, VR_ERR_INVALID_ARGUMENT, "insufficient parameter!");
  for (size_t f = 0; f < n_frames; ++f)
    if (!out_images[f]) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: null output");
  VR_TRY(set_device(c));
  const size_t bytes = (size_t)res[0] * res[1] * 3 * sizeof(float);
  if (bytes == 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "render: bad resolution");
  if (!c->pipe_open) {
    VR_TRY(collect(c));  // a synchronous render may still have counters pending
    c->pipe_open = true, c->pipe_batches = 0;
    c->pipe_total = vr_stats{};
  }
  const int k = (int)(c->pipe_batches & 1);
  LaunchSlot &S = c->slot[k];
  VR_TRY(harvest(c, S));  // batch k-2: its frames have left this slot's images (normally long ago)
  c->cur = k;
  std::vector<float *> d_outs;
  vr_status st = slot_images(c, S, bytes, n_frames, d_outs);
  if (st == VR_OK) st = render_common(c, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, d_outs.data(), c->stream);
  if (st != VR_OK) {
    c->cur = 0;
    return st;
  }
  S.launches = c->stats.kernel_launches;
  VR_CUDA(c, cudaEventRecord(S.ev_rendered, c->stream));
  VR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, S.ev_rendered, 0));
  for (size_t f = 0; f < n_frames; ++f) VR_TRY(enqueue_download(c, out_images[f], d_outs[f], res[1], res[0], c->copy_stream));
  VR_CUDA(c, cudaEventRecord(S.ev_copied, c->copy_stream));
  // (the next launch into this slot's images is batch k+2, whose begin waits for these copies in harvest() above)
  S.in_flight = true;
  c->pipe_batches++;
  c->cur = 0;
  return VR_OK;
}
}  // namespace
}  // extern "C++"

// Pipelined movie frames (examples/example2.m:59-66 renders 360 frames of one scene): vr_render_batch_begin enqueues the
// launch of a batch and the device->host copies of its frames and returns at once; the next begin renders into a second
// set of device images while those copies are on the wire.  out_images must stay valid (and should be pinned: a pageable
// destination makes the copy synchronous) until vr_render_batch_end, which waits for everything begun and leaves the summed
// counters in vr_get_stats.  Any other call on the context ends the pipeline first.
vr_status vr_render_batch_begin(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination, const float factors[3],
                                const float es[3], const uint64_t res[2], size_t n_frames, const float *rotations, const float *props,
                                float thr, const float color[3], float *const *out_images) {
  VR_CHECK_HANDLE(c);
  if (c->leader) return fail(c, VR_ERR_INVALID_ARGUMENT, "this context is a member of a device group: call its leader");
  for (vr_ctx *m : group_of(c)) {
    const vr_status st = batch_begin_one(m, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, out_images);
    if (st != VR_OK) return m == c ? st : fail(c, st, "device %d: %s", m->device, m->err.c_str());
  }
  return VR_OK;
}
vr_status vr_render_batch_end(vr_ctx *c) {
  VR_CHECK_HANDLE(c);
  if (c->leader) return fail(c, VR_ERR_INVALID_ARGUMENT, "this context is a member of a device group: call its leader");
  for (vr_ctx *m : group_of(c)) {
    VR_TRY(set_device(m));
    const vr_status st = drain_pipeline(m);
    if (st != VR_OK) return m == c ? st : fail(c, st, "device %d: %s", m->device, m->err.c_str());
  }
  VR_TRY(set_device(c));
  c->stats.bytes_resident = resident_bytes(c);
  return VR_OK;
}

// The synchronous host-image renders are one pipelined batch, waited for.
vr_status vr_render_batch(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                          const float factors[3], const float es[3], const uint64_t res[2], size_t n_frames,
                          const float *rotations, const float *props, float thr, const float color[3], float *const *out_images) {
  VR_CHECK_HANDLE(c);
  if (c->leader) return fail(c, VR_ERR_INVALID_ARGUMENT, "this context is a member of a device group: call its leader");
  for (vr_ctx *m : group_of(c)) {
    VR_TRY(set_device(m));
    VR_TRY(collect(m));  // ends an open pipeline: this batch's counters stand alone
  }
  const vr_status st = vr_render_batch_begin(c, lights, n_lights, illumination, factors, es, res, n_frames, rotations, props, thr, color, out_images);
  const vr_status se = vr_render_batch_end(c);  // also after a failed begin: members that did start are waited for
  return st != VR_OK ? st : se;
}

vr_status vr_render(vr_ctx *c, const vr_light *lights, size_t n_lights, const vr_volume *illumination,
                    const float factors[3], const float es[3], const uint64_t res[2], const float R[9],
                    const float props[3], float thr, const float color[3], float *out) {
  if (!out) {
    VR_CHECK_HANDLE(c);
    return fail(c, VR_ERR_INVALID_ARGUMENT, "insufficient parameter!");
  }
  // cudaDeviceSynchronize + cudaMemcpy D2H (volumeRender.cpp:280-281)
  float *outs[1] = {out};
  return vr_render_batch(c, lights, n_lights, illumination, factors, es, res, 1, R, props, thr, color, outs);
}

vr_status vr_henyey_greenstein(uint32_t n, float g, float *out) {
  if (!out || n == 0 || n > 1024) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "HenyeyGreenstein: bad size");
  if (g > 1.0f || g < -1.0f) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "g must be in interval [-1,1]");  // HenyeyGreenstein.cc:44-45
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(nullptr, VR_ERR_NO_DEVICE, "no CUDA device; there is no CPU fallback");
  float *d = nullptr;
  const size_t bytes = (size_t)n * n * n * sizeof(float);
  VR_CUDA(nullptr, cudaMalloc(&d, bytes));
  cudaError_t e = launch_hg_lut(d, n, g, nullptr);
  if (e == cudaSuccess) e = cudaMemcpy(out, d, bytes, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return fail(nullptr, VR_ERR_CUDA, "HenyeyGreenstein: %s", cudaGetErrorString(e));
  return VR_OK;
}

// vr_henyey_greenstein without any allocation: the device scratch is a static array of the library (resident_kernels.cu).
vr_status vr_henyey_greenstein_static(uint32_t n, float g, float *out, void *cuda_stream) {
  static_assert(VR_HG_STATIC_MAX_N == HG_STATIC_MAX_N, "header and kernel file disagree on the scratch size");
  if (!out || n == 0) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "HenyeyGreenstein: bad size");
  if (n > VR_HG_STATIC_MAX_N)
    return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "vr_henyey_greenstein_static: n = %u exceeds the static scratch (n <= %d, %d^3 floats); use vr_henyey_greenstein",
                n, (int)VR_HG_STATIC_MAX_N, (int)VR_HG_STATIC_MAX_N);
  if (g > 1.0f || g < -1.0f) return fail(nullptr, VR_ERR_INVALID_ARGUMENT, "g must be in interval [-1,1]");  // HenyeyGreenstein.cc:44-45
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(nullptr, VR_ERR_NO_DEVICE, "no CUDA device; there is no CPU fallback");
  cudaStream_t s = (cudaStream_t)cuda_stream;
  float *d = nullptr;
  cudaError_t e = launch_hg_lut_static(n, g, s, &d);
  const size_t bytes = (size_t)n * n * n * sizeof(float);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, d, bytes, cudaMemcpyDefault, s);  // `out`: host or device memory of the caller
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return fail(nullptr, VR_ERR_CUDA, "vr_henyey_greenstein_static: %s", cudaGetErrorString(e));
  return VR_OK;
}

uint64_t vr_timestamp_ms(int truncate_like_reference) {
  const uint64_t ms = (uint64_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
  return truncate_like_reference ? (ms & 0xffffffffull) : ms;
}

const char *vr_last_error(const vr_ctx *c) {
  if (c && live(c)) return c->err.c_str();
  return tl_error.c_str();
}
const char *vr_status_string(vr_status s) {
  switch (s) {
  case VR_OK: return "VR_OK";
  case VR_ERR_INVALID_ARGUMENT: return "VR_ERR_INVALID_ARGUMENT";
  case VR_ERR_INVALID_HANDLE: return "VR_ERR_INVALID_HANDLE";
  case VR_ERR_INSUFFICIENT_VRAM: return "VR_ERR_INSUFFICIENT_VRAM";
  case VR_ERR_CUDA: return "VR_ERR_CUDA";
  case VR_ERR_NO_DEVICE: return "VR_ERR_NO_DEVICE";
  case VR_ERR_UNSUPPORTED: return "VR_ERR_UNSUPPORTED";
  }
  return "?";
}
int vr_abi_version(void) { return VR_ABI_VERSION; }

vr_status vr_set_option(vr_ctx *c, int option, int64_t value) {
  VR_CHECK_HANDLE(c);
  if (option < 0 || option >= VR_OPT_COUNT_) return fail(c, VR_ERR_INVALID_ARGUMENT, "unknown option %d", option);
  if (option == VR_OPT_FILTER_FRAC_BITS && value != 0 && value != 8) return fail(c, VR_ERR_INVALID_ARGUMENT, "filter frac bits must be 0 or 8");
  if (option == VR_OPT_ARITHMETIC && (value < 0 || value > 2)) return fail(c, VR_ERR_INVALID_ARGUMENT, "arithmetic flavour must be 0 (exact), 1 (reference) or 2 (fast)");
  if (option == VR_OPT_RAY_SLICES && (value < 1 || value > 16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "ray slices must be in [1, 16]");
  if (option == VR_OPT_VS_SLICES && (value < 0 || value > 16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "depth slices must be in [0, 16] (0 = automatic)");
  if (c->members.size() > 1) {
    if (option == VR_OPT_PART_RANK || option == VR_OPT_PART_NRANKS || option == VR_OPT_PART_MODE)
      return fail(c, VR_ERR_INVALID_ARGUMENT, "the partition of a device group is set by vr_set_devices");
    for (size_t i = 1; i < c->members.size(); ++i) c->members[i]->opt[option] = value;
  }
  c->opt[option] = value;
  return VR_OK;
}
vr_status vr_get_option(vr_ctx *c, int option, int64_t *value) {
  VR_CHECK_HANDLE(c);
  if (option < 0 || option >= VR_OPT_COUNT_ || !value) return fail(c, VR_ERR_INVALID_ARGUMENT, "unknown option %d", option);
  *value = c->opt[option];
  return VR_OK;
}
vr_status vr_get_stats(vr_ctx *c, vr_stats *out) {
  VR_CHECK_HANDLE(c);
  if (!out) return fail(c, VR_ERR_INVALID_ARGUMENT, "null stats");
  for (vr_ctx *m : group_of(c)) {
    VR_TRY(set_device(m));
    VR_TRY(collect(m));
    // a synchronisation point by contract: everything the context has enqueued is complete when the counters are returned
    // (an opacity pass counts into a slot nobody reads, so `collect` alone may find nothing to wait for)
    VR_CUDA(m, cudaStreamSynchronize(m->stream));
    m->stats.bytes_resident = resident_bytes(m);
  }
  VR_TRY(set_device(c));
  *out = merged_stats(c);
  return VR_OK;
}
vr_status vr_set_stream(vr_ctx *c, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  VR_TRY(collect(c));
  if (c->own_stream) cudaStreamDestroy(c->stream);
  c->own_stream = false;
  c->stream = (cudaStream_t)cuda_stream;
  return VR_OK;
}
vr_status vr_get_device(vr_ctx *c, int *device) {
  VR_CHECK_HANDLE(c);
  if (device) *device = c->device;
  return VR_OK;
}

vr_status vr_synth_volume(vr_ctx *c, int kind, uint32_t seed, const uint64_t dims[3], int32_t dtype, void *d_out) {
  VR_CHECK_HANDLE(c);
  if (!dims || !d_out || kind < 0 || kind > 1 || (dtype != VR_DTYPE_F32 && dtype != VR_DTYPE_BF16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "synth: bad argument");
  VR_TRY(set_device(c));
  VR_CUDA(c, launch_synth(d_out, kind, seed, (int)dims[0], (int)dims[1], (int)dims[2], dtype, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  return VR_OK;
}
vr_status vr_synth_volume_region(vr_ctx *c, int kind, uint32_t seed, const uint64_t dims[3], const uint64_t lo[3], const uint64_t hi[3],
                                 int32_t dtype, void *d_out) {
  VR_CHECK_HANDLE(c);
  if (!dims || !lo || !hi || !d_out || kind < 0 || kind > 1 || (dtype != VR_DTYPE_F32 && dtype != VR_DTYPE_BF16)) return fail(c, VR_ERR_INVALID_ARGUMENT, "synth: bad argument");
  for (int i = 0; i < 3; ++i)
    if (lo[i] >= hi[i] || hi[i] > dims[i]) return fail(c, VR_ERR_INVALID_ARGUMENT, "synth: bad box on axis %d", i);
  VR_TRY(set_device(c));
  const int d3[3] = {(int)dims[0], (int)dims[1], (int)dims[2]}, l3[3] = {(int)lo[0], (int)lo[1], (int)lo[2]}, h3[3] = {(int)hi[0], (int)hi[1], (int)hi[2]};
  VR_CUDA(c, launch_synth_region(d_out, kind, seed, d3, l3, h3, dtype, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  return VR_OK;
}
vr_status vr_volume_gradient(vr_ctx *c, const vr_volume *in, int32_t out_memspace, void *gx, void *gy, void *gz) {
  VR_CHECK_HANDLE(c);
  if (!in || !in->data || !gx || !gy || !gz) return fail(c, VR_ERR_INVALID_ARGUMENT, "gradient: null argument");
  if (in->dtype != VR_DTYPE_F32 && in->dtype != VR_DTYPE_BF16) return fail(c, VR_ERR_INVALID_ARGUMENT, "gradient: bad dtype");
  VR_TRY(set_device(c));
  const uint64_t w = in->dims[0], h = in->dims[1], d = in->dims[2];
  if (w == 0 || h == 0 || d == 0 || w > 0x7fffffff || h > 0x7fffffff || d > 0x7fffffff) return fail(c, VR_ERR_INVALID_ARGUMENT, "gradient: bad dims");
  const size_t bytes = (size_t)w * h * d * elem_size(in->dtype);
  void *tmp[4] = {nullptr, nullptr, nullptr, nullptr};
  auto cleanup = [&]() {
    for (void *p : tmp)
      if (p) cudaFree(p);
  };
  const void *d_in = in->data;
  void *d_out[3] = {gx, gy, gz};
  cudaError_t e = cudaSuccess;
  if (in->memspace == VR_MEM_HOST) {
    e = cudaMalloc(&tmp[0], bytes);
    if (e == cudaSuccess) e = cudaMemcpyAsync(tmp[0], in->data, bytes, cudaMemcpyHostToDevice, c->stream);
    d_in = tmp[0];
  }
  if (out_memspace == VR_MEM_HOST)
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) {
      e = cudaMalloc(&tmp[1 + i], bytes);
      d_out[i] = tmp[1 + i];
    }
  if (e == cudaSuccess) e = launch_gradient(d_in, (int)w, (int)h, (int)d, in->dtype, d_out[0], d_out[1], d_out[2], c->stream);
  if (out_memspace == VR_MEM_HOST) {
    void *host[3] = {gx, gy, gz};
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) e = cudaMemcpyAsync(host[i], d_out[i], bytes, cudaMemcpyDeviceToHost, c->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cleanup();
  if (e != cudaSuccess)
    return fail(c, e == cudaErrorMemoryAllocation ? VR_ERR_INSUFFICIENT_VRAM : VR_ERR_CUDA, "gradient: %s", cudaGetErrorString(e));
  return VR_OK;
}
vr_status vr_ipc_export(vr_ctx *c, void *d_ptr, unsigned char handle[VR_IPC_HANDLE_BYTES]) {
  VR_CHECK_HANDLE(c);
  if (!d_ptr || !handle) return fail(c, VR_ERR_INVALID_ARGUMENT, "ipc_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == VR_IPC_HANDLE_BYTES, "handle size");
  VR_TRY(set_device(c));
  cudaIpcMemHandle_t h;
  VR_CUDA(c, cudaIpcGetMemHandle(&h, d_ptr));
  memcpy(handle, &h, sizeof h);
  return VR_OK;
}
vr_status vr_ipc_open(vr_ctx *c, const unsigned char handle[VR_IPC_HANDLE_BYTES], void **d_ptr) {
  VR_CHECK_HANDLE(c);
  if (!d_ptr || !handle) return fail(c, VR_ERR_INVALID_ARGUMENT, "ipc_open: null argument");
  VR_TRY(set_device(c));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  VR_CUDA(c, cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return VR_OK;
}
vr_status vr_ipc_close(vr_ctx *c, void *d_ptr) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  VR_CUDA(c, cudaIpcCloseMemHandle(d_ptr));
  return VR_OK;
}
vr_status vr_compose_stereo_device(vr_ctx *c, const float *d_left, const float *d_right, const uint64_t res_wide[2], uint64_t delta, int mode,
                                   float *d_out, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  if (!d_left || !d_right || !d_out || !res_wide || (mode != 0 && mode != 1) || delta >= res_wide[1]) return fail(c, VR_ERR_INVALID_ARGUMENT, "compose_stereo: bad arguments");
  VR_TRY(set_device(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  VR_CUDA(c, launch_stereo(d_left, d_right, (int)res_wide[0], (int)res_wide[1], (int)delta, mode, d_out, s));
  return VR_OK;
}
vr_status vr_normalize_sequence_device(vr_ctx *c, float *const *d_images, uint32_t n_images, uint64_t n_floats_each, void *cuda_stream) {
  VR_CHECK_HANDLE(c);
  if (!d_images || n_images == 0 || n_floats_each == 0) return fail(c, VR_ERR_INVALID_ARGUMENT, "normalize: bad arguments");
  VR_TRY(set_device(c));
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->stream;
  VR_CUDA(c, launch_normalize(d_images, (int)n_images, (size_t)n_floats_each, reinterpret_cast<unsigned *>(c->d_counter + 8), s));
  return VR_OK;
}
vr_status vr_device_upload(vr_ctx *c, void *d_dst, const void *host_src, size_t bytes) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  VR_CUDA(c, cudaMemcpyAsync(d_dst, host_src, bytes, cudaMemcpyHostToDevice, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  return VR_OK;
}
vr_status vr_probe_gather(vr_ctx *c, int kind, int log2_edge, double *fetches_per_s, double *ms) {
  VR_CHECK_HANDLE(c);
  if (!fetches_per_s || !ms || kind < 0 || kind > 4 || log2_edge < 4 || log2_edge > 10) return fail(c, VR_ERR_INVALID_ARGUMENT, "probe_gather: bad arguments");
  VR_TRY(set_device(c));
  VR_CUDA(c, launch_probe_gather(kind, log2_edge, c->sm_count, c->stream, fetches_per_s, ms));
  return VR_OK;
}
vr_status vr_device_alloc(vr_ctx *c, size_t bytes, void **d_ptr) {
  VR_CHECK_HANDLE(c);
  if (!d_ptr) return fail(c, VR_ERR_INVALID_ARGUMENT, "null pointer");
  VR_TRY(set_device(c));
  VR_CUDA(c, cudaMalloc(d_ptr, bytes));
  return VR_OK;
}
vr_status vr_device_free(vr_ctx *c, void *d_ptr) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  VR_CUDA(c, cudaFree(d_ptr));
  return VR_OK;
}
vr_status vr_debug_warp_times(vr_ctx *c, uint64_t *host_pairs, size_t max_warps) {
  VR_CHECK_HANDLE(c);
#ifdef VR_TRACE_PIXEL
  if (max_warps > 4 * 8192) max_warps = 4 * 8192;
  memcpy(host_pairs, g_trace_host, max_warps * 16);  // no CUDA call: readable after a fault
  return VR_OK;
#endif
  if (!c->d_warp_times) return fail(c, VR_ERR_UNSUPPORTED, "set VR_WARP_TIMES=1 before vr_create");
  VR_TRY(set_device(c));
  if (max_warps > 4 * 8192) max_warps = 4 * 8192;  // (beyond 8192: the item records of a VR_ITEM_PROFILE build)
  VR_CUDA(c, cudaMemcpyAsync(host_pairs, c->d_warp_times, max_warps * 16, cudaMemcpyDeviceToHost, c->stream));
  VR_CUDA(c, cudaStreamSynchronize(c->stream));
  return VR_OK;
}
vr_status vr_device_download(vr_ctx *c, void *host_dst, const void *d_src, size_t bytes) {
  VR_CHECK_HANDLE(c);
  VR_TRY(set_device(c));
  if (!host_dst || !d_src) return fail(c, VR_ERR_INVALID_ARGUMENT, "null pointer");
  return download_frame(c, host_dst, d_src, bytes);
}

}  // extern "C"
