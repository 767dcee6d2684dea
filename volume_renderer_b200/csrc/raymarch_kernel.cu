// raymarch_kernel.cu - the ray-march kernel for sm_100a and its launcher.
//
// Replaces d_render<<<grid,16x16>>> (src/C/vr/volumeRender_kernel.cu:365-507, launched at :590).
// Design:
//  * persistent grid: (SMs x resident blocks) blocks of 8 warps; every warp pulls 8x4-pixel
//    work items from a global atomic counter, so long and short rays balance across the chip;
//  * per warp a "while-while" loop: lanes first run the cheap empty-space phase (serial
//    t/pos recurrences only) until they stand on a sample that can contribute, then all such
//    lanes execute the shading phase together; __any_sync ends the tile when every lane has
//    terminated (early ray termination or t > tfar);
//  * volumes are padded (bricked) arrays read through the read-only path; no tensor cores, the
//    path is gather + fp32 accumulate.
//
// This file is compiled once per gradient method and arithmetic flavour (Makefile): -DVR_GRAD=1 -> raymarch_lookup_a*.o
// (lookupGradient kernels), -DVR_GRAD=0 -> raymarch_compute_a*.o (computeGradient kernels); -DVR_ARITH=0/1/2 with
// -fmad=false for flavour 0 (raymarch_core.h).  Separate modules keep ptxas' allocation of one family independent of
// edits to the other and the build parallel.
#include <cuda_runtime.h>

#ifndef VR_LUT_ADJACENT
#define VR_LUT_ADJACENT VR_GRAD /* which form of the LUT fetch each family runs faster with (A/B on B200) */
#endif

#ifndef VR_EARLY_RE
#define VR_EARLY_RE VR_GRAD /* reflection taps issued with the emission taps: faster for lookupGradient only (A/B on B200) */
#endif

#ifndef VR_ARITH
#define VR_ARITH 0
#endif

#include "raymarch_core.h"
#include "raymarch_launch.h"

#ifndef VR_GRAD
#error "compile with -DVR_GRAD=0 (computeGradient) or -DVR_GRAD=1 (lookupGradient)"
#endif

namespace vrb {

// Block shape and register budget, measured on B200 (configs[1] / the same with computeGradient; ms per
// 2048^2 frame): the kernel is bound by instruction issue AND by the latency of its gathers, so
// resident warps are worth a few spilled registers.
//   lookupGradient, interleaved Gx,Gy,Gz: 256x3 (80 regs) 25.1 | 128x7 (72) 24.5 | 256x4 (64) 25.1
//   computeGradient                     : 256x3 (80 regs) 27.8 | 128x7 (72) 27.5 | 256x4 (64) 29.3 | 128x5 (96) 28.8
// -DVR_BLOCK_THREADS / -DVR_MIN_BLOCKS override both (A/B builds).
template <int GRAD, int LAYOUT>
struct Tune {
#if defined(VR_BLOCK_THREADS) && defined(VR_MIN_BLOCKS)
  static constexpr int threads = VR_BLOCK_THREADS, min_blocks = VR_MIN_BLOCKS;
#else
  static constexpr int threads = 128, min_blocks = 7;
#endif
};
// texture-unit kernels: no tap arithmetic to keep live, the fetch latency is longer -> more resident warps.
// Measured (ms per frame, configs[1] / computeGradient with 1 light): 128x6 17.6 / 17.0 | 128x7 16.3 / 16.0 |
// 128x8 (64 registers) 15.9 / 15.2 | 256x4 15.9 / 15.2 | 128x10 (48) 16.8 / 17.2 | 128x12 (40) 17.7 / 17.6
// On the final kernel (straight-line fetches): 128x8 15.14 / 14.44 | 128x9 (56 registers, 88 bytes of spills) 15.39 / 14.94.
// The jump thresholds below on the same kernel: 24 lanes / 64 iterations 15.14 | 16 / 32 15.29 | 28 / 128 15.18.
#ifndef VR_TEX_THREADS
#define VR_TEX_THREADS 128
#endif
#ifndef VR_TEX_MIN_BLOCKS
#define VR_TEX_MIN_BLOCKS 8
#endif
template <int GRAD>
struct Tune<GRAD, LAYOUT_TEXHW> {
  static constexpr int threads = VR_TEX_THREADS, min_blocks = VR_TEX_MIN_BLOCKS;
};
static constexpr unsigned kFull = 0xffffffffu;
#ifndef VR_JUMP_LANES
#define VR_JUMP_LANES 24
#endif
#ifndef VR_JUMP_WAIT
#define VR_JUMP_WAIT 64
#endif

// ARITH (= VR_ARITH of this translation unit, raymarch_core.h) only keeps the kernels of the arithmetic flavours apart by name
// SL = 1: the instantiation for volume-space passes cut into depth slices (KernelParams::vs_slices > 1); whole rays - every other
// launch - run the SL = 0 kernels, in which no slice code exists.
template <int GRAD, int LAYOUT, typename E, int UNI, int ARITH, int SL>
__global__ void __launch_bounds__(Tune<GRAD, LAYOUT>::threads, Tune<GRAD, LAYOUT>::min_blocks) k_raymarch(const __grid_constant__ KernelParams P) {
  const int lane = threadIdx.x & 31;
  const F3 bmin = ld3(P.boxmin), ext = ld3(P.boxmax) - bmin;
  const F3 bscale = mk3(1.0f / ext.x, 1.0f / ext.y, 1.0f / ext.z);
  const uint32_t wt_per_tile = (uint32_t)(P.wt_x * P.wt_y);
  unsigned long long tot_marched = 0, tot_shaded = 0, tot_hits = 0;
  unsigned long long t_start = 0;
  if (P.warp_times) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  // whole-ray volume-space passes can time their warps (KernelParams::measure_tail); compiled out of the plain texture-unit kernels
  constexpr bool kTimed = SL == 0 && !(VR_TEX_NO_VS && LAYOUT == LAYOUT_TEXHW && UNI == 0);
  unsigned long long warp_start = 0;
  if (kTimed && P.measure_tail) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(warp_start));

  for (;;) {
#ifdef VR_ITEM_PROFILE  // developer build (tools/build_ab.sh): the longest item of every warp, behind the warp times
    unsigned long long it_start = 0;
    unsigned it_loops = 0, it_jumps = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(it_start));
#endif
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(P.work_counter, 1u);
    w = __shfl_sync(kFull, w, 0);
    if (w >= P.num_work) break;
    // volume-space passes with depth slices: S consecutive work items are the S slices of one warp footprint
    const int vs_mode_w = (VR_TEX_NO_VS && LAYOUT == LAYOUT_TEXHW && UNI == 0) ? 0 : P.vs_mode;
    const uint32_t S = (SL && vs_mode_w && P.vs_slices > 1) ? (uint32_t)P.vs_slices : 1u;
    const uint32_t slice = S > 1 ? w % S : 0u;
    if (S > 1) w /= S;
    // work item -> frame, tile, pixel.  `entry` = frame * tiles_total + tile.
    const uint32_t owned = w / wt_per_tile, sub = w - owned * wt_per_tile;
    const uint32_t entry = P.tile_order ? VR_LDG(P.tile_order + owned) : (uint32_t)P.part_rank + owned * (uint32_t)P.part_nranks;
    uint32_t tile = entry;
    Camera cam = P.cam;
    if (P.n_frames > 1) {  // batch: the camera of this item's frame (uniform per warp, L1/L2 resident)
      const uint32_t frame = entry / P.tiles_total;
      tile = entry - frame * P.tiles_total;
      const Camera *cp = P.cams + frame;
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) cam.m[i][j] = VR_LDG(&cp->m[i][j]);
      cam.out = cp->out;
    }
    const F3 origin = ray_origin(cam);
    const uint32_t ty = tile / (uint32_t)P.ntx, tx = tile - ty * (uint32_t)P.ntx;
    const uint32_t sy = sub / (uint32_t)P.wt_x, sx = sub - sy * (uint32_t)P.wt_x;
    const uint32_t x = tx * P.tile_w + (sx << P.wshift) + (lane & ((1u << P.wshift) - 1u));
    const uint32_t y = ty * P.tile_h + sy * (32u >> P.wshift) + (lane >> P.wshift);

    Ray r;
    bool hit = (x < (uint32_t)P.W) && (y < (uint32_t)P.H) && ray_init(P, cam, x, y, origin, r);
    bool active = hit;
#ifdef VR_TRACE_PIXEL
    r.trace_slot = 0, r.trace_n = 0;
    if (P.warp_times && x >= VR_TRACE_X0 && x < VR_TRACE_X0 + 8 && y >= VR_TRACE_Y0 && y < VR_TRACE_Y0 + 8) r.trace_slot = 1 + (x - VR_TRACE_X0) + 8 * (y - VR_TRACE_Y0);
    if (hit) VR_TRACE(P, r, 1, P.vs_mode, slice);
#endif
    const int vs_mode = vs_mode_w;  // (TEXHW with UNI = 1: region contexts)
    if (S > 1 && hit) {
      // This item's slice [bound(slice), bound(slice + 1)) of the ray parameter.  Its END becomes the ray's far end - the largest
      // float below the bound, so that `t > tfar` (kernel.cu:489-490) is exactly `t >= bound` - and the march below is the
      // whole-ray march, instruction for instruction: the loop tail and the empty-space jump already stop at tfar.
      const float ts1 = vs_slice_bound(vs_slice_cut(P, origin, r, bmin, ext, (int)S), (int)slice + 1, (int)S);
      if (slice > 0) {  // its first sample, from k_slice_states
        const float4 st = __ldg(reinterpret_cast<const float4 *>(P.vs_slice_state) + (size_t)(slice - 1) * ((size_t)P.W * P.H) + ((size_t)x * P.H + y));
        r.t = st.x, r.pos = mk3(st.y, st.z, st.w);
        if (!(st.x < 3.0e38f)) active = false;  // the ray ends before this slice
      }
      if (ts1 < 3.0e38f) r.tfar = fminf(r.tfar, float_below(ts1));
      if (r.t > r.tfar) active = false;  // an empty slice
    }
    if (vs_mode == 2 && hit) {  // start from the opacity of the bricks in front; nothing to do behind the threshold
      float a_in = 0.0f;
      active = vs_opacity_in(P, origin, r, (size_t)x * P.H + y, a_in) && active;
      if (S > 1 && active) {  // ... and of the slices of this brick in front of this one
        float T = 1.0f - a_in;
        const size_t plane = (size_t)P.W * P.H, pixel = (size_t)x * P.H + y;
        for (uint32_t q = 0; q < slice; ++q) T = T * (1.0f - VR_LDG(P.vs_slice_opacity + q * plane + pixel));
        a_in = 1.0f - T;
      }
      r.sw = a_in;
      if (a_in > P.thr) active = false;
    }
    F3 dq = mk3(0, 0, 0), inv_dq = mk3(0, 0, 0);
    if (active && P.skip == 1) skip_setup(P, r, bscale, dq, inv_dq);

    int waited = 0;  // warp-uniform: iterations the oldest postponed jump has been waiting
    while (__any_sync(kFull, active)) {
#ifdef VR_ITEM_PROFILE
      ++it_loops;
#endif
      bool ready = active;  // shades a sample in this iteration
      if (P.skip == 2) {
        // The jump over empty space is ~290 instructions and is usually wanted by one or two lanes
        // at a time; neighbouring rays reach the same empty cell within a few samples of each other.
        // A lane that stands on an empty cell therefore sits out up to VR_JUMP_WAIT iterations (it
        // costs the warp nothing: the others shade meanwhile) so that the lanes jump TOGETHER.
        // Measured on configs[1] (ms/frame): no waiting 27.4 | 8 lanes or 4 iterations 26.8 | 16/8 26.6 |
        // 16/16 26.3 | 24/64 26.0 | 24/256 26.4 | only when every live lane wants to 29.8.
        const bool want = active && on_empty_cell(P, r, bmin, bscale);
        const unsigned wm = __ballot_sync(kFull, want);
        if (wm) {
          const unsigned am = __ballot_sync(kFull, active);
          if (__popc(wm) >= VR_JUMP_LANES || wm == am || waited >= VR_JUMP_WAIT) {
#ifdef VR_ITEM_PROFILE
            ++it_jumps;
#endif
            if (want) ready = active = SL ? skip_empty_jump_t<0>(P, r, bmin, bscale, 3.0e38f) : skip_empty_jump(P, r, bmin, bscale);
            waited = 0;
          } else {
            ++waited;
            if (want) ready = false;
          }
        }
      } else if (P.skip == 1) {
        if (active) ready = active = skip_empty_serial(P, r, bmin, bscale, dq, inv_dq);
      }
      if (ready) {
        VR_TRACE(P, r, 6, 0, 0);
        shade_sample<GRAD, LAYOUT, E, UNI>(P, origin, r);
        VR_TRACE(P, r, 7, __float_as_uint(r.sw), 0);
        active = ray_advance(P, r);
      }
    }
    if (x < (uint32_t)P.W && y < (uint32_t)P.H) {
      // column-major planar H x W x 3 (kernel.cu:496-506); lanes of a warp cover 4 consecutive y.
      // A ray that misses the box leaves the reference's pixel at its memset 0 (volumeRender.cpp:271):
      // the owner of the pixel writes that 0 itself, so the image needs no clear and cam.out may be
      // another GPU's memory (image-space direct send: each rank stores its tiles into the root's image).
      const size_t size = (size_t)P.W * P.H, k = (size_t)x * P.H + y;
      if (hit) VR_TRACE(P, r, 8, k, 0);
      if (vs_mode == 1) {
        if (S > 1) P.vs_slice_opacity[slice * size + k] = hit ? r.sw : 0.0f;
        else P.vs_opacity_out[k] = hit ? r.sw : 0.0f;
      } else {
        float *o = cam.out + (size_t)slice * 3 * size;  // (slice 0 of 1: the image itself)
        o[k] = hit ? r.sx : 0.0f;
        o[k + size] = hit ? r.sy : 0.0f;
        o[k + 2 * size] = hit ? r.sz : 0.0f;
      }
    }
    if (hit) VR_TRACE(P, r, 9, 0, 0);
    if (P.tile_cost) {
      // What this item cost: the next launch hands the tiles with the LONGEST items out first.  An item lasts as long
      // as its longest ray (lanes march in lockstep), and what must not start late is a long item - a tile half
      // outside the volume has a small sum but items as long as any (measured: ordered by the sum, a single 2048^2
      // frame waited 2.8 ms of 25.5 for items that were started after 60 % of the queue).
      const unsigned cost = __reduce_add_sync(kFull, hit ? r.n_shaded : 0u);
      const unsigned longest = __reduce_max_sync(kFull, hit ? r.n_shaded : 0u);
      if (lane == 0) {
        atomicAdd(P.tile_cost + tile, cost + 1u);
        atomicMax(P.tile_cost + P.tiles_total + tile, longest);
      }
    }
#ifdef VR_ITEM_PROFILE
    if (P.warp_times) {
      unsigned long long it_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(it_end));
      const unsigned ms = __reduce_max_sync(kFull, hit ? r.n_shaded : 0u), mm = __reduce_max_sync(kFull, hit ? r.n_marched : 0u);
      const unsigned ss = __reduce_add_sync(kFull, hit ? r.n_shaded : 0u);
      const size_t wid = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
      unsigned long long *rec = P.warp_times + 2 * 8192 + 6 * wid;
      if (lane == 0 && it_end - it_start > rec[0]) {
        rec[0] = it_end - it_start, rec[1] = (unsigned long long)x | ((unsigned long long)y << 16) | ((unsigned long long)slice << 32);
        rec[2] = ms | ((unsigned long long)ss << 32), rec[3] = mm, rec[4] = it_loops, rec[5] = it_jumps;
      }
    }
#endif
    if (hit) {
      tot_marched += r.n_marched;
      tot_shaded += r.n_shaded;
      tot_hits += slice == 0 ? 1 : 0;  // (a pixel's slices are one ray)
    }
  }
  if (kTimed && P.measure_tail && lane == 0) {
    // when the first warp started, when the first warp found the queue empty, when the last one finished: the launch's tail
    unsigned long long warp_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(warp_end));
    atomicMax(P.stats + 3, ~warp_start), atomicMax(P.stats + 4, ~warp_end), atomicMax(P.stats + 5, warp_end);
  }
  if (P.warp_times && lane == 0) {
    unsigned long long t_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
    const size_t wid = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    P.warp_times[2 * wid] = t_start, P.warp_times[2 * wid + 1] = t_end;
  }
  if (P.count) {
    for (int o = 16; o > 0; o >>= 1) {
      tot_marched += __shfl_xor_sync(kFull, tot_marched, o);
      tot_shaded += __shfl_xor_sync(kFull, tot_shaded, o);
      tot_hits += __shfl_xor_sync(kFull, tot_hits, o);
    }
    if (lane == 0) {
      atomicAdd(P.stats + 0, tot_marched);
      atomicAdd(P.stats + 1, tot_shaded);
      atomicAdd(P.stats + 2, tot_hits);
    }
  }
}

#define VR_CAT3_(a, b, c) a##b##c
#define VR_CAT3(a, b, c) VR_CAT3_(a, b, c)
#if !VR_GRAD
// Depth slices: the first lattice sample of slices 1 .. S-1 of every pixel (vr_types.h), one thread per pixel.  Compiled with
// the march kernels of its arithmetic flavour (ray_init must produce the very rays they march).
template <int ARITH>
__global__ void __launch_bounds__(128) k_slice_states(const __grid_constant__ KernelParams P) {
  const F3 bmin = ld3(P.boxmin), ext = ld3(P.boxmax) - bmin;
  const F3 origin = ray_origin(P.cam);
  const size_t plane = (size_t)P.W * P.H;
  const int S = P.vs_slices;
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < plane; k += (size_t)gridDim.x * blockDim.x) {
    const uint32_t x = (uint32_t)(k / (size_t)P.H), y = (uint32_t)(k - (size_t)x * P.H);
    if (P.part_nranks > 1) {  // an image-space partition: only this context's tiles are marched
      const uint32_t tx = x / (uint32_t)P.tile_w, ty = y / (uint32_t)P.tile_h;
      if ((P.part_columns ? tx : ty * (uint32_t)P.ntx + tx) % (uint32_t)P.part_nranks != (uint32_t)P.part_rank) continue;
    }
    Ray r;
    bool alive = ray_init(P, P.cam, x, y, origin, r);
    SliceCut cut;
    cut.a = cut.w = 0.0f, cut.inside = false;
    if (alive) cut = vs_slice_cut(P, origin, r, bmin, ext, S);
    alive = alive && cut.inside;
    for (int s = 1; s < S; ++s) {
      if (alive) alive = fast_forward_to(P, r, vs_slice_bound(cut, s, S));
      reinterpret_cast<float4 *>(P.vs_slice_state)[(size_t)(s - 1) * plane + k] = alive ? make_float4(r.t, r.pos.x, r.pos.y, r.pos.z) : make_float4(3.0e38f, 0.0f, 0.0f, 0.0f);
    }
  }
}
cudaError_t VR_CAT3(launch_slice_states, _a, VR_ARITH)(const KernelParams &P, cudaStream_t stream, int sm_count) {
  const size_t plane = (size_t)P.W * P.H;
  size_t blocks = (plane + 127) / 128;
  if (blocks > (size_t)sm_count * 16) blocks = (size_t)sm_count * 16;
  if (blocks < 1) blocks = 1;
  k_slice_states<VR_ARITH><<<(unsigned)blocks, 128, 0, stream>>>(P);
  return cudaGetLastError();
}
#endif

template <int GRAD, int LAYOUT, typename E, int UNI, int SL>
static cudaError_t launch_sl(const KernelParams &P, cudaStream_t stream, int sm_count) {
  constexpr int kBlockThreads = Tune<GRAD, LAYOUT>::threads;
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_raymarch<GRAD, LAYOUT, E, UNI, VR_ARITH, SL>, kBlockThreads, 0);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  // persistent grid: a whole number of waves, never more warps than work items
  long long blocks = (long long)sm_count * per_sm;
  const long long need = ((long long)P.num_work + (kBlockThreads / 32) - 1) / (kBlockThreads / 32);
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  k_raymarch<GRAD, LAYOUT, E, UNI, VR_ARITH, SL><<<(unsigned)blocks, kBlockThreads, 0, stream>>>(P);
  return cudaGetLastError();
}
template <int GRAD, int LAYOUT, typename E, int UNI>
static cudaError_t launch_one(const KernelParams &P, cudaStream_t stream, int sm_count) {
  if (P.vs_mode && P.vs_slices > 1) {
    // sliced passes exist for the kernels region contexts normally run (shared tap sets / region texture arrays)
    if (UNI == 0) return cudaErrorInvalidValue;
    return launch_sl<GRAD, LAYOUT, E, (UNI == 0 ? 1 : UNI), 1>(P, stream, sm_count);
  }
  return launch_sl<GRAD, LAYOUT, E, UNI, 0>(P, stream, sm_count);
}

template <int LAYOUT, typename E>
static cudaError_t launch_le(const KernelParams &P, cudaStream_t stream, int sm_count) {
  if (P.uniform == 2) {  // 64-bit z terms: bricked layout only (context.cu refuses the other combinations)
    if (LAYOUT == LAYOUT_BRICK) return launch_one<VR_GRAD, LAYOUT_BRICK, E, 2>(P, stream, sm_count);
    return cudaErrorInvalidValue;
  }
  return P.uniform ? launch_one<VR_GRAD, LAYOUT, E, 1>(P, stream, sm_count) : launch_one<VR_GRAD, LAYOUT, E, 0>(P, stream, sm_count);
}

#if VR_GRAD
cudaError_t VR_CAT3(launch_raymarch_lookup, _a, VR_ARITH)(const KernelParams &P, int layout, int dtype, cudaStream_t stream, int sm_count) {
#else
cudaError_t VR_CAT3(launch_raymarch_compute, _a, VR_ARITH)(const KernelParams &P, int layout, int dtype, cudaStream_t stream, int sm_count) {
#endif
  // texture-unit launches: fp32 arrays, per-texture fetches (there are no tap sets to share)
  // (the arrays are fp32 whatever the storage type; uniform == 1 selects the instantiation for contexts that own a region)
  if (layout == LAYOUT_TEXHW) {
    if (P.uniform == 0) return launch_one<VR_GRAD, LAYOUT_TEXHW, ElemF32, 0>(P, stream, sm_count);
    if (P.uniform == 1) return launch_one<VR_GRAD, LAYOUT_TEXHW, ElemF32, 1>(P, stream, sm_count);
    return cudaErrorInvalidValue;
  }
  if (dtype == 0) return layout == LAYOUT_BRICK ? launch_le<LAYOUT_BRICK, ElemF32>(P, stream, sm_count) : launch_le<LAYOUT_LINEAR, ElemF32>(P, stream, sm_count);
  return layout == LAYOUT_BRICK ? launch_le<LAYOUT_BRICK, ElemBF16>(P, stream, sm_count) : launch_le<LAYOUT_LINEAR, ElemBF16>(P, stream, sm_count);
}

}  // namespace vrb
