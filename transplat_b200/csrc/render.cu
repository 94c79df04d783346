// render.cu — per-16x16-tile alpha compositing, forward and backward.
//
// One CTA per (view, tile); grid = N*T, so every view of the call renders in one launch.  The
// tile's slice of the depth-sorted record arrays (3 x float4 per entry, written by
// ranges_gather_kernel) is contiguous in HBM and is streamed through a shared-memory RING of
// kStages stages x kStageEntries entries by a dedicated producer warp with 1-D TMA bulk copies
// (cp.async.bulk -> mbarrier complete_tx).  The 8 consumer warps (one 8x4 pixel block each) run
// DECOUPLED: each waits on a stage's `full` barrier, culls the stage's entries against its own
// pixel block (minimum of the quadratic over the block, exact in effect), composites the survivors, and arrives on the
// stage's `empty` barrier; the producer refills a stage once all 8 warps have released it.  Depth
// order makes consecutive entries spatially clustered, so per-batch CTA-wide barriers left warps
// idle ~40% of the time; the ring lets them drift up to kStages-1 stages apart.
// Colour (3 channels) and the depth feature (4th channel) are composited in the same pass (the
// reference renders depth with a second full pass, cuda_splatting.py:404).
#include <cuda_fp16.h>

#include "common.cuh"

namespace tsplat {

#ifndef TSPLAT_STAGE_ENTRIES
#define TSPLAT_STAGE_ENTRIES 64
#endif
#ifndef TSPLAT_STAGES
#define TSPLAT_STAGES 8
#endif
constexpr int kStageEntries = TSPLAT_STAGE_ENTRIES;  // list entries per ring stage (cull rounds of 32)
constexpr int kStages = TSPLAT_STAGES;               // ring depth: warps of a tile may drift kStages-1 stages apart
// A tile's 8 pixel blocks (one warp each) can be split over kSplit CTAs that stream the same list: smaller work
// items for the tail of the launch at the price of one more producer warp per part
#ifndef TSPLAT_TILE_SPLIT
#define TSPLAT_TILE_SPLIT 1
#endif
constexpr int kSplit = TSPLAT_TILE_SPLIT;
constexpr int kConsumerWarps = kTilePix / 32 / kSplit;        // 8 (or 4)
constexpr int kRenderThreads = kConsumerWarps * 32 + 32;      // + the producer warp

#ifndef TSPLAT_FWD_MIN_BLOCKS
#define TSPLAT_FWD_MIN_BLOCKS 4  // 54 registers, 4 CTAs per SM: 1 -> 82 registers (0.200 ms), 5 -> 40 (0.166), 7 -> 32 + spills (0.181), 4 -> 0.160
#endif
struct RingSmem {
  float4 r0[kStages][kStageEntries];  // x, y, conic.x, conic.y
  float4 r1[kStages][kStageEntries];  // conic.z, opacity, r, g
  float4 r2[kStages][kStageEntries];  // b, depth feature, bits(Gaussian id), cull threshold tau_s
  uint64_t full[kStages];
  uint64_t empty[kStages];
  int done_warps;
  float red[2 * kConsumerWarps];  // fused MSE loss and PSNR: per-warp partial sums
  unsigned short hitq[kConsumerWarps][kStageEntries];  // per warp: byte offsets (16 j) of the batch's entries that pass the cull
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ int ld_volatile_smem(const int* p) {
  int v;
  asm volatile("ld.volatile.shared::cta.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}

// exp and reciprocal of the hot loops without the denormal guards of __expf / __fdividef (3 and 5 instructions per
// visit).  ex2.approx.ftz differs from __expf only where the result is denormal (< 1.2e-38): alpha is then below
// 1/255 either way; forward and backward use the same function, so they classify every (pixel, entry) alike.
// 1 - alpha lies in [0.01, 1]: rcp.approx.ftz needs no range reduction.
__device__ __forceinline__ float exp_hot(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
__device__ __forceinline__ float rcp_hot(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// shared-memory accesses of the backward's visit loop by 32-bit shared address: through generic pointers the
// compiler rebuilt the shared window base (S2R SR_CgaCtaId ...) on every visit
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ float2 lds_f2(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void red_shared_add(uint32_t a, float x) {
  asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(a), "f"(x) : "memory");
}

template <class Smem>
__device__ __forceinline__ void issue_stage(Smem& s, int stage, const Workspace& ws, unsigned int first,
                                            unsigned int cnt) {
  const uint32_t bytes = cnt * 16u;
  mbar_expect_tx(&s.full[stage], 3u * bytes);
  bulk_g2s(s.r0[stage], ws.rec0 + first, bytes, &s.full[stage]);
  bulk_g2s(s.r1[stage], ws.rec1 + first, bytes, &s.full[stage]);
  bulk_g2s(s.r2[stage], ws.rec2 + first, bytes, &s.full[stage]);
}

// Pixel <-> thread mapping: warp w owns an 8x4 pixel block of the 16x16 tile (2 block columns x 4
// block rows), lane l the pixel (l & 7, l >> 3) of it.  A compact block makes the per-warp culling
// effective (the mapping itself is invisible in the results).
__device__ __forceinline__ void pixel_of_thread(int tid, int part, int& bx, int& by, int& lx, int& ly) {
  const int warp = part * kConsumerWarps + (tid >> 5), lane = tid & 31;
  bx = (warp & 1) * 8; by = (warp >> 1) * 4;
  lx = lane & 7; ly = lane >> 3;
}

// Per-warp culling: bit l of the result is set iff entry j (one per lane) can pass the alpha test on some pixel of
// this warp's 8x4 block.  The test itself (minimum of the conic quadratic over the block against tau_s) is done once
// per (tile, Gaussian) pair by the gather kernels, which leave one bit per block in rec2.w (cull.cuh:
// cull_block_mask); here a lane only looks at its warp's bit.
__device__ __forceinline__ unsigned int cull_round(const float4* __restrict__ r2, int j, bool in_range, int block) {
  bool hit = false;
  if (in_range) hit = (__float_as_uint(r2[j].w) >> block) & 1u;
  return __ballot_sync(0xffffffffu, hit);
}

// ---------------------------------------------------------------------------------------------
// Forward
// ---------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(kRenderThreads, TSPLAT_FWD_MIN_BLOCKS) render_forward_kernel(Workspace ws, int H, int W,
                                                                        const TsplatCamera* __restrict__ cams,
                                                                        float* __restrict__ out_color,
                                                                        float* __restrict__ out_depth,
                                                                        const float* __restrict__ target,
                                                                        float* __restrict__ loss_sum, float loss_scale,
                                                                        float* __restrict__ psnr) {
  __shared__ __align__(128) RingSmem s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slot = (int)blockIdx.x / kSplit, part = (int)blockIdx.x % kSplit;
  const int bid = ws.use_order ? (int)ws.tile_order[slot] : slot;  // longest lists first
  const int n = bid / ws.T, tile = bid - n * ws.T;
  const uint2 range = ws.ranges[bid];
  const int total = (int)(range.y - range.x);
  const int nb = (total + kStageEntries - 1) / kStageEntries;

  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < kStages; i++) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], kConsumerWarps); }
    s.done_warps = 0;
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ---------------- producer: one lane streams the tile's list through the ring ----------------
    if (lane != 0) return;
    int issued = 0;
    for (int b = 0; b < nb; b++) {
      const int st = b % kStages, k = b / kStages;
      bool stop = false;
      if (b >= kStages) {  // wait until all 8 warps released the batch that occupied this stage
        while (!mbar_try_wait(&s.empty[st], (unsigned)(k - 1) & 1u))
          if (ld_volatile_smem(&s.done_warps) == kConsumerWarps) { stop = true; break; }
      }
      if (stop || ld_volatile_smem(&s.done_warps) == kConsumerWarps) break;  // every pixel saturated
      issue_stage(s, st, ws, range.x + (unsigned)b * kStageEntries,
                  (unsigned)min(kStageEntries, total - b * kStageEntries));
      issued = b + 1;
    }
    // never leave the CTA with a bulk copy still in flight into its shared memory
    for (int b = max(0, issued - kStages); b < issued; b++) mbar_wait(&s.full[b % kStages], (unsigned)(b / kStages) & 1u);
    return;
  }

  // ---------------- consumers ----------------
  const int tx = tile % ws.tiles_x, ty = tile / ws.tiles_x;
  int bx, by, lx, ly;
  pixel_of_thread(tid, part, bx, by, lx, ly);
  const int px = tx * kTile + bx + lx, py = ty * kTile + by + ly;
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const int block = part * kConsumerWarps + warp;  // this warp's bit in the pair's block mask (rec2.w)

  float T = 1.0f;
  float C[CH];
#pragma unroll
  for (int ch = 0; ch < CH; ch++) C[ch] = 0.0f;
  unsigned int last_contributor = 0;
  bool done = !inside;
  bool warp_done = __all_sync(0xffffffffu, done);
  if (warp_done && lane == 0) atomicAdd(&s.done_warps, 1);
  const unsigned int lt_mask = (1u << lane) - 1u;

  for (int b = 0; b < nb; b++) {
    const int st = b % kStages;
    const unsigned int par = (unsigned)(b / kStages) & 1u;
    if (!warp_done) {
      mbar_wait(&s.full[st], par);  // this warp still needs data, so the producer keeps going
    } else {
      bool quit = false;
      while (!mbar_try_wait(&s.full[st], par))
        if (ld_volatile_smem(&s.done_warps) == kConsumerWarps) { quit = true; break; }
      if (quit) break;
    }
    if (!warp_done) {
      const int cnt = min(kStageEntries, total - b * kStageEntries);
      const float4* __restrict__ r0 = s.r0[st];
      const float4* __restrict__ r1 = s.r1[st];
      const float4* __restrict__ r2 = s.r2[st];
      // survivors of the cull, compacted into this warp's queue (list order)
      int nhit = 0;
#pragma unroll
      for (int i = 0; i < kStageEntries / 32; i++) {
        const int j = i * 32 + lane;
        const unsigned int m = cull_round(r2, j, j < cnt, block);
        if ((m >> lane) & 1u) s.hitq[warp][nhit + __popc(m & lt_mask)] = (unsigned short)(j * 16);
        nhit += __popc(m);
      }
      __syncwarp();
      uint32_t rb = smem_u32(r0);
      const uint32_t hqb = smem_u32(s.hitq[warp]);
      asm volatile("" : "+r"(rb));
      const unsigned int base = (unsigned int)(b * kStageEntries) + 1u;  // 1-based list position of entry 0
      for (int k = 0; k < nhit; k++) {
        const uint32_t off = lds_u16(hqb + 2u * (uint32_t)k);  // 16 j
        const float4 a0 = lds_f4(rb + off);
        const float4 a1 = lds_f4(rb + off + (uint32_t)sizeof(s.r0));
        const float dx = a0.x - pxf, dy = a0.y - pyf;
        const float power = -0.5f * (a0.z * dx * dx + a1.x * dy * dy) - a0.w * dx * dy;
        if (done || power > 0.0f) continue;
        const float alpha = fminf(0.99f, a1.y * exp_hot(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1.0f - alpha);
        if (test_T < 0.0001f) { done = true; continue; }
        const float2 a2 = lds_f2(rb + off + 2u * (uint32_t)sizeof(s.r0));
        const float w = alpha * T;
        C[0] += a1.z * w;
        C[1] += a1.w * w;
        C[2] += a2.x * w;
        if (CH == 4) C[3] += a2.y * w;
        T = test_T;
        last_contributor = base + (off >> 4);
      }
      if (__all_sync(0xffffffffu, done)) {
        warp_done = true;
        if (lane == 0) atomicAdd(&s.done_warps, 1);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&s.empty[st]);
    if (warp_done && ld_volatile_smem(&s.done_warps) == kConsumerWarps) break;
  }
  {  // the backward never needs the list behind the tile's deepest contributor (28 % of a C2 list on average)
    const unsigned int wm = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && wm) atomicMax(&ws.tile_last[bid], wm);
  }
  float psq = 0.0f;  // per-view PSNR (metrics.py:12-19): sum over channels of (clip(target) - clip(colour))^2
  float sq = 0.0f;  // fused MSE (loss_mse.py:30): this pixel's sum over channels of (colour - target)^2
  if (inside) {
    const size_t pix = (size_t)py * W + px;
    const size_t P = (size_t)H * W;
    const float* bg = cams[n].bg;
    ws.final_T[n * P + pix] = T;
    ws.n_contrib[n * P + pix] = last_contributor;
    const float c0 = C[0] + T * __ldg(bg + 0), c1 = C[1] + T * __ldg(bg + 1), c2 = C[2] + T * __ldg(bg + 2);
    out_color[(n * 3 + 0) * P + pix] = c0;
    out_color[(n * 3 + 1) * P + pix] = c1;
    out_color[(n * 3 + 2) * P + pix] = c2;
    if (CH == 4) {  // mean over three identical channels (cuda_splatting.py:417)
      const float x = C[3];
      out_depth[n * P + pix] = __fdiv_rn(__fadd_rn(__fadd_rn(x, x), x), 3.0f);
    }
    if (target != nullptr) {
      const float t0 = target[(n * 3 + 0) * P + pix], t1 = target[(n * 3 + 1) * P + pix], t2 = target[(n * 3 + 2) * P + pix];
      const float d0 = c0 - t0, d1 = c1 - t1, d2 = c2 - t2;
      sq = d0 * d0 + d1 * d1 + d2 * d2;
      if (psnr != nullptr) {  // compute_psnr clips both images to [0,1] first (src/evaluation/metrics.py:16-17)
        const float e0 = __saturatef(t0) - __saturatef(c0), e1 = __saturatef(t1) - __saturatef(c1),
                    e2 = __saturatef(t2) - __saturatef(c2);
        psq = e0 * e0 + e1 * e1 + e2 * e2;
      }
    }
  }
  if (target != nullptr) {
    // CTA partial in a fixed order; the LAST CTA to finish adds all partials in index order: the loss value does
    // not depend on the order in which the tiles complete
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sq += __shfl_xor_sync(0xffffffffu, sq, o);
      psq += __shfl_xor_sync(0xffffffffu, psq, o);
    }
    if (lane == 0) { s.red[warp] = sq; s.red[kConsumerWarps + warp] = psq; }
    asm volatile("bar.sync 1, %0;" ::"r"(kConsumerWarps * 32) : "memory");  // the consumer warps (the producer has left)
    if (warp == 0) {
      unsigned int last = 0;
      float* psnr_partial = ws.loss_partial + (size_t)ws.N * ws.T * kSplit;  // indexed by (view, tile, part)
      if (lane == 0) {
        float t = 0.0f, u = 0.0f;
#pragma unroll
        for (int i = 0; i < kConsumerWarps; i++) { t += s.red[i]; u += s.red[kConsumerWarps + i]; }
        ws.loss_partial[blockIdx.x] = t;
        if (psnr != nullptr) psnr_partial[bid * kSplit + part] = u;
        __threadfence();
        last = atomicAdd(&ws.ctl->loss_done, 1u) == gridDim.x - 1 ? 1u : 0u;
      }
      last = __shfl_sync(0xffffffffu, last, 0);
      if (last) {
        __threadfence();
        float t = 0.0f;
        for (unsigned int i = lane; i < gridDim.x; i += 32) t += __ldcg(ws.loss_partial + i);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) *loss_sum = t * loss_scale;
        if (psnr != nullptr) {
          const int per_view = ws.T * kSplit;
          for (int v = 0; v < ws.N; v++) {
            float u = 0.0f;
            for (int i = lane; i < per_view; i += 32) u += __ldcg(psnr_partial + (size_t)v * per_view + i);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
            // -10 log10(mean over c,h,w); a perfect match gives +inf as in the reference
            if (lane == 0) psnr[v] = -10.0f * log10f(u / (3.0f * (float)H * (float)W));
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward: back-to-front over the same list through the same kind of ring.  Per entry the 10
// gradient components (mean2D.xy, conic.xyw, opacity, colour[CH]) are pre-reduced over the 32
// pixels of each warp with a 12-shuffle transpose-reduction and added into the stage's
// shared-memory accumulator (red.shared).  When all 8 warps have released a stage, the PRODUCER
// warp flushes its accumulators with three 16-B vector reductions per (tile, Gaussian)
// (red.global.add.v4.f32) and refills the stage: the flush is off the consumers' critical path, and
// global reductions are one per (tile, Gaussian) instead of one atomic per (pixel, Gaussian,
// component) as in the reference.
// ---------------------------------------------------------------------------------------------
#ifndef TSPLAT_BWD_MIN_BLOCKS
#define TSPLAT_BWD_MIN_BLOCKS 4  // 4 CTAs per SM by shared memory; 5 or 6 with fewer stages and capped registers are slower
#endif
struct RingBwdSmem {
  float4 r0[kStages][kStageEntries];
  float4 r1[kStages][kStageEntries];
  float4 r2[kStages][kStageEntries];
  float acc[kStages][kStageEntries][kAccFloats];
  uint64_t full[kStages];
  uint64_t empty[kStages];
  unsigned short hitq[kConsumerWarps][kStageEntries];  // per warp: byte offsets (16 j) of the batch's entries that pass the cull
};

// one 16-B vector reduction into global memory (sm_90+): 4 fp32 adds, one L2 operation
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// producer-side flush of one released stage: lane handles entries lane, lane+32
__device__ __forceinline__ void flush_stage(RingBwdSmem& s, int st, int cnt, int lane, float* __restrict__ grad_acc,
                                            size_t view_base) {
#pragma unroll
  for (int j = lane; j < kStageEntries; j += 32) {
    if (j < cnt) {
      const unsigned int id = __float_as_uint(s.r2[st][j].z);
      float4* a4 = reinterpret_cast<float4*>(&s.acc[st][j][0]);
      const float4 q0 = a4[0], q1 = a4[1], q2 = a4[2];
      const bool any = (q0.x != 0.0f) | (q0.y != 0.0f) | (q0.z != 0.0f) | (q0.w != 0.0f) | (q1.x != 0.0f) |
                       (q1.y != 0.0f) | (q1.z != 0.0f) | (q1.w != 0.0f) | (q2.x != 0.0f) | (q2.y != 0.0f);
      if (any) {
        float* dst = grad_acc + (view_base + id) * kAccFloats;
#ifdef TSPLAT_EXP_NOFLUSH
        if (q0.x == 123.456f)
#endif
        {
          red_add_v4(dst, q0.x, q0.y, q0.z, q0.w);
          red_add_v4(dst + 4, q1.x, q1.y, q1.z, q1.w);
          red_add_v4(dst + 8, q2.x, q2.y, 0.0f, 0.0f);
        }
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        a4[0] = z; a4[1] = z; a4[2] = z;
      }
    }
  }
}

template <int CH>
__global__ void __launch_bounds__(kRenderThreads, TSPLAT_BWD_MIN_BLOCKS) render_backward_kernel(Workspace ws, int H, int W, int G,
                                                                         const TsplatCamera* __restrict__ cams,
                                                                         const float* __restrict__ dL_dcolor,
                                                                         const float* __restrict__ dL_ddepth,
                                                                         TsplatMse mse) {
  extern __shared__ __align__(128) unsigned char bsmem_raw[];
  RingBwdSmem& s = *reinterpret_cast<RingBwdSmem*>(bsmem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slot = (int)blockIdx.x / kSplit, part = (int)blockIdx.x % kSplit;
  const int bid = ws.use_order ? (int)ws.tile_order[slot] : slot;  // longest lists first
  const int n = bid / ws.T, tile = bid - n * ws.T;
  uint2 range = ws.ranges[bid];
  range.y = range.x + min(range.y - range.x, ws.tile_last[bid]);  // nothing contributes behind this position
  const int total = (int)(range.y - range.x);
  const int nb = (total + kStageEntries - 1) / kStageEntries;

  // batch b (b = 0 is the END of the list) covers sorted entries [lo_b, hi_b)
  auto batch_hi = [&](int b) { return (int)range.y - b * kStageEntries; };
  auto batch_lo = [&](int b) { return max((int)range.x, (int)range.y - (b + 1) * kStageEntries); };

  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < kStages; i++) { mbar_init(&s.full[i], 1); mbar_init(&s.empty[i], kConsumerWarps); }
    mbar_fence_init();
  }
  for (int i = tid; i < kStages * kStageEntries * kAccFloats; i += kRenderThreads) (&s.acc[0][0][0])[i] = 0.0f;
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ---------------- producer + flusher (whole warp) ----------------
    const size_t view_base = (size_t)n * G;
    for (int b = 0; b < nb; b++) {
      const int st = b % kStages;
      if (b >= kStages) {
        const int pb = b - kStages;  // the batch that occupied this stage
        mbar_wait(&s.empty[st], (unsigned)(pb / kStages) & 1u);
        flush_stage(s, st, batch_hi(pb) - batch_lo(pb), lane, ws.grad_acc, view_base);
        __syncwarp();  // every lane has read its record ids before the stage is overwritten
      }
      if (lane == 0) issue_stage(s, st, ws, (unsigned)batch_lo(b), (unsigned)(batch_hi(b) - batch_lo(b)));
    }
    for (int pb = max(0, nb - kStages); pb < nb; pb++) {
      const int st = pb % kStages;
      mbar_wait(&s.empty[st], (unsigned)(pb / kStages) & 1u);
      flush_stage(s, st, batch_hi(pb) - batch_lo(pb), lane, ws.grad_acc, view_base);
    }
    return;
  }

  // ---------------- consumers ----------------
  const int tx = tile % ws.tiles_x, ty = tile / ws.tiles_x;
  int bx, by, lx, ly;
  pixel_of_thread(tid, part, bx, by, lx, ly);
  const int px = tx * kTile + bx + lx, py = ty * kTile + by + ly;
  const bool inside = px < W && py < H;
  const float pxf = (float)px, pyf = (float)py;
  const int block = part * kConsumerWarps + warp;  // this warp's bit in the pair's block mask (rec2.w)
  const size_t P = (size_t)H * W;
  const size_t pix = (size_t)py * W + px;

  const float T_final = inside ? ws.final_T[n * P + pix] : 0.0f;
  float T = T_final;
  const int last_contributor = inside ? (int)ws.n_contrib[n * P + pix] : 0;
  // The reference keeps accum_rec[ch] (colour behind the current entry) per channel and forms
  // dL/dalpha = sum_ch (c[ch] - accum_rec[ch]) dL_dpix[ch]; dL_dpix is constant along the list, so the recurrence
  // runs on the SCALAR accum_dl = sum_ch accum_rec[ch] dL_dpix[ch] (same arithmetic up to summation order).
  float dL_dpix[CH];
  float accum_dl = 0.0f, w_last = 0.0f, oml_last = 1.0f;
#pragma unroll
  for (int ch = 0; ch < CH; ch++) dL_dpix[ch] = 0.0f;
  if (inside) {
    if (dL_dcolor != nullptr) {
#pragma unroll
      for (int ch = 0; ch < 3; ch++) dL_dpix[ch] = dL_dcolor[(n * 3 + ch) * P + pix];
    }
    if (mse.target != nullptr) {  // fused MSE: dL/dcolor += dL/dloss * 2 w / numel * (colour - target)
      const float gl = __ldg(mse.dL_dloss) * mse.scale;
#pragma unroll
      for (int ch = 0; ch < 3; ch++)
        dL_dpix[ch] += gl * (mse.color[(n * 3 + ch) * P + pix] - mse.target[(n * 3 + ch) * P + pix]);
    }
    if (CH == 4) dL_dpix[3] = dL_ddepth[n * P + pix];
  }
  const float* bgp = cams[n].bg;
  // bg of the fused depth channel is 0 (render_depth_cuda passes zeros, cuda_splatting.py:410)
  const float bg_dot = __ldg(bgp) * dL_dpix[0] + __ldg(bgp + 1) * dL_dpix[1] + __ldg(bgp + 2) * dL_dpix[2];
  const float bgT = -T_final * bg_dot;
  // ---- reduce_plan: which lane ends up with which of the 10 per-entry sums ----
  // Lane bits b4 b3 b2 pick the component, bits b1 b0 are summed away.  Slots of the accumulator record:
  //   0 S_x  1 S_y  2 S_xx  3 S_xy  4 S_yy  5 dL/d opacity  6..9 dL/d colour (r, g, b, depth feature).
  // The swaps of a transposing butterfly cost two selects per exchanged pair; here they are pushed into operands that
  // are constant along the list: the colour sums are alpha T dL_dpix[c], so each lane multiplies by ITS permutation
  // of dL_dpix (LAk, LAs, LBk, LBs) and rounds 1 and 2 need no select for them; S_x|S_y and S_xx|S_yy share the
  // swapped offsets (dk, ds).  S_xy | opacity run as a sixth value through plain sums.  9 selects instead of 22.
  const bool b4i = lane & 16, b3i = lane & 8;
  const float dl0 = dL_dpix[0], dl1 = dL_dpix[1], dl2 = dL_dpix[2], dl3 = CH == 4 ? dL_dpix[CH - 1] : 0.0f;
  float LAk = b3i ? (b4i ? dl3 : dl2) : (b4i ? dl1 : dl0), LAs = b3i ? (b4i ? dl2 : dl3) : (b4i ? dl0 : dl1);
  float LBk = b3i ? (b4i ? dl1 : dl0) : (b4i ? dl3 : dl2), LBs = b3i ? (b4i ? dl0 : dl1) : (b4i ? dl2 : dl3);
  const bool g2_writer = (lane & 15) == 1;  // lane 1: S_xy, lane 17: opacity
  const int my_comp = g2_writer ? (b4i ? 5 : 3)
                                : ((lane & 4) ? 6 + (b3i ? 2 : 0) + (b4i ? 1 : 0) : (b3i ? (b4i ? 4 : 2) : (b4i ? 1 : 0)));
  const bool my_red = g2_writer || ((lane & 3) == 0 && (CH == 4 || my_comp != 9));
  // opaque to ptxas, which would otherwise rebuild the four permuted factors and the lane bits on every visit
  int lane_bits = lane | (g2_writer ? 32 : 0);
  asm volatile("" : "+f"(LAk), "+f"(LAs), "+f"(LBk), "+f"(LBs), "+r"(lane_bits));
  // list positions >= the warp's deepest contributor are never visited by this warp
  const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
  const unsigned int lt_mask = (1u << lane) - 1u;

  for (int b = 0; b < nb; b++) {
    const int st = b % kStages;
    const int lo = batch_lo(b), cnt = batch_hi(b) - lo;
    mbar_wait(&s.full[st], (unsigned)(b / kStages) & 1u);
    const float4* __restrict__ r0 = s.r0[st];
    const float4* __restrict__ r1 = s.r1[st];
    const float4* __restrict__ r2 = s.r2[st];
    const int pos0 = lo - (int)range.x;  // 0-based list position of entry 0 of this batch
    if (pos0 < warp_last) {
      // survivors of the cull, compacted into this warp's queue (ascending list order): walking a queue costs
      // 5 instructions per visit, scanning the ballot masks bit by bit cost 15
      unsigned short* hq = s.hitq[warp];
      int nhit = 0;
#pragma unroll
      for (int i = 0; i < kStageEntries / 32; i++) {
        const int j = i * 32 + lane;
        const unsigned int m = cull_round(r2, j, j < cnt && pos0 + j < warp_last, block);
        if ((m >> lane) & 1u) hq[nhit + __popc(m & lt_mask)] = (unsigned short)(j * 16);
        nhit += __popc(m);
      }
      __syncwarp();
      uint32_t rb = smem_u32(r0);
      uint32_t accb = smem_u32(&s.acc[st][0][0]) + (uint32_t)(my_comp * 4);
      const uint32_t hqb = smem_u32(hq);
      asm volatile("" : "+r"(rb), "+r"(accb));  // keep them in registers: ptxas otherwise rebuilds both on every visit
      // entries at list positions >= last_contributor are skipped (clamped: no overflow for any list length)
      const int lim = max(min(last_contributor - pos0, kStageEntries), 0) * 16;
      for (int k = nhit - 1; k >= 0; k--) {
        const uint32_t off = lds_u16(hqb + 2u * (uint32_t)k);  // 16 j
        const float4 a0 = lds_f4(rb + off);
        const float4 a1 = lds_f4(rb + off + (uint32_t)sizeof(s.r0));
        const float dx = a0.x - pxf, dy = a0.y - pyf;
        const float power = -0.5f * (a0.z * dx * dx + a1.x * dy * dy) - a0.w * dx * dy;
        const float Gv = exp_hot(power);
        const float alpha = fminf(0.99f, a1.y * Gv);
        const bool valid = ((int)off < lim) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        if (!__any_sync(0xffffffffu, valid)) continue;
        // No branch on `valid`.  A lane that does not contribute runs the same arithmetic with alpha = 0: that leaves
        // T unchanged (T / (1 - 0)), makes every accumulated product 0, and turns the suffix recurrence into the
        // identity for the next step (w = 0, 1 - alpha = 1), exactly as if the entry had been skipped.
        const float2 a2 = lds_f2(rb + off + 2u * (uint32_t)sizeof(s.r0));
        const float Gm = valid ? Gv : 0.0f, am = valid ? alpha : 0.0f;
        const float col[4] = {a1.z, a1.w, a2.x, a2.y};
        const float oma = 1.0f - am;
        const float inv1ma = rcp_hot(oma);
        const float Tn = T * inv1ma;
        const float dchannel_dcolor = am * Tn;
        float cdl = 0.0f;
#pragma unroll
        for (int ch = 0; ch < CH; ch++) cdl += col[ch] * dL_dpix[ch];
        // accum_dl = last_alpha * last_cdl + (1 - last_alpha) * accum_dl, with w_last = last_alpha * last_cdl and
        // oml_last = 1 - last_alpha carried as state
        const float acc_new = w_last + oml_last * accum_dl;
        const float dL_dalpha = (cdl - acc_new) * Tn + bgT * inv1ma;
        T = Tn; accum_dl = acc_new; w_last = am * cdl; oml_last = oma;
        // Accumulated per (pixel, entry): the MOMENTS of u = dL/dG * G over the pixel offsets.  Everything that is
        // constant per (view, Gaussian) - the conic in dL/dmean2D = -(A S_x + B S_y, C S_y + B S_x), the -1/2 of
        // dL/dconic = -1/2 (S_xx, S_xy, S_yy), the pixel scale W/2, H/2 - is applied once in preprocess_backward.
        const float v5 = Gm * dL_dalpha;  // dL/d opacity
        const float u = a1.y * v5;
        // ---- transposing warp reduction of the 10 sums (see reduce_plan above the kernel) ----
        const unsigned int full = 0xffffffffu;
        const bool b4 = lane_bits & 16, b3 = lane_bits & 8, b2 = lane_bits & 4;
        const float dk = b4 ? dy : dx, ds = b4 ? dx : dy;
        const float m1k = u * dk, m1s = u * ds;           // S_x | S_y
        const float m2k = m1k * dk, m2s = m1s * ds;       // S_xx | S_yy
        const float v3 = m1k * ds;                        // S_xy
        const float g2k = b4 ? v5 : v3, g2s = b4 ? v3 : v5;
        // round 1 (xor 16)
        float G0 = m1k + __shfl_xor_sync(full, m1s, 16);
        float G1 = m2k + __shfl_xor_sync(full, m2s, 16);
        float G2 = g2k + __shfl_xor_sync(full, g2s, 16);
        float CA = dchannel_dcolor * LAk + __shfl_xor_sync(full, dchannel_dcolor * LAs, 16);
        float CB = dchannel_dcolor * LBk + __shfl_xor_sync(full, dchannel_dcolor * LBs, 16);
        // round 2 (xor 8)
        const float Cx = CA + __shfl_xor_sync(full, CB, 8);
        const float gk = b3 ? G1 : G0, gs = b3 ? G0 : G1;
        const float Gx = gk + __shfl_xor_sync(full, gs, 8);
        G2 += __shfl_xor_sync(full, G2, 8);
        // round 3 (xor 4)
        const float mk = b2 ? Cx : Gx, msd = b2 ? Gx : Cx;
        float Mx = mk + __shfl_xor_sync(full, msd, 4);
        G2 += __shfl_xor_sync(full, G2, 4);
        // rounds 4, 5: plain sums
        Mx += __shfl_xor_sync(full, Mx, 2);
        G2 += __shfl_xor_sync(full, G2, 2);
        Mx += __shfl_xor_sync(full, Mx, 1);
        G2 += __shfl_xor_sync(full, G2, 1);
        const float r = (lane_bits & 32) ? G2 : Mx;
#ifdef TSPLAT_EXP_NOATOM
        if (r == 123.456f) s.acc[st][0][0] = 1.0f;
#else
        if (my_red) red_shared_add(accb + 3u * off, r);
#endif
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&s.empty[st]);
  }
}

cudaError_t launch_render_forward(const TsplatProblem& p, const Workspace& ws, float* out_color, float* out_depth,
                                  const TsplatMse* mse, cudaStream_t st) {
  const int grid = ws.N * ws.T * kSplit;
  if (grid == 0) return cudaSuccess;
  const float* target = mse ? mse->target : nullptr;
  float* loss_sum = mse ? mse->loss_sum : nullptr;
  const float loss_scale = mse ? mse->loss_scale : 0.0f;
  float* psnr = mse ? mse->psnr : nullptr;
  if (p.depth_mode != TSPLAT_DEPTH_NONE)
    render_forward_kernel<4><<<grid, kRenderThreads, 0, st>>>(ws, p.H, p.W, p.cams, out_color, out_depth, target, loss_sum, loss_scale, psnr);
  else
    render_forward_kernel<3><<<grid, kRenderThreads, 0, st>>>(ws, p.H, p.W, p.cams, out_color, out_depth, target, loss_sum, loss_scale, psnr);
  return cudaGetLastError();
}

cudaError_t launch_render_backward(const TsplatProblem& p, const Workspace& ws, const float* dL_dcolor,
                                   const float* dL_ddepth, const TsplatMse* mse_in, cudaStream_t st) {
  const int grid = ws.N * ws.T * kSplit;
  if (grid == 0) return cudaSuccess;
  TsplatMse mse{};
  if (mse_in) mse = *mse_in;
  const int smem = (int)sizeof(RingBwdSmem);
  cudaError_t e;
  if (p.depth_mode != TSPLAT_DEPTH_NONE) {
    e = cudaFuncSetAttribute(render_backward_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    render_backward_kernel<4><<<grid, kRenderThreads, smem, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  } else {
    e = cudaFuncSetAttribute(render_backward_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    render_backward_kernel<3><<<grid, kRenderThreads, smem, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  }
  return cudaGetLastError();
}

}  // namespace tsplat
