// render.cu — alpha compositing, forward and backward, one WARP per 8x4 pixel block.
//
// The unit of work is a (view, tile, block) job: one of the 8 pixel blocks of a 16x16 tile and ITS list - the entries
// of the tile's depth-sorted list that can touch the block, as 8-byte references (Gaussian id, position in the tile
// list) written contiguously by blocklist.cu.  Both kernels are persistent: CTAs of 8 independent warps.  render_forward
// deals the jobs heaviest tile first: a warp's first job is static (its index in the launch), every further one comes
// from a global counter when the warp has finished the previous one.  render_backward's units are the lists cut at the
// block's deepest contributor and then into 64-reference segments (independent thanks to per-pixel checkpoints the
// forward leaves), emitted by the forward and dealt round-robin.
//
// A warp streams its references through its own shared-memory ring with 1-D TMA bulk copies (cp.async.bulk -> mbarrier
// complete_tx; one elected lane issues, the whole warp waits on the stage's mbarrier), ACROSS job boundaries: while the
// last batch of a job is composited the first references of the next job are already in flight.  One batch ahead of the
// compositing, lane l reads reference l of the next batch and gathers that Gaussian's 48-B record from the L2-resident
// per-Gaussian array into a double-buffered shared-memory tile with asynchronous 16-B copies (cp.async); all lanes then
// read each record by broadcast.  There is no cull test, no survivor queue, no
// barrier between warps and no waiting for a tile's slowest warp - what round 1's shared per-tile ring spent a quarter
// of its stall samples on.  Colour (3 channels) and the depth feature (4th channel) are composited in the same pass
// (the reference renders depth with a second full pass, cuda_splatting.py:404).
#include "common.cuh"

namespace tsplat {

#ifndef TSPLAT_WSTAGES
#define TSPLAT_WSTAGES 4
#endif
constexpr int kWStages = TSPLAT_WSTAGES;     // reference-ring depth per warp
constexpr int kWBatch = 32;                  // references per stage = records per batch (one per lane)
constexpr int kRefStageBytes = kWBatch * 8;  // 256
constexpr int kRecBytes = 16 * kRecFloat4;   // 48
constexpr int kRecTileBytes = kWBatch * kRecBytes;
constexpr int kRenderWarps = 8;
constexpr int kRenderThreads = kRenderWarps * 32;
#ifndef TSPLAT_FWD_UNROLL
#define TSPLAT_FWD_UNROLL 4  // 1 / 2 / 4 visits in flight: 0.131 / 0.125 / 0.122 ms at C2
#endif
constexpr int kFwdUnroll = TSPLAT_FWD_UNROLL;  // visits of render_forward in flight per warp
#ifndef TSPLAT_BWD_UNROLL
#define TSPLAT_BWD_UNROLL 1
#endif
constexpr int kBwdUnroll = TSPLAT_BWD_UNROLL;
#ifndef TSPLAT_FWD_MIN_BLOCKS
#define TSPLAT_FWD_MIN_BLOCKS 4
#endif
#ifndef TSPLAT_BWD_MIN_BLOCKS
#define TSPLAT_BWD_MIN_BLOCKS 4
#endif
#ifndef TSPLAT_BWD_TWOPASS
#define TSPLAT_BWD_TWOPASS 1  // render_backward2_kernel (entry-per-lane accumulation) instead of the butterfly kernel
#endif
#ifndef TSPLAT_BWD2_MIN_BLOCKS
#define TSPLAT_BWD2_MIN_BLOCKS 3  // 16-row matrix: 71 KB of shared memory per CTA, 80 registers
#endif
#ifndef TSPLAT_BWD2_UNROLL
#define TSPLAT_BWD2_UNROLL 2
#endif

struct WarpSmem {
  float4 rec[kRenderWarps][2][kWBatch * kRecFloat4];  // the batch being composited / the one stored next
  uint2 ref[kRenderWarps][kWStages][kWBatch];
  uint64_t full[kRenderWarps][kWStages];
};

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
// shared-memory accesses of the hot loops by 32-bit shared address: through generic pointers the compiler rebuilt the
// shared window base on every visit
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ uint2 lds_u2(uint32_t a) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f4(uint32_t a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void red_global_add(float* addr, float x) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(x) : "memory");
}

// One job record (blocklist.cu): first reference, reference count, (view*T + tile)*8 + block.
struct JobRec {
  unsigned int start, count, id;
};
__device__ __forceinline__ JobRec load_job(const Workspace& ws, const uint4* __restrict__ jobs, unsigned int idx,
                                           unsigned int njobs, int H, int W) {
  JobRec r{0u, 0u, 0u};
  if (idx < njobs) {
    const uint4 q = __ldcg(jobs + idx);
    r.start = q.x; r.id = q.z;
    {
      // a block that lies outside the image (partial tiles at the right / bottom edge) has nothing to composite
      const int blk = (int)(q.z & 7u), tile = (int)(q.z >> 3) % ws.T, ty = tile / ws.tiles_x, tx = tile - ty * ws.tiles_x;
      const bool outside = tx * kTile + (blk & 1) * 8 >= W || ty * kTile + (blk >> 1) * 4 >= H;
      r.count = outside ? 0u : q.y;
    }
  }
  return r;
}

// Pixel <-> lane mapping: block b of a tile covers columns (b & 1) * 8 .. +7, rows (b >> 1) * 4 .. +3; lane l is the
// pixel (l & 7, l >> 3) of it (cull.cuh: cull_block_mask uses the same numbering).
__device__ __forceinline__ void pixel_of_lane(int blk, int lane, int tx, int ty, int& px, int& py) {
  px = tx * kTile + (blk & 1) * 8 + (lane & 7);
  py = ty * kTile + (blk >> 1) * 4 + (lane >> 3);
}

// ---------------------------------------------------------------------------------------------
// The warp's stream of batches (<= 32 references each), in the order the warp consumes them, across job boundaries.
// All lanes keep the same cursor values; only lane 0 issues bulk copies.
// ---------------------------------------------------------------------------------------------
struct Stream {
  uint32_t ref_addr;            // shared address of this warp's reference stage 0
  uint32_t bar_addr;            // shared address of this warp's full[0]
  uint32_t rec_addr;            // shared address of this warp's record tile 0
  uint32_t phase;               // bit s: parity the next wait on reference stage s uses
  unsigned int issued, waited;  // reference batches issued / consumed so far (stage = sequence number % kWStages)
  int p_next;                   // the producer cursor is in the current job (0) or already in the next one (1)
  int p_b;                      // next batch of that job to issue
  int pf_job, pf_b;             // which batch is being gathered into the spare record tile: job (0 current, 1 next, -1 none), batch
  int tile_sel;                 // record tile the gather in flight (or the next one) fills
};
__device__ __forceinline__ int batches_of(const JobRec& j) { return (int)((j.count + kWBatch - 1) / kWBatch); }
// Batches are the list's aligned chunks of 32 references (a list starts at an even reference index, so every chunk
// starts 16-B aligned as the bulk copy requires); the backward consumes them last chunk first.
template <bool kBack>
__device__ __forceinline__ int batch_lo(const JobRec& j, int b) {
  return (kBack ? batches_of(j) - 1 - b : b) * kWBatch;
}
template <bool kBack>
__device__ __forceinline__ int batch_cnt(const JobRec& j, int b) {
  return min(kWBatch, (int)j.count - batch_lo<kBack>(j, b));
}
// keep the reference ring full: the current job's remaining batches first, then the next job's
template <bool kBack>
__device__ __forceinline__ void pump(Stream& s, const JobRec& cur, const JobRec& nxt, const uint2* __restrict__ bref, int lane) {
  while (s.issued - s.waited < (unsigned int)kWStages) {
    const JobRec& j = s.p_next ? nxt : cur;
    if (s.p_b < batches_of(j)) {
      if (lane == 0) {
        const uint32_t stage = s.issued % kWStages, bar = s.bar_addr + 8u * stage;
        // whole 16-B units: an odd batch also copies the reference behind it (padding or the next list's first one)
        const uint32_t bytes = (uint32_t)((batch_cnt<kBack>(j, s.p_b) + 1) & ~1) * 8u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         s.ref_addr + stage * (uint32_t)kRefStageBytes),
                     "l"(bref + (size_t)j.start + batch_lo<kBack>(j, s.p_b)), "r"(bytes), "r"(bar)
                     : "memory");
      }
      s.p_b++;
      s.issued++;
    } else if (s.p_next == 0) {
      s.p_next = 1;
      s.p_b = 0;
    } else {
      break;
    }
  }
}
// waits for the next reference batch of the stream; returns the shared address of its first reference
__device__ __forceinline__ uint32_t next_refs(Stream& s) {
  const uint32_t stage = s.waited % kWStages, bar = s.bar_addr + 8u * stage, parity = (s.phase >> stage) & 1u;
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(kMbarSuspendNs)
        : "memory");
  } while (!ok);
  s.phase ^= 1u << stage;
  s.waited++;
  return s.ref_addr + stage * (uint32_t)kRefStageBytes;
}
// Gather the records of batch (job j, batch b) into the spare record tile: lane l reads reference l and copies that
// Gaussian's 48-B record from the L2-resident per-Gaussian array straight into shared memory with asynchronous 16-B
// copies (cp.async: no registers held across the compositing loop, no store instructions); the record's last word (the
// cull threshold, dead after blocklist.cu) is overwritten with the reference's position in the tile list.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_n(uint32_t dst, const void* src, int n) {  // n = 8 or 4
  if (n == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
template <bool kBack>
__device__ __forceinline__ void prefetch_records(Stream& s, const Workspace& ws, int G, const JobRec& j, int which, int b,
                                                 int lane) {
  const uint32_t ra = next_refs(s);
  if (lane < batch_cnt<kBack>(j, b)) {
    const uint2 ref = lds_u2(ra + 8u * (uint32_t)lane);
    const int view = (int)(j.id >> 3) / ws.T;
    TS_CHECK(ref.x < (unsigned int)G && ref.y >= 1u && view < ws.N);
    TS_CHECK((size_t)j.start + batch_lo<kBack>(j, b) + lane < ws.e_cap);
    const char* gr = reinterpret_cast<const char*>(ws.grec + ((size_t)view * G + ref.x) * kRecFloat4);
    const uint32_t dst = s.rec_addr + (uint32_t)(s.tile_sel * kRecTileBytes + lane * kRecBytes);
    cp_async16(dst, gr);
    cp_async16(dst + 16u, gr + 16);
    cp_async_n(dst + 32u, gr + 32, 8);
    cp_async_n(dst + 40u, gr + 40, 4);
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(dst + 44u), "r"(ref.y) : "memory");  // 1-based position in the tile list
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  s.pf_job = which; s.pf_b = b;
  __syncwarp();  // every lane has read its reference: the stage may be refilled
}
// the tile being gathered becomes the batch being composited; returns its shared address
__device__ __forceinline__ uint32_t publish_records(Stream& s) {
  const uint32_t base = s.rec_addr + (uint32_t)(s.tile_sel * kRecTileBytes);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  s.tile_sel ^= 1;
  s.pf_job = -1;
  __syncwarp();  // every lane's copies have landed, and every lane has left the tile the next gather overwrites
  return base;
}
// Makes batch b of the current job ready in shared memory and starts the gather of the batch after it (the current
// job's next one, else the next job's first).  Returns the shared address of the batch's first record.
template <bool kBack>
__device__ __forceinline__ uint32_t stage_batch(Stream& s, const Workspace& ws, int G, const JobRec& cur, const JobRec& nxt,
                                                int b, int lane) {
  if (!(s.pf_job == 0 && s.pf_b == b)) prefetch_records<kBack>(s, ws, G, cur, 0, b, lane);  // exposed: first batch after an idle ring
  const uint32_t base = publish_records(s);
  if (b + 1 < batches_of(cur)) prefetch_records<kBack>(s, ws, G, cur, 0, b + 1, lane);
  else if (batches_of(nxt) > 0) prefetch_records<kBack>(s, ws, G, nxt, 1, 0, lane);
  pump<kBack>(s, cur, nxt, ws.bref, lane);
  return base;
}
// the current job ends early, after batch b: drop what the ring and the registers hold for it
__device__ __forceinline__ void abandon_job(Stream& s, const JobRec& cur, int b) {
  int consumed = b + 1;
  if (s.pf_job == 0) { consumed = s.pf_b + 1; s.pf_job = -1; }
  const int issued_cur = s.p_next ? batches_of(cur) : s.p_b;
  for (int i = consumed; i < issued_cur; i++) (void)next_refs(s);
  if (s.p_next == 0) s.p_b = batches_of(cur);
}
// the next job becomes the current one
__device__ __forceinline__ void advance_job(Stream& s, JobRec& cur, JobRec& nxt, const JobRec& fut) {
  cur = nxt;
  nxt = fut;
  if (s.p_next) s.p_next = 0;   // the producer was already there: its batch cursor stands
  else s.p_b = 0;
  if (s.pf_job == 1) s.pf_job = 0;
}
__device__ __forceinline__ Stream make_stream(WarpSmem& sm, int warp, int lane) {
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < kWStages; i++) mbar_init(&sm.full[warp][i], 1);
    mbar_fence_init();
  }
  __syncwarp();
  Stream s;
  s.ref_addr = smem_u32(&sm.ref[warp][0][0]);
  s.bar_addr = smem_u32(&sm.full[warp][0]);
  s.rec_addr = smem_u32(&sm.rec[warp][0][0]);
  s.phase = 0u; s.issued = 0u; s.waited = 0u; s.p_next = 0; s.p_b = 0; s.pf_job = -1; s.pf_b = 0; s.tile_sel = 0;
  return s;
}
// Longest-processing-time-first dealing: the w-th warp of the launch starts with the w-th longest job; every further job
// is the next one of the sorted table, taken from a global counter by lane 0 (one job ahead of its use).
__device__ __forceinline__ unsigned int take_job(unsigned int* counter, unsigned int first_wave, int lane) {
  unsigned int j = 0;
  if (lane == 0) j = atomicAdd(counter, 1u);
  return first_wave + __shfl_sync(0xffffffffu, j, 0);
}

#ifdef TSPLAT_TIMELINE
// development only (scripts/exp_variants.sh): per-warp start / end time, jobs and references processed by render_forward
__device__ unsigned long long g_timeline[8192 * 4];
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#endif

// ---------------------------------------------------------------------------------------------
// Forward
// ---------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(kRenderThreads, TSPLAT_FWD_MIN_BLOCKS) render_forward_kernel(
    Workspace ws, int H, int W, int G, const TsplatCamera* __restrict__ cams, float* __restrict__ out_color,
    float* __restrict__ out_depth, const float* __restrict__ target, float* __restrict__ loss_sum, float loss_scale,
    float* __restrict__ psnr) {
  __shared__ __align__(128) WarpSmem sm;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Stream s = make_stream(sm, warp, lane);
  const unsigned int njobs = (unsigned int)(ws.N * ws.T) * kBlocksPerTile;
  const size_t P = (size_t)H * W;
  const unsigned int NT = (unsigned int)(ws.N * ws.T);
  // loss partials: [0, 8 NT) per job, [8 NT, 9 NT) per tile; the PSNR partials follow with the same layout
  float* const job_part = ws.loss_partial;
  float* const tile_part = ws.loss_partial + (size_t)NT * kBlocksPerTile;
  float* const pjob_part = tile_part + NT;
  float* const ptile_part = pjob_part + (size_t)NT * kBlocksPerTile;

  const unsigned int W_all = gridDim.x * kRenderWarps, gw = blockIdx.x * kRenderWarps + warp;
  // No look-ahead on the job counter: a warp claims its next job only when it has finished the current one.  Early
  // termination makes a job's cost unpredictable (45 % of the listed references are never visited at C2) and there are
  // fewer than two jobs per warp: with jobs claimed ahead, the first warps to reach the counter emptied it at once and
  // the rest ran a single job each (measured: warps ended between 30 and 161 us).  The claim's latency (counter ->
  // job record -> references -> records) is covered by the other warps of the SM.
  unsigned int cur_idx = gw;
  JobRec cur = load_job(ws, ws.block_jobs, cur_idx, njobs, H, W);
  const JobRec nxt{0u, 0u, 0u};
#ifdef TSPLAT_TIMELINE
  const unsigned long long t_start = gtime();
  unsigned long long n_jobs = 0, n_refs = 0;
#endif
  while (cur_idx < njobs) {
    pump<false>(s, cur, nxt, ws.bref, lane);
    const int blk = (int)(cur.id & 7u);
    const int bid = (int)(cur.id >> 3);
    const int n = bid / ws.T, tile = bid - n * ws.T;
    const int ty = tile / ws.tiles_x, tx = tile - ty * ws.tiles_x;
    int px, py;
    pixel_of_lane(blk, lane, tx, ty, px, py);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;

    float T = 1.0f;
    float C[CH];
#pragma unroll
    for (int ch = 0; ch < CH; ch++) C[ch] = 0.0f;
    unsigned int last_contributor = 0;  // 1-based position in the TILE list (n_contrib)
    int last_entry = 0;                 // 1-based position in this block's list
    bool done = !inside;
    const int nb = batches_of(cur);
    for (int b = 0; b < nb; b++) {
      uint32_t ra = stage_batch<false>(s, ws, G, cur, nxt, b, lane);
      const int cnt = batch_cnt<false>(cur, b);
      const int pos0 = b * kWBatch + 1;
      if (b > 0 && (b * kWBatch) % kSegLen == 0) {
        // per-pixel state in front of reference b * 32 of this list: render_backward restarts from it, so that the
        // list's segments are independent work units (blocklist.cu: segment_jobs_kernel)
        TS_CHECK((size_t)(cur.start / kSegLen) + (size_t)(b * kWBatch / kSegLen) < ws.e_cap / kSegLen + 2);
        float* ck = ws.ckpt + ((size_t)(cur.start / kSegLen) + (size_t)(b * kWBatch / kSegLen)) * kCkptFloats + lane;
        ck[0] = T; ck[32] = C[0]; ck[64] = C[1]; ck[96] = C[2]; ck[128] = CH == 4 ? C[CH - 1] : 0.0f;
      }
      // Branch-free visit (a lane that does not contribute runs the same arithmetic with weight 0), two visits in flight:
      // with the per-lane `continue`s of the reference a quarter of the issue slots went to divergent paths and the
      // dependent chain load -> exp -> test could not overlap between entries (0.132 -> 0.126 ms at C2)
#pragma unroll kFwdUnroll
      for (int k = 0; k < cnt; k++) {
        const uint32_t rk = ra + (uint32_t)(k * kRecBytes);
        const float4 a0 = lds_f4(rk);
        const float4 a1 = lds_f4(rk + 16u);
        const float dx = a0.x - pxf, dy = a0.y - pyf;
        const float power = -0.5f * (a0.z * dx * dx + a1.x * dy * dy) - a0.w * dx * dy;
        const float alpha = fminf(0.99f, a1.y * exp_hot(power));
        bool v = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
        const float test_T = T * (1.0f - alpha);
        const bool stop = v && test_T < 0.0001f;
        done = done || stop;
        v = v && !stop;
        if (__any_sync(0xffffffffu, v)) {
          const float4 a2 = lds_f4(rk + 32u);
          const float w = v ? alpha * T : 0.0f;
          C[0] += a1.z * w;
          C[1] += a1.w * w;
          C[2] += a2.x * w;
          if (CH == 4) C[3] += a2.y * w;
          T = v ? test_T : T;
          last_contributor = v ? __float_as_uint(a2.w) : last_contributor;
          last_entry = v ? pos0 + k : last_entry;
        }
      }
#ifdef TSPLAT_TIMELINE
      n_refs += cnt;
#endif
      if (__all_sync(0xffffffffu, done)) {  // warp-vote early termination
        abandon_job(s, cur, b);
        break;
      }
    }
#ifdef TSPLAT_TIMELINE
    n_jobs++;
#endif

    {
      // render_backward's work units: this list, cut at the block's deepest contributor (nothing behind it contributes:
      // half of a C2 block list on average), in segments of kSegLen references - independent thanks to the checkpoints
      // above.  Full segments are appended from the front of the table, the partial last one from its back.
      const unsigned int len = min(cur.count, (unsigned int)__reduce_max_sync(0xffffffffu, last_entry));
      const unsigned int nfull = len / kSegLen, rem = len % kSegLen;
      unsigned int at_full = 0, at_part = 0;
      if (lane == 0) {
        if (nfull) at_full = atomicAdd(&ws.ctl->seg_full, nfull);
        if (rem) at_part = atomicAdd(&ws.ctl->seg_part, 1u);
      }
      at_full = __shfl_sync(0xffffffffu, at_full, 0);
      TS_CHECK(at_full + nfull + at_part < ws.seg_cap);
      for (unsigned int k = lane; k < nfull; k += 32) ws.segs[at_full + k] = make_uint4(cur.start, len, cur.id, k);
      if (lane == 0 && rem) ws.segs[ws.seg_cap - 1 - at_part] = make_uint4(cur.start, len, cur.id, nfull);
    }
    float psq = 0.0f;  // per-view PSNR (metrics.py:12-19): sum over channels of (clip(target) - clip(colour))^2
    float sq = 0.0f;   // fused MSE (loss_mse.py:30): this pixel's sum over channels of (colour - target)^2
    if (inside) {
      const size_t pix = (size_t)py * W + px;
      const float* bg = cams[n].bg;
      ws.final_T[n * P + pix] = T;
      ws.n_contrib[n * P + pix] = last_contributor;
      {  // the composited sums themselves (no background, raw depth channel): the backward's segment restarts need them
        float* cf = ws.c_final + (size_t)n * 4 * P + pix;
        cf[0] = C[0]; cf[P] = C[1]; cf[2 * P] = C[2]; cf[3 * P] = CH == 4 ? C[CH - 1] : 0.0f;
      }
      float c0 = C[0] + T * __ldg(bg + 0);
      const float c1 = C[1] + T * __ldg(bg + 1), c2 = C[2] + T * __ldg(bg + 2);
      // a call whose pair list or block lists overflowed their capacity is incomplete: poison it where nobody can
      // miss it (first colour value of the call; the fused loss below)
      if (cur.id == 0u && lane == 0 && ws.ctl->overflow) c0 = __int_as_float(0x7fc00000);
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
      // Deterministic loss: job partial (fixed shuffle tree) -> the warp that completes a tile adds its 8 job partials
      // in block order -> the warp that completes the last tile adds all tile partials in index order.  Which warp
      // ran which job, and in which order jobs finished, does not enter the value.
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
        psq += __shfl_xor_sync(0xffffffffu, psq, o);
      }
      unsigned int final_pass = 0;
      if (lane == 0) {
        job_part[cur.id] = sq;
        if (psnr != nullptr) pjob_part[cur.id] = psq;
        __threadfence();
        if (atomicAdd(&ws.tile_jobs_done[bid], 1u) == kBlocksPerTile - 1) {
          __threadfence();
          float t = 0.0f, u = 0.0f;
#pragma unroll
          for (int i = 0; i < kBlocksPerTile; i++) {
            t += __ldcg(job_part + (size_t)bid * kBlocksPerTile + i);
            if (psnr != nullptr) u += __ldcg(pjob_part + (size_t)bid * kBlocksPerTile + i);
          }
          tile_part[bid] = t;
          if (psnr != nullptr) ptile_part[bid] = u;
          __threadfence();
          final_pass = atomicAdd(&ws.ctl->loss_done, 1u) == NT - 1 ? 1u : 0u;
        }
      }
      final_pass = __shfl_sync(0xffffffffu, final_pass, 0);
      if (final_pass) {
        __threadfence();
        float t = 0.0f;
        for (unsigned int i = lane; i < NT; i += 32) t += __ldcg(tile_part + i);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) *loss_sum = ws.ctl->overflow ? __int_as_float(0x7fc00000) : t * loss_scale;
        if (psnr != nullptr) {
          for (int v = 0; v < ws.N; v++) {
            float u = 0.0f;
            for (int i = lane; i < ws.T; i += 32) u += __ldcg(ptile_part + (size_t)v * ws.T + i);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
            // -10 log10(mean over c,h,w); a perfect match gives +inf as in the reference
            if (lane == 0) psnr[v] = -10.0f * log10f(u / (3.0f * (float)H * (float)W));
          }
        }
      }
    }
    cur_idx = take_job(&ws.ctl->fwd_next, W_all, lane);
    cur = load_job(ws, ws.block_jobs, cur_idx, njobs, H, W);
    s.p_next = 0; s.p_b = 0; s.pf_job = -1;  // the stream starts afresh (nothing of the finished job is in flight)
  }
#ifdef TSPLAT_TIMELINE
  if (lane == 0 && gw < 8192) {
    g_timeline[gw * 4 + 0] = t_start; g_timeline[gw * 4 + 1] = gtime(); unsigned int smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    g_timeline[gw * 4 + 2] = n_jobs | ((unsigned long long)smid << 32); g_timeline[gw * 4 + 3] = n_refs;
  }
#endif
}

// ---------------------------------------------------------------------------------------------
// Backward: the warp walks its block list back to front from the block's deepest contributor.  Per entry the 10
// gradient sums (moments of dL/dG G over the pixel offsets, dL/d opacity, dL/d colour[CH]) are reduced over the 32
// pixels with a transposing butterfly (14 shuffles) that leaves one sum per writer lane, and the 10 writer lanes add
// them to the Gaussian's screen-space accumulator with ONE warp-wide red.global.add.f32 (40 contiguous bytes: two L2
// sectors).  The reference issues one global atomic per (pixel, Gaussian, component).
// ---------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(kRenderThreads, TSPLAT_BWD_MIN_BLOCKS) render_backward_kernel(
    Workspace ws, int H, int W, int G, const TsplatCamera* __restrict__ cams, const float* __restrict__ dL_dcolor,
    const float* __restrict__ dL_ddepth, TsplatMse mse) {
  __shared__ __align__(128) WarpSmem sm;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Stream s = make_stream(sm, warp, lane);
  const unsigned int njobs = (unsigned int)(ws.N * ws.T) * kBlocksPerTile;
  const size_t P = (size_t)H * W;

  // ---- reduce_plan: which lane ends up with which of the 10 per-entry sums (lane-invariant across jobs) ----
  // Lane bits b4 b3 b2 pick the component, bits b1 b0 are summed away.  Slots of the accumulator record:
  //   0 S_x  1 S_y  2 S_xx  3 S_xy  4 S_yy  5 dL/d opacity  6..9 dL/d colour (r, g, b, depth feature).
  // The swaps of a transposing butterfly cost two selects per exchanged pair; here they are pushed into operands that
  // are constant along the list: the colour sums are alpha T dL_dpix[c], so each lane multiplies by ITS permutation
  // of dL_dpix (LAk, LAs, LBk, LBs) and rounds 1 and 2 need no select for them; S_x|S_y and S_xx|S_yy share the
  // swapped offsets (dk, ds).  S_xy | opacity run as a sixth value through plain sums.  9 selects instead of 22.
  const bool b4i = lane & 16, b3i = lane & 8;
  const bool g2_writer = (lane & 15) == 1;  // lane 1: S_xy, lane 17: opacity
  const int my_comp = g2_writer ? (b4i ? 5 : 3)
                                : ((lane & 4) ? 6 + (b3i ? 2 : 0) + (b4i ? 1 : 0) : (b3i ? (b4i ? 4 : 2) : (b4i ? 1 : 0)));
  const bool my_red = g2_writer || ((lane & 3) == 0 && (CH == 4 || my_comp != 9));

  // work units: list segments (emitted by render_forward), dealt round-robin: all full segments (equal work) first, then
  // the partial ones - every warp gets the same number of full segments +- 1 and of partial ones +- 1
  const unsigned int nfull_all = ws.ctl->seg_full, nseg = nfull_all + ws.ctl->seg_part;
  const unsigned int W_all = gridDim.x * kRenderWarps, gw = blockIdx.x * kRenderWarps + warp;
  auto load_seg = [&](unsigned int idx, unsigned int& list_start, unsigned int& list_len, unsigned int& k) {
    JobRec r{0u, 0u, 0u};
    list_start = 0u; list_len = 0u; k = 0u;
    if (idx < nseg) {
      const uint4 q = __ldcg(ws.segs + (idx < nfull_all ? idx : ws.seg_cap - 1 - (idx - nfull_all)));
      list_start = q.x; list_len = q.y; k = q.w;
      r.start = q.x + q.w * kSegLen;
      r.count = min((unsigned int)kSegLen, q.y - q.w * kSegLen);
      r.id = q.z;
    }
    return r;
  };
  unsigned int cur_ls, cur_ll, cur_k, nxt_ls, nxt_ll, nxt_k;
  JobRec cur = load_seg(gw, cur_ls, cur_ll, cur_k), nxt = load_seg(gw + W_all, nxt_ls, nxt_ll, nxt_k);
  for (unsigned int idx = gw; idx < nseg; idx += W_all) {
    unsigned int fut_ls, fut_ll, fut_k;
    const JobRec fut = load_seg(idx + 2 * W_all, fut_ls, fut_ll, fut_k);  // in flight while this unit runs
    pump<true>(s, cur, nxt, ws.bref, lane);
    const int count = (int)cur.count;  // > 0 by construction
    {
      const int blk = (int)(cur.id & 7u);
      const int bid = (int)(cur.id >> 3);
      const int n = bid / ws.T, tile = bid - n * ws.T;
      const int ty = tile / ws.tiles_x, tx = tile - ty * ws.tiles_x;
      int px, py;
      pixel_of_lane(blk, lane, tx, ty, px, py);
      const bool inside = px < W && py < H;
      const float pxf = (float)px, pyf = (float)py;
      const size_t pix = (size_t)py * W + px;

      const float T_final = inside ? ws.final_T[n * P + pix] : 0.0f;
      float T = T_final;
      const unsigned int last_contributor = inside ? ws.n_contrib[n * P + pix] : 0u;
      // this unit ends at reference `hi` of its list; unless that is the list's (cut) end, the walk restarts from the
      // checkpoint render_forward left in front of reference hi
      const unsigned int hi = cur_k * kSegLen + (unsigned int)count;
      const bool restart = hi < cur_ll;
      TS_CHECK(!restart || (size_t)(cur_ls / kSegLen) + hi / kSegLen < ws.e_cap / kSegLen + 2);
      const float* ck = ws.ckpt + ((size_t)(cur_ls / kSegLen) + hi / kSegLen) * kCkptFloats + lane;
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
        if (restart) {
          // colour behind reference hi as seen from there: (C_final - C_prefix) / T_prefix per channel, contracted with
          // dL/dpixel - the value the reference's accum_rec recurrence has reached at this point of its walk
          const float Tc = ck[0];
          const float* cf = ws.c_final + (size_t)n * 4 * P + pix;
          float a = 0.0f;
#pragma unroll
          for (int ch = 0; ch < CH; ch++) a += dL_dpix[ch] * (cf[(size_t)ch * P] - ck[32 * (ch + 1)]);
          accum_dl = a * rcp_hot(Tc) ;
          T = Tc;
        }
      }
      const float* bgp = cams[n].bg;
      // bg of the fused depth channel is 0 (render_depth_cuda passes zeros, cuda_splatting.py:410)
      const float bg_dot = __ldg(bgp) * dL_dpix[0] + __ldg(bgp + 1) * dL_dpix[1] + __ldg(bgp + 2) * dL_dpix[2];
      const float bgT = -T_final * bg_dot;
      const float dl0 = dL_dpix[0], dl1 = dL_dpix[1], dl2 = dL_dpix[2], dl3 = CH == 4 ? dL_dpix[CH - 1] : 0.0f;
      float LAk = b3i ? (b4i ? dl3 : dl2) : (b4i ? dl1 : dl0), LAs = b3i ? (b4i ? dl2 : dl3) : (b4i ? dl0 : dl1);
      float LBk = b3i ? (b4i ? dl1 : dl0) : (b4i ? dl3 : dl2), LBs = b3i ? (b4i ? dl0 : dl1) : (b4i ? dl2 : dl3);
      // opaque to ptxas, which would otherwise rebuild the four permuted factors and the lane bits on every visit
      int lane_bits = lane | (g2_writer ? 32 : 0);
      asm volatile("" : "+f"(LAk), "+f"(LAs), "+f"(LBk), "+f"(LBs), "+r"(lane_bits));
      // this lane's slot of the accumulator record of Gaussian 0 of the view
      float* const acc_lane = ws.grad_acc + (size_t)n * G * kAccFloats + my_comp;

      // batch b (b = 0 is the END of the list) is chunk nb - 1 - b of the block list
      const int nb = batches_of(cur);
      for (int b = 0; b < nb; b++) {
        const int cnt = batch_cnt<true>(cur, b);
        uint32_t ra = stage_batch<true>(s, ws, G, cur, nxt, b, lane) + (uint32_t)((cnt - 1) * kRecBytes);
#pragma unroll kBwdUnroll
        for (int k = cnt - 1; k >= 0; k--) {
          const uint32_t rk = ra - (uint32_t)((cnt - 1 - k) * kRecBytes);
          const float4 a0 = lds_f4(rk);
          const float4 a1 = lds_f4(rk + 16u);
          const float4 a2 = lds_f4(rk + 32u);
          const float dx = a0.x - pxf, dy = a0.y - pyf;
          const float power = -0.5f * (a0.z * dx * dx + a1.x * dy * dy) - a0.w * dx * dy;
          const float Gv = exp_hot(power);
          const float alpha = fminf(0.99f, a1.y * Gv);
          // entries behind this pixel's last contributor are skipped (tile-list positions, 1-based)
          const bool valid = (__float_as_uint(a2.w) <= last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
          if (!__any_sync(0xffffffffu, valid)) continue;
          // No branch on `valid`.  A lane that does not contribute runs the same arithmetic with alpha = 0: that leaves
          // T unchanged (T / (1 - 0)), makes every accumulated product 0, and turns the suffix recurrence into the
          // identity for the next step (w = 0, 1 - alpha = 1), exactly as if the entry had been skipped.
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
          // ---- transposing warp reduction of the 10 sums (see reduce_plan above) ----
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
          TS_CHECK(__float_as_uint(a2.z) < (unsigned int)G);
          if (my_red) red_global_add(acc_lane + (size_t)__float_as_uint(a2.z) * kAccFloats, r);
        }
      }
    }
    advance_job(s, cur, nxt, fut);
    cur_ls = nxt_ls; cur_ll = nxt_ll; cur_k = nxt_k;
    nxt_ls = fut_ls; nxt_ll = fut_ll; nxt_k = fut_k;
  }
}

// ---------------------------------------------------------------------------------------------
// render_backward, two passes per batch (the default).  The butterfly above costs 37 of a visit's 95 instructions
// because every (warp, entry) visit reduces its 10 sums over the 32 PIXEL lanes.  Here the warp changes its lane
// mapping in the middle of a batch instead:
//   pass 1, lane = pixel: the sequential part - alpha, the transmittance and suffix recurrences, dL/dalpha - for the
//     batch's entries back to front; per (entry, pixel) only two numbers leave: v = G dL/dalpha and w = alpha T, stored
//     into a [entry][pixel] matrix in shared memory (row stride 33 float2: both passes conflict-free);
//   pass 2, lane = ENTRY: lane k reads row k of the matrix, its own record (centre, opacity, id) and the block's 32
//     dL/dpixel vectors (broadcast reads), and accumulates its Gaussian's 10 sums over the pixels in its own registers:
//     no cross-lane reduction at all.  Three vector red.global.add per lane and batch deliver them.
// Per visit: ~47 (pass 1) + 16 (pass 2) instructions instead of 111.  kBwdRows = 16 halves the matrix (more resident
// warps): lane = (entry, half of the pixels), one shuffle per sum joins the halves.
// ---------------------------------------------------------------------------------------------
#ifndef TSPLAT_BWD_ROWS
#define TSPLAT_BWD_ROWS 16  // C2, render_backward: 32 rows (2 CTAs per SM) 0.1946 / 0.1935 ms at unroll 1 / 2; 16 rows (3 CTAs) 0.1935 / 0.1919
#endif
constexpr int kBwd2Unroll = TSPLAT_BWD2_UNROLL;    // pass-1 visits in flight
constexpr int kBwdRows = TSPLAT_BWD_ROWS;          // entries per pass-1 / pass-2 round: 32 or 16
constexpr int kBwdParts = 32 / kBwdRows;           // lanes per entry in pass 2
constexpr int kBwdPixPerPart = 32 / kBwdParts;     // pixels each of them sums
constexpr int kMatRowBytes = 33 * 8;

struct BwdSmem {
  WarpSmem w;
  float2 mat[kRenderWarps][kBwdRows][33];
  float4 dl[kRenderWarps][32];
};

__device__ __forceinline__ void sts_f2(uint32_t a, float x, float y) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ float2 lds_f2(uint32_t a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds_f1(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void red_global_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_global_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

template <int CH>
__global__ void __launch_bounds__(kRenderThreads, TSPLAT_BWD2_MIN_BLOCKS) render_backward2_kernel(
    Workspace ws, int H, int W, int G, const TsplatCamera* __restrict__ cams, const float* __restrict__ dL_dcolor,
    const float* __restrict__ dL_ddepth, TsplatMse mse) {
  extern __shared__ __align__(128) unsigned char bwd_smem_raw[];
  BwdSmem& sm = *reinterpret_cast<BwdSmem*>(bwd_smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Stream s = make_stream(sm.w, warp, lane);
  const size_t P = (size_t)H * W;
  const uint32_t mat_addr = smem_u32(&sm.mat[warp][0][0]);
  const uint32_t dl_addr = smem_u32(&sm.dl[warp][0]);
  // pass 2: this lane's entry of the round and its share of the block's pixels
  const int row = lane & (kBwdRows - 1), part = lane / kBwdRows;
  const uint32_t my_row = mat_addr + (uint32_t)(row * kMatRowBytes + part * kBwdPixPerPart * 8);
  const uint32_t my_dl = dl_addr + (uint32_t)(part * kBwdPixPerPart * 16);

  // work units: list segments (emitted by render_forward), dealt round-robin: all full segments (equal work) first, then
  // the partial ones
  const unsigned int nfull_all = ws.ctl->seg_full, nseg = nfull_all + ws.ctl->seg_part;
  const unsigned int W_all = gridDim.x * kRenderWarps, gw = blockIdx.x * kRenderWarps + warp;
  auto load_seg = [&](unsigned int idx, unsigned int& list_start, unsigned int& list_len, unsigned int& k) {
    JobRec r{0u, 0u, 0u};
    list_start = 0u; list_len = 0u; k = 0u;
    if (idx < nseg) {
      const uint4 q = __ldcg(ws.segs + (idx < nfull_all ? idx : ws.seg_cap - 1 - (idx - nfull_all)));
      list_start = q.x; list_len = q.y; k = q.w;
      r.start = q.x + q.w * kSegLen;
      r.count = min((unsigned int)kSegLen, q.y - q.w * kSegLen);
      r.id = q.z;
    }
    return r;
  };
  unsigned int cur_ls, cur_ll, cur_k, nxt_ls, nxt_ll, nxt_k;
  JobRec cur = load_seg(gw, cur_ls, cur_ll, cur_k), nxt = load_seg(gw + W_all, nxt_ls, nxt_ll, nxt_k);
  for (unsigned int idx = gw; idx < nseg; idx += W_all) {
    unsigned int fut_ls, fut_ll, fut_k;
    const JobRec fut = load_seg(idx + 2 * W_all, fut_ls, fut_ll, fut_k);  // in flight while this unit runs
    pump<true>(s, cur, nxt, ws.bref, lane);
    const int count = (int)cur.count;  // > 0 by construction
    {
      const int blk = (int)(cur.id & 7u);
      const int bid = (int)(cur.id >> 3);
      const int n = bid / ws.T, tile = bid - n * ws.T;
      const int ty = tile / ws.tiles_x, tx = tile - ty * ws.tiles_x;
      int px, py;
      pixel_of_lane(blk, lane, tx, ty, px, py);
      const bool inside = px < W && py < H;
      const float pxf = (float)px, pyf = (float)py;
      const size_t pix = (size_t)py * W + px;
      // origin of this lane's pixels in pass 2 (pixel p of the block = column p & 7, row p >> 3)
      const float ox = (float)(px - (lane & 7)), oy = (float)(py - (lane >> 3) + part * (kBwdPixPerPart / 8));

      const float T_final = inside ? ws.final_T[n * P + pix] : 0.0f;
      float T = T_final;
      const unsigned int last_contributor = inside ? ws.n_contrib[n * P + pix] : 0u;
      const unsigned int hi = cur_k * kSegLen + (unsigned int)count;
      const bool restart = hi < cur_ll;  // see render_backward_kernel
      TS_CHECK(!restart || (size_t)(cur_ls / kSegLen) + hi / kSegLen < ws.e_cap / kSegLen + 2);
      const float* ck = ws.ckpt + ((size_t)(cur_ls / kSegLen) + hi / kSegLen) * kCkptFloats + lane;
      float dL_dpix[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      float accum_dl = 0.0f, w_last = 0.0f, oml_last = 1.0f;
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
        if (restart) {
          const float Tc = ck[0];
          const float* cf = ws.c_final + (size_t)n * 4 * P + pix;
          float a = 0.0f;
#pragma unroll
          for (int ch = 0; ch < CH; ch++) a += dL_dpix[ch] * (cf[(size_t)ch * P] - ck[32 * (ch + 1)]);
          accum_dl = a * rcp_hot(Tc);
          T = Tc;
        }
      }
      __syncwarp();  // the previous unit's pass 2 has read dl
      sts_f4(dl_addr + (uint32_t)lane * 16u, make_float4(dL_dpix[0], dL_dpix[1], dL_dpix[2], dL_dpix[3]));
      const float* bgp = cams[n].bg;
      // bg of the fused depth channel is 0 (render_depth_cuda passes zeros, cuda_splatting.py:410)
      const float bg_dot = __ldg(bgp) * dL_dpix[0] + __ldg(bgp + 1) * dL_dpix[1] + __ldg(bgp + 2) * dL_dpix[2];
      const float bgT = -T_final * bg_dot;
      float* const acc_view = ws.grad_acc + (size_t)n * G * kAccFloats;

      // batch b (b = 0 is the END of the list) is chunk nb - 1 - b of the block list
      const int nb = batches_of(cur);
      for (int b = 0; b < nb; b++) {
        const int cnt = batch_cnt<true>(cur, b);
        const uint32_t base = stage_batch<true>(s, ws, G, cur, nxt, b, lane);
        for (int k_hi = cnt; k_hi > 0; k_hi -= kBwdRows) {
          const int k_lo = k_hi > kBwdRows ? k_hi - kBwdRows : 0;
          // ---- pass 1: lane = pixel ----
          unsigned int rowmask = 0u;
          uint32_t rk = base + (uint32_t)((k_hi - 1) * kRecBytes);
          uint32_t mk = mat_addr + (uint32_t)((k_hi - 1 - k_lo) * kMatRowBytes + lane * 8);
#pragma unroll kBwd2Unroll
          for (int k = k_hi - 1; k >= k_lo; k--, rk -= kRecBytes, mk -= kMatRowBytes) {
            const float4 a0 = lds_f4(rk);
            const float4 a1 = lds_f4(rk + 16u);
            const float4 a2 = lds_f4(rk + 32u);
            const float dx = a0.x - pxf, dy = a0.y - pyf;
            const float power = -0.5f * (a0.z * dx * dx + a1.x * dy * dy) - a0.w * dx * dy;
            const float Gv = exp_hot(power);
            const float alpha = fminf(0.99f, a1.y * Gv);
            // entries behind this pixel's last contributor are skipped (tile-list positions, 1-based)
            const bool valid = (__float_as_uint(a2.w) <= last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(0xffffffffu, valid)) continue;
            rowmask |= 1u << (k - k_lo);
            // no branch on `valid`: a lane that does not contribute runs the same arithmetic with alpha = 0
            const float Gm = valid ? Gv : 0.0f, am = valid ? alpha : 0.0f;
            const float col[4] = {a1.z, a1.w, a2.x, a2.y};
            const float oma = 1.0f - am;
            const float inv1ma = rcp_hot(oma);
            const float Tn = T * inv1ma;
            float cdl = 0.0f;
#pragma unroll
            for (int ch = 0; ch < CH; ch++) cdl += col[ch] * dL_dpix[ch];
            const float acc_new = w_last + oml_last * accum_dl;
            const float dL_dalpha = (cdl - acc_new) * Tn + bgT * inv1ma;
            T = Tn; accum_dl = acc_new; w_last = am * cdl; oml_last = oma;
            sts_f2(mk, Gm * dL_dalpha, am * Tn);
          }
          __syncwarp();
          // ---- pass 2: lane = entry (x part of the pixels) ----
          if (rowmask != 0u) {
            const uint32_t rr = base + (uint32_t)((k_lo + row) * kRecBytes);
            const float2 cxy = lds_f2(rr);
            const float opac = lds_f1(rr + 20u);
            const unsigned int gid = __float_as_uint(lds_f1(rr + 40u));
            const float gx = cxy.x - ox, gy = cxy.y - oy;
            float Sx = 0.f, Sy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f, Sop = 0.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, C3 = 0.f;
            float dxc[8];
#pragma unroll
            for (int c = 0; c < 8; c++) dxc[c] = gx - (float)c;
#pragma unroll
            for (int i = 0; i < kBwdPixPerPart; i++) {
              const float2 vw = lds_f2(my_row + (uint32_t)i * 8u);
              const float4 dl = lds_f4(my_dl + (uint32_t)i * 16u);
              const float dx = dxc[i & 7], dy = gy - (float)(i >> 3);
              const float u = opac * vw.x;
              const float ux = u * dx, uy = u * dy;
              Sx += ux; Sy += uy;
              Sxx += ux * dx; Sxy += ux * dy; Syy += uy * dy;
              Sop += vw.x;
              C0 += vw.y * dl.x; C1 += vw.y * dl.y; C2 += vw.y * dl.z;
              if (CH == 4) C3 += vw.y * dl.w;
            }
            if (kBwdParts > 1) {  // join the parts of an entry: lanes row, row + kBwdRows, ...
              const unsigned int full = 0xffffffffu;
#pragma unroll
              for (int off = kBwdRows; off < 32; off <<= 1) {
                Sx += __shfl_xor_sync(full, Sx, off); Sy += __shfl_xor_sync(full, Sy, off);
                Sxx += __shfl_xor_sync(full, Sxx, off); Sxy += __shfl_xor_sync(full, Sxy, off);
                Syy += __shfl_xor_sync(full, Syy, off); Sop += __shfl_xor_sync(full, Sop, off);
                C0 += __shfl_xor_sync(full, C0, off); C1 += __shfl_xor_sync(full, C1, off);
                C2 += __shfl_xor_sync(full, C2, off);
                if (CH == 4) C3 += __shfl_xor_sync(full, C3, off);
              }
            }
            if (part == 0 && ((rowmask >> row) & 1u)) {
              TS_CHECK(gid < (unsigned int)G);
              float* acc = acc_view + (size_t)gid * kAccFloats;
              red_global_add_v4(acc, Sx, Sy, Sxx, Sxy);
              red_global_add_v4(acc + 4, Syy, Sop, C0, C1);
              if (CH == 4) red_global_add_v2(acc + 8, C2, C3);
              else red_global_add(acc + 8, C2);
            }
          }
          __syncwarp();  // the next round's pass 1 rewrites the matrix
        }
      }
    }
    advance_job(s, cur, nxt, fut);
    cur_ls = nxt_ls; cur_ll = nxt_ll; cur_k = nxt_k;
    nxt_ls = fut_ls; nxt_ll = fut_ll; nxt_k = fut_k;
  }
}

// persistent grids: resident CTAs per SM x SMs of the device, queried once per kernel
template <class K>
static int persistent_grid(K kernel, int dyn_smem = 0) {
  int dev = 0, sms = 148, per_sm = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kRenderThreads, dyn_smem) != cudaSuccess || per_sm < 1) per_sm = 1;
#ifdef TSPLAT_RENDER_CTAS_PER_SM
  if (per_sm > TSPLAT_RENDER_CTAS_PER_SM) per_sm = TSPLAT_RENDER_CTAS_PER_SM;
#endif
  return sms * per_sm;
}

cudaError_t launch_render_forward(const TsplatProblem& p, const Workspace& ws, float* out_color, float* out_depth,
                                  const TsplatMse* mse, cudaStream_t st) {
  const int njobs = ws.N * ws.T * kBlocksPerTile;
  if (njobs == 0) return cudaSuccess;
  const float* target = mse ? mse->target : nullptr;
  float* loss_sum = mse ? mse->loss_sum : nullptr;
  const float loss_scale = mse ? mse->loss_scale : 0.0f;
  float* psnr = mse ? mse->psnr : nullptr;
  const int cap = (njobs + kRenderWarps - 1) / kRenderWarps;
  if (p.depth_mode != TSPLAT_DEPTH_NONE) {
    static const int per = persistent_grid(render_forward_kernel<4>);
    render_forward_kernel<4><<<min(per, cap), kRenderThreads, 0, st>>>(ws, p.H, p.W, p.G, p.cams, out_color, out_depth, target, loss_sum, loss_scale, psnr);
  } else {
    static const int per = persistent_grid(render_forward_kernel<3>);
    render_forward_kernel<3><<<min(per, cap), kRenderThreads, 0, st>>>(ws, p.H, p.W, p.G, p.cams, out_color, out_depth, target, loss_sum, loss_scale, psnr);
  }
  return cudaGetLastError();
}

cudaError_t launch_render_backward(const TsplatProblem& p, const Workspace& ws, const float* dL_dcolor,
                                   const float* dL_ddepth, const TsplatMse* mse_in, cudaStream_t st) {
  const int njobs = ws.N * ws.T * kBlocksPerTile;
  if (njobs == 0) return cudaSuccess;
  TsplatMse mse{};
  if (mse_in) mse = *mse_in;
  const int cap = (njobs + kRenderWarps - 1) / kRenderWarps;
#if TSPLAT_BWD_TWOPASS
  const int smem = (int)sizeof(BwdSmem);
  if (p.depth_mode != TSPLAT_DEPTH_NONE) {
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(render_backward2_kernel<4>), smem);
    if (e != cudaSuccess) return e;
    static const int per = persistent_grid(render_backward2_kernel<4>, smem);
    render_backward2_kernel<4><<<min(per, cap), kRenderThreads, smem, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  } else {
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(render_backward2_kernel<3>), smem);
    if (e != cudaSuccess) return e;
    static const int per = persistent_grid(render_backward2_kernel<3>, smem);
    render_backward2_kernel<3><<<min(per, cap), kRenderThreads, smem, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  }
#else
  if (p.depth_mode != TSPLAT_DEPTH_NONE) {
    static const int per = persistent_grid(render_backward_kernel<4>);
    render_backward_kernel<4><<<min(per, cap), kRenderThreads, 0, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  } else {
    static const int per = persistent_grid(render_backward_kernel<3>);
    render_backward_kernel<3><<<min(per, cap), kRenderThreads, 0, st>>>(ws, p.H, p.W, p.G, p.cams, dL_dcolor, dL_ddepth, mse);
  }
#endif
  return cudaGetLastError();
}

}  // namespace tsplat

#ifdef TSPLAT_TIMELINE
extern "C" int tsplat_debug_timeline(unsigned long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, tsplat::g_timeline, sizeof(unsigned long long) * n);
}
#endif
