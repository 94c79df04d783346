// blocklist.cu — per-8x4-pixel-block lists and the render job tables: the last binning stage, shared by both sort paths.
//
// A 16x16 tile is rendered by 8 warps, one 8x4 pixel block each.  In round 1 the 8 warps of a tile streamed the SAME
// tile list and each skipped the entries its block cannot see (an 8-bit block mask per pair), shared one ring and waited
// for the tile's slowest warp.  Here every block gets its OWN compact list: the entries of the tile's depth-sorted list
// whose mask has the block's bit, in list order, as 8-byte references (Gaussian id, 1-based position in the TILE list).
// A render warp then walks exactly what it composites (render.cu): no cull, no queue, no ring shared between warps,
// 8 N T independent jobs.  The tile-list position is what n_contrib stores and what the parity contract fixes (the tile
// lists themselves - keys, values, ranges - stay as they are and bit-exact).  A pair is listed once per block it can
// touch: E = 2.2 D references at C2, which is why the lists hold references and the render warps gather the 48-B records
// from the L2-resident per-Gaussian array themselves (copying the records E times cost 118 MB of DRAM writes per step).
//
// block_lists_kernel: one CTA per (view, tile); each of its 16 warps owns a contiguous run of the tile list.  Pass 1:
// per pair, the per-Gaussian record's cull inputs -> block mask (cull.cuh), masks stored, per-(warp, block) counts; one
// exchange of those counts and one returning atomic per tile reserve the tile's 8 lists as one contiguous run of the
// reference buffer.  Pass 2: per chunk of 32 pairs, ballot ranks + the warp's running per-block positions (registers)
// place every (pair, block) reference - no shared memory, no CTA barrier.  The CTA of slot s of the heaviest-first tile order also writes the records of render jobs
// 8 s .. 8 s + 7: render_forward deals them in that order (a separate counting sort of the jobs by their own list length
// cost 10 us and bought 6).
#include "cull.cuh"

namespace tsplat {

#ifndef TSPLAT_BL_THREADS
#define TSPLAT_BL_THREADS 512
#endif
#ifndef TSPLAT_BL_MIN_BLOCKS
#define TSPLAT_BL_MIN_BLOCKS 4  // 32 registers, 4 CTAs per SM: 2 / 3 / 4 -> 53.7 / 53.9 / 49.5 us at C2
#endif
constexpr int kBlThreads = TSPLAT_BL_THREADS;
constexpr int kBlWarps = kBlThreads / 32;

__global__ void __launch_bounds__(kBlThreads, TSPLAT_BL_MIN_BLOCKS) block_lists_kernel(Workspace ws, int G, const unsigned int* __restrict__ vals) {
  __shared__ unsigned int s_start[kBlocksPerTile];             // first entry of each block's list (clipped to e_cap)
  __shared__ unsigned int s_lim[kBlocksPerTile];               // one past its last usable entry
  __shared__ unsigned int s_wcnt[kBlWarps][kBlocksPerTile];    // per-warp counts -> exclusive offsets within the block's list
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slot = (int)blockIdx.x;                            // position in the heaviest-first order = render job order
  const int bid = ws.use_order ? (int)ws.tile_order[slot] : slot;
  const uint2 range = ws.ranges[bid];
  const int n = (int)(range.y - range.x);
  uint4* out_jobs = ws.block_jobs + (size_t)slot * kBlocksPerTile;
  if (n == 0 || n >= (1 << 24)) {  // (16-bit per-thread counters below: a single tile with 16 M pairs is not supported)
    if (tid < kBlocksPerTile) out_jobs[tid] = make_uint4(0u, 0u, (unsigned int)bid * kBlocksPerTile + tid, 0u);
    if (n != 0 && tid == 0) ws.ctl->overflow = 1u;
    return;
  }
  const int view = bid / ws.T, tile = bid - view * ws.T;
  const int ty = tile / ws.tiles_x, tx = tile - ty * ws.tiles_x;
  const float tileX0 = (float)(tx * kTile), tileY0 = (float)(ty * kTile);
  const float4* __restrict__ grec = ws.grec + (size_t)view * G * kRecFloat4;
  // Warp w owns a contiguous run of 32-pair chunks of the tile list: list order = (warp, chunk, lane) order, so after
  // one exchange of the per-warp counts the placement pass needs neither shared memory nor CTA barriers.
  const int chunks = (n + 31) >> 5;
  const int c_lo = (int)(((long long)chunks * warp) / kBlWarps), c_hi = (int)(((long long)chunks * (warp + 1)) / kBlWarps);

  // ---- pass 1: block masks and per-block counts ----
  unsigned int c01 = 0, c23 = 0, c45 = 0, c67 = 0;  // pairs of 16-bit counters (a lane sees at most n / 512 + 1 pairs)
  for (int c = c_lo; c < c_hi; c++) {
    const int j = (c << 5) + lane;
    if (j < n) {
      const unsigned int id = vals[range.x + j];
      TS_CHECK(id < (unsigned int)G && range.x + (unsigned int)j < ws.d_cap);
      const float4* __restrict__ gr = grec + (size_t)id * kRecFloat4;
      const float4 g0 = gr[0];
      const float cq = gr[1].x, tau = gr[2].w;
      const unsigned int m = cull_block_mask(g0, cq, tau, tileX0, tileY0);
      ws.pair_mask[range.x + j] = (unsigned char)m;
      c01 += (m & 1u) | ((m & 2u) << 15);
      c23 += ((m >> 2) & 1u) | ((m & 8u) << 13);
      c45 += ((m >> 4) & 1u) | ((m & 32u) << 11);
      c67 += ((m >> 6) & 1u) | ((m & 128u) << 9);
    }
  }
  {
    const unsigned int c[8] = {c01 & 0xffffu, c01 >> 16, c23 & 0xffffu, c23 >> 16, c45 & 0xffffu, c45 >> 16, c67 & 0xffffu, c67 >> 16};
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const unsigned int t = __reduce_add_sync(0xffffffffu, c[b]);
      if (lane == 0) s_wcnt[warp][b] = t;
    }
  }
  __syncthreads();
  if (warp == 0) {
    // lane b < 8: exclusive scan of block b's per-warp counts; then lane 0 reserves the tile's 8 lists as one run
    unsigned int total_b = 0;
    if (lane < kBlocksPerTile) {
#pragma unroll
      for (int w = 0; w < kBlWarps; w++) {
        const unsigned int t = s_wcnt[w][lane];
        s_wcnt[w][lane] = total_b;
        total_b += t;
      }
    }
    // every list starts at an even reference index (16-B aligned: the render warps stream it with bulk copies)
    const unsigned int padded = (total_b + 1u) & ~1u;
    unsigned int incl = padded;
#pragma unroll
    for (int d = 1; d < kBlocksPerTile; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    const unsigned int total = __shfl_sync(0xffffffffu, incl, kBlocksPerTile - 1);
    unsigned long long base = 0;
    if (lane == 0) {
      base = atomicAdd(&ws.ctl->num_block_entries, (unsigned long long)total);
      if (base + total > ws.e_cap) ws.ctl->overflow = 1u;  // entries beyond the capacity are dropped, never written
    }
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < kBlocksPerTile) {
      const unsigned long long off = base + (incl - padded);
      const unsigned long long st = off < ws.e_cap ? off : ws.e_cap;
      const unsigned long long en = off + total_b < ws.e_cap ? off + total_b : ws.e_cap;
      s_start[lane] = (unsigned int)st;
      s_lim[lane] = (unsigned int)en;
      out_jobs[lane] = make_uint4((unsigned int)st, (unsigned int)(en - st), (unsigned int)bid * kBlocksPerTile + lane, 0u);
    }
  }
  __syncthreads();

  // ---- pass 2: place the entries; the warp's running position in each block's list lives in registers ----
  unsigned int pos_b[8];
#pragma unroll
  for (int b = 0; b < 8; b++) pos_b[b] = s_start[b] + s_wcnt[warp][b];
  const unsigned int e_cap = (unsigned int)ws.e_cap;  // a position at or beyond the capacity is dropped (overflow is flagged)
  const unsigned int lt = (1u << lane) - 1u;
  for (int c = c_lo; c < c_hi; c++) {
    const int j = (c << 5) + lane;
    const unsigned int m = j < n ? (unsigned int)ws.pair_mask[range.x + j] : 0u;  // this thread's own store of pass 1
    // (Gaussian id, 1-based position in the tile list: what n_contrib records)
    const uint2 ref = make_uint2(m ? vals[range.x + j] : 0u, (unsigned int)(j + 1));
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const unsigned int bal = __ballot_sync(0xffffffffu, (m >> b) & 1u);
      if ((m >> b) & 1u) {
        const unsigned int pos = pos_b[b] + __popc(bal & lt);
        TS_CHECK(pos < s_lim[b] || s_lim[b] == e_cap);
        if (pos < e_cap) ws.bref[pos] = ref;
      }
      pos_b[b] += __popc(bal);
    }
  }
}

__global__ void empty_jobs_kernel(Workspace ws, int njobs) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < njobs) ws.block_jobs[j] = make_uint4(0u, 0u, (unsigned int)j, 0u);
}

cudaError_t launch_empty_jobs(const Workspace& ws, cudaStream_t st) {
  const int njobs = ws.N * ws.T * kBlocksPerTile;
  if (njobs == 0) return cudaSuccess;
  empty_jobs_kernel<<<(njobs + 255) / 256, 256, 0, st>>>(ws, njobs);
  return cudaGetLastError();
}

cudaError_t launch_block_lists(const TsplatProblem& p, const Workspace& ws, cudaStream_t st) {
  const int NT = ws.N * ws.T;
  if (NT == 0) return cudaSuccess;
  block_lists_kernel<<<NT, kBlThreads, 0, st>>>(ws, p.G, ws.vals[ws.sort_passes & 1]);
  return cudaGetLastError();
}

}  // namespace tsplat
