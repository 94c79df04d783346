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
// block_lists_kernel: one CTA per (view, tile).  Pass 1: per pair, the per-Gaussian record's cull inputs -> block mask
// (cull.cuh), masks stored, per-block counts reduced; one returning atomic per tile reserves the tile's 8 lists as one
// contiguous run of the reference buffer.  Pass 2: per chunk of pairs, ballot ranks + running per-block bases place every
// (pair, block) reference.  The CTA of slot s of the heaviest-first tile order also writes the records of render jobs
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
  __shared__ unsigned int s_cnt[kBlocksPerTile];               // pass 1: entries per block
  __shared__ unsigned int s_start[kBlocksPerTile];             // first entry of each block's list (clipped to e_cap)
  __shared__ unsigned int s_lim[kBlocksPerTile];               // one past its last usable entry
  __shared__ unsigned int s_run[kBlocksPerTile];               // pass 2: entries placed by earlier chunks
  __shared__ unsigned int s_wcnt[kBlWarps][kBlocksPerTile];    // pass 2: per-warp counts of the current chunk -> exclusive offsets
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
  if (tid < kBlocksPerTile) { s_cnt[tid] = 0u; s_run[tid] = 0u; }
  __syncthreads();

  // ---- pass 1: block masks and per-block counts ----
  unsigned int c01 = 0, c23 = 0, c45 = 0, c67 = 0;  // pairs of 16-bit counters (a thread sees at most n / 1024 + 1 pairs)
  for (int j = tid; j < n; j += kBlThreads) {
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
  {
    unsigned int c[8] = {c01 & 0xffffu, c01 >> 16, c23 & 0xffffu, c23 >> 16, c45 & 0xffffu, c45 >> 16, c67 & 0xffffu, c67 >> 16};
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const unsigned int t = __reduce_add_sync(0xffffffffu, c[b]);
      if (lane == 0 && t) atomicAdd(&s_cnt[b], t);
    }
  }
  __syncthreads();
  if (tid == 0) {
    // every list starts at an even reference index (16-B aligned: the render warps stream it with bulk copies)
    unsigned int total = 0;
#pragma unroll
    for (int b = 0; b < 8; b++) total += (s_cnt[b] + 1u) & ~1u;
    const unsigned long long base = atomicAdd(&ws.ctl->num_block_entries, (unsigned long long)total);
    if (base + total > ws.e_cap) ws.ctl->overflow = 1u;  // entries beyond the capacity are dropped, never written
    unsigned long long off = base;
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const unsigned long long st = off < ws.e_cap ? off : ws.e_cap;
      const unsigned long long en = off + s_cnt[b] < ws.e_cap ? off + s_cnt[b] : ws.e_cap;
      s_start[b] = (unsigned int)st;
      s_lim[b] = (unsigned int)en;
      out_jobs[b] = make_uint4((unsigned int)st, (unsigned int)(en - st), (unsigned int)bid * kBlocksPerTile + b, 0u);
      off += (s_cnt[b] + 1u) & ~1u;
    }
  }
  __syncthreads();

  // ---- pass 2: place the entries, chunk by chunk ----
  const unsigned int lt = (1u << lane) - 1u;
  for (int base = 0; base < n; base += kBlThreads) {
    const int j = base + tid;
    const unsigned int m = j < n ? (unsigned int)ws.pair_mask[range.x + j] : 0u;  // this thread's own store of pass 1
    unsigned int rank[8];
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const unsigned int bal = __ballot_sync(0xffffffffu, (m >> b) & 1u);
      rank[b] = __popc(bal & lt);
      if (lane == 0) s_wcnt[warp][b] = __popc(bal);
    }
    // (Gaussian id, 1-based position in the tile list: what n_contrib records)
    const uint2 ref = make_uint2(m ? vals[range.x + j] : 0u, (unsigned int)(j + 1));
    __syncthreads();
    if (warp < kBlocksPerTile) {  // warp b turns the per-warp counts of block b into exclusive offsets
      const unsigned int c = lane < kBlWarps ? s_wcnt[lane][warp] : 0u;
      unsigned int incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      const unsigned int run = s_run[warp];
      __syncwarp();
      if (lane < kBlWarps) s_wcnt[lane][warp] = run + incl - c;
      if (lane == 31) s_run[warp] = run + incl;
    }
    __syncthreads();
    if (m) {
#pragma unroll
      for (int b = 0; b < 8; b++) {
        if ((m >> b) & 1u) {
          const unsigned int pos = s_start[b] + s_wcnt[warp][b] + rank[b];
          TS_CHECK(pos < s_lim[b] || s_lim[b] == ws.e_cap);
          if (pos < s_lim[b]) ws.bref[pos] = ref;
        }
      }
    }
    __syncthreads();  // s_wcnt is rewritten by the next chunk
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
