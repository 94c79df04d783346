// bucket.cu — the default binning path: count -> scan -> emit -> per-tile sort (+ record gather).
// Produces exactly what the onesweep path (binning.cu) produces — the pair list sorted by
// ((view*T + tile) << 32 | depth bits) with ties in Gaussian order, the tile ranges, the
// sorted-order records — but without sorting the tile word at all:
//
//   1. the preprocess kernel counts pairs per (view, tile) with fire-and-forget reductions
//      (red.global.add.u32) while it computes the tile rectangles            [preprocess.cu]
//   2. bucket_scan_kernel: exclusive scan of the counts = the tile ranges (identifyTileRanges
//      for free), D, the overflow flag, the heaviest-first tile order, the list of big tiles
//   3. bucket_emit_kernel: every (Gaussian, tile) pair takes a slot of its tile's segment from an
//      atomic cursor and stores (depth bits << 32 | Gaussian id)
//   4. tile_sort_gather_kernel: one CTA per tile sorts its segment by that 64-bit word in SHARED
//      MEMORY (bitonic network; the word is unique, so the result equals the stable sort by depth
//      of the Gaussian-ordered list) and writes keys, values and the gathered 48-B records in
//      sorted order; tiles above kSmallTileCap go to a second launch with a 128-KB buffer, tiles
//      above kBigTileCap are sorted in place in global memory by one CTA.
//
// Why: at C2 (1.6 M pairs, 42 key bits) six LSD passes were latency-bound at ~26 us each plus a
// 38-us scan/emit and a 38-us range/gather pass (272 us before render); each pair now crosses HBM
// once in each direction and is sorted on chip.  The onesweep path stays available
// (TSPLAT_FLAG_SORT_ONESWEEP) and is held to the same bit-exact tests.
#include "cull.cuh"
#include "geom.cuh"

namespace tsplat {

// ---------------------------------------------------------------------------------------------
// 2. scan of the per-tile counts (one CTA; NT = views*tiles is 1024 at C2, 32 K at C5)
// ---------------------------------------------------------------------------------------------
constexpr int kScanCta = 1024;

__global__ void __launch_bounds__(kScanCta) bucket_scan_kernel(Workspace ws) {
  __shared__ unsigned long long s_warp[kScanCta / 32];
  __shared__ unsigned long long s_carry;
  __shared__ unsigned int s_bucket[1024];
  __shared__ unsigned int s_nbig;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int NT = ws.N * ws.T;
  if (tid == 0) { s_carry = 0; s_nbig = 0; }
  for (int i = tid; i < 1024; i += kScanCta) s_bucket[i] = 0;
  __syncthreads();
  for (int base = 0; base < NT; base += kScanCta) {
    const int i = base + tid;
    const unsigned int c = i < NT ? ws.tile_count[i] : 0u;
    unsigned long long incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned long long pre = s_carry;
    for (int w = 0; w < warp; w++) pre += s_warp[w];
    const unsigned long long start = pre + incl - c, end = pre + incl;
    if (i < NT) {
      // segments are clipped to the capacity: pairs beyond d_cap are dropped, never written
      const unsigned long long cs = start < ws.d_cap ? start : ws.d_cap, ce = end < ws.d_cap ? end : ws.d_cap;
      ws.ranges[i] = make_uint2((unsigned int)cs, (unsigned int)ce);
      const unsigned int len = (unsigned int)(ce - cs);
      if (len > (unsigned int)kSmallTileCap) ws.big_list[atomicAdd(&s_nbig, 1u)] = (unsigned int)i;
      if (ws.use_order) atomicAdd(&s_bucket[1023u - min(1023u, len >> 4)], 1u);
    }
    __syncthreads();
    if (tid == kScanCta - 1) s_carry = end;
    __syncthreads();
  }
  if (tid == 0) {
    const unsigned long long D = s_carry;
    ws.ctl->num_rendered = D;
    ws.ctl->overflow = D > ws.d_cap ? 1u : 0u;
    ws.ctl->num_big = s_nbig;
  }
  if (!ws.use_order) return;
  // heaviest-first tile order: counting sort on length/16 (bucket 0 = longest lists)
  {
    const unsigned int c = s_bucket[tid];
    unsigned int incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    __shared__ unsigned int s_w2[kScanCta / 32];
    if (lane == 31) s_w2[warp] = incl;
    __syncthreads();
    unsigned int pre = incl - c;
    for (int w = 0; w < warp; w++) pre += s_w2[w];
    s_bucket[tid] = pre;
  }
  __syncthreads();
  for (int i = tid; i < NT; i += kScanCta) {
    const uint2 r = ws.ranges[i];
    ws.tile_order[atomicAdd(&s_bucket[1023u - min(1023u, (r.y - r.x) >> 4)], 1u)] = (unsigned int)i;
  }
}

// ---------------------------------------------------------------------------------------------
// 3. emission: CTA = 256 Gaussians of ONE view.  Slots are reserved per (CTA, tile): the CTA counts
//    its pairs per tile in shared memory, takes one range per non-empty tile from the tile's global
//    cursor (one returning atomic per (CTA, tile) instead of one per pair: same-address atomics
//    serialise at ~75 ns), then deals the range out with shared-memory atomics.  Order inside a tile
//    segment is arbitrary; the per-tile sort makes it canonical.  Images with more than kEmitMaxTiles
//    tiles per view take the direct global-cursor path.
// ---------------------------------------------------------------------------------------------
constexpr int kEmitMaxTiles = 4096;

template <bool kSmemBins>
__global__ void __launch_bounds__(256) bucket_emit_kernel(Workspace ws, int G, unsigned long long* __restrict__ pairs) {
  extern __shared__ unsigned int s_bins[];  // [2][T]: counts / local cursors, reserved bases
  const int n = blockIdx.y, g = blockIdx.x * 256 + threadIdx.x, tid = threadIdx.x;
  const int T = ws.T;
  unsigned int* s_cnt = s_bins;
  unsigned int* s_base = s_bins + T;
  if (kSmemBins) {
    for (int i = tid; i < T; i += 256) s_cnt[i] = 0u;
    __syncthreads();
  }
  const size_t it = (size_t)n * G + g;
  const bool live = g < G && ws.tiles_touched[it] != 0u;
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
  unsigned long long word = 0;
  if (live) {
    const float2 c = ws.xy[it];
    get_rect(c.x, c.y, ws.radii[it], ws.tiles_x, ws.tiles_y, x0, y0, x1, y1);
    word = ((unsigned long long)__float_as_uint(ws.depths[it]) << 32) | (unsigned int)g;
  }
  const uint2* __restrict__ ranges = ws.ranges + (size_t)n * T;
  unsigned int* cursor = ws.tile_cursor + (size_t)n * T;
  if (!kSmemBins) {
    for (int y = y0; y < y1; y++)
      for (int x = x0; x < x1; x++) {
        const int b = y * ws.tiles_x + x;
        const uint2 r = ranges[b];
        const unsigned int slot = r.x + atomicAdd(&cursor[b], 1u);
        if (slot < r.y) pairs[slot] = word;
      }
    return;
  }
  for (int y = y0; y < y1; y++)
    for (int x = x0; x < x1; x++) atomicAdd(&s_cnt[y * ws.tiles_x + x], 1u);
  __syncthreads();
  for (int b = tid; b < T; b += 256) {
    const unsigned int c = s_cnt[b];
    if (c) {
      s_base[b] = ranges[b].x + atomicAdd(&cursor[b], c);
      s_cnt[b] = 0u;
    }
  }
  __syncthreads();
  for (int y = y0; y < y1; y++)
    for (int x = x0; x < x1; x++) {
      const int b = y * ws.tiles_x + x;
      const unsigned int slot = s_base[b] + atomicAdd(&s_cnt[b], 1u);
      if (slot < ranges[b].y) pairs[slot] = word;
    }
}

// ---------------------------------------------------------------------------------------------
// 4. per-tile sort + gather.  Bitonic network in its ascending-only form (first step of every merge
//    mirrors the partner), so a list of any length n sorts without padding: a comparator whose upper
//    index is >= n would compare against +inf and is skipped.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void compare_exchange(unsigned long long* s, int i, int l) {
  const unsigned long long a = s[i], b = s[l];
  if (a > b) { s[i] = b; s[l] = a; }
}

// s: the n words (shared or global memory, private to this CTA).  All threads of the CTA call.
__device__ __forceinline__ void bitonic_sort_words(unsigned long long* s, int n, int tid, int nthreads) {
  if (n < 2) return;
  int np = 2;
  while (np < n) np <<= 1;
  const int half = np >> 1;
  for (int k = 2; k <= np; k <<= 1) {
    const int hk = k >> 1;
    for (int t = tid; t < half; t += nthreads) {  // mirrored step
      const int blk = t / hk, off = t - blk * hk;
      const int i = blk * k + off, l = blk * k + (k - 1 - off);
      if (l < n) compare_exchange(s, i, l);
    }
    __syncthreads();
    for (int j = hk >> 1; j > 0; j >>= 1) {
      for (int t = tid; t < half; t += nthreads) {
        const int i = 2 * t - (t & (j - 1)), l = i + j;
        if (l < n) compare_exchange(s, i, l);
      }
      __syncthreads();
    }
  }
}

// write keys / values / records of one sorted tile segment
__device__ __forceinline__ void write_sorted_tile(const Workspace& ws, const unsigned long long* s, int n, unsigned int start,
                                                  int bid, int G, unsigned long long* __restrict__ keys_out,
                                                  unsigned int* __restrict__ vals_out, int tid, int nthreads) {
  const size_t vbase = (size_t)(bid / ws.T) * G;
  for (int j = tid; j < n; j += nthreads) {
    const unsigned long long w = s[j];
    const unsigned int id = (unsigned int)w;
    keys_out[start + j] = ((unsigned long long)(unsigned int)bid << 32) | (w >> 32);
    vals_out[start + j] = id;
    const size_t o = vbase + id;
    const float2 c = ws.xy[o];
    const float4 co = ws.conic_opacity[o];
    const float4 f = ws.rgbd[o];
    ws.rec0[start + j] = make_float4(c.x, c.y, co.x, co.y);
    ws.rec1[start + j] = make_float4(co.z, co.w, f.x, f.y);
    ws.rec2[start + j] = make_float4(f.z, f.w, __uint_as_float(id), __uint_as_float(cull_extent(co.x, co.y, co.z, co.w)));
  }
}

__global__ void __launch_bounds__(256) tile_sort_gather_kernel(Workspace ws, int G,
                                                               const unsigned long long* __restrict__ pairs,
                                                               unsigned long long* __restrict__ keys_out,
                                                               unsigned int* __restrict__ vals_out) {
  __shared__ unsigned long long s[kSmallTileCap];
  const int bid = ws.use_order ? (int)ws.tile_order[blockIdx.x] : (int)blockIdx.x;
  const uint2 r = ws.ranges[bid];
  const int n = (int)(r.y - r.x);
  if (n == 0 || n > kSmallTileCap) return;
  const int tid = threadIdx.x;
  for (int i = tid; i < n; i += 256) s[i] = pairs[r.x + i];
  __syncthreads();
  bitonic_sort_words(s, n, tid, 256);
  write_sorted_tile(ws, s, n, r.x, bid, G, keys_out, vals_out, tid, 256);
}

// tiles above kSmallTileCap: a few persistent CTAs with a 128-KB buffer; above kBigTileCap in place in HBM
__global__ void __launch_bounds__(1024) big_tile_sort_gather_kernel(Workspace ws, int G, unsigned long long* __restrict__ pairs,
                                                                    unsigned long long* __restrict__ keys_out,
                                                                    unsigned int* __restrict__ vals_out) {
  extern __shared__ __align__(16) unsigned long long sbig[];
  const int tid = threadIdx.x;
  const unsigned int nbig = ws.ctl->num_big;
  for (unsigned int k = blockIdx.x; k < nbig; k += gridDim.x) {
    const int bid = (int)ws.big_list[k];
    const uint2 r = ws.ranges[bid];
    const int n = (int)(r.y - r.x);
    if (n <= kBigTileCap) {
      for (int i = tid; i < n; i += 1024) sbig[i] = pairs[r.x + i];
      __syncthreads();
      bitonic_sort_words(sbig, n, tid, 1024);
      write_sorted_tile(ws, sbig, n, r.x, bid, G, keys_out, vals_out, tid, 1024);
    } else {
      unsigned long long* seg = pairs + r.x;  // this CTA is the only one that touches the segment
      __syncthreads();
      bitonic_sort_words(seg, n, tid, 1024);
      write_sorted_tile(ws, seg, n, r.x, bid, G, keys_out, vals_out, tid, 1024);
    }
    __syncthreads();  // the buffer is reused by the next tile of this CTA
  }
}

cudaError_t launch_bucket_scan_emit(const TsplatProblem& p, const Workspace& ws, cudaStream_t st) {
  bucket_scan_kernel<<<1, kScanCta, 0, st>>>(ws);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  // unsorted words go to the buffer the sorted keys do NOT end in
  unsigned long long* pairs = ws.keys[(ws.sort_passes & 1) ^ 1];
  const dim3 grid((p.G + 255) / 256, ws.N);
  if (ws.T <= kEmitMaxTiles)
    bucket_emit_kernel<true><<<grid, 256, 2 * sizeof(unsigned int) * ws.T, st>>>(ws, p.G, pairs);
  else
    bucket_emit_kernel<false><<<grid, 256, 0, st>>>(ws, p.G, pairs);
  return cudaGetLastError();
}

cudaError_t launch_bucket_sort_gather(const TsplatProblem& p, const Workspace& ws, cudaStream_t st) {
  unsigned long long* pairs = ws.keys[(ws.sort_passes & 1) ^ 1];
  unsigned long long* keys_out = ws.keys[ws.sort_passes & 1];
  unsigned int* vals_out = ws.vals[ws.sort_passes & 1];
  const int NT = ws.N * ws.T;
  tile_sort_gather_kernel<<<NT, 256, 0, st>>>(ws, p.G, pairs, keys_out, vals_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int smem = kBigTileCap * (int)sizeof(unsigned long long);
  e = cudaFuncSetAttribute(big_tile_sort_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  big_tile_sort_gather_kernel<<<148, 1024, smem, st>>>(ws, p.G, pairs, keys_out, vals_out);
  return cudaGetLastError();
}

}  // namespace tsplat
