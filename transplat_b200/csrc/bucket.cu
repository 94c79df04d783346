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
//      MEMORY (the word is unique, so the result equals the stable sort by depth of the
//      Gaussian-ordered list): a one-pass distribution sort over 4096 depth buckets, with a stable
//      8-bit LSD radix sort over the id and depth digits as the fallback for clustered / equal depths;
//      writes keys and values in sorted order; tiles above kSmallTileCap go to a second launch
//      (bitonic network in a 128-KB buffer), tiles above kBigTileCap are sorted in place in global
//      memory by one CTA.
//   5. block_lists_kernel (blocklist.cu, shared with the onesweep path): the per-8x4-block lists the render warps stream.
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
    const float2 c = *reinterpret_cast<const float2*>(ws.grec + it * 3);
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
      TS_CHECK(slot < ranges[b].y || ranges[b].y == ws.d_cap);
      if (slot < ranges[b].y) pairs[slot] = word;
    }
}

// ---------------------------------------------------------------------------------------------
// 4. per-tile sort + gather.  Bitonic network in its ascending-only form (first step of every merge
//    mirrors the partner), so a list of any length n sorts without padding: a comparator whose upper
//    index is >= n would compare against +inf and is skipped.
// ---------------------------------------------------------------------------------------------
// One pass over the array executing R consecutive comparator steps (xor masks m[0..R-1], whose top
// bits sit at consecutive positions p+R-1 .. p) on groups of 2^R elements held in registers: the
// array crosses shared memory once per pass instead of once per step (the plain network is
// shared-memory-bandwidth bound: 66 steps for 2048 words).  Elements at index >= n are +inf
// (never stored); every comparator leaves the minimum at the lower index.
template <int R>
__device__ __forceinline__ void bitonic_pass(unsigned long long* s, int n, int np, int p, const int (&m)[3], int tid,
                                             int nthreads) {
  constexpr int E = 1 << R;
  for (int t = tid; t < (np >> R); t += nthreads) {
    const int base = ((t >> p) << (p + R)) | (t & ((1 << p) - 1));
    int idx[E];
    unsigned long long v[E];
#pragma unroll
    for (int a = 0; a < E; a++) {
      int x = base;
#pragma unroll
      for (int r = 0; r < R; r++)
        if (a & (1 << (R - 1 - r))) x ^= m[r];
      idx[a] = x;
      v[a] = x < n ? s[x] : ~0ull;
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
      const int ab = 1 << (R - 1 - r);
#pragma unroll
      for (int a = 0; a < E; a++) {
        if (a & ab) continue;
        const unsigned long long x = v[a], y = v[a | ab];
        const bool a_low = idx[a] < idx[a | ab];  // the minimum goes to the lower index
        if (a_low ? (x > y) : (x < y)) { v[a] = y; v[a | ab] = x; }
      }
    }
#pragma unroll
    for (int a = 0; a < E; a++)
      if (idx[a] < n) s[idx[a]] = v[a];
  }
  __syncthreads();
}

// s: the n words (shared or global memory, private to this CTA).  All threads of the CTA call.
// Bitonic network in ascending-only form: level k = [mirror step: i <-> i ^ (k-1)], then strides
// k/4, k/8, ..., 1; steps are taken three at a time.
__device__ __forceinline__ void bitonic_sort_words(unsigned long long* s, int n, int tid, int nthreads) {
  if (n < 2) return;
  int lg = 1;
  while ((1 << lg) < n) lg++;
  const int np = 1 << lg;
  for (int mlev = 1; mlev <= lg; mlev++) {  // level k = 2^mlev has mlev steps: masks k-1, k/4, k/8, ..., 1
    int step = 0;
    while (step < mlev) {
      const int R = min(3, mlev - step);
      int m[3] = {0, 0, 0};
      for (int r = 0; r < R; r++) {
        const int st = step + r;
        m[r] = st == 0 ? (1 << mlev) - 1 : 1 << (mlev - 1 - st);
      }
      const int p = mlev - step - R;  // lowest top-bit position of the chunk
      if (R == 3) bitonic_pass<3>(s, n, np, p, m, tid, nthreads);
      else if (R == 2) bitonic_pass<2>(s, n, np, p, m, tid, nthreads);
      else bitonic_pass<1>(s, n, np, p, m, tid, nthreads);
      step += R;
    }
  }
}

// write keys / values of one sorted tile segment (the per-block record lists follow in block_lists_kernel)
__device__ __forceinline__ void write_sorted_tile(const Workspace& ws, const unsigned long long* s, int n, unsigned int start,
                                                  int bid, int G, unsigned long long* __restrict__ keys_out,
                                                  unsigned int* __restrict__ vals_out, int tid, int nthreads) {
  const unsigned long long hiw = (unsigned long long)(unsigned int)bid << 32;
  for (int j = tid; j < n; j += nthreads) {
    const unsigned long long w = s[j];
    keys_out[start + j] = hiw | (w >> 32);
    vals_out[start + j] = (unsigned int)w;
  }
}

// Fallback of the small-tile sort: CTA-wide stable LSD radix sort of the words (8-bit digits, in place in shared
// memory: a pass holds its words in registers between the load and the scatter).
constexpr int kTileSortThreads = 512;
constexpr int kTileSortWarps = kTileSortThreads / 32;

// 2048 depth buckets (74 KB of shared memory per CTA) and 40 registers let THREE 512-thread CTAs share an SM: the kernel
// is bound by shared-memory latency and barriers, not by issue.  C2: 4096 buckets x 2 CTAs 34.8 us, 2048 x 2 33.3,
// 2048 x 3 29.4.
#ifndef TSPLAT_DEPTH_BUCKET_BITS
#define TSPLAT_DEPTH_BUCKET_BITS 11
#endif
#ifndef TSPLAT_TS_MIN_BLOCKS
#define TSPLAT_TS_MIN_BLOCKS 3
#endif
constexpr int kDepthBucketBits = TSPLAT_DEPTH_BUCKET_BITS;
constexpr int kDepthBuckets = 1 << kDepthBucketBits;  // depth buckets of the distribution sort (at most 2 words per bucket on average)
constexpr int kMaxBucketFill = 48;   // a fuller bucket (clustered or equal depths) sends the tile to the radix path

struct TileSortSmem {
  unsigned long long words[kSmallTileCap];
  union {
    unsigned long long tmp[kSmallTileCap];       // distribution sort: words grouped by bucket
    unsigned int whist[kTileSortWarps][kRadix];  // radix path: per-warp digit counts -> exclusive prefixes over warps
  };
  unsigned int bucket[kDepthBuckets + 1];        // counts -> exclusive bases (bucket[kDepthBuckets] = n)
  unsigned int dexcl[kRadix];                  // exclusive offset of each digit in the CTA
  unsigned int wscan[kTileSortWarps];
  unsigned int dmin, dmax, maxfill;
};

// Distribution sort of the tile's words (depth bits << 32 | Gaussian id, all distinct) - the default.  The depth
// range of the tile is cut into kDepthBuckets equal intervals of the BIT PATTERN (monotone for positive floats):
//   count per bucket (shared-memory atomics) -> exclusive scan -> scatter by bucket -> every word finds its rank
//   among the (typically 1-3) words of its bucket by comparing whole words.
// The result is the ascending order of the words, i.e. the stable depth sort of the Gaussian-ordered list, with
// one pass over the data instead of the radix sort's four.  Returns false (nothing modified) when some bucket
// holds more than kMaxBucketFill words; the caller then takes the radix path.
__device__ __forceinline__ bool distribution_sort_tile(TileSortSmem& s, int n, int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int KMAX = kSmallTileCap / kTileSortThreads;  // 8 words per thread at most
  for (int i = tid; i <= kDepthBuckets; i += kTileSortThreads) s.bucket[i] = 0u;
  unsigned int lo = 0xffffffffu, hi = 0u;
  for (int i = tid; i < n; i += kTileSortThreads) {
    const unsigned int d = (unsigned int)(s.words[i] >> 32);
    lo = min(lo, d); hi = max(hi, d);
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if (lane == 0) { atomicMin(&s.dmin, lo); atomicMax(&s.dmax, hi); }
  __syncthreads();
  const unsigned int dmin = s.dmin, range = s.dmax - dmin;
  const int bits = 32 - __clz(range);  // 0 when every depth is equal
  const int shift = max(0, bits - kDepthBucketBits);
  unsigned short rank[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; k++) {
    const int i = tid + k * kTileSortThreads;
    if (i < n) {
      const unsigned int b = ((unsigned int)(s.words[i] >> 32) - dmin) >> shift;
      rank[k] = (unsigned short)atomicAdd(&s.bucket[b], 1u);
    }
  }
  __syncthreads();
  {  // exclusive scan of the counts: thread t owns buckets [8t, 8t+8)
    constexpr int PER = kDepthBuckets / kTileSortThreads;
    unsigned int c[PER], sum = 0, mx = 0;
#pragma unroll
    for (int k = 0; k < PER; k++) { c[k] = s.bucket[tid * PER + k]; sum += c[k]; mx = max(mx, c[k]); }
    unsigned int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 31) s.wscan[warp] = incl;
    if (lane == 0) atomicMax(&s.maxfill, mx);
    __syncthreads();
    unsigned int pre = incl - sum;
    for (int wv = 0; wv < warp; wv++) pre += s.wscan[wv];
#pragma unroll
    for (int k = 0; k < PER; k++) { s.bucket[tid * PER + k] = pre; pre += c[k]; }
    if (tid == 0) s.bucket[kDepthBuckets] = (unsigned int)n;
  }
  __syncthreads();
  if (s.maxfill > (unsigned int)kMaxBucketFill) return false;
#pragma unroll
  for (int k = 0; k < KMAX; k++) {
    const int i = tid + k * kTileSortThreads;
    if (i < n) {
      const unsigned long long w = s.words[i];
      const unsigned int b = ((unsigned int)(w >> 32) - dmin) >> shift;
      s.tmp[s.bucket[b] + rank[k]] = w;
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += kTileSortThreads) {
    const unsigned long long w = s.tmp[i];
    const unsigned int b = ((unsigned int)(w >> 32) - dmin) >> shift;
    const unsigned int b0 = s.bucket[b], b1 = s.bucket[b + 1];
    unsigned int r = 0;
    for (unsigned int q = b0; q < b1; q++) r += s.tmp[q] < w ? 1u : 0u;
    s.words[b0 + r] = w;
  }
  __syncthreads();
  return true;
}

template <int KPT>
__device__ __forceinline__ void radix_sort_tile_by_depth(TileSortSmem& s, int n, int tid, int id_passes) {
  const int lane = tid & 31, warp = tid >> 5;
  const int wbase = warp * (32 * KPT);
  // stable LSD over the whole word: the Gaussian-id digits first (id_passes = ceil(bits(G) / 8)), then the four depth
  // digits - equal depths come out in Gaussian order with no fix-up, whatever the emission order was
#pragma unroll 1
  for (int pass = 0; pass < id_passes + 4; pass++) {
    const int shift = pass < id_passes ? 8 * pass : 32 + 8 * (pass - id_passes);
    for (int i = tid; i < kTileSortWarps * kRadix; i += kTileSortThreads) (&s.whist[0][0])[i] = 0u;
    unsigned long long w[KPT];
#pragma unroll
    for (int i = 0; i < KPT; i++) {
      const int li = wbase + i * 32 + lane;
      w[i] = li < n ? s.words[li] : ~0ull;
    }
    __syncthreads();
    unsigned short rank[KPT];
#pragma unroll
    for (int i = 0; i < KPT; i++) {  // warp-private ranking: step i ascending, lane ascending == current order
      const bool valid = wbase + i * 32 + lane < n;
      const unsigned int dg = (unsigned int)(w[i] >> shift) & 255u;
      // peers with the same digit, from 8 ballots (MATCH.ANY is microcoded and slow on sm_100: it was 30 % of
      // this kernel's stall samples)
      unsigned int same = __ballot_sync(0xffffffffu, valid);
#pragma unroll
      for (int bit = 0; bit < 8; bit++) {
        const unsigned int v = __ballot_sync(0xffffffffu, (dg >> bit) & 1u);
        same &= ((dg >> bit) & 1u) ? v : ~v;
      }
      if (!valid) same = 1u << lane;
      const int leader = __ffs(same) - 1;
      unsigned int old = 0;
      if (valid && lane == leader) {
        old = s.whist[warp][dg];
        s.whist[warp][dg] = old + __popc(same);
      }
      old = __shfl_sync(0xffffffffu, old, leader);
      rank[i] = (unsigned short)(old + __popc(same & ((1u << lane) - 1u)));
      __syncwarp();
    }
    __syncthreads();
    unsigned int dcount = 0, incl = 0;
    if (tid < kRadix) {  // thread d owns digit d
#pragma unroll
      for (int wv = 0; wv < kTileSortWarps; wv++) {
        const unsigned int c = s.whist[wv][tid];
        s.whist[wv][tid] = dcount;
        dcount += c;
      }
      incl = dcount;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
      }
      if (lane == 31) s.wscan[warp] = incl;
    }
    __syncthreads();
    if (tid < kRadix) {
      unsigned int pre = 0;
      for (int wv = 0; wv < warp; wv++) pre += s.wscan[wv];
      s.dexcl[tid] = pre + incl - dcount;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < KPT; i++) {
      if (wbase + i * 32 + lane < n) {
        const unsigned int dg = (unsigned int)(w[i] >> shift) & 255u;
        s.words[s.dexcl[dg] + s.whist[warp][dg] + rank[i]] = w[i];
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kTileSortThreads, TSPLAT_TS_MIN_BLOCKS) tile_sort_gather_kernel(Workspace ws, int G,
                                                                            const unsigned long long* __restrict__ pairs,
                                                                            unsigned long long* __restrict__ keys_out,
                                                                            unsigned int* __restrict__ vals_out) {
  extern __shared__ __align__(16) unsigned char ts_raw[];
  TileSortSmem& s = *reinterpret_cast<TileSortSmem*>(ts_raw);
  const int bid = ws.use_order ? (int)ws.tile_order[blockIdx.x] : (int)blockIdx.x;
  const uint2 r = ws.ranges[bid];
  const int n = (int)(r.y - r.x);
  if (n == 0 || n > kSmallTileCap) return;
  const int tid = threadIdx.x;
  for (int i = tid; i < n; i += kTileSortThreads) s.words[i] = pairs[r.x + i];
  if (tid == 0) { s.dmin = 0xffffffffu; s.dmax = 0u; s.maxfill = 0u; }
  __syncthreads();
  const bool sorted = distribution_sort_tile(s, n, tid);
  if (!sorted) {  // clustered / equal depths: stable 8-bit LSD radix sort over the id digits, then the depth digits
    int id_passes = 1;
    while (id_passes < 4 && (G - 1) >> (8 * id_passes)) id_passes++;
    if (n <= 1 * kTileSortThreads) radix_sort_tile_by_depth<1>(s, n, tid, id_passes);
    else if (n <= 2 * kTileSortThreads) radix_sort_tile_by_depth<2>(s, n, tid, id_passes);
    else if (n <= 4 * kTileSortThreads) radix_sort_tile_by_depth<4>(s, n, tid, id_passes);
    else radix_sort_tile_by_depth<8>(s, n, tid, id_passes);
  }
  write_sorted_tile(ws, s.words, n, r.x, bid, G, keys_out, vals_out, tid, kTileSortThreads);
}

// tiles above kSmallTileCap: a few persistent CTAs with a 128-KB buffer; above kBigTileCap in place in HBM
__global__ void __launch_bounds__(512) big_tile_sort_gather_kernel(Workspace ws, int G, unsigned long long* __restrict__ pairs,
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
      for (int i = tid; i < n; i += 512) sbig[i] = pairs[r.x + i];
      __syncthreads();
      bitonic_sort_words(sbig, n, tid, 512);
      write_sorted_tile(ws, sbig, n, r.x, bid, G, keys_out, vals_out, tid, 512);
    } else {
      unsigned long long* seg = pairs + r.x;  // this CTA is the only one that touches the segment
      __syncthreads();
      bitonic_sort_words(seg, n, tid, 512);
      write_sorted_tile(ws, seg, n, r.x, bid, G, keys_out, vals_out, tid, 512);
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
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(tile_sort_gather_kernel), (int)sizeof(TileSortSmem));
  if (e != cudaSuccess) return e;
  tile_sort_gather_kernel<<<NT, kTileSortThreads, sizeof(TileSortSmem), st>>>(ws, p.G, pairs, keys_out, vals_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int smem = kBigTileCap * (int)sizeof(unsigned long long);
  e = ensure_dynamic_smem(reinterpret_cast<const void*>(big_tile_sort_gather_kernel), smem);
  if (e != cudaSuccess) return e;
  big_tile_sort_gather_kernel<<<148, 512, smem, st>>>(ws, p.G, pairs, keys_out, vals_out);
  return cudaGetLastError();
}

}  // namespace tsplat
