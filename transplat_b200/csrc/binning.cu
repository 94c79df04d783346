// binning.cu — tile|depth key duplication, onesweep LSD radix sort, tile ranges, record gather
// (SURVEY 8a X4-X7).  All V views of all B scenes of a call are binned and sorted TOGETHER: the
// high key word is view*T + tile, so one sort orders every view's list and one render launch
// covers every (view, tile).  Per view the key is bit-identical to the reference's
// (tile << 32) | float_bits(depth) after subtracting view*T from the high word.
//
// No host synchronisation: the pair count D lives in device memory (Control::num_rendered);
// grids are sized by the capacity d_cap and surplus CTAs exit.
#include "cull.cuh"
#include "geom.cuh"

namespace tsplat {

// ---------------------------------------------------------------------------------------------
// 1. bin_emit: exclusive scan of tiles_touched (single pass, decoupled look-back) fused with key
//    emission and with the onesweep digit histograms.  Replaces cub::DeviceScan::InclusiveSum +
//    the blocking D2H copy of num_rendered + duplicateWithKeys (X4, X5).
//    Emission is load-balanced: the CTA's output slots are dealt to threads round-robin and each
//    slot finds its (Gaussian, tile) by binary search, so writes are coalesced and a Gaussian
//    covering 200 tiles costs the same per slot as one covering 2.
// ---------------------------------------------------------------------------------------------
constexpr unsigned long long kFlagAgg = 1ull << 62, kFlagIncl = 2ull << 62, kValMask = (1ull << 62) - 1;

__global__ void __launch_bounds__(kScanThreads) bin_emit_kernel(Workspace ws, int G, int total_items) {
  __shared__ unsigned int s_off[kScanTile + 1];   // local exclusive offsets
  __shared__ unsigned int s_rect[kScanTile];      // x0 | y0 << 10 | width << 20
  __shared__ unsigned int s_depth[kScanTile];
  __shared__ unsigned int s_warp[kScanThreads / 32];
  __shared__ unsigned int s_hist[kMaxPasses][kRadix];
  __shared__ unsigned int s_tile;
  __shared__ unsigned long long s_base;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(&ws.ctl->scan_tile_counter, 1u);
  for (int i = tid; i < kMaxPasses * kRadix; i += kScanThreads) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const int tile = (int)s_tile;
  const int item0 = tile * kScanTile + tid * kScanItems;

  unsigned int cnt[kScanItems];
  unsigned int tsum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    const int it = item0 + i;
    cnt[i] = it < total_items ? ws.tiles_touched[it] : 0u;
    tsum += cnt[i];
  }
  // block exclusive scan of tsum
  unsigned int incl = tsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  unsigned int wpre = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; w++) {
    const unsigned int c = s_warp[w];
    if (w < warp) wpre += c;
    block_total += c;
  }
  unsigned int run = wpre + incl - tsum;
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    const int li = tid * kScanItems + i;
    s_off[li] = run;
    run += cnt[i];
    const int it = item0 + i;
    if (cnt[i] != 0) {
      const float2 c = *reinterpret_cast<const float2*>(ws.grec + it * 3);
      int x0, y0, x1, y1;
      get_rect(c.x, c.y, ws.radii[it], ws.tiles_x, ws.tiles_y, x0, y0, x1, y1);
      s_rect[li] = (unsigned)x0 | ((unsigned)y0 << 10) | ((unsigned)(x1 - x0) << 20);
      const unsigned int db = __float_as_uint(ws.depths[it]);
      s_depth[li] = db;
    }
  }
  if (tid == kScanThreads - 1) s_off[kScanTile] = run;

  // decoupled look-back over the tile aggregates (warp 0)
  if (warp == 0) {
    if (lane == 0) {
      const unsigned long long w = (tile == 0 ? kFlagIncl : kFlagAgg) | (unsigned long long)block_total;
      st_relaxed_u64(&ws.scan_state[tile], w);
    }
    unsigned long long excl = 0;
    int look = tile - 1 - lane;
    bool fin = (tile == 0);
    while (!fin) {
      unsigned long long s = kFlagIncl;  // virtual predecessor of tile 0: inclusive 0
      if (look >= 0) {
        do { s = ld_relaxed_u64(&ws.scan_state[look]); } while ((s >> 62) == 0);
      }
      const unsigned int incl_mask = __ballot_sync(0xffffffffu, (s >> 62) == 2);
      const int first = incl_mask ? (__ffs(incl_mask) - 1) : 32;
      unsigned long long v = (lane <= first) ? (s & kValMask) : 0ull;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
      excl += v;
      fin = incl_mask != 0;
      look -= 32;
    }
    if (lane == 0) {
      if (tile != 0) st_relaxed_u64(&ws.scan_state[tile], kFlagIncl | (excl + block_total));
      s_base = excl;
      if (tile == ws.num_scan_tiles - 1) {
        const unsigned long long D = excl + block_total;
        ws.ctl->num_rendered = D;
        ws.ctl->overflow = D > ws.d_cap ? 1u : 0u;
      }
    }
  }
  __syncthreads();

  const unsigned long long base = s_base;
  const int T = ws.T, gx = ws.tiles_x;
  // low-word digits are shared by all of a Gaussian's keys: count them once per Gaussian, limited
  // to the slots that fit under d_cap so the histograms always sum to min(D, d_cap)
#pragma unroll
  for (int i = 0; i < kScanItems; i++) {
    if (cnt[i] != 0) {
      const int li = tid * kScanItems + i;
      const unsigned long long start = base + s_off[li];
      const unsigned long long room = start < ws.d_cap ? ws.d_cap - start : 0ull;
      const unsigned int c = (unsigned int)(room < cnt[i] ? room : cnt[i]);
      const unsigned int db = s_depth[li];
      if (c != 0) {
#pragma unroll
        for (int ps = 0; ps < 4; ps++) atomicAdd(&s_hist[ps][(db >> (8 * ps)) & 255u], c);
      }
    }
  }
  const int first_item = tile * kScanTile;
  for (unsigned int j = tid; j < block_total; j += kScanThreads) {
    // last local item with s_off <= j
    int lo = 0, hi = kScanTile;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_off[mid] <= j) lo = mid; else hi = mid;
    }
    const unsigned int k = j - s_off[lo];
    const unsigned int r = s_rect[lo];
    const unsigned int w = r >> 20;
    const unsigned int x = (r & 1023u) + k % w, y = ((r >> 10) & 1023u) + k / w;
    const int it = first_item + lo;
    const int n = it / G;
    const unsigned int hiw = (unsigned int)(n * T) + y * gx + x;
    const unsigned long long out = base + j;
    if (out < ws.d_cap) {
      ws.keys[0][out] = ((unsigned long long)hiw << 32) | s_depth[lo];
      ws.vals[0][out] = (unsigned int)(it - n * G);
      // high-word digits: warp-aggregated (neighbouring slots share tiles/views)
#pragma unroll
      for (int ps = 4; ps < kMaxPasses; ps++) {
        if (ps >= ws.sort_passes) break;
        const unsigned int dg = (hiw >> (8 * (ps - 4))) & 255u;
        const unsigned int act = __activemask();
        const unsigned int same = __match_any_sync(act, dg);
        if (lane == __ffs(same) - 1) atomicAdd(&s_hist[ps][dg], (unsigned)__popc(same));
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < ws.sort_passes * kRadix; i += kScanThreads) {
    const unsigned int c = (&s_hist[0][0])[i];
    if (c) atomicAdd(&ws.ctl->hist[0][0] + i, c);
  }
}

// ---------------------------------------------------------------------------------------------
// 2. onesweep pass: one LSD radix pass (8-bit digit) over (u64 key, u32 value) pairs with
//    chained-scan decoupled look-back per digit; stable.  Replaces cub::DeviceRadixSort::SortPairs
//    (X6).  Tile = 256 threads x 12 keys, warp-striped so that rank order == input order.
// ---------------------------------------------------------------------------------------------
constexpr unsigned int kSFlagAgg = 1u << 30, kSFlagIncl = 2u << 30, kSValMask = (1u << 30) - 1;

__global__ void __launch_bounds__(kSortThreads) onesweep_pass_kernel(Workspace ws, int pass) {
  __shared__ unsigned int s_whist[kSortThreads / 32][kRadix];
  __shared__ unsigned int s_dexcl[kRadix];     // block-level exclusive offset of each digit
  __shared__ unsigned int s_goff[kRadix];      // global offset of digit d minus s_dexcl[d]
  __shared__ unsigned long long s_keys[kSortTile];
  __shared__ unsigned int s_vals[kSortTile];
  __shared__ unsigned int s_tile;
  __shared__ unsigned int s_scan[kSortThreads / 32];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned long long Dfull = ws.ctl->num_rendered;
  const unsigned int D = (unsigned int)(Dfull < ws.d_cap ? Dfull : ws.d_cap);
  if (tid == 0) s_tile = atomicAdd(&ws.ctl->sort_tile_counter[pass], 1u);
  for (int i = tid; i < (kSortThreads / 32) * kRadix; i += kSortThreads) (&s_whist[0][0])[i] = 0;
  __syncthreads();
  const unsigned int tile = s_tile;
  if ((unsigned long long)tile * kSortTile >= D) return;
  const unsigned int base = tile * kSortTile;
  const unsigned int n = min((unsigned int)kSortTile, D - base);
  const int shift = pass * kRadixBits;
  const unsigned long long* __restrict__ kin = ws.keys[pass & 1];
  const unsigned int* __restrict__ vin = ws.vals[pass & 1];
  unsigned long long* __restrict__ kout = ws.keys[(pass & 1) ^ 1];
  unsigned int* __restrict__ vout = ws.vals[(pass & 1) ^ 1];

  unsigned long long key[kSortKpt];
  unsigned int val[kSortKpt];
  unsigned short rank[kSortKpt];
  const unsigned int wbase = warp * (kSortKpt * 32);
#pragma unroll
  for (int i = 0; i < kSortKpt; i++) {
    const unsigned int li = wbase + i * 32 + lane;
    if (li < n) { key[i] = kin[base + li]; val[i] = vin[base + li]; }
    else { key[i] = ~0ull; val[i] = 0; }
  }
  // warp-private ranking: step i ascending, lane ascending == input order (stable)
#pragma unroll
  for (int i = 0; i < kSortKpt; i++) {
    const unsigned int li = wbase + i * 32 + lane;
    const bool valid = li < n;
    const unsigned int dg = (unsigned int)(key[i] >> shift) & 255u;
    // peers with the same digit from 8 ballots (MATCH.ANY is microcoded and slow on sm_100)
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
      old = s_whist[warp][dg];
      s_whist[warp][dg] = old + __popc(same);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[i] = (unsigned short)(old + __popc(same & ((1u << lane) - 1u)));
    __syncwarp();
  }
  __syncthreads();
  // thread d owns digit d: prefix over warps, block digit scan, look-back
  unsigned int dcount = 0;
  {
    const int d = tid;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; w++) {
      const unsigned int c = s_whist[w][d];
      s_whist[w][d] = dcount;
      dcount += c;
    }
  }
  unsigned int incl = dcount;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_scan[warp] = incl;
  // global digit base = exclusive scan over digits of the upfront histogram
  const unsigned int hcount = ws.ctl->hist[pass][tid];
  unsigned int hincl = hcount;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, hincl, d);
    if (lane >= d) hincl += t;
  }
  __shared__ unsigned int s_hscan[kSortThreads / 32];
  if (lane == 31) s_hscan[warp] = hincl;
  __syncthreads();
  unsigned int wpre = 0, hpre = 0;
#pragma unroll
  for (int w = 0; w < kSortThreads / 32; w++) {
    if (w < warp) { wpre += s_scan[w]; hpre += s_hscan[w]; }
  }
  const unsigned int dexcl = wpre + incl - dcount;
  const unsigned int hexcl = hpre + hincl - hcount;
  // look-back for digit tid
  unsigned int* state = ws.sort_state + ((size_t)pass * ws.num_sort_tiles) * kRadix;
  st_relaxed_u32(&state[(size_t)tile * kRadix + tid], (tile == 0 ? kSFlagIncl : kSFlagAgg) | dcount);
  // Look-back, kLook predecessors per round trip: the loads of a window are independent, so the chain
  // advances kLook tiles per L2 latency instead of one (with ~500 resident tiles and one hop per
  // latency the serial walk WAS the pass: ~24 us of a 27 us launch at C2).
  unsigned int gpre = 0;
  if (tile != 0) {
    constexpr int kLook = 8;
    int look = (int)tile - 1;
    bool fin = false;
    while (!fin) {
      unsigned int w[kLook];
#pragma unroll
      for (int i = 0; i < kLook; i++) {
        const int t = look - i;
        w[i] = t >= 0 ? ld_relaxed_u32(&state[(size_t)t * kRadix + tid]) : kSFlagIncl;  // virtual tile -1: inclusive 0
      }
#pragma unroll
      for (int i = 0; i < kLook; i++) {
        if (fin) break;
        if ((w[i] >> 30) == 0) break;  // not published yet: retry the window from here
        gpre += w[i] & kSValMask;
        look--;
        if ((w[i] >> 30) == 2) fin = true;
      }
    }
    st_relaxed_u32(&state[(size_t)tile * kRadix + tid], kSFlagIncl | (gpre + dcount));
  }
  s_dexcl[tid] = dexcl;
  s_goff[tid] = hexcl + gpre - dexcl;
  __syncthreads();
  // scatter into shared memory in rank order
#pragma unroll
  for (int i = 0; i < kSortKpt; i++) {
    const unsigned int li = wbase + i * 32 + lane;
    if (li < n) {
      const unsigned int dg = (unsigned int)(key[i] >> shift) & 255u;
      const unsigned int pos = s_dexcl[dg] + s_whist[warp][dg] + rank[i];
      s_keys[pos] = key[i];
      s_vals[pos] = val[i];
    }
  }
  __syncthreads();
  for (unsigned int s = tid; s < n; s += kSortThreads) {
    const unsigned long long k = s_keys[s];
    const unsigned int dg = (unsigned int)(k >> shift) & 255u;
    const unsigned int out = s_goff[dg] + s;
    kout[out] = k;
    vout[out] = s_vals[s];
  }
}

// ---------------------------------------------------------------------------------------------
// 3. identifyTileRanges (X7) and the heaviest-first tile order; the per-block record lists the render kernels
//    stream follow in block_lists_kernel (blocklist.cu), as in the bucket path.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ranges_gather_kernel(Workspace ws, int G) {
  const unsigned long long Dfull = ws.ctl->num_rendered;
  const unsigned int D = (unsigned int)(Dfull < ws.d_cap ? Dfull : ws.d_cap);
  const unsigned long long* __restrict__ keys = ws.keys[ws.sort_passes & 1];
  for (unsigned int j = blockIdx.x * blockDim.x + threadIdx.x; j < D; j += gridDim.x * blockDim.x) {
    const unsigned int tile = (unsigned int)(keys[j] >> 32);
    if (j == 0) ws.ranges[tile].x = 0;
    else {
      const unsigned int prev = (unsigned int)(keys[j - 1] >> 32);
      if (tile != prev) { ws.ranges[prev].y = j; ws.ranges[tile].x = j; }
    }
    if (j == D - 1) ws.ranges[tile].y = D;
  }
  if (!ws.use_order) return;
  // The last CTA to finish orders the (view, tile) ids by list length, longest first (counting sort
  // on length/16): the render kernels take their tile from this table, so the hardware CTA
  // scheduler hands out the long lists first and the short ones fill the tail of the launch.
  __shared__ unsigned int s_bucket[1024];
  __shared__ unsigned int s_warpsum[8];
  __shared__ bool s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&ws.ctl->gather_done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int NT = ws.N * ws.T, tid = threadIdx.x;
  for (int i = tid; i < 1024; i += 256) s_bucket[i] = 0;
  __syncthreads();
  auto bucket_of = [&](int i) {
    const uint2 r = __ldcg(&ws.ranges[i]);
    const unsigned int len = r.y - r.x;
    return 1023u - min(1023u, len >> 4);  // bucket 0 = longest lists
  };
  for (int i = tid; i < NT; i += 256) atomicAdd(&s_bucket[bucket_of(i)], 1u);
  __syncthreads();
  {  // exclusive scan of the 1024 bucket counts: 4 per thread
    unsigned int c[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) { c[k] = s_bucket[tid * 4 + k]; sum += c[k]; }
    unsigned int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, d);
      if ((tid & 31) >= d) incl += t;
    }
    if ((tid & 31) == 31) s_warpsum[tid >> 5] = incl;
    __syncthreads();
    unsigned int pre = incl - sum;
    for (int w = 0; w < (tid >> 5); w++) pre += s_warpsum[w];
#pragma unroll
    for (int k = 0; k < 4; k++) { s_bucket[tid * 4 + k] = pre; pre += c[k]; }
  }
  __syncthreads();
  for (int i = tid; i < NT; i += 256) ws.tile_order[atomicAdd(&s_bucket[bucket_of(i)], 1u)] = (unsigned int)i;
}

cudaError_t launch_bin_emit(const TsplatProblem& p, const Workspace& ws, cudaStream_t st) {
  const int total = ws.N * p.G;
  bin_emit_kernel<<<ws.num_scan_tiles, kScanThreads, 0, st>>>(ws, p.G, total);
  return cudaGetLastError();
}

cudaError_t launch_sort(const TsplatProblem&, const Workspace& ws, cudaStream_t st) {
  for (int pass = 0; pass < ws.sort_passes; pass++) {
    onesweep_pass_kernel<<<ws.num_sort_tiles, kSortThreads, 0, st>>>(ws, pass);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_ranges_gather(const TsplatProblem& p, const Workspace& ws, cudaStream_t st) {
  ranges_gather_kernel<<<148 * 8, 256, 0, st>>>(ws, p.G);
  return cudaGetLastError();
}

}  // namespace tsplat
