// common.cuh — shared declarations for libtsplat.so (sm_100a only).
// Internal header: the public boundary is include/tsplat.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tsplat.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libtsplat is written for sm_100a (B200) only"
#endif

// Checked builds (make EXTRA=-DTSPLAT_CHECKED; scripts/exp_variants.sh): device-side bounds assertions on every index the
// binning and render kernels derive from data.  compute-sanitizer is closed on the B200 pool this was developed on, so
// this is the memory check that can be run there (profiles/r02_checked_build.md).
#ifdef TSPLAT_CHECKED
#include <assert.h>
#define TS_CHECK(cond) assert(cond)
#else
#define TS_CHECK(cond) ((void)0)
#endif

namespace tsplat {

constexpr int kTile = TSPLAT_TILE;          // 16x16 pixels per tile (parity contract)
constexpr int kTilePix = kTile * kTile;     // 256
#ifndef TSPLAT_PRE_BLOCK
#define TSPLAT_PRE_BLOCK 128
#endif
constexpr int kPreBlock = TSPLAT_PRE_BLOCK;
#ifndef TSPLAT_PRE_BLOCK_BWD
#define TSPLAT_PRE_BLOCK_BWD 64
#endif
constexpr int kPreBwdBlock = TSPLAT_PRE_BLOCK_BWD;  // Gaussians per preprocess-backward CTA              // Gaussians per preprocess CTA
constexpr int kScanThreads = 256;
constexpr int kScanItems = 4;               // (view, Gaussian) items per thread in bin_emit
constexpr int kScanTile = kScanThreads * kScanItems;
constexpr int kSortThreads = 256;
constexpr int kSortKpt = 12;                // keys per thread per onesweep tile
constexpr int kSortTile = kSortThreads * kSortKpt;
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
constexpr int kMaxPasses = 8;
constexpr int kBlocksPerTile = 8;            // 8x4-pixel blocks of a 16x16 tile = warp jobs of the render kernels
constexpr int kBlockEntryRatio = 2;          // capacity of the per-block lists in entries per pair-list slot (e_cap = 2 d_cap)
constexpr int kRecFloat4 = 3;                // one per-Gaussian record = 3 x float4 = 48 B
constexpr int kSegLen = 64;                  // render_backward work unit: references per list segment (a multiple of 32);
                                             // render_forward leaves a per-pixel checkpoint at every multiple of it
constexpr int kCkptFloats = 5 * 32;          // one checkpoint: T and the 4 channel prefixes of the block's 32 pixels
constexpr int kSmallTileCap = 4096;          // bucket path: pairs per tile sorted by the small-tile kernel (32 KB smem)
constexpr int kBigTileCap = 16384;           // ... by the big-tile kernel in shared memory (128 KB); beyond: in global memory
constexpr int kOrderMaxTiles = 8192;         // order (view, tile) CTAs heaviest-first only up to this many
constexpr int kAccFloats = 12;              // screen-space gradient accumulators per (view, Gaussian)

// Device-side control block at the start of the workspace (superset of TsplatStatus).
struct Control {
  unsigned long long num_rendered;  // D (uncapped)
  unsigned int overflow;
  unsigned int reserved;
  unsigned long long num_block_entries;  // entries of all per-block lists (uncapped; block_lists_kernel's allocation cursor)
  unsigned long long reserved1;          // (the first 32 bytes are the public TsplatStatus)
  unsigned int scan_tile_counter;   // dynamic tile ids for bin_emit
  unsigned int sort_tile_counter[kMaxPasses];
  unsigned int gather_done;         // CTAs of ranges_gather that have finished (last one orders the tiles)
  unsigned int num_big;             // bucket path: tiles whose list does not fit the small-tile sort kernel
  unsigned int loss_done;           // tiles of render_forward whose MSE partial is complete (fused loss)
  unsigned int fwd_next;            // render_forward: jobs handed out beyond each warp's first one
  unsigned int seg_full, seg_part;  // render_backward work units emitted by render_forward: full / partial list segments
  unsigned int hist[kMaxPasses][kRadix];  // global digit histograms (onesweep: one upfront sweep)
};

// Resolved workspace pointers (host-side struct passed to kernels by value).
struct Workspace {
  Control* ctl;
  float* depths;
  int* radii;
  float4* grec;  // per (view, Gaussian) record, 3 x float4 in the layout of rec0/rec1/rec2 below: the gather into
                 // sorted order is a plain 48-B copy, and the cull threshold is computed once per Gaussian
  unsigned int* tiles_touched;
  unsigned char* clamped;
  unsigned long long* keys[2];
  unsigned int* vals[2];
  uint2* ranges;
  float* final_T;
  unsigned int* n_contrib;
  float* grad_acc;
  float* loss_partial;             // fused MSE loss / PSNR partial sums: [2][N*T*8] per (tile, block) job, then [2][N*T] per tile
  unsigned int* tile_jobs_done;    // per (view, tile): block jobs of render_forward that have written their loss partial
  // per-block lists (blocklist.cu): the entries of a tile's sorted list that can touch one 8x4 pixel block, in list order,
  // as references (Gaussian id, 1-based position in the TILE list) - contiguous per block: a render warp streams its
  // own list with 1-D TMA bulk copies and gathers the 48-B records from grec
  uint2* bref;
  // render job records: first reference, reference count, (view*T + tile) * 8 + block, 0
  uint4* block_jobs;               // in the order render_forward deals them: heaviest tile first (block_lists_kernel)
  // render_backward work units (emitted by render_forward): a block list cut at its deepest contributor and then into
  // segments of kSegLen references: first reference of the LIST, cut length, job id, segment index.  Full segments grow
  // from the front of the table, partial ones from its back; dealt round-robin over the warps of the launch.
  uint4* segs;
  unsigned int seg_cap;
  float* ckpt;                     // [e_cap / kSegLen + 2][5][32] per-pixel state in front of reference k * kSegLen of a list
  float* c_final;                  // [N][4][H*W] the composited sums (colour without background, raw depth channel)
  unsigned char* pair_mask;        // per sorted pair: its 8-bit block mask (cull.cuh), between the two passes of block_lists
  unsigned long long e_cap;        // capacity of brec in entries
  unsigned int* tile_count;        // bucket path: pairs per (view, tile), counted by the preprocess kernel
  unsigned int* tile_cursor;       // bucket path: emission cursors per (view, tile)
  unsigned int* big_list;          // bucket path: (view, tile) ids with more than kSmallTileCap pairs
  int bucket;                      // 1: bucket path (count -> scan -> emit -> per-tile sort), 0: onesweep LSD path
  unsigned int* tile_order;        // (view, tile) ids, heaviest list first (render launch order)
  int use_order;                   // 0: identity order (very many tiles: waves balance themselves)
  unsigned long long* scan_state;  // decoupled look-back words for bin_emit
  unsigned int* sort_state;        // [passes][sort tiles][256]
  unsigned long long d_cap;
  int tiles_x, tiles_y, T, N;      // tiles per view, views per call
  int sort_bits, sort_passes;
  int num_scan_tiles, num_sort_tiles;
};

struct LayoutFull {
  TsplatLayout pub;
  uint64_t keys_alt, vals_alt, bref, block_jobs, segs, ckpt, c_final, pair_mask, tile_jobs_done, loss_partial, scan_state, sort_state,
      tile_order, tile_count, tile_cursor, big_list, e_cap,
      zero_begin, zero_mid, zero_end;  // [zero_begin, zero_mid) is all the bucket path needs cleared
  int num_scan_tiles, num_sort_tiles;
};

// Raises a kernel's dynamic shared-memory limit once per (device, kernel, size) instead of on every launch
// (cudaFuncSetAttribute is a driver call; the attribute is per device).  Thread-safe.
cudaError_t ensure_dynamic_smem(const void* kernel, int bytes);

int compute_layout(const TsplatProblem& p, uint64_t d_cap, LayoutFull* out);
Workspace resolve(const LayoutFull& L, void* base, uint64_t d_cap, int N);
void set_error(const char* fmt, ...);

// stage launchers (each only enqueues on `st`)
cudaError_t launch_preprocess(const TsplatProblem& p, const Workspace& ws, int* out_radii, void* zero_ptr,
                              uint64_t zero_bytes, cudaStream_t st);
cudaError_t launch_bucket_scan_emit(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_bucket_sort_gather(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_bin_emit(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_sort(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_ranges_gather(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_block_lists(const TsplatProblem& p, const Workspace& ws, cudaStream_t st);
cudaError_t launch_empty_jobs(const Workspace& ws, cudaStream_t st);  // G == 0: every render job has an empty list
cudaError_t launch_render_forward(const TsplatProblem& p, const Workspace& ws, float* out_color,
                                  float* out_depth, const TsplatMse* mse, cudaStream_t st);
cudaError_t launch_render_backward(const TsplatProblem& p, const Workspace& ws, const float* dL_dcolor,
                                   const float* dL_ddepth, const TsplatMse* mse, cudaStream_t st);
cudaError_t launch_preprocess_backward(const TsplatProblem& p, const Workspace& ws, float* dL_dmeans,
                                       float* dL_dcov, float* dL_dsh, float* dL_dopacity,
                                       float* dL_dmeans2D, int g_begin, int g_count, bool multimem, cudaStream_t st);
cudaError_t launch_pack_cameras(int N, const float* extrinsics, const float* intrinsics, const float* near_,
                                const float* far_, const float* bg, int bg_stride, int scale_invariant,
                                TsplatCamera* out, cudaStream_t st);
cudaError_t launch_adapter_forward(const TsplatAdapterProblem& p, float* means, float* cov, float* harmonics,
                                   float* scales, float* rotations, cudaStream_t st, int* bad);
cudaError_t launch_adapter_backward(const TsplatAdapterProblem& p, const float* g_means, const float* g_cov,
                                    const float* g_harm, float* g_coords, float* g_depths, float* g_raw,
                                    cudaStream_t st, int* bad);
cudaError_t launch_cov3d_forward(int G, const float* scales, const float* rot, float mod, float* cov6, cudaStream_t st);
cudaError_t launch_cov3d_backward(int G, const float* scales, const float* rot, float mod, const float* dL_dcov6,
                                  float* dL_dscales, float* dL_drot, cudaStream_t st);
void launch_status_store(const void* workspace, void* host_alias, cudaStream_t st);
cudaError_t launch_mark_visible(int G, const float* means, const float* view, uint8_t* present,
                                cudaStream_t st);

// ---------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + 1-D bulk async copy (TMA, SASS UBLKCP) — sm_90+/sm_100a.
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
// suspend-time hint (ns): a waiting warp sleeps in hardware instead of spinning on the issue port
constexpr uint32_t kMbarSuspendNs = 2000;
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(kMbarSuspendNs)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// (Measured on B200: replacing this hardware-assisted wait by test_wait + __nanosleep back-off removes the
// TRYWAIT/NANOSLEEP.SYNCS loop instructions - 17 % of the render kernels' issued instructions - but does not
// shorten the kernels: those issue slots were idle anyway; profiles/r01_experiments.md.)
// global -> shared::cta bulk copy; bytes % 16 == 0, both addresses 16-B aligned.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// shared::cta -> global bulk copy (bulk async-group completion)
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// make generic-proxy smem writes visible to the async proxy before a bulk store
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_relaxed_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
#endif  // __CUDACC__

}  // namespace tsplat
