// api.cu — the C-ABI of libtsplat.so (include/tsplat.h): workspace carving and stage orchestration.
// Replaces rasterize_points.cu + CudaRasterizer::Rasterizer::{forward,backward} of the reference's
// external extension.  Nothing here allocates device memory or synchronises.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace tsplat {

cudaError_t ensure_dynamic_smem(const void* kernel, int bytes) {
  struct Entry { int dev; const void* fn; int bytes; };
  static std::mutex mu;
  static std::vector<Entry> seen;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  for (Entry& en : seen)
    if (en.dev == dev && en.fn == kernel) {
      if (bytes <= en.bytes) return cudaSuccess;
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
      if (e == cudaSuccess) en.bytes = bytes;
      return e;
    }
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) seen.push_back({dev, kernel, bytes});
  return e;
}

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int bits_for(uint64_t n) {  // smallest b with n <= 2^b
  int b = 0;
  while (((uint64_t)1 << b) < n) b++;
  return b;
}

static int validate(const TsplatProblem* p, uint64_t d_cap) {
  if (!p) { set_error("problem is NULL"); return TSPLAT_EINVAL; }
  if (p->B < 1 || p->V < 1 || p->G < 0 || p->H < 1 || p->W < 1) {
    set_error("bad sizes B=%d G=%d V=%d H=%d W=%d", p->B, p->G, p->V, p->H, p->W);
    return TSPLAT_EINVAL;
  }
  if (p->sh_degree < 0 || p->sh_degree > 4) { set_error("sh_degree %d not in 0..4", p->sh_degree); return TSPLAT_EINVAL; }
  if (p->sh_coeffs < 0 || p->sh_coeffs > 25) { set_error("sh_coeffs %d not in 0..25", p->sh_coeffs); return TSPLAT_EINVAL; }
  if (p->sh_coeffs > 0) {
    const int eff = (p->sh_degree >= 4 && !p->sh_l4) ? 3 : p->sh_degree;
    if ((eff + 1) * (eff + 1) > p->sh_coeffs) {
      set_error("sh_degree %d needs %d coefficients, got %d", p->sh_degree, (eff + 1) * (eff + 1), p->sh_coeffs);
      return TSPLAT_EINVAL;
    }
  }
  if (p->sh_layout != TSPLAT_SH_GM3 && p->sh_layout != TSPLAT_SH_G3M) { set_error("bad sh_layout"); return TSPLAT_EINVAL; }
  if (p->cov_layout != TSPLAT_COV_TRIU6 && p->cov_layout != TSPLAT_COV_FULL9) { set_error("bad cov_layout"); return TSPLAT_EINVAL; }
  if (p->depth_mode < TSPLAT_DEPTH_NONE || p->depth_mode > TSPLAT_DEPTH_LOG) { set_error("bad depth_mode"); return TSPLAT_EINVAL; }
  if (d_cap < 1 || d_cap >= (1ull << 30)) { set_error("d_cap %llu not in [1, 2^30)", (unsigned long long)d_cap); return TSPLAT_EINVAL; }
  if ((uint64_t)p->B * p->V * (uint64_t)p->G >= (1ull << 31)) { set_error("B*V*G too large"); return TSPLAT_EINVAL; }
  const int tx = (p->W + kTile - 1) / kTile, ty = (p->H + kTile - 1) / kTile;
  if (tx > 1023 || ty > 1023) { set_error("image too large (more than 1023 tiles per axis)"); return TSPLAT_EINVAL; }
  if ((uint64_t)p->B * p->V * tx * ty >= (1ull << 31)) { set_error("too many (view, tile) pairs"); return TSPLAT_EINVAL; }
  return TSPLAT_OK;
}

int compute_layout(const TsplatProblem& p, uint64_t d_cap, LayoutFull* out) {
  memset(out, 0, sizeof(*out));
  const uint64_t N = (uint64_t)p.B * p.V, G = (uint64_t)p.G, NG = N * G;
  const int tx = (p.W + kTile - 1) / kTile, ty = (p.H + kTile - 1) / kTile;
  const uint64_t T = (uint64_t)tx * ty, P = (uint64_t)p.H * p.W;
  uint64_t off = 0;
  auto take = [&](uint64_t bytes) { const uint64_t o = off; off = (off + bytes + 255) & ~(uint64_t)255; return o; };
  TsplatLayout& L = out->pub;
  out->num_scan_tiles = (int)((NG + kScanTile - 1) / kScanTile);
  if (out->num_scan_tiles < 1) out->num_scan_tiles = 1;
  out->num_sort_tiles = (int)((d_cap + kSortTile - 1) / kSortTile);
  L.tiles_x = tx; L.tiles_y = ty;
  L.sort_bits = 32 + (bits_for(N * T) > 0 ? bits_for(N * T) : 1);
  L.sort_passes = (L.sort_bits + kRadixBits - 1) / kRadixBits;
  // ---- region zeroed at the start of every forward ----
  out->zero_begin = off;
  L.status = take(sizeof(Control));
  L.ranges = take(8ull * N * T);
  out->tile_cursor = take(4ull * N * T);
  out->tile_jobs_done = take(4ull * N * T);
  // tile_count is accumulated by the preprocess kernel itself (bucket path), so that kernel cannot also be the one
  // that clears it: it sits last in the region and is cleared by a memset node ahead of the launch
  out->tile_count = take(4ull * N * T);
  out->zero_mid = off;
  out->scan_state = take(8ull * out->num_scan_tiles);
  out->sort_state = take(4ull * kRadix * out->num_sort_tiles * L.sort_passes);
  out->zero_end = off;
  out->tile_order = take(4ull * N * T);
  out->big_list = take(4ull * N * T);
  // ---- per (view, Gaussian) state ----
  L.depths = take(4 * NG);
  L.radii = take(4 * NG);
  // one 48-B record per (view, Gaussian): x, y, conic.xyz, opacity, r, g, b, depth feature, id bits, cull threshold
  const uint64_t grec = take(48 * NG);
  L.xy = grec; L.conic_opacity = grec + 8; L.rgbd = grec + 24;
  L.rec_stride = 48; L.reserved0 = 0;
  L.tiles_touched = take(4 * NG);
  L.clamped = take(NG);
  // ---- pair list (double-buffered for the radix passes) ----
  const uint64_t k0 = take(8 * d_cap), k1 = take(8 * d_cap);
  const uint64_t v0 = take(4 * d_cap), v1 = take(4 * d_cap);
  const bool odd = L.sort_passes & 1;
  L.keys = odd ? k1 : k0; out->keys_alt = odd ? k0 : k1;
  L.vals = odd ? v1 : v0; out->vals_alt = odd ? v0 : v1;
  // ---- per-block lists ----
  out->e_cap = (uint64_t)kBlockEntryRatio * d_cap;
  out->bref = take(8ull * (out->e_cap + 2));  // + the padding reference an odd last batch copies along
  L.block_refs = out->bref; L.e_cap = out->e_cap;
  out->pair_mask = take(d_cap);
  out->block_jobs = take(16ull * kBlocksPerTile * N * T);
  out->segs = take(16ull * (out->e_cap / kSegLen + kBlocksPerTile * N * T + 1));
  out->ckpt = take(4ull * kCkptFloats * (out->e_cap / kSegLen + 2));
  out->c_final = take(16ull * N * P);
  L.block_jobs = out->block_jobs;
  // ---- per pixel ----
  L.final_T = take(4 * N * P);
  L.n_contrib = take(4 * N * P);
  // ---- backward ----
  L.grad_acc = take(4ull * kAccFloats * NG);
  out->loss_partial = take(4ull * N * T * 2 * (kBlocksPerTile + 1));  // loss and PSNR: one partial each per (tile, block) job and per tile
  L.total_bytes = off;
  return TSPLAT_OK;
}

Workspace resolve(const LayoutFull& F, void* base, uint64_t d_cap, int N) {
  const TsplatLayout& L = F.pub;
  char* b = static_cast<char*>(base);
  Workspace w;
  w.ctl = reinterpret_cast<Control*>(b + L.status);
  w.depths = reinterpret_cast<float*>(b + L.depths);
  w.radii = reinterpret_cast<int*>(b + L.radii);
  w.grec = reinterpret_cast<float4*>(b + L.xy);
  w.tiles_touched = reinterpret_cast<unsigned int*>(b + L.tiles_touched);
  w.clamped = reinterpret_cast<unsigned char*>(b + L.clamped);
  const bool odd = L.sort_passes & 1;  // buffer 0 is where bin_emit writes; sorted result lands in [passes & 1]
  w.keys[0] = reinterpret_cast<unsigned long long*>(b + (odd ? F.keys_alt : L.keys));
  w.keys[1] = reinterpret_cast<unsigned long long*>(b + (odd ? L.keys : F.keys_alt));
  w.vals[0] = reinterpret_cast<unsigned int*>(b + (odd ? F.vals_alt : L.vals));
  w.vals[1] = reinterpret_cast<unsigned int*>(b + (odd ? L.vals : F.vals_alt));
  w.ranges = reinterpret_cast<uint2*>(b + L.ranges);
  w.final_T = reinterpret_cast<float*>(b + L.final_T);
  w.n_contrib = reinterpret_cast<unsigned int*>(b + L.n_contrib);
  w.grad_acc = reinterpret_cast<float*>(b + L.grad_acc);
  w.loss_partial = reinterpret_cast<float*>(b + F.loss_partial);
  w.bref = reinterpret_cast<uint2*>(b + F.bref);
  w.block_jobs = reinterpret_cast<uint4*>(b + F.block_jobs);
  w.segs = reinterpret_cast<uint4*>(b + F.segs);
  w.seg_cap = (unsigned int)(F.e_cap / kSegLen + (uint64_t)kBlocksPerTile * N * (L.tiles_x * L.tiles_y) + 1);
  w.ckpt = reinterpret_cast<float*>(b + F.ckpt);
  w.c_final = reinterpret_cast<float*>(b + F.c_final);
  w.pair_mask = reinterpret_cast<unsigned char*>(b + F.pair_mask);
  w.tile_jobs_done = reinterpret_cast<unsigned int*>(b + F.tile_jobs_done);
  w.e_cap = F.e_cap;
  w.tile_order = reinterpret_cast<unsigned int*>(b + F.tile_order);
  w.tile_count = reinterpret_cast<unsigned int*>(b + F.tile_count);
  w.tile_cursor = reinterpret_cast<unsigned int*>(b + F.tile_cursor);
  w.big_list = reinterpret_cast<unsigned int*>(b + F.big_list);
  w.bucket = 1;
  w.use_order = (uint64_t)N * (uint64_t)(L.tiles_x * L.tiles_y) <= (uint64_t)kOrderMaxTiles ? 1 : 0;
  w.scan_state = reinterpret_cast<unsigned long long*>(b + F.scan_state);
  w.sort_state = reinterpret_cast<unsigned int*>(b + F.sort_state);
  w.d_cap = d_cap;
  w.tiles_x = L.tiles_x; w.tiles_y = L.tiles_y; w.T = L.tiles_x * L.tiles_y; w.N = N;
  w.sort_bits = L.sort_bits; w.sort_passes = L.sort_passes;
  w.num_scan_tiles = F.num_scan_tiles; w.num_sort_tiles = F.num_sort_tiles;
  return w;
}

// ---- optional per-stage timing (cudaEvents on the caller's stream; off by default) ----
enum Stage { kStZero = 0, kStPre, kStBin, kStSort, kStRanges, kStRender, kStZeroB, kStRenderB, kStPreB, kStCount };
constexpr int kProfSlots = 256;  // ring of recorded calls; a slot = one forward + its backward
static bool g_prof = false;
static cudaEvent_t g_ev[kProfSlots][kStCount][2];
static bool g_ev_made = false;
static bool g_ev_used[kProfSlots][kStCount];
static long long g_prof_calls = 0;  // forwards seen since enable
struct StageTimer {
  int s, slot; cudaStream_t st; bool on;
  StageTimer(int s_, cudaStream_t st_) : s(s_), st(st_), on(g_prof && g_prof_calls > 0) {
    if (!on) return;
    slot = (int)((g_prof_calls - 1) % kProfSlots);
    cudaEventRecord(g_ev[slot][s][0], st);
  }
  ~StageTimer() { if (on) { cudaEventRecord(g_ev[slot][s][1], st); g_ev_used[slot][s] = true; } }
};
static void prof_begin_forward() {
  if (!g_prof) return;
  if (!g_ev_made) {
    for (auto& sl : g_ev) for (auto& p : sl) { cudaEventCreate(&p[0]); cudaEventCreate(&p[1]); }
    g_ev_made = true;
  }
  g_prof_calls++;
  const int slot = (int)((g_prof_calls - 1) % kProfSlots);
  for (int s = 0; s < kStCount; s++) g_ev_used[slot][s] = false;
}

static int cuda_fail(const char* what, cudaError_t e) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return TSPLAT_ECUDA;
}

// debug=True of the reference: CHECK_CUDA(..., debug) = cudaDeviceSynchronize + throw after every kernel
static cudaError_t debug_sync(const TsplatProblem* p, cudaStream_t st, cudaError_t e) {
  if (e != cudaSuccess || !(p->flags & TSPLAT_FLAG_DEBUG)) return e;
  return cudaStreamSynchronize(st);
}

static int prepare(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                   LayoutFull* F, Workspace* ws) {
  int rc = validate(p, d_cap);
  if (rc) return rc;
  if (!workspace) { set_error("workspace is NULL"); return TSPLAT_EINVAL; }
  if (reinterpret_cast<uintptr_t>(workspace) & 255u) { set_error("workspace must be 256-B aligned"); return TSPLAT_EINVAL; }
  if (!p->means || !p->cov || !p->sh || !p->opacity || !p->cams) { set_error("NULL input pointer"); return TSPLAT_EINVAL; }
  compute_layout(*p, d_cap, F);
  if (F->pub.total_bytes > workspace_bytes) {
    set_error("workspace too small: need %llu bytes, have %llu", (unsigned long long)F->pub.total_bytes,
              (unsigned long long)workspace_bytes);
    return TSPLAT_ENOSPC;
  }
  *ws = resolve(*F, workspace, d_cap, p->B * p->V);
  return TSPLAT_OK;
}

}  // namespace tsplat

using namespace tsplat;

extern "C" {

int tsplat_version(void) { return TSPLAT_VERSION; }
const char* tsplat_last_error(void) { return g_err; }

int tsplat_workspace_layout(const TsplatProblem* p, uint64_t d_cap, TsplatLayout* out) {
  int rc = validate(p, d_cap);
  if (rc) return rc;
  if (!out) { set_error("out is NULL"); return TSPLAT_EINVAL; }
  LayoutFull F;
  compute_layout(*p, d_cap, &F);
  *out = F.pub;
  return TSPLAT_OK;
}

static int forward_impl(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                        float* out_color, float* out_depth, int32_t* out_radii, const TsplatMse* mse,
                        void* cuda_stream) {
  LayoutFull F;
  Workspace ws;
  int rc = prepare(p, workspace, workspace_bytes, d_cap, &F, &ws);
  if (rc) return rc;
  if (!out_color) { set_error("out_color is NULL"); return TSPLAT_EINVAL; }
  if (p->depth_mode != TSPLAT_DEPTH_NONE && !out_depth) { set_error("depth_mode set but out_depth is NULL"); return TSPLAT_EINVAL; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e;
  prof_begin_forward();
  e = cudaSuccess;
  if (p->G == 0) {
    // nothing to draw: background only.  num_rendered stays 0; the render kernel sees empty ranges.
    StageTimer t(kStZero, st);
    e = cudaMemsetAsync(static_cast<char*>(workspace) + F.zero_begin, 0, F.zero_end - F.zero_begin, st);
    if (e != cudaSuccess) return cuda_fail("memset(control)", e);
    // every block list empty; the tile id of job j is j / 8 (identity order)
    e = launch_empty_jobs(ws, st);
    if (e != cudaSuccess) return cuda_fail("empty_jobs", e);
    ws.use_order = 0;
  } else {
    // the preprocess kernel also zeroes the control / cursors / ranges region (nothing in it is touched before the
    // binning kernels) and the screen-space gradient accumulators of the Gaussians it keeps.  The per-tile pair
    // counters of the bucket path are the exception: the same launch accumulates them, and nothing orders one CTA's
    // clear against another CTA's reduction, so they are cleared by a memset node ahead of the launch.
    ws.bucket = (p->flags & TSPLAT_FLAG_SORT_ONESWEEP) ? 0 : 1;
    if (ws.bucket) {
      StageTimer t(kStZero, st);
      e = cudaMemsetAsync(static_cast<char*>(workspace) + F.tile_count, 0, F.zero_mid - F.tile_count, st);
      if (e != cudaSuccess) return cuda_fail("memset(tile_count)", e);
    }
    {
      StageTimer t(kStPre, st);
      e = debug_sync(p, st, launch_preprocess(*p, ws, out_radii, static_cast<char*>(workspace) + F.zero_begin,
                                              (ws.bucket ? F.tile_count : F.zero_end) - F.zero_begin, st));
    }
    if (e != cudaSuccess) return cuda_fail("preprocess", e);
    if (ws.bucket) {
      { StageTimer t(kStBin, st); e = debug_sync(p, st, launch_bucket_scan_emit(*p, ws, st)); }
      if (e != cudaSuccess) return cuda_fail("bucket_scan_emit", e);
      { StageTimer t(kStSort, st); e = debug_sync(p, st, launch_bucket_sort_gather(*p, ws, st)); }
      if (e != cudaSuccess) return cuda_fail("bucket_sort_gather", e);
      { StageTimer t(kStRanges, st); e = debug_sync(p, st, launch_block_lists(*p, ws, st)); }
      if (e != cudaSuccess) return cuda_fail("block_lists", e);
    } else {
      { StageTimer t(kStBin, st); e = debug_sync(p, st, launch_bin_emit(*p, ws, st)); }
      if (e != cudaSuccess) return cuda_fail("bin_emit", e);
      { StageTimer t(kStSort, st); e = debug_sync(p, st, launch_sort(*p, ws, st)); }
      if (e != cudaSuccess) return cuda_fail("sort", e);
      {
        StageTimer t(kStRanges, st);
        e = debug_sync(p, st, launch_ranges_gather(*p, ws, st));
        if (e == cudaSuccess) e = debug_sync(p, st, launch_block_lists(*p, ws, st));
      }
      if (e != cudaSuccess) return cuda_fail("ranges / block_lists", e);
    }
  }
  { StageTimer t(kStRender, st); e = debug_sync(p, st, launch_render_forward(*p, ws, out_color, out_depth, mse, st)); }
  if (e != cudaSuccess) return cuda_fail("render_forward", e);
  return TSPLAT_OK;
}

int tsplat_forward(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                   float* out_color, float* out_depth, int32_t* out_radii, void* cuda_stream) {
  return forward_impl(p, workspace, workspace_bytes, d_cap, out_color, out_depth, out_radii, nullptr, cuda_stream);
}

int tsplat_forward_mse(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                       float* out_color, float* out_depth, int32_t* out_radii, const TsplatMse* mse,
                       void* cuda_stream) {
  if (!mse || !mse->target || !mse->loss_sum) { set_error("forward_mse: target and loss_sum are required"); return TSPLAT_EINVAL; }
  return forward_impl(p, workspace, workspace_bytes, d_cap, out_color, out_depth, out_radii, mse, cuda_stream);
}

static int backward_impl(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                         const float* dL_dcolor, const float* dL_ddepth, float* dL_dmeans, float* dL_dcov,
                         float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, const TsplatMse* mse,
                         void* cuda_stream) {
  LayoutFull F;
  Workspace ws;
  int rc = prepare(p, workspace, workspace_bytes, d_cap, &F, &ws);
  if (rc) return rc;
  if ((!dL_dcolor && !mse) || !dL_dmeans || !dL_dcov || !dL_dsh || !dL_dopacity) { set_error("NULL gradient pointer"); return TSPLAT_EINVAL; }
  if (p->depth_mode != TSPLAT_DEPTH_NONE && !dL_ddepth) { set_error("depth_mode set but dL_ddepth is NULL"); return TSPLAT_EINVAL; }
  if (p->G == 0) return TSPLAT_OK;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e;
  // grad_acc is zero after the forward (its preprocess clears the kept Gaussians' accumulators); a repeated
  // backward over the same forward state asks for a fresh clear
  if ((p->flags & TSPLAT_FLAG_REZERO_GRAD_ACC) && !(p->flags & TSPLAT_FLAG_BWD_GAUSSIANS_ONLY)) {
    StageTimer t(kStZeroB, st);
    e = cudaMemsetAsync(ws.grad_acc, 0, sizeof(float) * kAccFloats * (uint64_t)p->B * p->V * p->G, st);
    if (e != cudaSuccess) return cuda_fail("memset(grad_acc)", e);
  }
  if (!(p->flags & TSPLAT_FLAG_BWD_GAUSSIANS_ONLY)) {
    StageTimer t(kStRenderB, st);
    e = debug_sync(p, st, launch_render_backward(*p, ws, dL_dcolor, dL_ddepth, mse, st));
    if (e != cudaSuccess) return cuda_fail("render_backward", e);
  }
  if (p->flags & TSPLAT_FLAG_BWD_RENDER_ONLY) return TSPLAT_OK;
  int g_begin = 0, g_count = p->G;
  if ((p->flags & TSPLAT_FLAG_BWD_GAUSSIANS_ONLY) && p->g_count > 0) {
    g_begin = p->g_begin; g_count = p->g_count;
    if (g_begin < 0 || g_begin % kPreBwdBlock != 0 || (int64_t)g_begin + g_count > p->G) {
      set_error("backward: bad Gaussian range [%d, +%d) (G = %d, g_begin must be a multiple of %d)", g_begin, g_count, p->G, kPreBwdBlock);
      return TSPLAT_EINVAL;
    }
  }
  {
    StageTimer t(kStPreB, st);
    const bool multimem = (p->flags & TSPLAT_FLAG_BWD_MULTIMEM) != 0;
    if (multimem && dL_dmeans2D) { set_error("backward: TSPLAT_FLAG_BWD_MULTIMEM does not cover dL_dmeans2D"); return TSPLAT_EINVAL; }
    e = debug_sync(p, st, launch_preprocess_backward(*p, ws, dL_dmeans, dL_dcov, dL_dsh, dL_dopacity, dL_dmeans2D,
                                                     g_begin, g_count, multimem, st));
  }
  if (e != cudaSuccess) return cuda_fail("preprocess_backward", e);
  return TSPLAT_OK;
}

int tsplat_backward(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                    const float* dL_dcolor, const float* dL_ddepth, float* dL_dmeans, float* dL_dcov,
                    float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, void* cuda_stream) {
  return backward_impl(p, workspace, workspace_bytes, d_cap, dL_dcolor, dL_ddepth, dL_dmeans, dL_dcov, dL_dsh,
                       dL_dopacity, dL_dmeans2D, nullptr, cuda_stream);
}

int tsplat_backward_mse(const TsplatProblem* p, void* workspace, uint64_t workspace_bytes, uint64_t d_cap,
                        const float* dL_dcolor, const float* dL_ddepth, float* dL_dmeans, float* dL_dcov,
                        float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, const TsplatMse* mse,
                        void* cuda_stream) {
  if (!mse || !mse->target || !mse->color || !mse->dL_dloss) {
    set_error("backward_mse: target, color and dL_dloss are required");
    return TSPLAT_EINVAL;
  }
  return backward_impl(p, workspace, workspace_bytes, d_cap, dL_dcolor, dL_ddepth, dL_dmeans, dL_dcov, dL_dsh,
                       dL_dopacity, dL_dmeans2D, mse, cuda_stream);
}

int tsplat_mark_visible(int32_t G, const float* means, const float* viewmatrix, const float* projmatrix,
                        uint8_t* present, void* cuda_stream) {
  (void)projmatrix;  // the reference's frustum test only uses the view matrix (SURVEY A.2 in_frustum)
  if (G < 0 || !means || !viewmatrix || !present) { set_error("bad mark_visible arguments"); return TSPLAT_EINVAL; }
  cudaError_t e = launch_mark_visible(G, means, viewmatrix, present, static_cast<cudaStream_t>(cuda_stream));
  if (e != cudaSuccess) return cuda_fail("mark_visible", e);
  return TSPLAT_OK;
}

int tsplat_cov3d_forward(int32_t G, const float* scales, const float* rotations, float scale_modifier, float* cov6,
                         void* cuda_stream) {
  if (G < 0 || !scales || !rotations || !cov6 || (reinterpret_cast<uintptr_t>(rotations) & 15u)) {
    set_error("bad cov3d_forward arguments (rotations must be 16-B aligned)");
    return TSPLAT_EINVAL;
  }
  cudaError_t e = launch_cov3d_forward(G, scales, rotations, scale_modifier, cov6, static_cast<cudaStream_t>(cuda_stream));
  if (e != cudaSuccess) return cuda_fail("cov3d_forward", e);
  return TSPLAT_OK;
}

int tsplat_cov3d_backward(int32_t G, const float* scales, const float* rotations, float scale_modifier,
                          const float* dL_dcov6, float* dL_dscales, float* dL_drotations, void* cuda_stream) {
  if (G < 0 || !scales || !rotations || !dL_dcov6 || !dL_dscales || !dL_drotations ||
      (reinterpret_cast<uintptr_t>(rotations) & 15u) || (reinterpret_cast<uintptr_t>(dL_drotations) & 15u)) {
    set_error("bad cov3d_backward arguments (rotations and dL_drotations must be 16-B aligned)");
    return TSPLAT_EINVAL;
  }
  cudaError_t e = launch_cov3d_backward(G, scales, rotations, scale_modifier, dL_dcov6, dL_dscales, dL_drotations,
                                        static_cast<cudaStream_t>(cuda_stream));
  if (e != cudaSuccess) return cuda_fail("cov3d_backward", e);
  return TSPLAT_OK;
}

int tsplat_pack_cameras(int32_t N, const float* extrinsics, const float* intrinsics, const float* near_plane,
                        const float* far_plane, const float* bg, int32_t bg_stride, int32_t scale_invariant,
                        TsplatCamera* cams, void* cuda_stream) {
  if (N < 0 || !extrinsics || !intrinsics || !near_plane || !far_plane || !bg || !cams ||
      (bg_stride != 0 && bg_stride != 3)) {
    set_error("bad pack_cameras arguments");
    return TSPLAT_EINVAL;
  }
  cudaError_t e = launch_pack_cameras(N, extrinsics, intrinsics, near_plane, far_plane, bg, bg_stride,
                                      scale_invariant, cams, static_cast<cudaStream_t>(cuda_stream));
  if (e != cudaSuccess) return cuda_fail("pack_cameras", e);
  return TSPLAT_OK;
}

int tsplat_profile_enable(int on) {
  g_prof = on != 0;
  g_prof_calls = 0;
  memset(g_ev_used, 0, sizeof(g_ev_used));
  return TSPLAT_OK;
}

int tsplat_profile_read(float* ms_host, int n) {
  if (!ms_host || n < TSPLAT_NUM_STAGES + 1) { set_error("profile_read needs room for %d floats", TSPLAT_NUM_STAGES + 1); return TSPLAT_EINVAL; }
  for (int s = 0; s <= kStCount; s++) ms_host[s] = 0.0f;
  const long long calls = g_prof_calls < kProfSlots ? g_prof_calls : kProfSlots;
  ms_host[kStCount] = (float)calls;
  if (!g_ev_made) return TSPLAT_OK;
  for (int slot = 0; slot < calls; slot++)
    for (int s = 0; s < kStCount; s++) {
      if (!g_ev_used[slot][s]) continue;
      cudaError_t e = cudaEventSynchronize(g_ev[slot][s][1]);
      if (e != cudaSuccess) return cuda_fail("profile_read", e);
      float ms = 0.0f;
      cudaEventElapsedTime(&ms, g_ev[slot][s][0], g_ev[slot][s][1]);
      ms_host[s] += ms;
    }
  return TSPLAT_OK;
}

int tsplat_status_read(const void* workspace, TsplatStatus* status_host, void* cuda_stream) {
  if (!workspace || !status_host) { set_error("bad status_read arguments"); return TSPLAT_EINVAL; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  // Pinned (or registered) host memory is written by a one-warp kernel through its device alias: a 16-byte
  // cudaMemcpyAsync would go through a copy engine and queue behind whatever bulk D2H transfer another stream
  // has in flight there (measured: it stalled the backward kernels ~0.7 ms per step in the host-buffer pipeline).
  cudaPointerAttributes attr;
  cudaError_t e = cudaPointerGetAttributes(&attr, status_host);
  if (e == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer != nullptr) {
    tsplat::launch_status_store(workspace, attr.devicePointer, st);
    e = cudaGetLastError();
  } else {
    (void)cudaGetLastError();  // pageable memory: clear the attribute query's error, fall back to a copy
    e = cudaMemcpyAsync(status_host, workspace, sizeof(TsplatStatus), cudaMemcpyDeviceToHost, st);
  }
  if (e != cudaSuccess) return cuda_fail("status_read", e);
  return TSPLAT_OK;
}

int tsplat_adapter_forward(const TsplatAdapterProblem* p, float* means, float* covariances, float* harmonics,
                           float* scales, float* rotations, void* cuda_stream) {
  if (!p) { set_error("adapter_forward: null problem"); return TSPLAT_EINVAL; }
  int bad = 0;
  cudaError_t e = tsplat::launch_adapter_forward(*p, means, covariances, harmonics, scales, rotations,
                                                 static_cast<cudaStream_t>(cuda_stream), &bad);
  if (bad) { set_error("adapter_forward: bad shape (sh_degree 0..4, H, W > 0) or null pointer"); return TSPLAT_EINVAL; }
  if (e != cudaSuccess) return cuda_fail("adapter_forward", e);
  return TSPLAT_OK;
}

int tsplat_adapter_backward(const TsplatAdapterProblem* p, const float* dL_dmeans, const float* dL_dcovariances,
                            const float* dL_dharmonics, float* dL_dcoordinates, float* dL_ddepths, float* dL_draw,
                            void* cuda_stream) {
  if (!p) { set_error("adapter_backward: null problem"); return TSPLAT_EINVAL; }
  int bad = 0;
  cudaError_t e = tsplat::launch_adapter_backward(*p, dL_dmeans, dL_dcovariances, dL_dharmonics, dL_dcoordinates,
                                                  dL_ddepths, dL_draw, static_cast<cudaStream_t>(cuda_stream), &bad);
  if (bad) { set_error("adapter_backward: bad shape (sh_degree 0..4, H, W > 0) or null pointer"); return TSPLAT_EINVAL; }
  if (e != cudaSuccess) return cuda_fail("adapter_backward", e);
  return TSPLAT_OK;
}

}  // extern "C"
