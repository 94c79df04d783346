// preprocess.cu — per-Gaussian forward preprocess for all V views of a scene in one
// launch, plus markVisible.  COMPILED WITH -fmad=false (see geom.cuh: the results feed the sort
// keys, radii and tile rectangles, which must be bit-exact against the oracle).
//
// B200 design: one CTA = 32 Gaussians of one scene x all V views, one WARP per view (8 warps at most; more views are
// walked in rounds).  The CTA's slabs of the AoS inputs (means 12 B, covariance 24/36 B, SH 300 B, opacity 4 B per
// Gaussian - contiguous per CTA) are pulled into shared memory once with 1-D TMA bulk copies (cp.async.bulk, SASS
// UBLKCP) that complete on an mbarrier; lane l of every warp then reads Gaussian l from shared memory (odd word
// strides: conflict-free) and writes its view's SoA state coalesced.  SH is read from HBM once per scene instead of once
// per view (the reference re-reads and re-copies it per view: decoder_splatting_cuda.py:59-62, cuda_splatting.py:84).
// Round 1 walked the V views serially in one thread (128 Gaussians per CTA, 61 KB of shared memory, 12 warps per SM,
// 1.06 TB/s of DRAM traffic): 11 KB per CTA and a thread per (Gaussian, view) keep 4x as many loads in flight.
// HBM-bound: 340 B read once + 61 B written per (view, Gaussian).
#include "cull.cuh"
#include "geom.cuh"

namespace tsplat {

constexpr int kPreG = 32;          // Gaussians per CTA = lanes of a warp
constexpr int kPreMaxWarps = 8;    // views handled concurrently by one CTA
constexpr int kAggMaxArea = 16;    // tile rectangles up to this area are counted warp-aggregated

struct PreArgs {
  TsplatProblem p;
  int cov_floats;  // 6 or 9
  int sh_floats;   // 3*M, or 3 for colors_precomp
  int bulk_ok;     // all slabs 16-B aligned
  uint4* zero_ptr;              // control / look-back / ranges region to clear (replaces a memset launch)
  unsigned long long zero_vec;  // its size in 16-B words
};

// Pair counts per (view, tile) for the bucket path (blocklist / bucket.cu need them before any pair is emitted).  The 32
// Gaussians of a warp are neighbours in a context image, so their tile rectangles overlap: the warp walks the union of
// the small rectangles and adds ONE count per tile (ballot + popc) instead of one reduction per (Gaussian, tile) -
// same-address global reductions serialise in L2.  Big rectangles are counted by their own lane.
__device__ __forceinline__ void count_tiles(unsigned int* __restrict__ cnt, int gx, bool touched, int x0, int y0, int x1,
                                            int y1, int lane) {
  const int area = (x1 - x0) * (y1 - y0);
  const bool small = touched && area <= kAggMaxArea;
  if (touched && !small) {
    for (int y = y0; y < y1; y++)
      for (int x = x0; x < x1; x++) atomicAdd(cnt + y * gx + x, 1u);
  }
  if (__any_sync(0xffffffffu, small)) {
    const int ux0 = __reduce_min_sync(0xffffffffu, small ? x0 : 0x7fffffff), uy0 = __reduce_min_sync(0xffffffffu, small ? y0 : 0x7fffffff);
    const int ux1 = __reduce_max_sync(0xffffffffu, small ? x1 : 0), uy1 = __reduce_max_sync(0xffffffffu, small ? y1 : 0);
    if ((ux1 - ux0) * (uy1 - uy0) <= 64) {
      for (int y = uy0; y < uy1; y++)
        for (int x = ux0; x < ux1; x++) {
          const unsigned int m = __ballot_sync(0xffffffffu, small && x >= x0 && x < x1 && y >= y0 && y < y1);
          if (lane == 0 && m) atomicAdd(cnt + y * gx + x, (unsigned int)__popc(m));
        }
    } else if (small) {  // Gaussians of this warp are scattered over the image (depth discontinuity): count one by one
      for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) atomicAdd(cnt + y * gx + x, 1u);
    }
  }
}

#ifndef TSPLAT_PRE_MIN_BLOCKS
#define TSPLAT_PRE_MIN_BLOCKS 4  // 64 registers: 8 CTAs of 4 warps per SM (3 -> 80 registers, 39.6 us; 4 -> 37.3 us at C2)
#endif
template <int LAYOUT>
__global__ void __launch_bounds__(kPreG * kPreMaxWarps, TSPLAT_PRE_MIN_BLOCKS) preprocess_forward_kernel(PreArgs a, Workspace ws, int* out_radii) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  float* s_op = reinterpret_cast<float*>(smem_raw + 16);
  float* s_mean = s_op + kPreG;
  float* s_cov = s_mean + 3 * kPreG;
  float* s_sh = s_cov + a.cov_floats * kPreG;

  const TsplatProblem& p = a.p;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x, nwarps = nthreads >> 5;
  const int b = blockIdx.y;
  const int g0 = blockIdx.x * kPreG;
  const int cnt = min(kPreG, p.G - g0);
  const size_t gbase = (size_t)b * p.G + g0;
  const float* g_mean = p.means + gbase * 3;
  const float* g_cov = p.cov + gbase * a.cov_floats;
  const float* g_sh = p.sh + gbase * a.sh_floats;
  const float* g_op = p.opacity + gbase;

  const bool bulk = a.bulk_ok && cnt == kPreG;
  if (bulk) {
    if (tid == 0) {
      mbar_init(bar, 1);
      mbar_fence_init();
      const uint32_t bytes = (uint32_t)(kPreG * (4 + 12 + 4 * a.cov_floats + 4 * a.sh_floats));
      mbar_expect_tx(bar, bytes);
      bulk_g2s(s_op, g_op, kPreG * 4, bar);
      bulk_g2s(s_mean, g_mean, kPreG * 12, bar);
      bulk_g2s(s_cov, g_cov, kPreG * 4 * a.cov_floats, bar);
      bulk_g2s(s_sh, g_sh, kPreG * 4 * a.sh_floats, bar);
    }
  }
  {  // grid-stride clear of the region bin_emit / sort / ranges expect to be zero (while the slabs are in flight)
    const unsigned long long nthr = (unsigned long long)gridDim.x * gridDim.y * nthreads;
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (unsigned long long i = ((unsigned long long)blockIdx.y * gridDim.x + blockIdx.x) * nthreads + tid; i < a.zero_vec;
         i += nthr)
      a.zero_ptr[i] = z;
  }
  if (bulk) {
    __syncthreads();  // barrier init visible before anyone waits
    mbar_wait(bar, 0);
  } else {
    for (int i = tid; i < cnt; i += nthreads) s_op[i] = g_op[i];
    for (int i = tid; i < cnt * 3; i += nthreads) s_mean[i] = g_mean[i];
    for (int i = tid; i < cnt * a.cov_floats; i += nthreads) s_cov[i] = g_cov[i];
    for (int i = tid; i < cnt * a.sh_floats; i += nthreads) s_sh[i] = g_sh[i];
    __syncthreads();
  }
  const bool active = lane < cnt;
  const int g = g0 + lane;
  const float mu0 = s_mean[3 * lane], mu1 = s_mean[3 * lane + 1], mu2 = s_mean[3 * lane + 2];
  float craw[6];
  if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
    for (int k = 0; k < 6; k++) craw[k] = s_cov[6 * lane + k];
  } else {
    const float* c = s_cov + 9 * lane;
    craw[0] = c[0]; craw[1] = c[1]; craw[2] = c[2]; craw[3] = c[4]; craw[4] = c[5]; craw[5] = c[8];
  }
  const float opac = s_op[lane];
  const float* my_sh = s_sh + (size_t)lane * a.sh_floats;
  const int M = p.sh_coeffs;
  const int gx = ws.tiles_x, gy = ws.tiles_y;

  for (int v = warp; v < p.V; v += nwarps) {  // warp-uniform
    const int n = b * p.V + v;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    bool counted = false;
    if (active) {
    const float* cam = reinterpret_cast<const float*>(p.cams + n);
    const float* vm = cam;        // view[16]
    const float* pj = cam + 16;   // proj[16]
    const float tanx = __ldg(cam + 35), tany = __ldg(cam + 36);
    const float ms = __ldg(cam + 40), cs = __ldg(cam + 41);
    const size_t o = (size_t)n * p.G + g;

    float depth = 0.0f, px = 0.0f, py = 0.0f;
    float4 con_o = make_float4(0.f, 0.f, 0.f, 0.f), rgbd = make_float4(0.f, 0.f, 0.f, 0.f);
    int radius = 0;
    unsigned int touched = 0;
    unsigned char cl = 0;

    const float3 m = make_float3(mu0 * ms, mu1 * ms, mu2 * ms);
    const float3 pv = transform4x3(m, vm);
    if (pv.z > 0.2f) {  // in_frustum
      const float4 ph = transform4x4(m, pj);
      const float pw = 1.0f / (ph.w + 0.0000001f);
      const float ppx = ph.x * pw, ppy = ph.y * pw;
      float c6[6];
#pragma unroll
      for (int k = 0; k < 6; k++) c6[k] = craw[k] * cs;
      const float fx = (float)p.W / (2.0f * tanx), fy = (float)p.H / (2.0f * tany);
      Cov2D q;
      compute_cov2d(m, fx, fy, tanx, tany, c6, vm, q);
      const float det = q.a * q.c - q.b * q.b;
      if (det != 0.0f) {
        const float det_inv = 1.0f / det;
        const float mid = 0.5f * (q.a + q.c);
        const float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
        const float l1 = mid + sq, l2 = mid - sq;
        const int rad = (int)ceilf(3.0f * sqrtf(fmaxf(l1, l2)));
        const float fpx = ndc2pix(ppx, p.W), fpy = ndc2pix(ppy, p.H);
        get_rect(fpx, fpy, rad, gx, gy, x0, y0, x1, y1);
        const int area = (x1 - x0) * (y1 - y0);
        if (area != 0) {
          depth = pv.z; radius = rad; px = fpx; py = fpy; touched = (unsigned int)area;
          counted = true;
          con_o = make_float4(q.c * det_inv, -q.b * det_inv, q.a * det_inv, opac);
          if (M > 0) {  // computeColorFromSH (A.5)
            float dx = m.x - __ldg(cam + 32), dy = m.y - __ldg(cam + 33), dz = m.z - __ldg(cam + 34);
            const float len = sqrtf(dx * dx + dy * dy + dz * dz);
            dx /= len; dy /= len; dz /= len;
            float bs[25];
            const int nb = sh_basis<false>(p.sh_degree, p.sh_l4, dx, dy, dz, bs, nullptr);
            float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 25; k++) {
              if (k < nb) {
#pragma unroll
                for (int ch = 0; ch < 3; ch++) acc[ch] += bs[k] * my_sh[sh_idx(LAYOUT, M, k, ch)];
              }
            }
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
              acc[ch] += 0.5f;
              if (acc[ch] < 0.0f) cl |= (unsigned char)(1u << ch);
              acc[ch] = fmaxf(acc[ch], 0.0f);
            }
            rgbd.x = acc[0]; rgbd.y = acc[1]; rgbd.z = acc[2];
          } else {
            rgbd.x = my_sh[0]; rgbd.y = my_sh[1]; rgbd.z = my_sh[2];
          }
          if (p.depth_mode != TSPLAT_DEPTH_NONE) {  // fused depth feature (A.7)
            const float z = ((__ldg(cam + 42) * mu0 + __ldg(cam + 43) * mu1) + __ldg(cam + 44) * mu2) + __ldg(cam + 45);
            float dfdz;
            const float f = depth_feature(p.depth_mode, z, __ldg(cam + 46), __ldg(cam + 47), dfdz);
            float fd = TS_SH_C0 * f + 0.5f;
            if (fd < 0.0f) cl |= 8u;
            rgbd.w = fmaxf(fd, 0.0f);
          }
        }
      }
    }
    ws.depths[o] = depth;
    ws.radii[o] = radius;
    {
      // the per-Gaussian record the render warps gather (render.cu); its last word is the cull threshold of cull.cuh
      float4* gr = ws.grec + o * 3;
      const float tau = radius > 0 ? cull_tau(con_o.x, con_o.y, con_o.z, con_o.w) : 0.0f;
      gr[0] = make_float4(px, py, con_o.x, con_o.y);
      gr[1] = make_float4(con_o.z, con_o.w, rgbd.x, rgbd.y);
      gr[2] = make_float4(rgbd.z, rgbd.w, __uint_as_float((unsigned int)g), tau);
    }
    ws.tiles_touched[o] = touched;
    ws.clamped[o] = cl;
    if (out_radii) out_radii[o] = radius;
    if (radius > 0) {  // render_backward accumulates into these; preprocess_backward reads them iff radius > 0
      float4* ga = reinterpret_cast<float4*>(ws.grad_acc + o * kAccFloats);
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      ga[0] = z; ga[1] = z; ga[2] = z;
    }
    }  // active
    // bucket path: count the pairs of every (view, tile) while the rectangles are at hand
    if (ws.bucket) count_tiles(ws.tile_count + (size_t)n * ws.T, gx, counted, x0, y0, x1, y1, lane);
  }
}

__global__ void mark_visible_kernel(int G, const float* __restrict__ means, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G) return;
  const float3 m = make_float3(means[3 * i], means[3 * i + 1], means[3 * i + 2]);
  present[i] = transform4x3(m, view).z > 0.2f;
}

cudaError_t launch_preprocess(const TsplatProblem& p, const Workspace& ws, int* out_radii, void* zero_ptr,
                              uint64_t zero_bytes, cudaStream_t st) {
  PreArgs a;
  a.p = p;
  a.zero_ptr = static_cast<uint4*>(zero_ptr);
  a.zero_vec = zero_bytes / 16;  // the region is carved in 256-B units
  a.cov_floats = p.cov_layout == TSPLAT_COV_TRIU6 ? 6 : 9;
  a.sh_floats = p.sh_coeffs > 0 ? 3 * p.sh_coeffs : 3;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  // per-scene strides must keep every CTA slab 16-B aligned: G*4, G*12, ... multiples of 16
  const bool stride_ok = p.B == 1 || (p.G % 4 == 0);
  a.bulk_ok = al16(p.means) && al16(p.cov) && al16(p.sh) && al16(p.opacity) && stride_ok &&
              ((kPreG * 4 * a.sh_floats) % 16 == 0);
  const size_t smem = 16 + sizeof(float) * kPreG * (1 + 3 + a.cov_floats + a.sh_floats);
  dim3 grid((p.G + kPreG - 1) / kPreG, p.B);
  const int warps = p.V < kPreMaxWarps ? p.V : kPreMaxWarps;
  auto go = [&](auto kernel) -> cudaError_t {
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(kernel), (int)smem);
    if (e != cudaSuccess) return e;
    kernel<<<grid, 32 * warps, smem, st>>>(a, ws, out_radii);
    return cudaGetLastError();
  };
  return p.sh_layout == TSPLAT_SH_GM3 ? go(preprocess_forward_kernel<TSPLAT_SH_GM3>) : go(preprocess_forward_kernel<TSPLAT_SH_G3M>);
}

cudaError_t launch_mark_visible(int G, const float* means, const float* view, uint8_t* present, cudaStream_t st) {
  if (G <= 0) return cudaSuccess;
  mark_visible_kernel<<<(G + 255) / 256, 256, 0, st>>>(G, means, view, present);
  return cudaGetLastError();
}

}  // namespace tsplat
