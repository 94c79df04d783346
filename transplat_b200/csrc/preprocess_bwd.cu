// preprocess_bwd.cu — backward of the per-Gaussian preprocess (SURVEY 8a X10 + X11 fused):
// dL/dconic -> dL/dcov3D and dL/dmean (cov2D path), projection path, SH path (dL/dsh and the
// view-direction path into dL/dmean), fused depth-feature path, host-side scaling chain.
//
// One CTA = 32 Gaussians of one scene x all V views, one WARP per view (8 at most; more views in rounds), inputs
// staged with TMA bulk loads exactly like the forward kernel.  The views' gradients are SUMMED on chip - the reference
// writes one gradient set per view and lets autograd add the V copies (decoder_splatting_cuda.py:59-62):
//   phase A  thread (Gaussian, view): cov2D / projection / depth-feature backward in registers; the SH part leaves the
//            view's 25 basis values and 3 masked colour gradients in shared memory and accumulates dL/d(direction) on the
//            fly (no 25 x 3 derivative table in registers: round 1 needed 167 registers and ran at 14 % occupancy);
//   phase B  the CTA's 32 x 75 dL/dsh values: a warp per channel (and half of the coefficients), a lane per Gaussian: sum
//            over views of basis x colour gradient in registers (no shared-memory read-modify-write per view), staged
//            once and written to HBM with one bulk store per CTA;
//   phase C  the 10 remaining per-Gaussian sums (mean 3, covariance 6, opacity 1) over the views through shared memory.
// HBM-bound: reads 340 + V (48 + 5) B, writes 340 B per Gaussian.
#include "geom.cuh"

namespace tsplat {

constexpr int kBwdG = 32;         // Gaussians per CTA = lanes of a warp
constexpr int kBwdMaxWarps = 8;   // views handled concurrently by one CTA
constexpr int kScr = 29;          // scratch floats per (view, Gaussian): 25 basis values, 3 colour gradients, 1 pad (odd stride)
constexpr int kPart = 11;         // partial sums per (view, Gaussian): mean 3, cov 6, opacity 1, 1 pad (odd stride)

struct PreBwdArgs {
  TsplatProblem p;
  int cov_floats, sh_floats, bulk_ok, out_bulk_ok;
  int multimem;        // outputs are multicast addresses: add with multimem.red instead of storing (view sharding)
  int g_begin, g_end;  // Gaussian range of this launch (every scene); g_begin is a multiple of kPreBwdBlock
  float* dL_dmeans;
  float* dL_dcov;
  float* dL_dsh;
  float* dL_dopacity;
  float* dL_dmeans2D;
};

// multimem.red: one instruction adds the value(s) into EVERY GPU's replica of a multicast-mapped buffer; the NVSwitch
// performs the reduction (NVLS).  sm_90+ / PTX 8.1.
__device__ __forceinline__ void mc_red_add(float* mc_addr, float v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(mc_addr), "f"(v) : "memory");
}
__device__ __forceinline__ void mc_red_add_v4(float* mc_addr, float4 v) {
  asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// SH basis values (stored through `store(k, value)`) and dL/d(direction) += sum_k dbasis_k/d(direction) * w(k), where
// w(k) = sum_ch sh[k][ch] * dL/dcolour[ch] is produced on demand - same polynomials as geom.cuh: sh_basis.
template <class WF, class SF>
__device__ __forceinline__ int sh_basis_backward(int deg, int sh_l4, float x, float y, float z, WF w, SF store, float& gx,
                                                 float& gy, float& gz) {
  int n = 1;
  store(0, TS_SH_C0);
  if (deg > 0) {
    n = 4;
    store(1, -TS_SH_C1 * y); store(2, TS_SH_C1 * z); store(3, -TS_SH_C1 * x);
    gy += -TS_SH_C1 * w(1); gz += TS_SH_C1 * w(2); gx += -TS_SH_C1 * w(3);
    if (deg > 1) {
      n = 9;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      store(4, kC2[0] * xy);
      store(5, kC2[1] * yz);
      store(6, kC2[2] * (2.0f * zz - xx - yy));
      store(7, kC2[3] * xz);
      store(8, kC2[4] * (xx - yy));
      { const float t = kC2[0] * w(4); gx += t * y; gy += t * x; }
      { const float t = kC2[1] * w(5); gy += t * z; gz += t * y; }
      { const float t = kC2[2] * w(6); gx += t * 2.0f * -x; gy += t * 2.0f * -y; gz += t * 4.0f * z; }
      { const float t = kC2[3] * w(7); gx += t * z; gz += t * x; }
      { const float t = kC2[4] * w(8); gx += t * 2.0f * x; gy += t * 2.0f * -y; }
      if (deg > 2) {
        n = 16;
        store(9, kC3[0] * y * (3.0f * xx - yy));
        store(10, kC3[1] * xy * z);
        store(11, kC3[2] * y * (4.0f * zz - xx - yy));
        store(12, kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy));
        store(13, kC3[4] * x * (4.0f * zz - xx - yy));
        store(14, kC3[5] * z * (xx - yy));
        store(15, kC3[6] * x * (xx - 3.0f * yy));
        { const float t = kC3[0] * w(9); gx += t * 6.0f * xy; gy += t * 3.0f * (xx - yy); }
        { const float t = kC3[1] * w(10); gx += t * yz; gy += t * xz; gz += t * xy; }
        { const float t = kC3[2] * w(11); gx += t * -2.0f * xy; gy += t * (-3.0f * yy + 4.0f * zz - xx); gz += t * 8.0f * yz; }
        { const float t = kC3[3] * w(12); gx += t * -6.0f * xz; gy += t * -6.0f * yz; gz += t * 3.0f * (2.0f * zz - xx - yy); }
        { const float t = kC3[4] * w(13); gx += t * (-3.0f * xx + 4.0f * zz - yy); gy += t * -2.0f * xy; gz += t * 8.0f * xz; }
        { const float t = kC3[5] * w(14); gx += t * 2.0f * xz; gy += t * -2.0f * yz; gz += t * (xx - yy); }
        { const float t = kC3[6] * w(15); gx += t * 3.0f * (xx - yy); gy += t * -6.0f * xy; }
        if (deg > 3 && sh_l4) {
          n = 25;
          store(16, kC4[0] * xy * (xx - yy));
          store(17, kC4[1] * yz * (3.0f * xx - yy));
          store(18, kC4[2] * xy * (7.0f * zz - 1.0f));
          store(19, kC4[3] * yz * (7.0f * zz - 3.0f));
          store(20, kC4[4] * (zz * (35.0f * zz - 30.0f) + 3.0f));
          store(21, kC4[5] * xz * (7.0f * zz - 3.0f));
          store(22, kC4[6] * (xx - yy) * (7.0f * zz - 1.0f));
          store(23, kC4[7] * xz * (xx - 3.0f * yy));
          store(24, kC4[8] * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy)));
          { const float t = kC4[0] * w(16); gx += t * y * (3.0f * xx - yy); gy += t * x * (xx - 3.0f * yy); }
          { const float t = kC4[1] * w(17); gx += t * 6.0f * xy * z; gy += t * z * 3.0f * (xx - yy); gz += t * y * (3.0f * xx - yy); }
          { const float t = kC4[2] * w(18); gx += t * y * (7.0f * zz - 1.0f); gy += t * x * (7.0f * zz - 1.0f); gz += t * 14.0f * xy * z; }
          { const float t = kC4[3] * w(19); gy += t * z * (7.0f * zz - 3.0f); gz += t * y * (21.0f * zz - 3.0f); }
          { const float t = kC4[4] * w(20); gz += t * z * (140.0f * zz - 60.0f); }
          { const float t = kC4[5] * w(21); gx += t * z * (7.0f * zz - 3.0f); gz += t * x * (21.0f * zz - 3.0f); }
          { const float t = kC4[6] * w(22); gx += t * 2.0f * x * (7.0f * zz - 1.0f); gy += t * -2.0f * y * (7.0f * zz - 1.0f); gz += t * 14.0f * z * (xx - yy); }
          { const float t = kC4[7] * w(23); gx += t * z * 3.0f * (xx - yy); gy += t * -6.0f * xy * z; gz += t * x * (xx - 3.0f * yy); }
          { const float t = kC4[8] * w(24); gx += t * 4.0f * x * (xx - 3.0f * yy); gy += t * -4.0f * y * (3.0f * xx - yy); }
        }
      }
    }
  }
  return n;
}

// resident CTAs per SM (the register budget): 4-warp CTAs at C2: 4 / 5 / 6 -> 65.7 / 52.1 / 55.0 us (128 / 102 / 80
// registers per thread, the last with spills)
#ifndef TSPLAT_PREB_MIN_BLOCKS4
#define TSPLAT_PREB_MIN_BLOCKS4 5
#endif
#ifndef TSPLAT_PREB_MIN_BLOCKS8
#define TSPLAT_PREB_MIN_BLOCKS8 3  // 8-warp CTAs (more than 4 views per scene) at C5: 2 / 3 -> 2.38 / 2.01 ms
#endif
// NW warps per CTA (4 or 8: at least 4 even for fewer views, so that phase B's 2400 outputs are at most 19 per thread and
// stay in registers)
template <int NW, int LAYOUT>
__global__ void __launch_bounds__(kBwdG * NW, NW == 4 ? TSPLAT_PREB_MIN_BLOCKS4 : TSPLAT_PREB_MIN_BLOCKS8) preprocess_backward_kernel(PreBwdArgs a, Workspace ws) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  float* s_mean = reinterpret_cast<float*>(smem_raw + 16);
  float* s_cov = s_mean + 3 * kBwdG;
  float* s_sh = s_cov + a.cov_floats * kBwdG;
  const TsplatProblem& p = a.p;
  constexpr int nwarps = NW, nthreads = 32 * NW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float* s_scr = s_sh + a.sh_floats * kBwdG;          // [nwarps][32][kScr]
  float* s_part = s_scr + nwarps * kBwdG * kScr;      // [nwarps][32][kPart]

  const int b = blockIdx.y;
  const int g0 = a.g_begin + blockIdx.x * kBwdG;
  const int cnt = min(kBwdG, a.g_end - g0);
  const size_t gbase = (size_t)b * p.G + g0;
  const float* g_mean = p.means + gbase * 3;
  const float* g_cov = p.cov + gbase * a.cov_floats;
  const float* g_sh = p.sh + gbase * a.sh_floats;

  const bool bulk = a.bulk_ok && cnt == kBwdG;
  if (bulk) {
    if (tid == 0) {
      mbar_init(bar, 1);
      mbar_fence_init();
      mbar_expect_tx(bar, (uint32_t)(kBwdG * (12 + 4 * a.cov_floats + 4 * a.sh_floats)));
      bulk_g2s(s_mean, g_mean, kBwdG * 12, bar);
      bulk_g2s(s_cov, g_cov, kBwdG * 4 * a.cov_floats, bar);
      bulk_g2s(s_sh, g_sh, kBwdG * 4 * a.sh_floats, bar);
    }
    __syncthreads();
    mbar_wait(bar, 0);
  } else {
    for (int i = tid; i < cnt * 3; i += nthreads) s_mean[i] = g_mean[i];
    for (int i = tid; i < cnt * a.cov_floats; i += nthreads) s_cov[i] = g_cov[i];
    for (int i = tid; i < cnt * a.sh_floats; i += nthreads) s_sh[i] = g_sh[i];
    __syncthreads();
  }

  const int M = p.sh_coeffs;
  const bool active = lane < cnt;
  const int g = g0 + lane;
  float gmean[3] = {0.f, 0.f, 0.f}, gcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, gop = 0.0f;
  float mu0 = 0.f, mu1 = 0.f, mu2 = 0.f, craw[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (active) {
    mu0 = s_mean[3 * lane]; mu1 = s_mean[3 * lane + 1]; mu2 = s_mean[3 * lane + 2];
    if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
      for (int k = 0; k < 6; k++) craw[k] = s_cov[6 * lane + k];
    } else {
      const float* c = s_cov + 9 * lane;
      craw[0] = c[0]; craw[1] = c[1]; craw[2] = c[2]; craw[3] = c[4]; craw[4] = c[5]; craw[5] = c[8];
    }
  }
  const float* my_sh = s_sh + (size_t)lane * a.sh_floats;
  float* my_scr = s_scr + ((size_t)warp * kBwdG + lane) * kScr;

  // phase B's outputs: warp w sums channel w % 3 (and, with 8 warps, half w / 3 of the coefficients) for the CTA's 32
  // Gaussians, one per lane: the colour gradient is read once per view and every basis value feeds one FFMA; the
  // sums stay in registers over all rounds.  Warps beyond 3 (6) have no phase-B work.
  constexpr int kSplit = NW / 4;                        // 1 or 2
  constexpr int kKPer = (25 + kSplit - 1) / kSplit;     // 25 or 13 coefficients per thread
  constexpr int kShStride = LAYOUT == TSPLAT_SH_GM3 ? 3 : 1;
  const bool b_on = warp < 3 * kSplit;
  const int b_ch = warp % 3, b_k0 = (warp / 3) * kKPer;
  float dsh[kKPer];
#pragma unroll
  for (int i = 0; i < kKPer; i++) dsh[i] = 0.0f;
  // this lane's coefficients of the three channels (phase A): immediate offsets for every constant k
  const float* sh_c0 = my_sh;
  const float* sh_c1 = LAYOUT == TSPLAT_SH_GM3 ? my_sh + 1 : my_sh + M;
  const float* sh_c2 = LAYOUT == TSPLAT_SH_GM3 ? my_sh + 2 : my_sh + 2 * M;

  const int rounds = (p.V + nwarps - 1) / nwarps;
  for (int r = 0; r < rounds; r++) {
    const int v = r * nwarps + warp;  // warp-uniform
    int nb = 0;
    float dc0 = 0.f, dc1 = 0.f, dc2 = 0.f;
    if (v < p.V && active) {
      const int n = b * p.V + v;
      const size_t o = (size_t)n * p.G + g;
      // all of the view's inputs are requested at once (the accumulators of a culled Gaussian are never initialised
      // and never used): nothing else hides the L2 latency at the head of this long dependent chain
      const int radius = ws.radii[o];
      const unsigned char cl = ws.clamped[o];
      const float4* ga = reinterpret_cast<const float4*>(ws.grad_acc + o * kAccFloats);
      const float4 q0 = __ldcg(ga), q1 = __ldcg(ga + 1), q2 = __ldcg(ga + 2);
      if (radius > 0) {
        gop += q1.y;
        float dcol[4] = {q1.z, q1.w, q2.x, q2.y};

        const float* cam = reinterpret_cast<const float*>(p.cams + n);
        const float* vm = cam;
        const float* pj = cam + 16;
        const float tanx = __ldg(cam + 35), tany = __ldg(cam + 36);
        const float ms = __ldg(cam + 40), cs = __ldg(cam + 41);
        const float3 m = make_float3(mu0 * ms, mu1 * ms, mu2 * ms);
        float c6[6];
#pragma unroll
        for (int k = 0; k < 6; k++) c6[k] = craw[k] * cs;
        const float fx = (float)p.W / (2.0f * tanx), fy = (float)p.H / (2.0f * tany);

        // ---- cov2D backward (A.4) ----
        Cov2D q;
        compute_cov2d(m, fx, fy, tanx, tany, c6, vm, q);
        const float aa = q.a, bb = q.b, cc = q.c;
        const float denom = aa * cc - bb * bb;
        // render_backward leaves the moments S = sum over pixels of (dL/dG G) (dx, dy, dx^2, dx dy, dy^2) in slots 0..4
        // (render.cu): with the conic (A, B, C) of the forward (preprocess.cu: c/det, -b/det, a/det)
        //   dL/dmean2D = -(A S_x + B S_y) W/2, -(C S_y + B S_x) H/2      dL/dconic = -1/2 (S_xx, S_xy, S_yy)
        const float det_inv = 1.0f / denom;
        const float cA = cc * det_inv, cB = -bb * det_inv, cC = aa * det_inv;
        const float dmx = -(cA * q0.x + cB * q0.y) * (0.5f * (float)p.W);
        const float dmy = -(cC * q0.y + cB * q0.x) * (0.5f * (float)p.H);
        const float dcon[3] = {-0.5f * q0.z, -0.5f * q0.w, -0.5f * q1.x};
        if (a.dL_dmeans2D) {
          float* d2 = a.dL_dmeans2D + o * 3;
          d2[0] = dmx; d2[1] = dmy; d2[2] = 0.0f;
        }
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        const float(*T)[3] = q.T;
        if (denom2inv != 0.0f) {
          dL_da = denom2inv * (-cc * cc * dcon[0] + 2 * bb * cc * dcon[1] + (denom - aa * cc) * dcon[2]);
          dL_dc = denom2inv * (-aa * aa * dcon[2] + 2 * aa * bb * dcon[1] + (denom - aa * cc) * dcon[0]);
          dL_db = denom2inv * 2 * (bb * cc * dcon[0] - (denom + 2 * bb * bb) * dcon[1] + aa * bb * dcon[2]);
          dcov[0] = T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc;
          dcov[3] = T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc;
          dcov[5] = T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc;
          dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
          dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
          dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
        }
        const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        float dT[2][3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const float t0 = T[0][0] * S[k][0] + T[0][1] * S[k][1] + T[0][2] * S[k][2];
          const float t1 = T[1][0] * S[k][0] + T[1][1] * S[k][1] + T[1][2] * S[k][2];
          dT[0][k] = 2 * t0 * dL_da + t1 * dL_db;
          dT[1][k] = 2 * t1 * dL_dc + t0 * dL_db;
        }
        const float dJ00 = vm[0] * dT[0][0] + vm[4] * dT[0][1] + vm[8] * dT[0][2];
        const float dJ02 = vm[2] * dT[0][0] + vm[6] * dT[0][1] + vm[10] * dT[0][2];
        const float dJ11 = vm[1] * dT[1][0] + vm[5] * dT[1][1] + vm[9] * dT[1][2];
        const float dJ12 = vm[2] * dT[1][0] + vm[6] * dT[1][1] + vm[10] * dT[1][2];
        const float tz = 1.0f / q.t.z, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = q.xmul * -fx * tz2 * dJ02;
        const float dty = q.ymul * -fy * tz2 * dJ12;
        const float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * q.t.x) * tz3 * dJ02 + (2 * fy * q.t.y) * tz3 * dJ12;
        float dmean[3];
        dmean[0] = vm[0] * dtx + vm[1] * dty + vm[2] * dtz;
        dmean[1] = vm[4] * dtx + vm[5] * dty + vm[6] * dtz;
        dmean[2] = vm[8] * dtx + vm[9] * dty + vm[10] * dtz;

        // ---- projection backward ----
        const float4 mh = transform4x4(m, pj);
        const float m_w = 1.0f / (mh.w + 0.0000001f);
        const float mul1 = (pj[0] * m.x + pj[4] * m.y + pj[8] * m.z + pj[12]) * m_w * m_w;
        const float mul2 = (pj[1] * m.x + pj[5] * m.y + pj[9] * m.z + pj[13]) * m_w * m_w;
        dmean[0] += (pj[0] * m_w - pj[3] * mul1) * dmx + (pj[1] * m_w - pj[3] * mul2) * dmy;
        dmean[1] += (pj[4] * m_w - pj[7] * mul1) * dmx + (pj[5] * m_w - pj[7] * mul2) * dmy;
        dmean[2] += (pj[8] * m_w - pj[11] * mul1) * dmx + (pj[9] * m_w - pj[11] * mul2) * dmy;

        // ---- SH backward (A.5) ----
        if (M > 0) {
#pragma unroll
          for (int ch = 0; ch < 3; ch++)
            if (cl & (1u << ch)) dcol[ch] = 0.0f;
          dc0 = dcol[0]; dc1 = dcol[1]; dc2 = dcol[2];
          const float ox = m.x - __ldg(cam + 32), oy = m.y - __ldg(cam + 33), oz = m.z - __ldg(cam + 34);
          const float sum2 = ox * ox + oy * oy + oz * oz;
          const float len = sqrtf(sum2);
          const float dxn = ox / len, dyn = oy / len, dzn = oz / len;
          float ddx = 0.f, ddy = 0.f, ddz = 0.f;
          nb = sh_basis_backward(
              p.sh_degree, p.sh_l4, dxn, dyn, dzn,
              [&](int k) { return sh_c0[k * kShStride] * dc0 + sh_c1[k * kShStride] * dc1 + sh_c2[k * kShStride] * dc2; },
              [&](int k, float val) { my_scr[k] = val; }, ddx, ddy, ddz);
          const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
          dmean[0] += ((sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv32;
          dmean[1] += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * inv32;
          dmean[2] += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * inv32;
        } else {
          dc0 = dcol[0]; dc1 = dcol[1]; dc2 = dcol[2];  // colors_precomp: the gradient passes through
        }
        gmean[0] += dmean[0] * ms; gmean[1] += dmean[1] * ms; gmean[2] += dmean[2] * ms;
#pragma unroll
        for (int k = 0; k < 6; k++) gcov[k] += dcov[k] * cs;

        // ---- fused depth feature (A.7) ----
        if (p.depth_mode != TSPLAT_DEPTH_NONE && !(cl & 8u)) {
          const float r0 = __ldg(cam + 42), r1 = __ldg(cam + 43), r2 = __ldg(cam + 44);
          const float z = ((r0 * mu0 + r1 * mu1) + r2 * mu2) + __ldg(cam + 45);
          float dfdz;
          (void)depth_feature(p.depth_mode, z, __ldg(cam + 46), __ldg(cam + 47), dfdz);
          const float dz = TS_SH_C0 * dcol[3] * dfdz;
          gmean[0] += dz * r0; gmean[1] += dz * r1; gmean[2] += dz * r2;
        }
      } else if (a.dL_dmeans2D) {
        float* d2 = a.dL_dmeans2D + o * 3;
        d2[0] = 0.0f; d2[1] = 0.0f; d2[2] = 0.0f;
      }
    }
    // basis values this view did not evaluate (culled Gaussian, lower degree, idle warp of the last round) are zero
    for (int k = nb; k < 25; k++) my_scr[k] = 0.0f;
    my_scr[25] = dc0; my_scr[26] = dc1; my_scr[27] = dc2;
    __syncthreads();
    // ---- phase B: dL/dsh[g][k][ch] += sum over this round's views of basis[k] * dcolour[ch] ----
    if (b_on) {
      const float* sc = s_scr + (size_t)lane * kScr;
#pragma unroll
      for (int w = 0; w < nwarps; w++, sc += kBwdG * kScr) {
        const float dc = sc[25 + b_ch];
        if (M == 0) {
          dsh[0] += dc;
        } else {
#pragma unroll
          for (int i = 0; i < kKPer; i++)
            if (b_k0 + i < 25) dsh[i] += sc[b_k0 + i] * dc;
        }
      }
    }
    __syncthreads();  // the scratch is rewritten by the next round
  }
  // ---- dL/dsh leaves through the (now idle) staging area of the SH inputs: one bulk store per CTA ----
  float* g_dsh = a.dL_dsh + gbase * a.sh_floats;
  if (b_on) {
    float* so = s_sh + (size_t)lane * a.sh_floats;
    if (M == 0) {
      if (b_k0 == 0) so[b_ch] = dsh[0];
    } else {
#pragma unroll
      for (int i = 0; i < kKPer; i++) {
        const int k = b_k0 + i;
        if (k < M && k < 25) so[LAYOUT == TSPLAT_SH_GM3 ? k * 3 + b_ch : b_ch * M + k] = dsh[i];
      }
    }
  }
  if (a.multimem) {
    // view sharding: this rank's sums go straight into every rank's gradient slab, reduced by the switch; Gaussians no
    // local view sees send nothing
    __syncthreads();
    const int nfl = cnt * a.sh_floats;
    if ((reinterpret_cast<uintptr_t>(g_dsh) & 15u) == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(s_sh);
      for (int i = tid; i < nfl / 4; i += nthreads) {
        const float4 v = s4[i];
        if (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f || v.w != 0.0f) mc_red_add_v4(g_dsh + 4 * i, v);
      }
      for (int i = (nfl & ~3) + tid; i < nfl; i += nthreads)
        if (s_sh[i] != 0.0f) mc_red_add(g_dsh + i, s_sh[i]);
    } else {
      for (int i = tid; i < nfl; i += nthreads)
        if (s_sh[i] != 0.0f) mc_red_add(g_dsh + i, s_sh[i]);
    }
  } else if (bulk && a.out_bulk_ok) {
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(g_dsh, s_sh, (uint32_t)(kBwdG * 4 * a.sh_floats));
      bulk_commit();
      bulk_wait_all();
    }
  } else {
    __syncthreads();
    for (int i = tid; i < cnt * a.sh_floats; i += nthreads) g_dsh[i] = s_sh[i];
  }
  // ---- phase C: mean / covariance / opacity gradients summed over the warps (views) ----
  {
    float* mp = s_part + ((size_t)warp * kBwdG + lane) * kPart;
    mp[0] = gmean[0]; mp[1] = gmean[1]; mp[2] = gmean[2];
#pragma unroll
    for (int k = 0; k < 6; k++) mp[3 + k] = gcov[k];
    mp[9] = gop;
  }
  __syncthreads();
  if (warp == 0 && active) {
    float sum[10];
#pragma unroll
    for (int k = 0; k < 10; k++) sum[k] = 0.0f;
    for (int w = 0; w < nwarps; w++) {
      const float* mp = s_part + ((size_t)w * kBwdG + lane) * kPart;
#pragma unroll
      for (int k = 0; k < 10; k++) sum[k] += mp[k];
    }
    const size_t og = (size_t)b * p.G + g;
    if (a.multimem) {
      auto add = [](float* q, float v) { if (v != 0.0f) mc_red_add(q, v); };
      add(a.dL_dmeans + og * 3 + 0, sum[0]); add(a.dL_dmeans + og * 3 + 1, sum[1]); add(a.dL_dmeans + og * 3 + 2, sum[2]);
      add(a.dL_dopacity + og, sum[9]);
      if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
        for (int k = 0; k < 6; k++) add(a.dL_dcov + og * 6 + k, sum[3 + k]);
      } else {
        float* c = a.dL_dcov + og * 9;
        add(c + 0, sum[3]); add(c + 1, sum[4]); add(c + 2, sum[5]); add(c + 4, sum[6]); add(c + 5, sum[7]); add(c + 8, sum[8]);
      }
    } else {
    a.dL_dmeans[og * 3 + 0] = sum[0]; a.dL_dmeans[og * 3 + 1] = sum[1]; a.dL_dmeans[og * 3 + 2] = sum[2];
    a.dL_dopacity[og] = sum[9];
    if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
      for (int k = 0; k < 6; k++) a.dL_dcov[og * 6 + k] = sum[3 + k];
    } else {
      float* c = a.dL_dcov + og * 9;
      c[0] = sum[3]; c[1] = sum[4]; c[2] = sum[5]; c[3] = 0.f; c[4] = sum[6]; c[5] = sum[7];
      c[6] = 0.f; c[7] = 0.f; c[8] = sum[8];
    }
    }
  }
  if (a.multimem) __threadfence_system();
}

cudaError_t launch_preprocess_backward(const TsplatProblem& p, const Workspace& ws, float* dL_dmeans, float* dL_dcov,
                                       float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, int g_begin, int g_count,
                                       bool multimem, cudaStream_t st) {
  if (g_count <= 0) return cudaSuccess;
  PreBwdArgs a;
  a.p = p;
  a.g_begin = g_begin; a.g_end = g_begin + g_count;
  a.cov_floats = p.cov_layout == TSPLAT_COV_TRIU6 ? 6 : 9;
  a.sh_floats = p.sh_coeffs > 0 ? 3 * p.sh_coeffs : 3;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  const bool stride_ok = p.B == 1 || (p.G % 4 == 0);
  a.bulk_ok = al16(p.means) && al16(p.cov) && al16(p.sh) && stride_ok && ((kBwdG * 4 * a.sh_floats) % 16 == 0);
  a.out_bulk_ok = al16(dL_dsh) && stride_ok;
  a.multimem = multimem ? 1 : 0;
  a.dL_dmeans = dL_dmeans; a.dL_dcov = dL_dcov; a.dL_dsh = dL_dsh; a.dL_dopacity = dL_dopacity;
  a.dL_dmeans2D = dL_dmeans2D;
  const int warps = p.V <= 4 ? 4 : kBwdMaxWarps;
  const size_t smem = 16 + sizeof(float) * kBwdG * (3 + a.cov_floats + a.sh_floats + warps * (kScr + kPart));
  dim3 grid((g_count + kBwdG - 1) / kBwdG, p.B);
  auto go = [&](auto kernel, int nw) -> cudaError_t {
    cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(kernel), (int)smem);
    if (e != cudaSuccess) return e;
    kernel<<<grid, 32 * nw, smem, st>>>(a, ws);
    return cudaSuccess;
  };
  cudaError_t e;
  if (p.sh_layout == TSPLAT_SH_GM3)
    e = warps == 4 ? go(preprocess_backward_kernel<4, TSPLAT_SH_GM3>, 4) : go(preprocess_backward_kernel<8, TSPLAT_SH_GM3>, 8);
  else
    e = warps == 4 ? go(preprocess_backward_kernel<4, TSPLAT_SH_G3M>, 4) : go(preprocess_backward_kernel<8, TSPLAT_SH_G3M>, 8);
  if (e != cudaSuccess) return e;
  return cudaGetLastError();
}

}  // namespace tsplat
