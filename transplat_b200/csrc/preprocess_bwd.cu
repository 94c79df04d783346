// preprocess_bwd.cu — backward of the per-Gaussian preprocess (SURVEY 8a X10 + X11 fused):
// dL/dconic -> dL/dcov3D and dL/dmean (cov2D path), projection path, SH path (dL/dsh and the
// view-direction path into dL/dmean), fused depth-feature path, host-side scaling chain.
//
// One CTA = 128 Gaussians of one scene, looping over the scene's V views and SUMMING the views'
// gradients on chip: dL/dsh (300 B/Gaussian, the dominant write) is accumulated in shared memory
// and written once per scene, where the reference writes it once per view and lets autograd add
// the V copies (decoder_splatting_cuda.py:59-62).  Inputs are staged with TMA bulk loads exactly
// like the forward kernel.  HBM-bound: reads 340 + V*(48+5) B, writes 340 B per Gaussian.
#include "geom.cuh"

namespace tsplat {

struct PreBwdArgs {
  TsplatProblem p;
  int cov_floats, sh_floats, bulk_ok;
  int g_begin, g_end;  // Gaussian range of this launch (every scene); g_begin is a multiple of kPreBwdBlock
  float* dL_dmeans;
  float* dL_dcov;
  float* dL_dsh;
  float* dL_dopacity;
  float* dL_dmeans2D;
};

__global__ void __launch_bounds__(kPreBwdBlock) preprocess_backward_kernel(PreBwdArgs a, Workspace ws) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  float* s_mean = reinterpret_cast<float*>(smem_raw + 16);
  float* s_cov = s_mean + 3 * kPreBwdBlock;
  float* s_sh = s_cov + a.cov_floats * kPreBwdBlock;
  float* s_dsh = s_sh + a.sh_floats * kPreBwdBlock;

  const TsplatProblem& p = a.p;
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int g0 = a.g_begin + blockIdx.x * kPreBwdBlock;
  const int cnt = min(kPreBwdBlock, a.g_end - g0);
  const size_t gbase = (size_t)b * p.G + g0;
  const float* g_mean = p.means + gbase * 3;
  const float* g_cov = p.cov + gbase * a.cov_floats;
  const float* g_sh = p.sh + gbase * a.sh_floats;

  const bool bulk = a.bulk_ok && cnt == kPreBwdBlock;
  if (bulk) {
    if (tid == 0) {
      mbar_init(bar, 1);
      mbar_fence_init();
      mbar_expect_tx(bar, (uint32_t)(kPreBwdBlock * (12 + 4 * a.cov_floats + 4 * a.sh_floats)));
      bulk_g2s(s_mean, g_mean, kPreBwdBlock * 12, bar);
      bulk_g2s(s_cov, g_cov, kPreBwdBlock * 4 * a.cov_floats, bar);
      bulk_g2s(s_sh, g_sh, kPreBwdBlock * 4 * a.sh_floats, bar);
    }
  } else {
    for (int i = tid; i < cnt * 3; i += kPreBwdBlock) s_mean[i] = g_mean[i];
    for (int i = tid; i < cnt * a.cov_floats; i += kPreBwdBlock) s_cov[i] = g_cov[i];
    for (int i = tid; i < cnt * a.sh_floats; i += kPreBwdBlock) s_sh[i] = g_sh[i];
  }
  for (int i = tid; i < kPreBwdBlock * a.sh_floats; i += kPreBwdBlock) s_dsh[i] = 0.0f;
  __syncthreads();
  if (bulk) mbar_wait(bar, 0);

  const int M = p.sh_coeffs;
  float gmean[3] = {0.f, 0.f, 0.f}, gcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, gop = 0.0f;
  if (tid < cnt) {
    const int g = g0 + tid;
    const float mu0 = s_mean[3 * tid], mu1 = s_mean[3 * tid + 1], mu2 = s_mean[3 * tid + 2];
    float craw[6];
    if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
      for (int k = 0; k < 6; k++) craw[k] = s_cov[6 * tid + k];
    } else {
      const float* c = s_cov + 9 * tid;
      craw[0] = c[0]; craw[1] = c[1]; craw[2] = c[2]; craw[3] = c[4]; craw[4] = c[5]; craw[5] = c[8];
    }
    const float* my_sh = s_sh + (size_t)tid * a.sh_floats;
    float* my_dsh = s_dsh + (size_t)tid * a.sh_floats;

    // per-view inputs are fetched one view AHEAD: the loop body is ~1500 dependent instructions per view and only
    // 8 warps fit on an SM (648 B of shared memory per Gaussian), so nothing else hides the L2/HBM latency
    struct ViewIn { float4 q0, q1, q2; int radius; unsigned char cl; };
    auto fetch = [&](int v) {
      ViewIn in;
      const size_t o = (size_t)(b * p.V + v) * p.G + g;
      in.radius = ws.radii[o];
      in.cl = ws.clamped[o];
      const float4* ga = reinterpret_cast<const float4*>(ws.grad_acc + o * kAccFloats);
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      in.q0 = in.radius > 0 ? ga[0] : z;  // accumulators of culled Gaussians are never initialised
      in.q1 = in.radius > 0 ? ga[1] : z;
      in.q2 = in.radius > 0 ? ga[2] : z;
      return in;
    };
    ViewIn nxt = fetch(0);
    for (int v = 0; v < p.V; v++) {
      const int n = b * p.V + v;
      const size_t o = (size_t)n * p.G + g;
      const ViewIn cur = nxt;
      if (v + 1 < p.V) nxt = fetch(v + 1);
      const bool vis = cur.radius > 0;
      if (!vis) {
        if (a.dL_dmeans2D) {
          float* d2 = a.dL_dmeans2D + o * 3;
          d2[0] = 0.0f; d2[1] = 0.0f; d2[2] = 0.0f;
        }
        continue;
      }
      const float4 q0 = cur.q0, q1 = cur.q1, q2 = cur.q2;
      gop += q1.y;
      float dcol[4] = {q1.z, q1.w, q2.x, q2.y};
      const unsigned char cl = cur.cl;

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
        const float ox = m.x - __ldg(cam + 32), oy = m.y - __ldg(cam + 33), oz = m.z - __ldg(cam + 34);
        const float sum2 = ox * ox + oy * oy + oz * oz;
        const float len = sqrtf(sum2);
        const float dxn = ox / len, dyn = oy / len, dzn = oz / len;
        float bs[25], db[25][3];
        const int nb = sh_basis<true>(p.sh_degree, p.sh_l4, dxn, dyn, dzn, bs, db);
        float ddir[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 25; k++) {
          if (k < nb) {
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
              const int si = sh_idx(p.sh_layout, M, k, ch);
              my_dsh[si] += bs[k] * dcol[ch];
              const float w = my_sh[si] * dcol[ch];
              ddir[0] += db[k][0] * w; ddir[1] += db[k][1] * w; ddir[2] += db[k][2] * w;
            }
          }
        }
        const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmean[0] += ((sum2 - ox * ox) * ddir[0] - oy * ox * ddir[1] - oz * ox * ddir[2]) * inv32;
        dmean[1] += (-ox * oy * ddir[0] + (sum2 - oy * oy) * ddir[1] - oz * oy * ddir[2]) * inv32;
        dmean[2] += (-ox * oz * ddir[0] - oy * oz * ddir[1] + (sum2 - oz * oz) * ddir[2]) * inv32;
      } else {
        my_dsh[0] += dcol[0]; my_dsh[1] += dcol[1]; my_dsh[2] += dcol[2];
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
    }
    // per-Gaussian outputs (small; strided stores merge in L2)
    const size_t og = (size_t)b * p.G + g;
    a.dL_dmeans[og * 3 + 0] = gmean[0]; a.dL_dmeans[og * 3 + 1] = gmean[1]; a.dL_dmeans[og * 3 + 2] = gmean[2];
    a.dL_dopacity[og] = gop;
    if (p.cov_layout == TSPLAT_COV_TRIU6) {
#pragma unroll
      for (int k = 0; k < 6; k++) a.dL_dcov[og * 6 + k] = gcov[k];
    } else {
      float* c = a.dL_dcov + og * 9;
      c[0] = gcov[0]; c[1] = gcov[1]; c[2] = gcov[2]; c[3] = 0.f; c[4] = gcov[3]; c[5] = gcov[4];
      c[6] = 0.f; c[7] = 0.f; c[8] = gcov[5];
    }
  }
  // dL/dsh slab: shared memory -> global, one contiguous write per CTA
  __syncthreads();
  float* g_dsh = a.dL_dsh + gbase * a.sh_floats;
  if (bulk) {
    fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(g_dsh, s_dsh, (uint32_t)(kPreBwdBlock * 4 * a.sh_floats));
      bulk_commit();
      bulk_wait_all();
    }
  } else {
    for (int i = tid; i < cnt * a.sh_floats; i += kPreBwdBlock) g_dsh[i] = s_dsh[i];
  }
}

cudaError_t launch_preprocess_backward(const TsplatProblem& p, const Workspace& ws, float* dL_dmeans, float* dL_dcov,
                                       float* dL_dsh, float* dL_dopacity, float* dL_dmeans2D, int g_begin, int g_count,
                                       cudaStream_t st) {
  if (g_count <= 0) return cudaSuccess;
  PreBwdArgs a;
  a.p = p;
  a.g_begin = g_begin; a.g_end = g_begin + g_count;
  a.cov_floats = p.cov_layout == TSPLAT_COV_TRIU6 ? 6 : 9;
  a.sh_floats = p.sh_coeffs > 0 ? 3 * p.sh_coeffs : 3;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  const bool stride_ok = p.B == 1 || (p.G % 4 == 0);
  a.bulk_ok = al16(p.means) && al16(p.cov) && al16(p.sh) && al16(dL_dsh) && stride_ok;
  a.dL_dmeans = dL_dmeans; a.dL_dcov = dL_dcov; a.dL_dsh = dL_dsh; a.dL_dopacity = dL_dopacity;
  a.dL_dmeans2D = dL_dmeans2D;
  const size_t smem = 16 + sizeof(float) * kPreBwdBlock * (3 + a.cov_floats + 2 * a.sh_floats);
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(preprocess_backward_kernel), (int)smem);
  if (e != cudaSuccess) return e;
  dim3 grid((g_count + kPreBwdBlock - 1) / kPreBwdBlock, p.B);
  preprocess_backward_kernel<<<grid, kPreBwdBlock, smem, st>>>(a, ws);
  return cudaGetLastError();
}

}  // namespace tsplat
