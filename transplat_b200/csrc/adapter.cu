// adapter.cu — the per-pixel Gaussian producer in front of the render path: one kernel for
// GaussianAdapter.forward (/root/reference/src/model/encoder/common/gaussian_adapter.py:50-98) and one for its
// backward.  The reference runs ~40 torch ops that materialise [.., 3, 3] rotation / scale / covariance chains
// (gaussians.py:32-43), broadcast world rays (projection.py:91-113) and an einsum per SH degree
// (sh_rotation_optimized.py:80-99); here a CTA takes 128 consecutive Gaussians of one context view:
//   * the raw feature slab [128, 7 + 3 d_sh] arrives in shared memory by ONE 1-D TMA bulk copy (mbarrier tx),
//   * a thread owns one Gaussian: activation of scales, quaternion -> R, covariance A diag(s^2) A^T with
//     A = C2W_rot * R (never forming the intermediates), world ray and mean, SH mask + Wigner-D rotation of the
//     three colour channels IN PLACE in the staged row (the per-view Wigner blocks sit in shared memory and are
//     read as broadcasts),
//   * harmonics leave through a coalesced copy of the staged rows; means / covariances / scales / rotations are
//     small per-thread stores.
// HBM-bound: 340 B read + 376 B written per Gaussian forward, 1 012 B read + 340 B written backward.
// The Wigner-D matrices are an INPUT (computed per context view by the caller, as in the reference's
// precompute_wigner_d_matrices): they depend on the camera only.
#include "common.cuh"

namespace tsplat {

constexpr int kAdBlock = 128;
constexpr int kWigFloats = 1 + 9 + 25 + 49 + 81;  // per-degree (2l+1)^2 blocks up to degree 4

struct AdapterArgs {
  TsplatAdapterProblem p;
  int d_sh, row, wig_floats, bulk_ok;
  float* means; float* cov; float* harmonics; float* scales; float* rotations;                  // forward outputs
  const float* g_means; const float* g_cov; const float* g_harm;                                // backward inputs
  float* g_coords; float* g_depths; float* g_raw;                                               // backward outputs
};

// per-view constants every thread of the CTA needs (computed redundantly per thread from 25 floats: cheaper
// than a broadcast through shared memory plus a barrier)
struct ViewConst {
  float C[3][3];     // camera-to-world rotation, extrinsics[:3,:3]
  float o[3];        // ray origin, extrinsics[:3,3]
  float Ki[3][3];    // inverse intrinsics
  float mult;        // get_scale_multiplier (gaussian_adapter.py:100-111)
};

__device__ __forceinline__ void load_view(const TsplatAdapterProblem& p, int n, ViewConst& v) {
  const float* e = p.extrinsics + (size_t)n * 16;
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) v.C[i][j] = __ldg(e + 4 * i + j);
    v.o[i] = __ldg(e + 4 * i + 3);
  }
  const float* k = p.intrinsics + (size_t)n * 9;
  float K[3][3];
#pragma unroll
  for (int i = 0; i < 9; i++) K[i / 3][i % 3] = __ldg(k + i);
  // 3x3 inverse by cofactors (the reference calls Tensor.inverse(): LU in binary32; both are within 2 ulp-ish of
  // the exact inverse for the near-triangular intrinsics this path sees)
  const float c00 = K[1][1] * K[2][2] - K[1][2] * K[2][1], c01 = K[1][2] * K[2][0] - K[1][0] * K[2][2],
              c02 = K[1][0] * K[2][1] - K[1][1] * K[2][0];
  const float det = K[0][0] * c00 + K[0][1] * c01 + K[0][2] * c02;
  const float id = 1.0f / det;
  v.Ki[0][0] = c00 * id; v.Ki[1][0] = c01 * id; v.Ki[2][0] = c02 * id;
  v.Ki[0][1] = (K[0][2] * K[2][1] - K[0][1] * K[2][2]) * id;
  v.Ki[1][1] = (K[0][0] * K[2][2] - K[0][2] * K[2][0]) * id;
  v.Ki[2][1] = (K[0][1] * K[2][0] - K[0][0] * K[2][1]) * id;
  v.Ki[0][2] = (K[0][1] * K[1][2] - K[0][2] * K[1][1]) * id;
  v.Ki[1][2] = (K[0][2] * K[1][0] - K[0][0] * K[1][2]) * id;
  v.Ki[2][2] = (K[0][0] * K[1][1] - K[0][1] * K[1][0]) * id;
  // multiplier = 0.1 * sum_i (inverse(K[:2,:2]) @ (1/w, 1/h))_i
  const float d2 = K[0][0] * K[1][1] - K[0][1] * K[1][0];
  const float pw = 1.0f / (float)p.W, ph = 1.0f / (float)p.H;
  const float m0 = (K[1][1] * pw - K[0][1] * ph) / d2, m1 = (-K[1][0] * pw + K[0][0] * ph) / d2;
  v.mult = 0.1f * m0 + 0.1f * m1;
}

// quaternion (i, j, k, r) -> rotation matrix exactly as gaussians.py:12-29 (two_s carries its own eps)
__device__ __forceinline__ void quat_to_R(float i, float j, float k, float r, float& t, float (&R)[3][3]) {
  t = 2.0f / ((i * i + j * j + k * k + r * r) + 1e-8f);
  R[0][0] = 1.0f - t * (j * j + k * k); R[0][1] = t * (i * j - k * r); R[0][2] = t * (i * k + j * r);
  R[1][0] = t * (i * j + k * r); R[1][1] = 1.0f - t * (i * i + k * k); R[1][2] = t * (j * k - i * r);
  R[2][0] = t * (i * k - j * r); R[2][1] = t * (j * k + i * r); R[2][2] = 1.0f - t * (i * i + j * j);
}

__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

// stage `rows` rows of `row` floats (contiguous in global memory) in shared memory
__device__ __forceinline__ void stage_rows(float* s_dst, const float* g_src, int cnt, int row, bool bulk, uint64_t* bar,
                                           int tid) {
  if (bulk) {
    if (tid == 0) {
      mbar_expect_tx(bar, (uint32_t)(kAdBlock * row * 4));
      bulk_g2s(s_dst, g_src, (uint32_t)(kAdBlock * row * 4), bar);
    }
  } else {
    for (int i = tid; i < cnt * row; i += kAdBlock) s_dst[i] = g_src[i];
  }
}

// out[c][i] = sum_j D_l[i][j] * (in[c][j] * mask[j]) for the degree-L block of the three channels, in place
template <int L>
__device__ __forceinline__ void rotate_block(float* sh, int d_sh, const float* __restrict__ sD,
                                             const float* __restrict__ mask) {
  constexpr int M = 2 * L + 1, O = L * L;
  float in[3][M];
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int j = 0; j < M; j++) in[c][j] = sh[c * d_sh + O + j] * mask[O + j];
#pragma unroll
  for (int i = 0; i < M; i++) {
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
#pragma unroll
    for (int j = 0; j < M; j++) {
      const float d = sD[i * M + j];
      a0 += d * in[0][j]; a1 += d * in[1][j]; a2 += d * in[2][j];
    }
    sh[0 * d_sh + O + i] = a0; sh[1 * d_sh + O + i] = a1; sh[2 * d_sh + O + i] = a2;
  }
}
// transpose of the above for the backward: g_in[c][j] = mask[j] * sum_i D_l[i][j] * g_out[c][i]; reads g (staged
// upstream gradient row), writes dst (the staged raw row's SH part, which becomes dL/draw)
template <int L>
__device__ __forceinline__ void rotate_block_T(const float* g, float* dst, int d_sh, const float* __restrict__ sD,
                                               const float* __restrict__ mask) {
  constexpr int M = 2 * L + 1, O = L * L;
  float go[3][M];
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int i = 0; i < M; i++) go[c][i] = g[c * d_sh + O + i];
#pragma unroll
  for (int j = 0; j < M; j++) {
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
#pragma unroll
    for (int i = 0; i < M; i++) {
      const float d = sD[i * M + j];
      a0 += d * go[0][i]; a1 += d * go[1][i]; a2 += d * go[2][i];
    }
    const float m = mask[O + j];
    dst[0 * d_sh + O + j] = a0 * m; dst[1 * d_sh + O + j] = a1 * m; dst[2 * d_sh + O + j] = a2 * m;
  }
}

__global__ void __launch_bounds__(kAdBlock) adapter_forward_kernel(AdapterArgs a) {
  extern __shared__ __align__(128) unsigned char ad_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(ad_raw);
  float* s_wig = reinterpret_cast<float*>(ad_raw + 16);
  float* s_mask = s_wig + kWigFloats;
  float* s_row = reinterpret_cast<float*>(ad_raw + 1024);  // header (barrier, Wigner blocks, mask) fits in 1 KB
  const TsplatAdapterProblem& p = a.p;
  const int tid = threadIdx.x, n = blockIdx.y;
  const int g0 = blockIdx.x * kAdBlock, cnt = min(kAdBlock, p.n - g0);
  const size_t gbase = (size_t)n * p.n + g0;
  const bool bulk = a.bulk_ok && cnt == kAdBlock;
  if (tid == 0 && bulk) { mbar_init(bar, 1); mbar_fence_init(); }
  for (int i = tid; i < a.wig_floats; i += kAdBlock) s_wig[i] = p.wigner[(size_t)n * a.wig_floats + i];
  if (tid < a.d_sh) s_mask[tid] = p.sh_mask[tid];
  __syncthreads();
  stage_rows(s_row, p.raw + gbase * a.row, cnt, a.row, bulk, bar, tid);
  ViewConst vc;
  load_view(p, n, vc);
  __syncthreads();
  if (bulk) mbar_wait(bar, 0);

  if (tid < cnt) {
    const size_t g = gbase + tid;
    float* r = s_row + (size_t)tid * a.row;
    const float depth = p.depths[g];
    const float2 xy = reinterpret_cast<const float2*>(p.coordinates)[g];
    // scales (gaussian_adapter.py:62-69)
    float s[3];
#pragma unroll
    for (int k = 0; k < 3; k++) s[k] = ((p.scale_min + (p.scale_max - p.scale_min) * sigmoidf(r[k])) * depth) * vc.mult;
    // rotation (:72), xyzw
    const float qn = sqrtf(r[3] * r[3] + r[4] * r[4] + r[5] * r[5] + r[6] * r[6]) + p.eps;
    const float qi = r[3] / qn, qj = r[4] / qn, qk = r[5] / qn, qr = r[6] / qn;
    float t, R[3][3], A[3][3];
    quat_to_R(qi, qj, qk, qr, t, R);
    // covariance = (C R) diag(s^2) (C R)^T (gaussians.py:36-43, gaussian_adapter.py:79-81)
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) A[i][j] = vc.C[i][0] * R[0][j] + vc.C[i][1] * R[1][j] + vc.C[i][2] * R[2][j];
    float* co = a.cov + g * 9;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++)
        co[3 * i + j] = A[i][0] * A[j][0] * (s[0] * s[0]) + A[i][1] * A[j][1] * (s[1] * s[1]) + A[i][2] * A[j][2] * (s[2] * s[2]);
    // world ray and mean (projection.py:74-113, gaussian_adapter.py:84-85)
    float d[3];
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = vc.Ki[i][0] * xy.x + vc.Ki[i][1] * xy.y + vc.Ki[i][2];
    const float dn = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] /= dn;
#pragma unroll
    for (int i = 0; i < 3; i++)
      a.means[g * 3 + i] = vc.o[i] + (vc.C[i][0] * d[0] + vc.C[i][1] * d[1] + vc.C[i][2] * d[2]) * depth;
    a.scales[g * 3 + 0] = s[0]; a.scales[g * 3 + 1] = s[1]; a.scales[g * 3 + 2] = s[2];
    reinterpret_cast<float4*>(a.rotations)[g] = make_float4(qi, qj, qk, qr);
    // SH: mask, then per-degree Wigner-D rotation of the three channels, in place in the staged row
    float* sh = r + 7;
    rotate_block<0>(sh, a.d_sh, s_wig, s_mask);
    if (p.sh_degree >= 1) rotate_block<1>(sh, a.d_sh, s_wig + 1, s_mask);
    if (p.sh_degree >= 2) rotate_block<2>(sh, a.d_sh, s_wig + 10, s_mask);
    if (p.sh_degree >= 3) rotate_block<3>(sh, a.d_sh, s_wig + 35, s_mask);
    if (p.sh_degree >= 4) rotate_block<4>(sh, a.d_sh, s_wig + 84, s_mask);
  }
  __syncthreads();
  const int hrow = 3 * a.d_sh;
  float* out = a.harmonics + gbase * hrow;
  for (int f = tid; f < cnt * hrow; f += kAdBlock) {
    const int gi = f / hrow, k = f - gi * hrow;
    out[f] = s_row[(size_t)gi * a.row + 7 + k];
  }
}

// Backward: only raw[0..6] (scales, quaternion) are needed again - harmonics are linear in the SH features - so the
// slab staged by TMA is the UPSTREAM gradient of the harmonics; it is rotated back (D^T, mask) in place and leaves,
// together with the 7 head gradients, as whole rows of dL/draw through a coalesced copy.
__global__ void __launch_bounds__(kAdBlock) adapter_backward_kernel(AdapterArgs a) {
  extern __shared__ __align__(128) unsigned char ad_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(ad_raw);
  float* s_wig = reinterpret_cast<float*>(ad_raw + 16);
  float* s_mask = s_wig + kWigFloats;
  float* s_g = reinterpret_cast<float*>(ad_raw + 1024);
  const TsplatAdapterProblem& p = a.p;
  const int hrow = 3 * a.d_sh;
  float* s_head = s_g + (size_t)kAdBlock * hrow;
  const int tid = threadIdx.x, n = blockIdx.y;
  const int g0 = blockIdx.x * kAdBlock, cnt = min(kAdBlock, p.n - g0);
  const size_t gbase = (size_t)n * p.n + g0;
  const bool bulk = a.bulk_ok && cnt == kAdBlock;
  if (tid == 0 && bulk) { mbar_init(bar, 1); mbar_fence_init(); }
  for (int i = tid; i < a.wig_floats; i += kAdBlock) s_wig[i] = p.wigner[(size_t)n * a.wig_floats + i];
  if (tid < a.d_sh) s_mask[tid] = p.sh_mask[tid];
  __syncthreads();
  stage_rows(s_g, a.g_harm + gbase * hrow, cnt, hrow, bulk, bar, tid);
  ViewConst vc;
  load_view(p, n, vc);
  __syncthreads();
  if (bulk) mbar_wait(bar, 0);

  if (tid < cnt) {
    const size_t g = gbase + tid;
    const float* r = p.raw + g * a.row;
    const float depth = p.depths[g];
    const float2 xy = reinterpret_cast<const float2*>(p.coordinates)[g];
    float gm[3], gc[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) gm[i] = a.g_means[g * 3 + i];
#pragma unroll
    for (int i = 0; i < 9; i++) gc[i / 3][i % 3] = a.g_cov[g * 9 + i];
    // ---- forward quantities again
    const float dsc = p.scale_max - p.scale_min;
    float sg[3], act[3], s[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      sg[k] = sigmoidf(r[k]);
      act[k] = p.scale_min + dsc * sg[k];
      s[k] = (act[k] * depth) * vc.mult;
    }
    const float q[4] = {r[3], r[4], r[5], r[6]};
    const float nrm = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const float qd = nrm + p.eps;
    const float qi = q[0] / qd, qj = q[1] / qd, qk = q[2] / qd, qr = q[3] / qd;
    float t, R[3][3], A[3][3];
    quat_to_R(qi, qj, qk, qr, t, R);
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) A[i][j] = vc.C[i][0] * R[0][j] + vc.C[i][1] * R[1][j] + vc.C[i][2] * R[2][j];
    // ---- mean path: mean = o + C (d/|d|) depth
    float d[3], dh[3], dw[3];
#pragma unroll
    for (int i = 0; i < 3; i++) d[i] = vc.Ki[i][0] * xy.x + vc.Ki[i][1] * xy.y + vc.Ki[i][2];
    const float dn = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
#pragma unroll
    for (int i = 0; i < 3; i++) dh[i] = d[i] / dn;
#pragma unroll
    for (int i = 0; i < 3; i++) dw[i] = vc.C[i][0] * dh[0] + vc.C[i][1] * dh[1] + vc.C[i][2] * dh[2];
    float g_depth = gm[0] * dw[0] + gm[1] * dw[1] + gm[2] * dw[2];
    float gdh[3];
#pragma unroll
    for (int j = 0; j < 3; j++) gdh[j] = (vc.C[0][j] * gm[0] + vc.C[1][j] * gm[1] + vc.C[2][j] * gm[2]) * depth;
    const float dot = dh[0] * gdh[0] + dh[1] * gdh[1] + dh[2] * gdh[2];
    float gd[3];
#pragma unroll
    for (int j = 0; j < 3; j++) gd[j] = (gdh[j] - dh[j] * dot) / dn;
    reinterpret_cast<float2*>(a.g_coords)[g] =
        make_float2(vc.Ki[0][0] * gd[0] + vc.Ki[1][0] * gd[1] + vc.Ki[2][0] * gd[2],
                    vc.Ki[0][1] * gd[0] + vc.Ki[1][1] * gd[1] + vc.Ki[2][1] * gd[2]);
    // ---- covariance path: cov = sum_k s_k^2 a_k a_k^T, a_k = column k of A = C R
    float B[3][3];  // (g_cov + g_cov^T) A
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int k = 0; k < 3; k++)
        B[i][k] = (gc[i][0] + gc[0][i]) * A[0][k] + (gc[i][1] + gc[1][i]) * A[1][k] + (gc[i][2] + gc[2][i]) * A[2][k];
    float gR[3][3];
    float* head = s_head + tid * 7;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float gs = s[k] * (A[0][k] * B[0][k] + A[1][k] * B[1][k] + A[2][k] * B[2][k]);  // dL/ds_k
      g_depth += gs * act[k] * vc.mult;
      head[k] = gs * depth * vc.mult * dsc * sg[k] * (1.0f - sg[k]);
      const float s2 = s[k] * s[k];
#pragma unroll
      for (int m = 0; m < 3; m++) gR[m][k] = (vc.C[0][m] * B[0][k] + vc.C[1][m] * B[1][k] + vc.C[2][m] * B[2][k]) * s2;
    }
    a.g_depths[g] = g_depth;
    // ---- quaternion path: R = I + t Q(qn), t = 2 / (qn.qn + 1e-8), qn = q / (|q| + eps)
    const float Q[3][3] = {{-(qj * qj + qk * qk), qi * qj - qk * qr, qi * qk + qj * qr},
                           {qi * qj + qk * qr, -(qi * qi + qk * qk), qj * qk - qi * qr},
                           {qi * qk - qj * qr, qj * qk + qi * qr, -(qi * qi + qj * qj)}};
    float gt = 0.0f;
#pragma unroll
    for (int m = 0; m < 3; m++)
#pragma unroll
      for (int k = 0; k < 3; k++) gt += gR[m][k] * Q[m][k];
    const float x00 = t * gR[0][0], x01 = t * gR[0][1], x02 = t * gR[0][2], x10 = t * gR[1][0], x11 = t * gR[1][1],
                x12 = t * gR[1][2], x20 = t * gR[2][0], x21 = t * gR[2][1], x22 = t * gR[2][2];
    const float c = -gt * t * t;  // dt/dqn = -t^2 qn
    float gqn[4];
    gqn[0] = (x01 + x10) * qj + (x02 + x20) * qk + (x21 - x12) * qr - 2.0f * qi * (x11 + x22) + c * qi;
    gqn[1] = (x01 + x10) * qi + (x12 + x21) * qk + (x02 - x20) * qr - 2.0f * qj * (x00 + x22) + c * qj;
    gqn[2] = (x02 + x20) * qi + (x12 + x21) * qj + (x10 - x01) * qr - 2.0f * qk * (x00 + x11) + c * qk;
    gqn[3] = (x10 - x01) * qk + (x02 - x20) * qj + (x21 - x12) * qi + c * qr;
    const float gq_dot_q = gqn[0] * q[0] + gqn[1] * q[1] + gqn[2] * q[2] + gqn[3] * q[3];
    const float w = nrm > 0.0f ? gq_dot_q / (qd * qd * nrm) : 0.0f;
#pragma unroll
    for (int k = 0; k < 4; k++) head[3 + k] = gqn[k] / qd - q[k] * w;
    // ---- SH path, in place in the staged upstream-gradient row
    float* gh = s_g + (size_t)tid * hrow;
    rotate_block_T<0>(gh, gh, a.d_sh, s_wig, s_mask);
    if (p.sh_degree >= 1) rotate_block_T<1>(gh, gh, a.d_sh, s_wig + 1, s_mask);
    if (p.sh_degree >= 2) rotate_block_T<2>(gh, gh, a.d_sh, s_wig + 10, s_mask);
    if (p.sh_degree >= 3) rotate_block_T<3>(gh, gh, a.d_sh, s_wig + 35, s_mask);
    if (p.sh_degree >= 4) rotate_block_T<4>(gh, gh, a.d_sh, s_wig + 84, s_mask);
  }
  __syncthreads();
  float* out = a.g_raw + gbase * a.row;
  for (int f = tid; f < cnt * a.row; f += kAdBlock) {
    const int gi = f / a.row, k = f - gi * a.row;
    out[f] = k < 7 ? s_head[gi * 7 + k] : s_g[(size_t)gi * hrow + (k - 7)];
  }
}

static int adapter_args(const TsplatAdapterProblem& p, AdapterArgs& a) {
  if (p.N < 0 || p.n < 0 || p.sh_degree < 0 || p.sh_degree > 4 || p.H <= 0 || p.W <= 0) return 1;
  if ((p.N > 0 && p.n > 0) && (!p.extrinsics || !p.intrinsics || !p.coordinates || !p.depths || !p.raw || !p.wigner || !p.sh_mask))
    return 1;
  a.p = p;
  a.d_sh = (p.sh_degree + 1) * (p.sh_degree + 1);
  a.row = 7 + 3 * a.d_sh;
  a.wig_floats = 0;
  for (int l = 0; l <= p.sh_degree; l++) a.wig_floats += (2 * l + 1) * (2 * l + 1);
  return 0;
}
static bool al16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; }

cudaError_t launch_adapter_forward(const TsplatAdapterProblem& p, float* means, float* cov, float* harmonics,
                                   float* scales, float* rotations, cudaStream_t st, int* bad) {
  AdapterArgs a{};
  *bad = adapter_args(p, a) || !means || !cov || !harmonics || !scales || !rotations;
  if (*bad || p.N == 0 || p.n == 0) return cudaSuccess;
  a.means = means; a.cov = cov; a.harmonics = harmonics; a.scales = scales; a.rotations = rotations;
  // a CTA's slab starts at (view*n + 128*k) rows: 16-B aligned iff every such offset is (row bytes * 128 always is)
  a.bulk_ok = al16(p.raw) && ((size_t)p.n * a.row * 4) % 16 == 0;
  const size_t smem = 1024 + sizeof(float) * kAdBlock * a.row;
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(adapter_forward_kernel), (int)smem);
  if (e != cudaSuccess) return e;
  dim3 grid((p.n + kAdBlock - 1) / kAdBlock, p.N);
  adapter_forward_kernel<<<grid, kAdBlock, smem, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_adapter_backward(const TsplatAdapterProblem& p, const float* g_means, const float* g_cov,
                                    const float* g_harm, float* g_coords, float* g_depths, float* g_raw,
                                    cudaStream_t st, int* bad) {
  AdapterArgs a{};
  *bad = adapter_args(p, a) || !g_means || !g_cov || !g_harm || !g_coords || !g_depths || !g_raw;
  if (*bad || p.N == 0 || p.n == 0) return cudaSuccess;
  a.g_means = g_means; a.g_cov = g_cov; a.g_harm = g_harm; a.g_coords = g_coords; a.g_depths = g_depths; a.g_raw = g_raw;
  a.bulk_ok = al16(g_harm) && ((size_t)p.n * 3 * a.d_sh * 4) % 16 == 0;
  const size_t smem = 1024 + sizeof(float) * kAdBlock * (3 * a.d_sh + 7);
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(adapter_backward_kernel), (int)smem);
  if (e != cudaSuccess) return e;
  dim3 grid((p.n + kAdBlock - 1) / kAdBlock, p.N);
  adapter_backward_kernel<<<grid, kAdBlock, smem, st>>>(a);
  return cudaGetLastError();
}

}  // namespace tsplat
