// cov3d.cu — the scales / rotations input mode of GaussianRasterizer.forward: computeCov3D of
// the reference's external extension and its backward, one thread per Gaussian.  TranSplat itself always passes
// cov3D_precomp (cuda_splatting.py:132), so this is call-surface coverage, HBM-bound: 28 B read + 24 B written per
// Gaussian forward, 52 B read + 28 B written backward.
//
//   s = scale_modifier * scale;  q = (r, x, y, z) USED AS GIVEN (the extension does not normalise it);
//   R = the quaternion's rotation-matrix formula;  M = R diag(s);  Sigma = M M^T, upper triangle xx xy xz yy yz zz.
#include "common.cuh"

namespace tsplat {

__device__ __forceinline__ void quat_matrix(float r, float x, float y, float z, float (&R)[3][3]) {
  R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - r * z); R[0][2] = 2.f * (x * z + r * y);
  R[1][0] = 2.f * (x * y + r * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - r * x);
  R[2][0] = 2.f * (x * z - r * y); R[2][1] = 2.f * (y * z + r * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
}

__global__ void __launch_bounds__(256) cov3d_forward_kernel(int G, const float* __restrict__ scales,
                                                            const float* __restrict__ rot, float mod,
                                                            float* __restrict__ cov6) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G) return;
  const float4 q = *reinterpret_cast<const float4*>(rot + 4 * (size_t)i);
  const float s[3] = {mod * scales[3 * (size_t)i], mod * scales[3 * (size_t)i + 1], mod * scales[3 * (size_t)i + 2]};
  float R[3][3], M[3][3];
  quat_matrix(q.x, q.y, q.z, q.w, R);
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = 0; b < 3; b++) M[a][b] = R[a][b] * s[b];
  auto dot = [&](int a, int b) { return M[a][0] * M[b][0] + M[a][1] * M[b][1] + M[a][2] * M[b][2]; };
  float* o = cov6 + 6 * (size_t)i;
  o[0] = dot(0, 0); o[1] = dot(0, 1); o[2] = dot(0, 2); o[3] = dot(1, 1); o[4] = dot(1, 2); o[5] = dot(2, 2);
}

// dL/dcov6 holds ONE value per off-diagonal pair: as a symmetric matrix the pair shares it half and half.
__global__ void __launch_bounds__(256) cov3d_backward_kernel(int G, const float* __restrict__ scales,
                                                             const float* __restrict__ rot, float mod,
                                                             const float* __restrict__ dL_dcov6,
                                                             float* __restrict__ dL_dscales, float* __restrict__ dL_drot) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G) return;
  const float4 q = *reinterpret_cast<const float4*>(rot + 4 * (size_t)i);
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const float s[3] = {mod * scales[3 * (size_t)i], mod * scales[3 * (size_t)i + 1], mod * scales[3 * (size_t)i + 2]};
  float R[3][3], M[3][3];
  quat_matrix(r, x, y, z, R);
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = 0; b < 3; b++) M[a][b] = R[a][b] * s[b];
  const float* g = dL_dcov6 + 6 * (size_t)i;
  const float Gs[3][3] = {{g[0], 0.5f * g[1], 0.5f * g[2]}, {0.5f * g[1], g[3], 0.5f * g[4]}, {0.5f * g[2], 0.5f * g[4], g[5]}};
  // Sigma = M M^T  =>  dL/dM = 2 Gs M
  float dM[3][3];
#pragma unroll
  for (int a = 0; a < 3; a++)
#pragma unroll
    for (int b = 0; b < 3; b++) dM[a][b] = 2.f * (Gs[a][0] * M[0][b] + Gs[a][1] * M[1][b] + Gs[a][2] * M[2][b]);
  float dR[3][3];
#pragma unroll
  for (int b = 0; b < 3; b++) {
    dL_dscales[3 * (size_t)i + b] = mod * (dM[0][b] * R[0][b] + dM[1][b] * R[1][b] + dM[2][b] * R[2][b]);
#pragma unroll
    for (int a = 0; a < 3; a++) dR[a][b] = dM[a][b] * s[b];
  }
  float4 dq;
  dq.x = 2.f * (-z * dR[0][1] + y * dR[0][2] + z * dR[1][0] - x * dR[1][2] - y * dR[2][0] + x * dR[2][1]);
  dq.y = 2.f * (y * dR[0][1] + z * dR[0][2] + y * dR[1][0] - r * dR[1][2] + z * dR[2][0] + r * dR[2][1]) -
         4.f * x * (dR[1][1] + dR[2][2]);
  dq.z = 2.f * (x * dR[0][1] + r * dR[0][2] + x * dR[1][0] + z * dR[1][2] - r * dR[2][0] + z * dR[2][1]) -
         4.f * y * (dR[0][0] + dR[2][2]);
  dq.w = 2.f * (-r * dR[0][1] + x * dR[0][2] + r * dR[1][0] + y * dR[1][2] + x * dR[2][0] + y * dR[2][1]) -
         4.f * z * (dR[0][0] + dR[1][1]);
  *reinterpret_cast<float4*>(dL_drot + 4 * (size_t)i) = dq;
}

cudaError_t launch_cov3d_forward(int G, const float* scales, const float* rot, float mod, float* cov6, cudaStream_t st) {
  if (G <= 0) return cudaSuccess;
  cov3d_forward_kernel<<<(G + 255) / 256, 256, 0, st>>>(G, scales, rot, mod, cov6);
  return cudaGetLastError();
}

cudaError_t launch_cov3d_backward(int G, const float* scales, const float* rot, float mod, const float* dL_dcov6,
                                  float* dL_dscales, float* dL_drot, cudaStream_t st) {
  if (G <= 0) return cudaSuccess;
  cov3d_backward_kernel<<<(G + 255) / 256, 256, 0, st>>>(G, scales, rot, mod, dL_dcov6, dL_dscales, dL_drot);
  return cudaGetLastError();
}

}  // namespace tsplat
