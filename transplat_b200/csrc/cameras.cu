// cameras.cu — device-side camera preparation: everything render_cuda does on the
// host per call before the per-view loop, for all N = B*V cameras in ONE launch and with no host
// synchronisation.  Replaces, for the batched decoder path,
//   cuda_splatting.py:73-80   scale invariance (extrinsics translation, near, far scaled by 1/near)
//   cuda_splatting.py:89-91   get_fov (projection.py:233-247) and tan(fov/2)
//   cuda_splatting.py:93-96   get_projection_matrix (:26-53), extrinsics.inverse(), view @ proj
//   cuda_splatting.py:111-112 the two .item() reads per view (tan fov stays on the device)
//   cuda_splatting.py:388-391 row 2 of the UNSCALED inverse extrinsics (depth feature)
// which in torch are ~100 tiny kernels (batched LU, triangular solves, elementwise ops).
//
// One CTA per camera (see the kernel).  The matrix inverses are Gauss-Jordan with partial pivoting carried in
// binary64 and rounded once to binary32 (torch's batched LU is binary32; the results agree to the
// last ulp or two, they cannot be bit-identical across BLAS back ends anyway); everything else is
// binary32 in the reference's operation order.  Compiled with -fmad=false like preprocess.cu.
#include "common.cuh"

namespace tsplat {

template <int N>
__device__ __forceinline__ bool invert_gj(const float* __restrict__ src, double (&inv)[N][N]) {
  double a[N][N];
#pragma unroll
  for (int r = 0; r < N; r++)
#pragma unroll
    for (int c = 0; c < N; c++) {
      a[r][c] = (double)src[r * N + c];
      inv[r][c] = r == c ? 1.0 : 0.0;
    }
  bool ok = true;
#pragma unroll
  for (int k = 0; k < N; k++) {
    int piv = k;
    double best = fabs(a[k][k]);
#pragma unroll
    for (int r = k + 1; r < N; r++) {
      const double v = fabs(a[r][k]);
      if (v > best) { best = v; piv = r; }
    }
    if (best == 0.0) ok = false;
#pragma unroll
    for (int r = k + 1; r < N; r++) {
      if (r == piv) {
#pragma unroll
        for (int c = 0; c < N; c++) {
          const double t0 = a[k][c]; a[k][c] = a[r][c]; a[r][c] = t0;
          const double t1 = inv[k][c]; inv[k][c] = inv[r][c]; inv[r][c] = t1;
        }
      }
    }
    const double d = 1.0 / a[k][k];
#pragma unroll
    for (int c = 0; c < N; c++) { a[k][c] *= d; inv[k][c] *= d; }
#pragma unroll
    for (int r = 0; r < N; r++) {
      if (r != k) {
        const double f = a[r][k];
#pragma unroll
        for (int c = 0; c < N; c++) { a[r][c] -= f * a[k][c]; inv[r][c] -= f * inv[k][c]; }
      }
    }
  }
  return ok;
}

// unit vector K^-1 * (x, y, 1)  (projection.py:236-239)
__device__ __forceinline__ void unit_ray(const float (&ki)[3][3], float x, float y, float (&o)[3]) {
#pragma unroll
  for (int r = 0; r < 3; r++) o[r] = (ki[r][0] * x + ki[r][1] * y) + ki[r][2] * 1.0f;
  const float n = sqrtf((o[0] * o[0] + o[1] * o[1]) + o[2] * o[2]);
#pragma unroll
  for (int r = 0; r < 3; r++) o[r] = o[r] / n;
}

// One CTA of three warps per camera, one live lane per warp: the unscaled inverse, the scaled inverse and the intrinsics
// chain (inverse, field of view, tangents) are independent, so they run side by side (three warps: lanes of one warp
// would serialise the three code paths) and meet in shared memory; the arithmetic of each is what one thread did
// before, so the results are bit-identical.  The kernel is a chain of dependent binary64 operations and nothing else;
// this takes about 2 us off the C2 step (0.4825 -> 0.4805 ms, graph replay).
__global__ void __launch_bounds__(96) pack_cameras_kernel(int N, const float* __restrict__ extrinsics,
                                                          const float* __restrict__ intrinsics,
                                                          const float* __restrict__ near_, const float* __restrict__ far_,
                                                          const float* __restrict__ bg, int bg_stride, int scale_invariant,
                                                          TsplatCamera* __restrict__ out) {
  __shared__ float s_inv_u_row2[4];
  __shared__ float s_tan[2];
  const int n = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (n >= N) return;
  float e[16];
  float near_u = 0.f, far_u = 0.f, scale = 1.0f, nr = 0.f, fr = 0.f;
  double inv_s[4][4];
  if (lane == 0 && warp < 2) {
#pragma unroll
    for (int i = 0; i < 16; i++) e[i] = extrinsics[(size_t)n * 16 + i];
    near_u = near_[n]; far_u = far_[n];
    nr = near_u; fr = far_u;
    if (warp == 0) {
      double inv_u[4][4];
      invert_gj<4>(e, inv_u);  // unscaled inverse: depth feature row (cuda_splatting.py:389)
#pragma unroll
      for (int j = 0; j < 4; j++) s_inv_u_row2[j] = (float)inv_u[2][j];
    } else {
      if (scale_invariant) {  // cuda_splatting.py:73-80
        scale = 1.0f / near_u;
        e[3] *= scale; e[7] *= scale; e[11] *= scale;
        nr = near_u * scale;
        fr = far_u * scale;
      }
      invert_gj<4>(e, inv_s);
    }
  }
  if (lane == 0 && warp == 2) {
    float ki[3][3];
    {
      double kd[3][3];
      float k[9];
#pragma unroll
      for (int i = 0; i < 9; i++) k[i] = intrinsics[(size_t)n * 9 + i];
      invert_gj<3>(k, kd);
#pragma unroll
      for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) ki[r][c] = (float)kd[r][c];
    }
    float l[3], r_[3], t[3], b[3];
    unit_ray(ki, 0.0f, 0.5f, l);
    unit_ray(ki, 1.0f, 0.5f, r_);
    unit_ray(ki, 0.5f, 0.0f, t);
    unit_ray(ki, 0.5f, 1.0f, b);
    const float fov_x = acosf((l[0] * r_[0] + l[1] * r_[1]) + l[2] * r_[2]);
    const float fov_y = acosf((t[0] * b[0] + t[1] * b[1]) + t[2] * b[2]);
    s_tan[0] = tanf(0.5f * fov_x);
    s_tan[1] = tanf(0.5f * fov_y);
  }
  __syncthreads();
  if (!(lane == 0 && warp == 1)) return;  // the warp that holds the scaled inverse assembles the record
  const float tan_x = s_tan[0], tan_y = s_tan[1];

  // get_projection_matrix (cuda_splatting.py:38-52)
  const float top = tan_y * nr, bottom = -top, right = tan_x * nr, left = -right;
  float P[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) P[i][j] = 0.0f;
  P[0][0] = (2.0f * nr) / (right - left);
  P[1][1] = (2.0f * nr) / (top - bottom);
  P[0][2] = (right + left) / (right - left);
  P[1][2] = (top + bottom) / (top - bottom);
  P[3][2] = 1.0f;
  P[2][2] = fr / (fr - nr);
  P[2][3] = -(fr * nr) / (fr - nr);

  TsplatCamera c;
  float vm[4][4];  // view_matrix = inverse(extrinsics)^T  (cuda_splatting.py:95)
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      vm[i][j] = (float)inv_s[j][i];
      c.view[4 * i + j] = vm[i][j];
    }
  // full_projection = view_matrix @ projection^T  (cuda_splatting.py:96), k ascending
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < 4; k++) acc = __fmaf_rn(vm[i][k], P[j][k], acc);
      c.proj[4 * i + j] = acc;
    }
  c.campos[0] = e[3]; c.campos[1] = e[7]; c.campos[2] = e[11];  // extrinsics[i, :3, 3] (:118)
  c.tanfovx = tan_x;
  c.tanfovy = tan_y;
  const float* bgp = bg + (size_t)n * bg_stride;
  c.bg[0] = bgp[0]; c.bg[1] = bgp[1]; c.bg[2] = bgp[2];
  c.mean_scale = scale;
  c.cov_scale = scale * scale;
#pragma unroll
  for (int j = 0; j < 4; j++) c.depth_row[j] = s_inv_u_row2[j];
  c.near_plane = near_u;
  c.far_plane = far_u;
#pragma unroll
  for (int i = 0; i < 16; i++) c.reserved[i] = 0.0f;
  out[n] = c;
}

cudaError_t launch_pack_cameras(int N, const float* extrinsics, const float* intrinsics, const float* near_,
                                const float* far_, const float* bg, int bg_stride, int scale_invariant,
                                TsplatCamera* out, cudaStream_t st) {
  if (N <= 0) return cudaSuccess;
  pack_cameras_kernel<<<N, 96, 0, st>>>(N, extrinsics, intrinsics, near_, far_, bg, bg_stride, scale_invariant, out);
  return cudaGetLastError();
}

// 16-byte status block -> pinned host memory through its device alias (tsplat_status_read)
__global__ void status_store_kernel(const uint4* __restrict__ status, uint4* host_alias) {
  if (threadIdx.x < 2) {  // the 32-byte TsplatStatus
    host_alias[threadIdx.x] = status[threadIdx.x];
    __threadfence_system();
  }
}
void launch_status_store(const void* workspace, void* host_alias, cudaStream_t st) {
  status_store_kernel<<<1, 32, 0, st>>>(static_cast<const uint4*>(workspace), static_cast<uint4*>(host_alias));
}

}  // namespace tsplat
