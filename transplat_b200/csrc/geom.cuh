// geom.cuh — per-Gaussian geometry and spherical-harmonics math shared by the preprocess
// forward and backward kernels.  Semantics: SURVEY.md Appendix A.2/A.3/A.5/A.6/A.7 (the published
// 3DGS rasterizer the reference reaches at cuda_splatting.py:122-133).
//
// Floating-point contract (parity with the oracle's integer decisions): every operation below is
// written as one IEEE binary32 operation in left-to-right order; the translation unit that feeds
// depth bits / radii / tile rectangles (preprocess.cu) is compiled with -fmad=false so ptxas
// cannot contract a*b+c, and without --use_fast_math so '/' and sqrtf are correctly rounded.
#pragma once
#include "common.cuh"

namespace tsplat {

#define TS_SH_C0 0.28209479177387814f
#define TS_SH_C1 0.4886025119029199f
static __constant__ float kC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                              -1.0925484305920792f, 0.5462742152960396f};
static __constant__ float kC3[7] = {-0.5900435899266435f, 2.890611442640554f,  -0.4570457994644658f,
                                              0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                              -0.5900435899266435f};
static __constant__ float kC4[9] = {2.5033429417967046f,  -1.7701307697799304f, 0.9461746957575601f,
                                              -0.6690465435572892f, 0.10578554691520431f, -0.6690465435572892f,
                                              0.47308734787878004f, -1.7701307697799304f, 0.6258357354491761f};

__device__ __forceinline__ float3 transform4x3(const float3 p, const float* __restrict__ m) {
  float3 o;
  o.x = ((m[0] * p.x + m[4] * p.y) + m[8] * p.z) + m[12];
  o.y = ((m[1] * p.x + m[5] * p.y) + m[9] * p.z) + m[13];
  o.z = ((m[2] * p.x + m[6] * p.y) + m[10] * p.z) + m[14];
  return o;
}
__device__ __forceinline__ float4 transform4x4(const float3 p, const float* __restrict__ m) {
  float4 o;
  o.x = ((m[0] * p.x + m[4] * p.y) + m[8] * p.z) + m[12];
  o.y = ((m[1] * p.x + m[5] * p.y) + m[9] * p.z) + m[13];
  o.z = ((m[2] * p.x + m[6] * p.y) + m[10] * p.z) + m[14];
  o.w = ((m[3] * p.x + m[7] * p.y) + m[11] * p.z) + m[15];
  return o;
}
// upstream evaluates ((v + 1.0) * S - 1.0) * 0.5 with double literals
__device__ __forceinline__ float ndc2pix(float v, int S) {
  return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5);
}
// getRect (A.2).  (int) is C truncation; CUDA's cvt.rzi saturates and maps NaN to 0 like the oracle.
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, int& x0, int& y0,
                                         int& x1, int& y1) {
  const float r = (float)radius;
  x0 = min(gx, max(0, (int)((px - r) / 16.0f)));
  y0 = min(gy, max(0, (int)((py - r) / 16.0f)));
  x1 = min(gx, max(0, (int)((((px + r) + 16.0f) - 1.0f) / 16.0f)));
  y1 = min(gy, max(0, (int)((((py + r) + 16.0f) - 1.0f) / 16.0f)));
}

struct Cov2D {
  float a, b, c;
  float T[2][3];
  float3 t;
  float xmul, ymul;
};

// computeCov2D (A.3): Sigma' = (J W) Sigma (J W)^T + 0.3 I ; c6 = xx,xy,xz,yy,yz,zz
__device__ __forceinline__ void compute_cov2d(const float3 mean, float fx, float fy, float tanx, float tany,
                                              const float* c6, const float* __restrict__ vm, Cov2D& o) {
  float3 t = transform4x3(mean, vm);
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  o.xmul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
  o.ymul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  o.t = t;
  const float J00 = fx / t.z;
  const float J02 = -(fx * t.x) / (t.z * t.z);
  const float J11 = fy / t.z;
  const float J12 = -(fy * t.y) / (t.z * t.z);
#pragma unroll
  for (int k = 0; k < 3; k++) {
    o.T[0][k] = vm[4 * k + 0] * J00 + vm[4 * k + 2] * J02;
    o.T[1][k] = vm[4 * k + 1] * J11 + vm[4 * k + 2] * J12;
  }
  const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
  float u[2][3];
#pragma unroll
  for (int a = 0; a < 2; a++)
#pragma unroll
    for (int j = 0; j < 3; j++) u[a][j] = (o.T[a][0] * S[j][0] + o.T[a][1] * S[j][1]) + o.T[a][2] * S[j][2];
  const float c00 = (u[0][0] * o.T[0][0] + u[0][1] * o.T[0][1]) + u[0][2] * o.T[0][2];
  const float c01 = (u[0][0] * o.T[1][0] + u[0][1] * o.T[1][1]) + u[0][2] * o.T[1][2];
  const float c11 = (u[1][0] * o.T[1][0] + u[1][1] * o.T[1][1]) + u[1][2] * o.T[1][2];
  o.a = c00 + 0.3f;
  o.b = c01;
  o.c = c11 + 0.3f;
}

// Real SH basis (A.5 for l<=3, A.6 for l=4).  Returns the number of active coefficients.
template <bool kDeriv>
__device__ __forceinline__ int sh_basis(int deg, int sh_l4, float x, float y, float z, float* b, float (*db)[3]) {
  int n = 1;
#pragma unroll
  for (int k = 0; k < 25; k++) {
    b[k] = 0.0f;
    if (kDeriv) db[k][0] = db[k][1] = db[k][2] = 0.0f;
  }
  b[0] = TS_SH_C0;
  if (deg > 0) {
    n = 4;
    b[1] = -TS_SH_C1 * y; b[2] = TS_SH_C1 * z; b[3] = -TS_SH_C1 * x;
    if (kDeriv) { db[1][1] = -TS_SH_C1; db[2][2] = TS_SH_C1; db[3][0] = -TS_SH_C1; }
    if (deg > 1) {
      n = 9;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = kC2[0] * xy;
      b[5] = kC2[1] * yz;
      b[6] = kC2[2] * (2.0f * zz - xx - yy);
      b[7] = kC2[3] * xz;
      b[8] = kC2[4] * (xx - yy);
      if (kDeriv) {
        db[4][0] = kC2[0] * y; db[4][1] = kC2[0] * x;
        db[5][1] = kC2[1] * z; db[5][2] = kC2[1] * y;
        db[6][0] = kC2[2] * 2.0f * -x; db[6][1] = kC2[2] * 2.0f * -y; db[6][2] = kC2[2] * 2.0f * 2.0f * z;
        db[7][0] = kC2[3] * z; db[7][2] = kC2[3] * x;
        db[8][0] = kC2[4] * 2.0f * x; db[8][1] = kC2[4] * 2.0f * -y;
      }
      if (deg > 2) {
        n = 16;
        b[9] = kC3[0] * y * (3.0f * xx - yy);
        b[10] = kC3[1] * xy * z;
        b[11] = kC3[2] * y * (4.0f * zz - xx - yy);
        b[12] = kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        b[13] = kC3[4] * x * (4.0f * zz - xx - yy);
        b[14] = kC3[5] * z * (xx - yy);
        b[15] = kC3[6] * x * (xx - 3.0f * yy);
        if (kDeriv) {
          db[9][0] = kC3[0] * 3.0f * 2.0f * xy; db[9][1] = kC3[0] * 3.0f * (xx - yy);
          db[10][0] = kC3[1] * yz; db[10][1] = kC3[1] * xz; db[10][2] = kC3[1] * xy;
          db[11][0] = kC3[2] * -2.0f * xy; db[11][1] = kC3[2] * (-3.0f * yy + 4.0f * zz - xx);
          db[11][2] = kC3[2] * 4.0f * 2.0f * yz;
          db[12][0] = kC3[3] * 3.0f * 2.0f * -xz; db[12][1] = kC3[3] * 3.0f * 2.0f * -yz;
          db[12][2] = kC3[3] * 3.0f * (2.0f * zz - xx - yy);
          db[13][0] = kC3[4] * (-3.0f * xx + 4.0f * zz - yy); db[13][1] = kC3[4] * -2.0f * xy;
          db[13][2] = kC3[4] * 4.0f * 2.0f * xz;
          db[14][0] = kC3[5] * 2.0f * xz; db[14][1] = kC3[5] * 2.0f * -yz; db[14][2] = kC3[5] * (xx - yy);
          db[15][0] = kC3[6] * 3.0f * (xx - yy); db[15][1] = kC3[6] * 3.0f * 2.0f * -xy;
        }
        if (deg > 3 && sh_l4) {
          n = 25;
          b[16] = kC4[0] * xy * (xx - yy);
          b[17] = kC4[1] * yz * (3.0f * xx - yy);
          b[18] = kC4[2] * xy * (7.0f * zz - 1.0f);
          b[19] = kC4[3] * yz * (7.0f * zz - 3.0f);
          b[20] = kC4[4] * (zz * (35.0f * zz - 30.0f) + 3.0f);
          b[21] = kC4[5] * xz * (7.0f * zz - 3.0f);
          b[22] = kC4[6] * (xx - yy) * (7.0f * zz - 1.0f);
          b[23] = kC4[7] * xz * (xx - 3.0f * yy);
          b[24] = kC4[8] * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
          if (kDeriv) {
            db[16][0] = kC4[0] * y * (3.0f * xx - yy); db[16][1] = kC4[0] * x * (xx - 3.0f * yy);
            db[17][0] = kC4[1] * 6.0f * xy * z; db[17][1] = kC4[1] * z * 3.0f * (xx - yy);
            db[17][2] = kC4[1] * y * (3.0f * xx - yy);
            db[18][0] = kC4[2] * y * (7.0f * zz - 1.0f); db[18][1] = kC4[2] * x * (7.0f * zz - 1.0f);
            db[18][2] = kC4[2] * 14.0f * xy * z;
            db[19][1] = kC4[3] * z * (7.0f * zz - 3.0f); db[19][2] = kC4[3] * y * (21.0f * zz - 3.0f);
            db[20][2] = kC4[4] * z * (140.0f * zz - 60.0f);
            db[21][0] = kC4[5] * z * (7.0f * zz - 3.0f); db[21][2] = kC4[5] * x * (21.0f * zz - 3.0f);
            db[22][0] = kC4[6] * 2.0f * x * (7.0f * zz - 1.0f); db[22][1] = kC4[6] * -2.0f * y * (7.0f * zz - 1.0f);
            db[22][2] = kC4[6] * 14.0f * z * (xx - yy);
            db[23][0] = kC4[7] * z * 3.0f * (xx - yy); db[23][1] = kC4[7] * -6.0f * xy * z;
            db[23][2] = kC4[7] * x * (xx - 3.0f * yy);
            db[24][0] = kC4[8] * 4.0f * x * (xx - 3.0f * yy); db[24][1] = kC4[8] * -4.0f * y * (3.0f * xx - yy);
          }
        }
      }
    }
  }
  return n;
}

// depth feature f(z) of render_depth_cuda (cuda_splatting.py:388-400) and df/dz
__device__ __forceinline__ float depth_feature(int mode, float z, float near_, float far_, float& dfdz) {
  const float eps = 1e-10f;
  switch (mode) {
    case TSPLAT_DEPTH_DEPTH: dfdz = 1.0f; return z;
    case TSPLAT_DEPTH_DISPARITY: { const float r = 1.0f / z; dfdz = -(r * r); return r; }
    case TSPLAT_DEPTH_RELATIVE_DISPARITY: {  // conversions.py:17-27
      const float disp_near = 1.0f / (near_ + eps), disp_far = 1.0f / (far_ + eps);
      const float disp = 1.0f / (z + eps);
      const float den = (disp_near - disp_far) + eps;
      dfdz = (disp * disp) / den;
      return 1.0f - (disp - disp_far) / den;
    }
    case TSPLAT_DEPTH_LOG:  // z.minimum(near).maximum(far).log(): log(far) whenever near < far
      dfdz = (fminf(z, near_) >= far_) ? ((z <= near_) ? 1.0f / z : 0.0f) : 0.0f;
      return logf(fmaxf(fminf(z, near_), far_));
    default: dfdz = 0.0f; return 0.0f;
  }
}

// index of coefficient k / channel ch of Gaussian-local SH block (floats), for both layouts
__device__ __forceinline__ int sh_idx(int layout, int M, int k, int ch) {
  return layout == TSPLAT_SH_GM3 ? k * 3 + ch : ch * M + k;
}

}  // namespace tsplat
