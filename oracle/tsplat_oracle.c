/*
 * tsplat_oracle.c — CPU restatement of the reference rasterizer.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's library.  The product (transplat_b200/) never does.
 *
 * PARITY UNPINNED.  The algorithm restated here lives in a third-party dependency that is absent
 * from /root/reference: `diff_gaussian_rasterization`, installed by the reference from
 * git+https://github.com/dcharatan/diff-gaussian-rasterization-modified (requirements.txt:17, no
 * commit pinned), a fork of graphdeco-inria/diff-gaussian-rasterization.  Its source is nowhere on
 * this machine and the reference holds no test, golden vector or fixture for the path
 * (SURVEY.md 8c).  What is restated is the published 3DGS rasterizer algorithm as reconstructed in
 * SURVEY.md Appendix A (A.1 constants, A.2 helpers, A.3 forward, A.4 backward, A.5 spherical
 * harmonics, A.6 the l=4 band, A.7 depth semantics), anchored on the reference's own call sites:
 *   /root/reference/src/model/decoder/cuda_splatting.py:108-133  (the call; argument meaning)
 *   /root/reference/src/model/decoder/cuda_splatting.py:73-96    (scale invariance, matrices)
 *   /root/reference/src/model/decoder/cuda_splatting.py:375-417  (depth = a degree-0 colour pass)
 *   /root/reference/src/model/encoder/matching/conversions.py:17-27 (relative disparity)
 * The oracle is pinned only by the closed-form known-answer tests of tests/kat.py (run through this file by
 * tests/test_oracle_kat.py and through the CUDA library by tests/test_gpu_kat.py) and by a float64 dense autograd
 * formulation of the same math that takes its own discrete decisions (oracle/torch_dense.py,
 * tests/test_oracle_grad.py).
 *
 * Floating-point contract: every quantity that feeds an integer decision (depth bits, radius,
 * tile rectangle, tiles_touched) is computed in IEEE binary32, one rounding per operation,
 * operations in the left-to-right order written below, no fused multiply-add (build with
 * -ffp-contract=off).  ndc2Pix is evaluated in binary64 as upstream writes it with double
 * literals.  The sm_100a kernels follow the same order with contraction disabled.
 *
 * Build: see oracle/Makefile.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/tsplat.h"

#define BLOCK 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};
/* the l=4 band of the fork (svox2 / plenoxels constants) */
static const float SH_C4[9] = {2.5033429417967046f,  -1.7701307697799304f, 0.9461746957575601f,
                               -0.6690465435572892f, 0.10578554691520431f, -0.6690465435572892f,
                               0.47308734787878004f, -1.7701307697799304f, 0.6258357354491761f};

/* ---------------------------------------------------------------- helpers ------ */

/* CUDA float->int conversion saturates and maps NaN to 0; C leaves both undefined. */
static int f2i_trunc(float v) {
  if (v != v) return 0;
  if (v >= 2147483648.0f) return 2147483647;
  if (v <= -2147483648.0f) return (-2147483647 - 1);
  return (int)v;
}
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

static void transform4x3(const float* p, const float* m, float* o) {
  o[0] = ((m[0] * p[0] + m[4] * p[1]) + m[8] * p[2]) + m[12];
  o[1] = ((m[1] * p[0] + m[5] * p[1]) + m[9] * p[2]) + m[13];
  o[2] = ((m[2] * p[0] + m[6] * p[1]) + m[10] * p[2]) + m[14];
}
static void transform4x4(const float* p, const float* m, float* o) {
  o[0] = ((m[0] * p[0] + m[4] * p[1]) + m[8] * p[2]) + m[12];
  o[1] = ((m[1] * p[0] + m[5] * p[1]) + m[9] * p[2]) + m[13];
  o[2] = ((m[2] * p[0] + m[6] * p[1]) + m[10] * p[2]) + m[14];
  o[3] = ((m[3] * p[0] + m[7] * p[1]) + m[11] * p[2]) + m[15];
}
/* upstream writes ((v + 1.0) * S - 1.0) * 0.5 with double literals */
static float ndc2pix(float v, int S) {
  return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5);
}
static void get_rect(float px, float py, int radius, int gx, int gy, int* rmin, int* rmax) {
  float r = (float)radius;
  rmin[0] = imin(gx, imax(0, f2i_trunc((px - r) / 16.0f)));
  rmin[1] = imin(gy, imax(0, f2i_trunc((py - r) / 16.0f)));
  rmax[0] = imin(gx, imax(0, f2i_trunc((((px + r) + 16.0f) - 1.0f) / 16.0f)));
  rmax[1] = imin(gy, imax(0, f2i_trunc((((py + r) + 16.0f) - 1.0f) / 16.0f)));
}

int oracle_higher_msb(uint32_t n) {
  uint32_t msb = 16, step = 16;
  while (step > 1) {
    step /= 2;
    if (n >> msb)
      msb += step;
    else
      msb -= step;
  }
  if (n >> msb) msb++;
  return (int)msb;
}

/* Scaled mean / covariance exactly as the reference host code produces them
 * (cuda_splatting.py:76-77: gaussian_covariances * scale**2, gaussian_means * scale). */
static void load_mean(const float* means, int i, float s, float* m) {
  m[0] = means[3 * i + 0] * s;
  m[1] = means[3 * i + 1] * s;
  m[2] = means[3 * i + 2] * s;
}
static void load_cov6(const float* cov, int layout, int i, float s2, float* c) {
  if (layout == TSPLAT_COV_TRIU6) {
    for (int k = 0; k < 6; k++) c[k] = cov[6 * i + k] * s2;
  } else { /* [3,3] -> (0,0) (0,1) (0,2) (1,1) (1,2) (2,2) (cuda_splatting.py:124) */
    static const int idx[6] = {0, 1, 2, 4, 5, 8};
    for (int k = 0; k < 6; k++) c[k] = cov[9 * i + idx[k]] * s2;
  }
}
static float load_sh(const float* sh, int layout, int M, int i, int k, int ch) {
  if (M == 0) return sh[3 * i + ch];
  return layout == TSPLAT_SH_GM3 ? sh[((size_t)i * M + k) * 3 + ch] : sh[((size_t)i * 3 + ch) * M + k];
}
static size_t sh_index(int layout, int M, int i, int k, int ch) {
  return layout == TSPLAT_SH_GM3 ? ((size_t)i * M + k) * 3 + ch : ((size_t)i * 3 + ch) * M + k;
}

/* 2-D covariance (computeCov2D).  Outputs a,b,c after the 0.3 dilation, the 2x3
 * matrix T = J*W (row a = screen axis), the clamped camera-space point t and the clamp masks. */
typedef struct {
  float a, b, c;
  float T[2][3];
  float t[3];
  float xmul, ymul;
  float J00, J02, J11, J12;
} Cov2D;

static void compute_cov2d(const float* mean, float fx, float fy, float tanx, float tany,
                          const float* c6, const float* vm, Cov2D* o) {
  float t[3];
  transform4x3(mean, vm, t);
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  const float txtz = t[0] / t[2], tytz = t[1] / t[2];
  o->xmul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
  o->ymul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
  t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
  t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
  o->t[0] = t[0]; o->t[1] = t[1]; o->t[2] = t[2];
  const float J00 = fx / t[2];
  const float J02 = -(fx * t[0]) / (t[2] * t[2]);
  const float J11 = fy / t[2];
  const float J12 = -(fy * t[1]) / (t[2] * t[2]);
  o->J00 = J00; o->J02 = J02; o->J11 = J11; o->J12 = J12;
  /* R[r][c] = vm[4c + r]; T[0][k] = J00 R[0][k] + J02 R[2][k]; T[1][k] = J11 R[1][k] + J12 R[2][k] */
  for (int k = 0; k < 3; k++) {
    o->T[0][k] = vm[4 * k + 0] * J00 + vm[4 * k + 2] * J02;
    o->T[1][k] = vm[4 * k + 1] * J11 + vm[4 * k + 2] * J12;
  }
  const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
  float u[2][3];
  for (int a = 0; a < 2; a++)
    for (int j = 0; j < 3; j++)
      u[a][j] = (o->T[a][0] * S[j][0] + o->T[a][1] * S[j][1]) + o->T[a][2] * S[j][2];
  const float c00 = (u[0][0] * o->T[0][0] + u[0][1] * o->T[0][1]) + u[0][2] * o->T[0][2];
  const float c01 = (u[0][0] * o->T[1][0] + u[0][1] * o->T[1][1]) + u[0][2] * o->T[1][2];
  const float c11 = (u[1][0] * o->T[1][0] + u[1][1] * o->T[1][1]) + u[1][2] * o->T[1][2];
  o->a = c00 + 0.3f;
  o->b = c01;
  o->c = c11 + 0.3f;
}

/* SH basis values and their plain partial derivatives. */
static int sh_basis(int deg, int sh_l4, float x, float y, float z, float* b, float (*db)[3]) {
  int n = 1;
  for (int k = 0; k < 25; k++) { b[k] = 0.0f; if (db) db[k][0] = db[k][1] = db[k][2] = 0.0f; }
  b[0] = SH_C0;
  if (deg > 0) {
    n = 4;
    b[1] = -SH_C1 * y; b[2] = SH_C1 * z; b[3] = -SH_C1 * x;
    if (db) { db[1][1] = -SH_C1; db[2][2] = SH_C1; db[3][0] = -SH_C1; }
    if (deg > 1) {
      n = 9;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      b[4] = SH_C2[0] * xy;
      b[5] = SH_C2[1] * yz;
      b[6] = SH_C2[2] * (2.0f * zz - xx - yy);
      b[7] = SH_C2[3] * xz;
      b[8] = SH_C2[4] * (xx - yy);
      if (db) {
        db[4][0] = SH_C2[0] * y; db[4][1] = SH_C2[0] * x;
        db[5][1] = SH_C2[1] * z; db[5][2] = SH_C2[1] * y;
        db[6][0] = SH_C2[2] * 2.0f * -x; db[6][1] = SH_C2[2] * 2.0f * -y; db[6][2] = SH_C2[2] * 2.0f * 2.0f * z;
        db[7][0] = SH_C2[3] * z; db[7][2] = SH_C2[3] * x;
        db[8][0] = SH_C2[4] * 2.0f * x; db[8][1] = SH_C2[4] * 2.0f * -y;
      }
      if (deg > 2) {
        n = 16;
        b[9] = SH_C3[0] * y * (3.0f * xx - yy);
        b[10] = SH_C3[1] * xy * z;
        b[11] = SH_C3[2] * y * (4.0f * zz - xx - yy);
        b[12] = SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
        b[13] = SH_C3[4] * x * (4.0f * zz - xx - yy);
        b[14] = SH_C3[5] * z * (xx - yy);
        b[15] = SH_C3[6] * x * (xx - 3.0f * yy);
        if (db) {
          db[9][0] = SH_C3[0] * 3.0f * 2.0f * xy; db[9][1] = SH_C3[0] * 3.0f * (xx - yy);
          db[10][0] = SH_C3[1] * yz; db[10][1] = SH_C3[1] * xz; db[10][2] = SH_C3[1] * xy;
          db[11][0] = SH_C3[2] * -2.0f * xy; db[11][1] = SH_C3[2] * (-3.0f * yy + 4.0f * zz - xx);
          db[11][2] = SH_C3[2] * 4.0f * 2.0f * yz;
          db[12][0] = SH_C3[3] * 3.0f * 2.0f * -xz; db[12][1] = SH_C3[3] * 3.0f * 2.0f * -yz;
          db[12][2] = SH_C3[3] * 3.0f * (2.0f * zz - xx - yy);
          db[13][0] = SH_C3[4] * (-3.0f * xx + 4.0f * zz - yy); db[13][1] = SH_C3[4] * -2.0f * xy;
          db[13][2] = SH_C3[4] * 4.0f * 2.0f * xz;
          db[14][0] = SH_C3[5] * 2.0f * xz; db[14][1] = SH_C3[5] * 2.0f * -yz; db[14][2] = SH_C3[5] * (xx - yy);
          db[15][0] = SH_C3[6] * 3.0f * (xx - yy); db[15][1] = SH_C3[6] * 3.0f * 2.0f * -xy;
        }
        if (deg > 3 && sh_l4) {
          n = 25;
          b[16] = SH_C4[0] * xy * (xx - yy);
          b[17] = SH_C4[1] * yz * (3.0f * xx - yy);
          b[18] = SH_C4[2] * xy * (7.0f * zz - 1.0f);
          b[19] = SH_C4[3] * yz * (7.0f * zz - 3.0f);
          b[20] = SH_C4[4] * (zz * (35.0f * zz - 30.0f) + 3.0f);
          b[21] = SH_C4[5] * xz * (7.0f * zz - 3.0f);
          b[22] = SH_C4[6] * (xx - yy) * (7.0f * zz - 1.0f);
          b[23] = SH_C4[7] * xz * (xx - 3.0f * yy);
          b[24] = SH_C4[8] * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
          if (db) {
            db[16][0] = SH_C4[0] * y * (3.0f * xx - yy); db[16][1] = SH_C4[0] * x * (xx - 3.0f * yy);
            db[17][0] = SH_C4[1] * 6.0f * xy * z; db[17][1] = SH_C4[1] * z * 3.0f * (xx - yy);
            db[17][2] = SH_C4[1] * y * (3.0f * xx - yy);
            db[18][0] = SH_C4[2] * y * (7.0f * zz - 1.0f); db[18][1] = SH_C4[2] * x * (7.0f * zz - 1.0f);
            db[18][2] = SH_C4[2] * 14.0f * xy * z;
            db[19][1] = SH_C4[3] * z * (7.0f * zz - 3.0f); db[19][2] = SH_C4[3] * y * (21.0f * zz - 3.0f);
            db[20][2] = SH_C4[4] * z * (140.0f * zz - 60.0f);
            db[21][0] = SH_C4[5] * z * (7.0f * zz - 3.0f); db[21][2] = SH_C4[5] * x * (21.0f * zz - 3.0f);
            db[22][0] = SH_C4[6] * 2.0f * x * (7.0f * zz - 1.0f); db[22][1] = SH_C4[6] * -2.0f * y * (7.0f * zz - 1.0f);
            db[22][2] = SH_C4[6] * 14.0f * z * (xx - yy);
            db[23][0] = SH_C4[7] * z * 3.0f * (xx - yy); db[23][1] = SH_C4[7] * -6.0f * xy * z;
            db[23][2] = SH_C4[7] * x * (xx - 3.0f * yy);
            db[24][0] = SH_C4[8] * 4.0f * x * (xx - 3.0f * yy); db[24][1] = SH_C4[8] * -4.0f * y * (3.0f * xx - yy);
          }
        }
      }
    }
  }
  return n;
}

/* depth feature f(z) of render_depth_cuda (cuda_splatting.py:388-400) and df/dz */
static float depth_feature(int mode, float z, float near_, float far_, float* dfdz) {
  const float eps = 1e-10f;
  switch (mode) {
    case TSPLAT_DEPTH_DEPTH: *dfdz = 1.0f; return z;
    case TSPLAT_DEPTH_DISPARITY: { float r = 1.0f / z; *dfdz = -(r * r); return r; }
    case TSPLAT_DEPTH_RELATIVE_DISPARITY: { /* conversions.py:17-27 */
      const float disp_near = 1.0f / (near_ + eps), disp_far = 1.0f / (far_ + eps);
      const float disp = 1.0f / (z + eps);
      const float den = (disp_near - disp_far) + eps;
      *dfdz = (disp * disp) / den;
      return 1.0f - (disp - disp_far) / den;
    }
    case TSPLAT_DEPTH_LOG: /* z.minimum(near).maximum(far).log()  (cuda_splatting.py:400) */
      *dfdz = (fminf(z, near_) >= far_) ? ((z <= near_) ? 1.0f / z : 0.0f) : 0.0f;
      return logf(fmaxf(fminf(z, near_), far_));
    default: *dfdz = 0.0f; return 0.0f;
  }
}

/* --------------------------------------------------------------- forward ------- */

typedef struct OracleView {
  int32_t G, H, W;
  int32_t sh_degree, sh_coeffs, sh_layout, cov_layout, depth_mode, sh_l4;
  const float* means;
  const float* cov;
  const float* sh;
  const float* opacity;
  const TsplatCamera* cam; /* HOST pointer here */
} OracleView;

/* preprocessCUDA.  Outputs are zero for culled Gaussians. */
int oracle_preprocess(const OracleView* v, float* depths, int32_t* radii, float* xy,
                      float* conic_opacity, float* rgbd, uint32_t* tiles_touched, uint8_t* clamped) {
  const TsplatCamera* cam = v->cam;
  const int G = v->G, W = v->W, H = v->H, M = v->sh_coeffs;
  const float fx = (float)W / (2.0f * cam->tanfovx), fy = (float)H / (2.0f * cam->tanfovy);
  const int gx = (W + BLOCK - 1) / BLOCK, gy = (H + BLOCK - 1) / BLOCK;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < G; i++) {
    depths[i] = 0.0f; radii[i] = 0; tiles_touched[i] = 0; clamped[i] = 0;
    xy[2 * i] = xy[2 * i + 1] = 0.0f;
    for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = rgbd[4 * i + k] = 0.0f;
    float m[3], c6[6], pv[3], ph[4];
    load_mean(v->means, i, cam->mean_scale, m);
    transform4x3(m, cam->view, pv);
    if (pv[2] <= 0.2f) continue; /* in_frustum */
    transform4x4(m, cam->proj, ph);
    const float pw = 1.0f / (ph[3] + 0.0000001f);
    const float ppx = ph[0] * pw, ppy = ph[1] * pw;
    load_cov6(v->cov, v->cov_layout, i, cam->cov_scale, c6);
    Cov2D q;
    compute_cov2d(m, fx, fy, cam->tanfovx, cam->tanfovy, c6, cam->view, &q);
    const float det = q.a * q.c - q.b * q.b;
    if (det == 0.0f) continue;
    const float det_inv = 1.0f / det;
    const float con[3] = {q.c * det_inv, -q.b * det_inv, q.a * det_inv};
    const float mid = 0.5f * (q.a + q.c);
    const float sq = sqrtf(fmaxf(0.1f, mid * mid - det));
    const float l1 = mid + sq, l2 = mid - sq;
    const float rad_f = ceilf(3.0f * sqrtf(fmaxf(l1, l2)));
    const int rad = f2i_trunc(rad_f);
    const float px = ndc2pix(ppx, W), py = ndc2pix(ppy, H);
    int rmin[2], rmax[2];
    get_rect(px, py, rad, gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;

    /* colour: computeColorFromSH (A.5) or colors_precomp */
    float rgb[3];
    uint8_t cl = 0;
    if (M > 0) {
      float d[3] = {m[0] - cam->campos[0], m[1] - cam->campos[1], m[2] - cam->campos[2]};
      const float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
      d[0] /= len; d[1] /= len; d[2] /= len;
      float b[25];
      const int n = sh_basis(v->sh_degree, v->sh_l4, d[0], d[1], d[2], b, NULL);
      for (int ch = 0; ch < 3; ch++) {
        float acc = 0.0f;
        for (int k = 0; k < n; k++) acc += b[k] * load_sh(v->sh, v->sh_layout, M, i, k, ch);
        acc += 0.5f;
        if (acc < 0.0f) cl |= (uint8_t)(1u << ch);
        rgb[ch] = fmaxf(acc, 0.0f);
      }
    } else {
      for (int ch = 0; ch < 3; ch++) rgb[ch] = v->sh[3 * i + ch];
    }
    /* fused depth channel (A.7): a degree-0 "colour" whose coefficient is f(z_cam) */
    float fd = 0.0f;
    if (v->depth_mode != TSPLAT_DEPTH_NONE) {
      const float* r = cam->depth_row;
      const float* mu = v->means + 3 * (size_t)i;
      const float z = ((r[0] * mu[0] + r[1] * mu[1]) + r[2] * mu[2]) + r[3];
      float dfdz;
      const float f = depth_feature(v->depth_mode, z, cam->near_plane, cam->far_plane, &dfdz);
      fd = SH_C0 * f + 0.5f;
      if (fd < 0.0f) cl |= 8u;
      fd = fmaxf(fd, 0.0f);
    }
    depths[i] = pv[2];
    radii[i] = rad;
    xy[2 * i] = px; xy[2 * i + 1] = py;
    conic_opacity[4 * i + 0] = con[0]; conic_opacity[4 * i + 1] = con[1];
    conic_opacity[4 * i + 2] = con[2]; conic_opacity[4 * i + 3] = v->opacity[i];
    rgbd[4 * i + 0] = rgb[0]; rgbd[4 * i + 1] = rgb[1]; rgbd[4 * i + 2] = rgb[2]; rgbd[4 * i + 3] = fd;
    tiles_touched[i] = (uint32_t)((rmax[1] - rmin[1]) * (rmax[0] - rmin[0]));
    clamped[i] = cl;
  }
  return 0;
}

/* cub::DeviceScan::InclusiveSum (X4); returns D */
uint64_t oracle_scan(int G, const uint32_t* tiles_touched, uint32_t* offsets) {
  uint64_t acc = 0;
  for (int i = 0; i < G; i++) { acc += tiles_touched[i]; offsets[i] = (uint32_t)acc; }
  return acc;
}

/* duplicateWithKeys (X5) */
void oracle_duplicate_keys(int G, int H, int W, const float* xy, const float* depths,
                           const int32_t* radii, const uint32_t* offsets, uint64_t* keys,
                           uint32_t* vals) {
  const int gx = (W + BLOCK - 1) / BLOCK, gy = (H + BLOCK - 1) / BLOCK;
  for (int i = 0; i < G; i++) {
    if (radii[i] <= 0) continue;
    uint32_t off = (i == 0) ? 0 : offsets[i - 1];
    int rmin[2], rmax[2];
    get_rect(xy[2 * i], xy[2 * i + 1], radii[i], gx, gy, rmin, rmax);
    uint32_t dbits;
    memcpy(&dbits, &depths[i], 4);
    for (int y = rmin[1]; y < rmax[1]; y++)
      for (int x = rmin[0]; x < rmax[0]; x++) {
        uint64_t key = (uint64_t)(uint32_t)(y * gx + x);
        key <<= 32;
        key |= dbits;
        keys[off] = key;
        vals[off] = (uint32_t)i;
        off++;
      }
  }
}

/* cub::DeviceRadixSort::SortPairs(keys, vals, D, 0, bits) (X6): stable LSD radix sort */
int oracle_sort_pairs(uint64_t D, uint64_t* keys, uint32_t* vals, int bits) {
  if (D == 0) return 0;
  uint64_t* k2 = (uint64_t*)malloc(D * sizeof(uint64_t));
  uint32_t* v2 = (uint32_t*)malloc(D * sizeof(uint32_t));
  if (!k2 || !v2) { free(k2); free(v2); return -1; }
  uint64_t *ka = keys, *kb = k2;
  uint32_t *va = vals, *vb = v2;
  for (int shift = 0; shift < bits; shift += 8) {
    const int nb = (bits - shift) < 8 ? (bits - shift) : 8;
    const uint64_t mask = ((uint64_t)1 << nb) - 1;
    uint64_t count[257];
    memset(count, 0, sizeof(count));
    for (uint64_t i = 0; i < D; i++) count[((ka[i] >> shift) & mask) + 1]++;
    for (int d = 0; d < 256; d++) count[d + 1] += count[d];
    for (uint64_t i = 0; i < D; i++) {
      const uint64_t d = (ka[i] >> shift) & mask;
      kb[count[d]] = ka[i];
      vb[count[d]] = va[i];
      count[d]++;
    }
    uint64_t* tk = ka; ka = kb; kb = tk;
    uint32_t* tv = va; va = vb; vb = tv;
  }
  if (ka != keys) { memcpy(keys, ka, D * sizeof(uint64_t)); memcpy(vals, va, D * sizeof(uint32_t)); }
  free(k2); free(v2);
  return 0;
}

/* identifyTileRanges (X7); ranges [T][2] pre-zeroed by this function */
void oracle_tile_ranges(uint64_t D, const uint64_t* keys, int T, uint32_t* ranges) {
  memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)T);
  for (uint64_t i = 0; i < D; i++) {
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[2 * t] = 0;
    else {
      const uint32_t p = (uint32_t)(keys[i - 1] >> 32);
      if (t != p) { ranges[2 * p + 1] = (uint32_t)i; ranges[2 * t] = (uint32_t)i; }
    }
    if (i == D - 1) ranges[2 * t + 1] = (uint32_t)D;
  }
}

/* renderCUDA forward (X8).  CH = 3, or 4 with the fused depth channel.  out is planar [CH,H,W]
 * except that with CH==4 channel 3 goes to out_depth [H,W] as ((x+x)+x)/3
 * (cuda_splatting.py:417 takes the mean of three identical channels). */
void oracle_render_forward(int H, int W, int CH, const uint32_t* ranges, const uint32_t* vals,
                           const float* xy, const float* conic_opacity, const float* rgbd,
                           const float* bg, float* out_color, float* out_depth, float* final_T,
                           uint32_t* n_contrib) {
  const int gx = (W + BLOCK - 1) / BLOCK, gy = (H + BLOCK - 1) / BLOCK;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK; ly++)
      for (int lx = 0; lx < BLOCK; lx++) {
        const int px = tx * BLOCK + lx, py = ty * BLOCK + ly;
        if (px >= W || py >= H) continue;
        const float pxf = (float)px, pyf = (float)py;
        float T = 1.0f, C[4] = {0, 0, 0, 0};
        uint32_t contributor = 0, last = 0;
        for (uint32_t e = r0; e < r1; e++) {
          contributor++;
          const uint32_t g = vals[e];
          const float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
          const float* co = conic_opacity + 4 * (size_t)g;
          const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0.0f) continue;
          const float alpha = fminf(0.99f, co[3] * expf(power));
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = T * (1.0f - alpha);
          if (test_T < 0.0001f) break;
          for (int ch = 0; ch < CH; ch++) C[ch] += rgbd[4 * (size_t)g + ch] * alpha * T;
          T = test_T;
          last = contributor;
        }
        const size_t pix = (size_t)py * W + px;
        final_T[pix] = T;
        n_contrib[pix] = last;
        for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pix] = C[ch] + T * bg[ch];
        if (CH == 4) { const float x = C[3]; out_depth[pix] = ((x + x) + x) / 3.0f; }
      }
  }
}

/* Pixels whose result hangs on a float comparison that sits within rounding distance of its threshold: the same walk
 * as oracle_render_forward, flag[pix] = 1 if for some evaluated entry
 *   power is within eps of 0 (relative to its terms), or alpha within eps of 1/255, or T (1 - alpha) within eps of 1e-4.
 * On such a pixel an implementation whose exp() differs in the last ulps (device ex2-based exp vs glibc expf) may
 * legitimately take the other branch: one contribution of alpha ~ 1/255 more or less.  The parity tests hold every
 * other pixel to the stated tolerance and these to a looser, explicit one (tests/parity.py). */
void oracle_render_borderline(int H, int W, const uint32_t* ranges, const uint32_t* vals, const float* xy,
                              const float* conic_opacity, float eps, uint8_t* flag) {
  const int gx = (W + BLOCK - 1) / BLOCK, gy = (H + BLOCK - 1) / BLOCK;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK; ly++)
      for (int lx = 0; lx < BLOCK; lx++) {
        const int px = tx * BLOCK + lx, py = ty * BLOCK + ly;
        if (px >= W || py >= H) continue;
        const float pxf = (float)px, pyf = (float)py;
        float T = 1.0f;
        uint8_t f = 0;
        for (uint32_t e = r0; e < r1; e++) {
          const uint32_t g = vals[e];
          const float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
          const float* co = conic_opacity + 4 * (size_t)g;
          const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          const float mag = 0.5f * (fabsf(co[0]) * dx * dx + fabsf(co[2]) * dy * dy) + fabsf(co[1] * dx * dy);
          if (fabsf(power) <= eps * mag && co[3] >= 1.0f / 255.0f) f = 1;
          if (power > 0.0f) continue;
          const float alpha = fminf(0.99f, co[3] * expf(power));
          /* two exp implementations, each within ~2 ulp, plus the rounding of the argument (an absolute error
           * ~6e-8 |power| in the exponent is a relative one in exp) and of o * exp: eps (1 + 0.3 |power|), eps = 4e-7 */
          const float rel = eps * (1.0f + 0.3f * fabsf(power));
          if (fabsf(alpha - 1.0f / 255.0f) <= rel * (1.0f / 255.0f)) f = 1;
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = T * (1.0f - alpha);
          /* T is a product of up to hundreds of factors (1 - alpha_i), each off by alpha_i rel / (1 - alpha_i):
           * with sum(alpha_i) ~ ln(1e4) at the threshold the products differ by up to ~50 eps */
          if (fabsf(test_T - 0.0001f) <= 50.0f * eps * 0.0001f) f = 1;
          if (test_T < 0.0001f) break;
          T = test_T;
        }
        flag[(size_t)py * W + px] = f;
      }
  }
}

/* --------------------------------------------------------------- backward ------ */

/* renderCUDA backward (X9).  Per-pixel arithmetic in binary32 as the reference; the per-Gaussian
 * sums (atomicAdd in the reference, order undefined) are accumulated here in binary64.
 * acc [G][10]: dmean2D.x, dmean2D.y, dconic.x, dconic.y, dconic.w, dopacity, dcolor[0..3]. */
void oracle_render_backward(int H, int W, int CH, const uint32_t* ranges, const uint32_t* vals,
                            const float* xy, const float* conic_opacity, const float* rgbd,
                            const float* bg, const float* final_T, const uint32_t* n_contrib,
                            const float* dL_dcolor, const float* dL_ddepth, double* acc) {
  const int gx = (W + BLOCK - 1) / BLOCK, gy = (H + BLOCK - 1) / BLOCK;
  const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < BLOCK; ly++)
      for (int lx = 0; lx < BLOCK; lx++) {
        const int px = tx * BLOCK + lx, py = ty * BLOCK + ly;
        if (px >= W || py >= H) continue;
        const size_t pix = (size_t)py * W + px;
        const float pxf = (float)px, pyf = (float)py;
        const float T_final = final_T[pix];
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const uint32_t last_contributor = n_contrib[pix];
        float accum_rec[4] = {0, 0, 0, 0}, last_color[4] = {0, 0, 0, 0}, dL_dpix[4] = {0, 0, 0, 0};
        float bg4[4] = {bg[0], bg[1], bg[2], 0.0f};
        for (int ch = 0; ch < 3; ch++) dL_dpix[ch] = dL_dcolor[(size_t)ch * H * W + pix];
        if (CH == 4) dL_dpix[3] = dL_ddepth[pix];
        float last_alpha = 0.0f;
        for (uint32_t e = r1; e-- > r0;) {
          contributor--;
          if (contributor >= last_contributor) continue;
          const uint32_t g = vals[e];
          const float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
          const float* co = conic_opacity + 4 * (size_t)g;
          const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0.0f) continue;
          const float Gv = expf(power);
          const float alpha = fminf(0.99f, co[3] * Gv);
          if (alpha < 1.0f / 255.0f) continue;
          T = T / (1.0f - alpha);
          const float dchannel_dcolor = alpha * T;
          float dL_dalpha = 0.0f;
          double* a = acc + 10 * (size_t)g;
          for (int ch = 0; ch < CH; ch++) {
            const float c = rgbd[4 * (size_t)g + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.0f - last_alpha) * accum_rec[ch];
            last_color[ch] = c;
            dL_dalpha += (c - accum_rec[ch]) * dL_dpix[ch];
            const double add = (double)(dchannel_dcolor * dL_dpix[ch]);
#pragma omp atomic
            a[6 + ch] += add;
          }
          dL_dalpha *= T;
          last_alpha = alpha;
          float bg_dot = 0.0f;
          for (int ch = 0; ch < CH; ch++) bg_dot += bg4[ch] * dL_dpix[ch];
          dL_dalpha += (-T_final / (1.0f - alpha)) * bg_dot;
          const float dL_dG = co[3] * dL_dalpha;
          const float gdx = Gv * dx, gdy = Gv * dy;
          const float dG_ddelx = -gdx * co[0] - gdy * co[1];
          const float dG_ddely = -gdy * co[2] - gdx * co[1];
          const double add0 = (double)(dL_dG * dG_ddelx * ddelx_dx);
          const double add1 = (double)(dL_dG * dG_ddely * ddely_dy);
          const double add2 = (double)(-0.5f * gdx * dx * dL_dG);
          const double add3 = (double)(-0.5f * gdx * dy * dL_dG);
          const double add4 = (double)(-0.5f * gdy * dy * dL_dG);
          const double add5 = (double)(Gv * dL_dalpha);
#pragma omp atomic
          a[0] += add0;
#pragma omp atomic
          a[1] += add1;
#pragma omp atomic
          a[2] += add2;
#pragma omp atomic
          a[3] += add3;
#pragma omp atomic
          a[4] += add4;
#pragma omp atomic
          a[5] += add5;
        }
      }
  }
}

/* computeCov2DCUDA + preprocessCUDA backward (X10, X11), one view.  Gradients of this view are
 * ADDED into the double accumulators so the caller can sum several views of one scene
 * (the reference sums them through autograd of `repeat`, decoder_splatting_cuda.py:59-62).
 *   g_means [G,3], g_cov [G,6] or [G,9], g_sh (layout of v->sh), g_opacity [G]  (binary64)
 *   g_mean2d [G,3] float or NULL: the screen-space gradient the reference returns for `means2D`. */
void oracle_preprocess_backward(const OracleView* v, const int32_t* radii, const uint8_t* clamped,
                                const double* acc, double* g_means, double* g_cov, double* g_sh,
                                double* g_opacity, float* g_mean2d) {
  const TsplatCamera* cam = v->cam;
  const int G = v->G, W = v->W, H = v->H, M = v->sh_coeffs;
  const float fx = (float)W / (2.0f * cam->tanfovx), fy = (float)H / (2.0f * cam->tanfovy);
  const float s = cam->mean_scale, s2 = cam->cov_scale;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < G; i++) {
    if (g_mean2d) g_mean2d[3 * i] = g_mean2d[3 * i + 1] = g_mean2d[3 * i + 2] = 0.0f;
    if (!(radii[i] > 0)) continue;
    const double* a = acc + 10 * (size_t)i;
    const float dmx = (float)a[0], dmy = (float)a[1];
    const float dcon[3] = {(float)a[2], (float)a[3], (float)a[4]};
    float dcol[4] = {(float)a[6], (float)a[7], (float)a[8], (float)a[9]};
    if (g_mean2d) { g_mean2d[3 * i] = dmx; g_mean2d[3 * i + 1] = dmy; }
    g_opacity[i] += a[5];

    float m[3], c6[6];
    load_mean(v->means, i, s, m);
    load_cov6(v->cov, v->cov_layout, i, s2, c6);
    /* ---- cov2D backward ---- */
    Cov2D q;
    compute_cov2d(m, fx, fy, cam->tanfovx, cam->tanfovy, c6, cam->view, &q);
    const float aa = q.a, bb = q.b, cc = q.c;
    const float denom = aa * cc - bb * bb;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float dcov[6] = {0, 0, 0, 0, 0, 0};
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
    for (int k = 0; k < 3; k++) {
      const float t0 = T[0][0] * S[k][0] + T[0][1] * S[k][1] + T[0][2] * S[k][2];
      const float t1 = T[1][0] * S[k][0] + T[1][1] * S[k][1] + T[1][2] * S[k][2];
      dT[0][k] = 2 * t0 * dL_da + t1 * dL_db;
      dT[1][k] = 2 * t1 * dL_dc + t0 * dL_db;
    }
    const float* vm = cam->view; /* R[r][c] = vm[4c+r] */
    const float dJ00 = vm[0] * dT[0][0] + vm[4] * dT[0][1] + vm[8] * dT[0][2];
    const float dJ02 = vm[2] * dT[0][0] + vm[6] * dT[0][1] + vm[10] * dT[0][2];
    const float dJ11 = vm[1] * dT[1][0] + vm[5] * dT[1][1] + vm[9] * dT[1][2];
    const float dJ12 = vm[2] * dT[1][0] + vm[6] * dT[1][1] + vm[10] * dT[1][2];
    const float tz = 1.0f / q.t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = q.xmul * -fx * tz2 * dJ02;
    const float dty = q.ymul * -fy * tz2 * dJ12;
    const float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * q.t[0]) * tz3 * dJ02 + (2 * fy * q.t[1]) * tz3 * dJ12;
    float dmean[3];
    dmean[0] = vm[0] * dtx + vm[1] * dty + vm[2] * dtz;
    dmean[1] = vm[4] * dtx + vm[5] * dty + vm[6] * dtz;
    dmean[2] = vm[8] * dtx + vm[9] * dty + vm[10] * dtz;

    /* ---- projection backward ---- */
    const float* pj = cam->proj;
    float mh[4];
    transform4x4(m, pj, mh);
    const float m_w = 1.0f / (mh[3] + 0.0000001f);
    const float mul1 = (pj[0] * m[0] + pj[4] * m[1] + pj[8] * m[2] + pj[12]) * m_w * m_w;
    const float mul2 = (pj[1] * m[0] + pj[5] * m[1] + pj[9] * m[2] + pj[13]) * m_w * m_w;
    dmean[0] += (pj[0] * m_w - pj[3] * mul1) * dmx + (pj[1] * m_w - pj[3] * mul2) * dmy;
    dmean[1] += (pj[4] * m_w - pj[7] * mul1) * dmx + (pj[5] * m_w - pj[7] * mul2) * dmy;
    dmean[2] += (pj[8] * m_w - pj[11] * mul1) * dmx + (pj[9] * m_w - pj[11] * mul2) * dmy;

    /* ---- SH backward (A.5) ---- */
    const uint8_t cl = clamped[i];
    if (M > 0) {
      for (int ch = 0; ch < 3; ch++) if (cl & (1u << ch)) dcol[ch] = 0.0f;
      const float dorig[3] = {m[0] - cam->campos[0], m[1] - cam->campos[1], m[2] - cam->campos[2]};
      const float len = sqrtf(dorig[0] * dorig[0] + dorig[1] * dorig[1] + dorig[2] * dorig[2]);
      const float d[3] = {dorig[0] / len, dorig[1] / len, dorig[2] / len};
      float b[25], db[25][3];
      const int n = sh_basis(v->sh_degree, v->sh_l4, d[0], d[1], d[2], b, db);
      float ddir[3] = {0, 0, 0};
      for (int k = 0; k < n; k++)
        for (int ch = 0; ch < 3; ch++) {
          g_sh[sh_index(v->sh_layout, M, i, k, ch)] += (double)(b[k] * dcol[ch]);
          const float shv = load_sh(v->sh, v->sh_layout, M, i, k, ch);
          for (int ax = 0; ax < 3; ax++) ddir[ax] += db[k][ax] * shv * dcol[ch];
        }
      const float sum2 = dorig[0] * dorig[0] + dorig[1] * dorig[1] + dorig[2] * dorig[2];
      const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmean[0] += ((sum2 - dorig[0] * dorig[0]) * ddir[0] - dorig[1] * dorig[0] * ddir[1] - dorig[2] * dorig[0] * ddir[2]) * inv32;
      dmean[1] += (-dorig[0] * dorig[1] * ddir[0] + (sum2 - dorig[1] * dorig[1]) * ddir[1] - dorig[2] * dorig[1] * ddir[2]) * inv32;
      dmean[2] += (-dorig[0] * dorig[2] * ddir[0] - dorig[1] * dorig[2] * ddir[1] + (sum2 - dorig[2] * dorig[2]) * ddir[2]) * inv32;
    } else {
      for (int ch = 0; ch < 3; ch++) g_sh[3 * (size_t)i + ch] += (double)dcol[ch];
    }

    /* chain through the host-side scaling (cuda_splatting.py:76-77) */
    double gm[3] = {(double)(dmean[0] * s), (double)(dmean[1] * s), (double)(dmean[2] * s)};

    /* ---- fused depth channel: dL/df = SH_C0 dL/dfd, then df/dz * dz/dmean (A.7) ---- */
    if (v->depth_mode != TSPLAT_DEPTH_NONE && !(cl & 8u)) {
      const float* r = cam->depth_row;
      const float* mu = v->means + 3 * (size_t)i;
      const float z = ((r[0] * mu[0] + r[1] * mu[1]) + r[2] * mu[2]) + r[3];
      float dfdz;
      (void)depth_feature(v->depth_mode, z, cam->near_plane, cam->far_plane, &dfdz);
      const float dz = SH_C0 * dcol[3] * dfdz;
      gm[0] += (double)(dz * r[0]); gm[1] += (double)(dz * r[1]); gm[2] += (double)(dz * r[2]);
    }
    for (int ax = 0; ax < 3; ax++) g_means[3 * (size_t)i + ax] += gm[ax];
    if (v->cov_layout == TSPLAT_COV_TRIU6) {
      for (int k = 0; k < 6; k++) g_cov[6 * (size_t)i + k] += (double)(dcov[k] * s2);
    } else {
      static const int idx[6] = {0, 1, 2, 4, 5, 8};
      for (int k = 0; k < 6; k++) g_cov[9 * (size_t)i + idx[k]] += (double)(dcov[k] * s2);
    }
  }
}

/* checkFrustum / markVisible */
void oracle_mark_visible(int G, const float* means, const float* view, uint8_t* present) {
  for (int i = 0; i < G; i++) {
    float pv[3];
    transform4x3(means + 3 * (size_t)i, view, pv);
    present[i] = pv[2] > 0.2f;
  }
}

/* OpenMP team size.  torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the timed CPU arm of bench.py
 * sets the count it wants explicitly instead of inheriting that. */
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  extern void omp_set_num_threads(int);
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  extern int omp_get_max_threads(void);
  return omp_get_max_threads();
#else
  return 1;
#endif
}
