// cull.cuh — per-pair cull threshold shared by the two binning paths (binning.cu,
// bucket.cu) and consumed by the render kernels (render.cu: cull_round).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace tsplat {

// ---------------------------------------------------------------------------------------------
// Exact per-warp culling (replaces the bounding-box extents in the render kernels): the record carries
// one float, tau_s >= ln(255 o) with the margins below, and the render kernels compare it with the
// MINIMUM of q over the warp's pixel rectangle (closed form for a convex quadratic over a box, see
// render.cu: cull_round).  An entry is skipped only if q > tau on the whole rectangle, i.e. if it fails
// the alpha test on every pixel of the warp.
//   margins: the kernels' binary32 q and the cull's binary32 q_min each carry a relative error
//   <= 12 eps kappa < delta = 2e-6 kappa  =>  tau_s = (ln(255 o) + 2e-4) * 1.003 / (1 - 2 delta);
//   +inf (never culled) for degenerate conics, -inf (always culled) for o < 1/255.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float cull_tau(float A, float B, float C, float o) {
  const float inf = __int_as_float(0x7f800000);
  if (o < 1.0f / 255.0f) return -inf;  // alpha <= o < 1/255 on every pixel
  const float det = (float)((double)A * (double)C - (double)B * (double)B);
  if (!(det > 0.0f) || !(A > 0.0f) || !(C > 0.0f)) return inf;
  const float kappa = __fdividef(A * C, det);
  const float delta2 = 4e-6f * kappa;
  if (!(delta2 < 0.5f)) return inf;
  const float tau = (__logf(255.0f * o) + 2e-4f) * __fdividef(1.003f, 1.0f - delta2);
  return (tau == tau) ? tau : inf;
}

// The cull itself, done ONCE per (tile, Gaussian) pair where the sorted records are written (gather kernels): bit w
// of the result is set iff the entry can pass the alpha test on some pixel of block w of the 16x16 tile at
// (tileX0, tileY0) - block w = the 8x4 pixels of consumer warp w of the render kernels, columns (w & 1) * 8, rows
// (w >> 1) * 4.  Exact in the continuous sense: the minimum of
//   q(dx, dy) = (A dx^2 + C dy^2)/2 + B dx dy      (power = -q)
// over the rectangle is compared with tau_s.  q is convex with its minimum at the centre, so over a box that does not
// contain the centre the minimum lies on an edge facing it: with (ax, by) the box point nearest the centre per axis,
// the candidates are the 1-D minima along dx = ax and along dy = by (if the centre is inside, both are 0).
// Evaluating q at a slightly different point of the box can only over-estimate the minimum by O(eps^2 q kappa),
// inside the margin of tau_s.  The render kernels only test their bit (they used to run this test per warp and
// batch: 160 instructions per 64 entries and warp, a third of the forward's instruction count).
__device__ __forceinline__ unsigned int cull_block_mask(float4 a0 /* x, y, A, B */, float Cq, float tau, float tileX0,
                                                        float tileY0) {
  unsigned int mask = 0;
  // Conservative bounding box of {q <= tau} first: half-widths sqrt(2 tau C / det), sqrt(2 tau A / det), widened by
  // 0.1 % + 0.02 px; only the block columns / rows it overlaps get the exact test below (2-4 of the 8 blocks for an
  // encoder-sized Gaussian).  Degenerate conics (tau = +inf, det <= 0, NaN) keep the full range.
  unsigned int cols = 3u, rows = 15u;
  if (!(tau > -3.0e38f)) return 0u;  // opacity below 1/255: alpha < 1/255 on every pixel
  {
    const float det = a0.z * Cq - a0.w * a0.w;
    if (det > 0.0f && tau < 3.0e38f) {
      float rdet;
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rdet) : "f"(det));
      const float s = 2.0f * fmaxf(tau, 0.0f) * rdet;
      const float ex = sqrtf(s * Cq) * 1.001f + 0.02f, ey = sqrtf(s * a0.z) * 1.001f + 0.02f;
      if (ex == ex && ey == ey) {
        const float lx = a0.x - ex - tileX0, hx = a0.x + ex - tileX0;  // the box in tile coordinates
        const float ly = a0.y - ey - tileY0, hy = a0.y + ey - tileY0;
        cols = (lx <= 7.0f && hx >= 0.0f ? 1u : 0u) | (lx <= 15.0f && hx >= 8.0f ? 2u : 0u);
        rows = (ly <= 3.0f && hy >= 0.0f ? 1u : 0u) | (ly <= 7.0f && hy >= 4.0f ? 2u : 0u) |
               (ly <= 11.0f && hy >= 8.0f ? 4u : 0u) | (ly <= 15.0f && hy >= 12.0f ? 8u : 0u);
      }
    }
  }
  // raw reciprocals: A, C of a kept Gaussian are normal positive numbers; anything else gives inf/NaN, which the
  // clamps turn back into a point of the box, and tau = +inf keeps a degenerate conic anyway
  float rA, rC;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rA) : "f"(a0.z));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rC) : "f"(Cq));
#pragma unroll
  for (int cx = 0; cx < 2; cx++) {
    if (!((cols >> cx) & 1u)) continue;
    const float a_lo = tileX0 + 8.0f * cx - a0.x, a_hi = a_lo + 7.0f;
    const float ax = fminf(fmaxf(0.0f, a_lo), a_hi);
    const float dy1_free = -a0.w * ax * rC;
    const float q1x = a0.z * ax * ax;
#pragma unroll
    for (int ry = 0; ry < 4; ry++) {
      if (!((rows >> ry) & 1u)) continue;
      const float b_lo = tileY0 + 4.0f * ry - a0.y, b_hi = b_lo + 3.0f;
      const float by = fminf(fmaxf(0.0f, b_lo), b_hi);
      const float dy1 = fminf(fmaxf(dy1_free, b_lo), b_hi);
      const float dx2 = fminf(fmaxf(-a0.w * by * rA, a_lo), a_hi);
      const float q1 = 0.5f * (q1x + Cq * dy1 * dy1) + a0.w * ax * dy1;
      const float q2 = 0.5f * (a0.z * dx2 * dx2 + Cq * by * by) + a0.w * dx2 * by;
      if (!(fminf(q1, q2) > tau)) mask |= 1u << (ry * 2 + cx);  // NaN (degenerate conic) => keep
    }
  }
  return mask;
}

}  // namespace tsplat
