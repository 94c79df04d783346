// cull.cuh — conservative per-pair cull extents shared by the two binning paths (binning.cu,
// bucket.cu) and consumed by the render kernels (render.cu: cull_round).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace tsplat {

// ---------------------------------------------------------------------------------------------
// Conservative pixel-space half extents (hx, hy) of the region where a list entry can pass the
// render kernels' test alpha = min(0.99, o*exp(power)) >= 1/255, packed as two fp16 rounded UP.
// The render kernels use it to skip, per warp, entries that cannot touch the warp's 8x4 pixels
// (exact culling: a skipped entry would have failed the alpha test on every pixel of the warp).
//   power = -q, q = (A dx^2 + C dy^2)/2 + B dx dy ; passes only if q <= tau = ln(255 o)
//   => |dx| <= sqrt(2 tau C/det), |dy| <= sqrt(2 tau A/det), det = AC - B^2.
// Margins: binary32 evaluation of q has relative error <= 12 eps kappa, kappa = AC/det (see
// DESIGN.md); the fast exp adds ~1e-6 to tau; coordinates compare with +0.01 px.  Evaluated in
// binary64 so that det itself carries no cancellation error.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int pack_extent(float hx, float hy) {
  const unsigned int lo = __half_as_ushort(__float2half_ru(hx)), hi = __half_as_ushort(__float2half_ru(hy));
  return lo | (hi << 16);
}
__device__ __forceinline__ unsigned int cull_extent(float A, float B, float C, float o) {
  const float inf = __int_as_float(0x7f800000);
  if (o < 1.0f / 255.0f) return pack_extent(-inf, -inf);  // alpha <= o < 1/255 on every pixel
  const double det = (double)A * (double)C - (double)B * (double)B;
  if (!(det > 0.0) || !(A > 0.0f) || !(C > 0.0f)) return pack_extent(inf, inf);
  const double kappa = ((double)A * (double)C) / det;
  const double delta = 2e-6 * kappa;
  if (!(delta < 0.5)) return pack_extent(inf, inf);
  const double tau = (log(255.0 * (double)o) + 1e-4) * 1.002 / (1.0 - delta);
  const double hx = sqrt(2.0 * tau * (double)C / det) + 0.01, hy = sqrt(2.0 * tau * (double)A / det) + 0.01;
  float fx = (float)hx * 1.0001f, fy = (float)hy * 1.0001f;
  if (!(fx == fx) || !(fy == fy)) return pack_extent(inf, inf);
  return pack_extent(fx, fy);
}

}  // namespace tsplat
