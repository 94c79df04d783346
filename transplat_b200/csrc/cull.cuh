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
// DESIGN.md); the fast exp/log add ~1e-6 to tau; coordinates compare with +0.01 px.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int pack_extent(float hx, float hy) {
  const unsigned int lo = __half_as_ushort(__float2half_ru(hx)), hi = __half_as_ushort(__float2half_ru(hy));
  return lo | (hi << 16);
}
__device__ __forceinline__ unsigned int cull_extent(float A, float B, float C, float o) {
  const float inf = __int_as_float(0x7f800000);
  if (o < 1.0f / 255.0f) return pack_extent(-inf, -inf);  // alpha <= o < 1/255 on every pixel
  // det in binary64 (the products are exact there, so det carries no cancellation error), the rest in
  // binary32: its ~1e-6 relative error is far inside the 2e-3 margin on tau
  const float det = (float)((double)A * (double)C - (double)B * (double)B);
  if (!(det > 0.0f) || !(A > 0.0f) || !(C > 0.0f)) return pack_extent(inf, inf);
  const float kappa = __fdividef(A * C, det);
  const float delta = 2e-6f * kappa;
  if (!(delta < 0.5f)) return pack_extent(inf, inf);
  const float tau = (__logf(255.0f * o) + 2e-4f) * __fdividef(1.003f, 1.0f - delta);
  const float s2 = __fdividef(2.0f * tau, det);
  const float hx = sqrtf(s2 * C) * 1.0002f + 0.01f, hy = sqrtf(s2 * A) * 1.0002f + 0.01f;
  if (!(hx == hx) || !(hy == hy)) return pack_extent(inf, inf);
  return pack_extent(hx, hy);
}

}  // namespace tsplat
