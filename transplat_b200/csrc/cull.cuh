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

}  // namespace tsplat
