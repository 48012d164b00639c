// cull.cuh — conservative per-warp splat culling shared by render_fwd / render_bwd.
//
// A warp owns an 8x4 pixel block.  A splat of the tile's list can only change a
// pixel p of that block if the per-pixel tests of the reference pass
// (CR/forward.cu:346-358):  power(p) <= 0  and  min(0.99, o*exp(power(p))) >= 1/255,
// i.e.  q(d) = 1/2 (a dx^2 + c dy^2) + b dx dy  <=  tau = ln(255 o),  d = mean - p.
// One lane evaluates the minimum of the convex quadratic q over the block's
// rectangle (it lies at d = 0 if the mean is inside, else on one of the 4 edges) and
// the warp skips the splat when even that minimum misses the threshold by a safety
// margin far above fp32 rounding.  Skipped splats would have been skipped by every
// pixel of the block, so images, n_contrib and gradients are unchanged; the per-pixel
// work drops from "every splat of the 16x16 tile" to "splats that reach this 8x4 block".
#pragma once
#include <cuda_runtime.h>

namespace b2gs {

// min over t in [t0,t1] of  A + B t + C t^2 (C > 0), with the sum of |terms| for the margin
__device__ __forceinline__ void edge_min(float A, float absA, float B, float C, float t0, float t1,
                                         float& q, float& s) {
  float t = fminf(fmaxf(-B / (2.f * C), t0), t1);
  const float bt = B * t, ct = C * t * t;
  q = A + bt + ct;
  s = absA + fabsf(bt) + fabsf(ct);
}

// false only if NO pixel of [x0,x1]x[y0,y1] can pass the reference's per-pixel tests.
__device__ __forceinline__ bool splat_may_touch_rect(float2 m, float4 co, float x0, float y0,
                                                     float x1, float y1) {
  const float a = co.x, b = co.y, c = co.z, o = co.w;
  if (o < 1.0f / 255.0f) return false;       // alpha <= o < 1/255 wherever power <= 0
  if (!(a > 0.f) || !(c > 0.f)) return true;  // not a proper conic (or NaN): let the pixels decide
  // d = mean - pixel ranges over [dx0,dx1] x [dy0,dy1]
  const float dx0 = m.x - x1, dx1 = m.x - x0, dy0 = m.y - y1, dy1 = m.y - y0;
  if (!(dx0 > 0.f || dx1 < 0.f || dy0 > 0.f || dy1 < 0.f)) return true;  // mean inside (or NaN)
  const float tau = __logf(255.0f * o) + 0.01f;  // __logf abs error << 0.01
  float qmin = 3.0e38f, smin = 0.f;
  float q, s;
  // edges dx = X: q = 1/2 a X^2 + (b X) dy + 1/2 c dy^2
  {
    float A = 0.5f * a * dx0 * dx0;
    edge_min(A, A, b * dx0, 0.5f * c, dy0, dy1, q, s);
    if (q < qmin) { qmin = q; smin = s; }
    A = 0.5f * a * dx1 * dx1;
    edge_min(A, A, b * dx1, 0.5f * c, dy0, dy1, q, s);
    if (q < qmin) { qmin = q; smin = s; }
  }
  // edges dy = Y
  {
    float A = 0.5f * c * dy0 * dy0;
    edge_min(A, A, b * dy0, 0.5f * a, dx0, dx1, q, s);
    if (q < qmin) { qmin = q; smin = s; }
    A = 0.5f * c * dy1 * dy1;
    edge_min(A, A, b * dy1, 0.5f * a, dx0, dx1, q, s);
    if (q < qmin) { qmin = q; smin = s; }
  }
  // keep unless the best case misses by more than the rounding margin (NaN -> keep)
  return !(qmin - (1.0e-5f * smin + 1.0e-3f) > tau);
}

}  // namespace b2gs
