// cull.cuh — conservative per-block splat culling for render_fwd / render_bwd.
//
// A warp of the render kernels owns an 8x4 pixel block of its 16x16 tile.  A splat of the
// tile's list can only change a pixel p of that block if the per-pixel tests of the reference
// pass (CR/forward.cu:346-358):  power(p) <= 0  and  min(0.99, o*exp(power(p))) >= 1/255,
// i.e.  q(d) = 1/2 (a dx^2 + c dy^2) + b dx dy  <=  tau = ln(255 o),  d = mean - p.
// The minimum of the convex quadratic q over the block's rectangle lies at d = 0 if the mean
// is inside, else on one of the 4 edges; the block is skipped when even that minimum misses the
// threshold by a safety margin far above fp32 rounding.  Skipped splats would have been skipped
// by every pixel of the block, so images, n_contrib and gradients are unchanged; the per-pixel
// work drops from "every splat of the 16x16 tile" to "splats that reach this 8x4 block".
//
// The test is evaluated ONCE per (tile, splat) instance for all eight blocks of the tile, by the
// instance scatter (binning.cu) where the splat's data is already in the owner lane's hands, and
// travels as the low byte of the per-tile sort key; both render kernels consume it from the
// sorted list (id << 8 | mask).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2gs {

struct SplatCull {
  float2 m;      // pixel-space mean
  float a, b, c; // conic
  float inv_a, inv_c;
  float tau;     // ln(255 o) + margin
  bool proper;   // a > 0 && c > 0 (else: let the pixels decide)
};

// min over t in [t0,t1] of  A + B t + C t^2 (C > 0, inv2C = 1/(2C)), with the sum of |terms|
// for the rounding margin
__device__ __forceinline__ void edge_min(float A, float B, float C, float inv2C, float t0, float t1,
                                         float& qmin, float& smin) {
  const float t = fminf(fmaxf(-B * inv2C, t0), t1);
  const float bt = B * t, ct = C * t * t;
  const float q = A + bt + ct;
  if (q < qmin) {
    qmin = q;
    smin = A + fabsf(bt) + fabsf(ct);
  }
}

// false only if NO pixel of [x0,x1]x[y0,y1] can pass the reference's per-pixel tests.
__device__ __forceinline__ bool splat_may_touch_rect(const SplatCull& s, float x0, float y0, float x1,
                                                     float y1) {
  if (!s.proper) return true;
  // d = mean - pixel ranges over [dx0,dx1] x [dy0,dy1]
  const float dx0 = s.m.x - x1, dx1 = s.m.x - x0, dy0 = s.m.y - y1, dy1 = s.m.y - y0;
  if (!(dx0 > 0.f || dx1 < 0.f || dy0 > 0.f || dy1 < 0.f)) return true;  // mean inside (or NaN)
  float qmin = 3.0e38f, smin = 0.f;
  const float ha = 0.5f * s.a, hc = 0.5f * s.c;
  // edges dx = X: q = 1/2 a X^2 + (b X) dy + 1/2 c dy^2
  edge_min(ha * dx0 * dx0, s.b * dx0, hc, s.inv_c, dy0, dy1, qmin, smin);
  edge_min(ha * dx1 * dx1, s.b * dx1, hc, s.inv_c, dy0, dy1, qmin, smin);
  // edges dy = Y
  edge_min(hc * dy0 * dy0, s.b * dy0, ha, s.inv_a, dx0, dx1, qmin, smin);
  edge_min(hc * dy1 * dy1, s.b * dy1, ha, s.inv_a, dx0, dx1, qmin, smin);
  // keep unless the best case misses by more than the rounding margin (NaN -> keep)
  return !(qmin - (1.0e-5f * smin + 1.0e-3f) > s.tau);
}

// Per-splat setup, done once per Gaussian: thresholds, reciprocals and the bounding box of the
// ellipse q(d) <= tau (half extents sqrt(2 tau c / det), sqrt(2 tau a / det)) for a cheap
// conservative pre-reject.
struct SplatCullBox {
  SplatCull s;
  float bx0, bx1, by0, by1;
  bool dead;     // opacity below 1/255: alpha < 1/255 wherever power <= 0
};

__device__ __forceinline__ SplatCullBox splat_cull_setup(float2 m, float4 co) {
  SplatCullBox r;
  const float a = co.x, b = co.y, c = co.z, o = co.w;
  r.dead = o < 1.0f / 255.0f;
  r.s.m = m; r.s.a = a; r.s.b = b; r.s.c = c;
  r.s.proper = a > 0.f && c > 0.f;
  r.s.inv_a = 1.0f / a;
  r.s.inv_c = 1.0f / c;
  r.s.tau = __logf(255.0f * fmaxf(o, 1.0e-30f)) + 0.01f;   // __logf abs error << 0.01
  float ex = 3.0e38f, ey = 3.0e38f;
  const float det = a * c - b * b;
  if (r.s.proper && det > 0.f) {
    const float inv = 2.f * r.s.tau / det;
    ex = sqrtf(fmaxf(inv * c, 0.f)) * 1.001f + 0.01f;
    ey = sqrtf(fmaxf(inv * a, 0.f)) * 1.001f + 0.01f;
  }
  r.bx0 = m.x - ex; r.bx1 = m.x + ex; r.by0 = m.y - ey; r.by1 = m.y + ey;
  return r;
}

// bit w set <=> the splat may touch warp w's 8x4 block of the 16x16 tile at (tx0, ty0):
// warp w owns columns (w & 1) * 8 .. +7 and rows (w >> 1) * 4 .. +3 (render_fwd / render_bwd).
__device__ __forceinline__ uint32_t splat_block_mask(const SplatCullBox& r, float tx0, float ty0) {
  if (r.dead) return 0u;
  uint32_t bits = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const float x0 = tx0 + (float)((w & 1) * 8), y0 = ty0 + (float)((w >> 1) * 4);
    const float x1 = x0 + 7.f, y1 = y0 + 3.f;
    const bool box = !(x0 > r.bx1 || x1 < r.bx0 || y0 > r.by1 || y1 < r.by0);   // NaN -> keep
    if (box && splat_may_touch_rect(r.s, x0, y0, x1, y1)) bits |= 1u << w;
  }
  return bits;
}

}  // namespace b2gs
