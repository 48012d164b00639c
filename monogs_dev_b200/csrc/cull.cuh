// cull.cuh — conservative per-block splat culling for render_fwd / render_bwd.
//
// A warp of the render kernels owns an 8x4 pixel block of its 16x16 tile.  A splat of the
// tile's list can only change a pixel p of that block if the per-pixel tests of the reference
// pass (CR/forward.cu:346-358):  power(p) <= 0  and  min(0.99, o*exp(power(p))) >= 1/255,
// i.e.  q(d) = 1/2 (a dx^2 + c dy^2) + b dx dy  <=  tau = ln(255 o),  d = mean - p.
// The block is skipped when the minimum of the convex quadratic q over the block's rectangle
// misses the threshold by a safety margin far above fp32 rounding.  Skipped splats would have
// been skipped by every pixel of the block, so images, n_contrib and gradients are unchanged;
// the per-pixel work drops from "every splat of the 16x16 tile" to "splats that reach this 8x4
// block".
//
// Where the minimum lies: if the mean is inside the rectangle, at d = 0.  Otherwise q decreases
// along the segment from the minimiser x* towards the mean, so that segment must leave the
// rectangle at once: x* sits on an edge that faces the mean — the vertical edge nearest in x
// (if the mean is outside the rectangle's x range) or the horizontal edge nearest in y.  With
// dxc = clamp(0, dx range) and dyc = clamp(0, dy range) the two candidates are the 1-D minima
// of q on the segments {dx = dxc} and {dy = dyc}; when dxc (dyc) is 0 the segment runs through
// the rectangle's interior and can only over-estimate, and when both are 0 both give q = 0.
// Two clamped parabola minima per block, no branches; the per-line terms are shared by the
// blocks of a column (row) of the tile.
//
// The test is evaluated ONCE per (tile, splat) instance for all eight blocks of the tile, by the
// instance scatter (binning.cu) where the splat's data is already in the owner lane's hands, and
// travels as the low byte of the per-tile sort key; both render kernels consume it from the
// sorted list (id << 8 | mask).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2gs {

// Per-splat setup, done once per Gaussian.
struct SplatCullBox {
  float2 m;        // pixel-space mean
  float ha, b, hc; // a/2, b, c/2
  float inv_a, inv_c;
  float thr;       // (ln(255 o) + margins) / (1 - 1e-5): cull when the margin-reduced minimum exceeds it
  uint32_t fixed;  // 0: evaluate; else 0x100 | mask to return as is (dead -> 0, improper -> 0xff)
};

__device__ __forceinline__ SplatCullBox splat_cull_setup(float2 m, float4 co) {
  SplatCullBox r;
  const float a = co.x, b = co.y, c = co.z, o = co.w;
  r.m = m; r.ha = 0.5f * a; r.b = b; r.hc = 0.5f * c;
  r.inv_a = 1.0f / a;
  r.inv_c = 1.0f / c;
  // __logf abs error << 0.01; + 1e-3 absolute slack, 1e-5 relative (applied to the minimum)
  r.thr = (__logf(255.0f * fmaxf(o, 1.0e-30f)) + 0.011f) * (1.0f / (1.0f - 1.0e-5f));
  // anything odd (non-positive diagonal, non-finite input or reciprocal): let the pixels decide
  const float chk = m.x + m.y + b + r.inv_a + r.inv_c + r.thr;
  const bool proper = a > 0.f && c > 0.f && fabsf(chk) < 3.0e38f;
  r.fixed = o < 1.0f / 255.0f ? 0x100u            // alpha < 1/255 wherever power <= 0
                              : (proper ? 0u : 0x1ffu);
  return r;
}

// margin-reduced minimum over t in [t0,t1] of  A + B t + C t^2  (C > 0, T = -B / (2C)):
// q - 1e-5 (|B t| - B t); the remaining relative term 1e-5 q is folded into `thr`.
__device__ __forceinline__ float edge_low(float A, float B, float C, float T, float t0, float t1) {
  const float t = fminf(fmaxf(T, t0), t1);
  const float q = fmaf(fmaf(C, t, B), t, A);
  const float bt = B * t;
  return fmaf(-1.0e-5f, fabsf(bt) - bt, q);
}

// bit w set <=> the splat may touch warp w's 8x4 block of the 16x16 tile at (tx0, ty0):
// warp w owns columns (w & 1) * 8 .. +7 and rows (w >> 1) * 4 .. +3 (render_fwd / render_bwd).
__device__ __forceinline__ uint32_t splat_block_mask(const SplatCullBox& r, float tx0, float ty0) {
  if (r.fixed) return r.fixed & 0xffu;
  // d = mean - pixel: block column j spans dx in [dxl, dxh], block row i spans dy in [dyl, dyh]
  float dxl[2], dxh[2], Av[2], Bv[2], Tv[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    dxh[j] = r.m.x - (tx0 + (float)(8 * j));
    dxl[j] = dxh[j] - 7.f;
    const float dxc = fminf(fmaxf(0.f, dxl[j]), dxh[j]);
    Bv[j] = r.b * dxc;
    Av[j] = r.ha * dxc * dxc;
    Tv[j] = -Bv[j] * r.inv_c;
  }
  uint32_t bits = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float dyh = r.m.y - (ty0 + (float)(4 * i)), dyl = dyh - 3.f;
    const float dyc = fminf(fmaxf(0.f, dyl), dyh);
    const float Bh = r.b * dyc, Ah = r.hc * dyc * dyc, Th = -Bh * r.inv_a;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float lv = edge_low(Av[j], Bv[j], r.hc, Tv[j], dyl, dyh);        // segment dx = dxc
      const float lh = edge_low(Ah, Bh, r.ha, Th, dxl[j], dxh[j]);           // segment dy = dyc
      if (!(fminf(lv, lh) > r.thr)) bits |= 1u << (2 * i + j);
    }
  }
  return bits;
}

}  // namespace b2gs
