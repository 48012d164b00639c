// loss_terms.cuh — the per-pixel terms of the mapping / tracking losses
// (utils/slam_utils.py:56-128), shared by the stand-alone loss kernels (loss.cu) and the
// forward render kernel's fused-loss epilogue (render_fwd.cu).
#pragma once
#include "common.cuh"

namespace b2gs {

struct PixelTerms {
  float d[3];   // image_ab*mask - gt*mask per channel (0 where masked)
  float dd;     // depth*mask - gt_depth*mask
  float o;      // opacity weight of the rgb term (1 in mapping)
  float m;      // rgb mask (0/1), incl. grad_mask in tracking
  float md;     // depth mask (0/1)
};

__device__ __forceinline__ float sgnf(float x) { return (float)(x > 0.f) - (float)(x < 0.f); }

// mode bit 0: tracking (opacity weighting, opacity>0.95 depth mask, grad_mask); bit 1: monocular.
// c0..c2 / dep / op: the rendered pixel (from memory in loss.cu, from registers in render_fwd.cu);
// gt: the ground-truth pixel (r, g, b, depth) — render_fwd.cu fetches it before its blend loop so
// that the loads are long done when the epilogue needs them.
__device__ __forceinline__ float4 load_gt(const B2gsLossArgs& a, int i) {
  const size_t N = (size_t)a.N;
  return make_float4(a.gt_image[i], a.gt_image[N + i], a.gt_image[2 * N + i],
                     (a.mode & 2) ? 0.f : a.gt_depth[i]);
}

__device__ __forceinline__ void prefetch_gt(const B2gsLossArgs& a, int i) {
  const size_t N = (size_t)a.N;
  asm volatile("prefetch.global.L2 [%0];" ::"l"(a.gt_image + i));
  asm volatile("prefetch.global.L2 [%0];" ::"l"(a.gt_image + N + i));
  asm volatile("prefetch.global.L2 [%0];" ::"l"(a.gt_image + 2 * N + i));
  if (!(a.mode & 2)) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.gt_depth + i));
}

__device__ __forceinline__ PixelTerms pixel_terms_of(const B2gsLossArgs& a, float ea, float eb, int i,
                                                    float c0, float c1, float c2, float dep, float op_in,
                                                    float4 gt) {
  PixelTerms t;
  const float g0 = gt.x, g1 = gt.y, g2 = gt.z;
  // gt_image.sum(dim=0) > threshold  (torch sums channel 0 + 1 + 2 in this order)
  float m = ((g0 + g1) + g2) > a.rgb_boundary_threshold ? 1.f : 0.f;
  const bool tracking = a.mode & 1;
  float op = 1.f;
  if (tracking) {
    op = op_in;
    if (a.grad_mask) m *= a.grad_mask[i];
  }
  t.m = m;
  t.o = op;
  const float i0 = ea * c0 + eb, i1 = ea * c1 + eb, i2 = ea * c2 + eb;
  t.d[0] = i0 * m - g0 * m;
  t.d[1] = i1 * m - g1 * m;
  t.d[2] = i2 * m - g2 * m;
  t.md = 0.f;
  t.dd = 0.f;
  if (!(a.mode & 2)) {
    const float gd = gt.w;
    float md = gd > 0.01f ? 1.f : 0.f;
    if (tracking) md *= (op > 0.95f ? 1.f : 0.f);
    t.md = md;
    t.dd = dep * md - gd * md;
  }
  return t;
}

__device__ __forceinline__ void exposure(const B2gsLossArgs& a, float& ea, float& eb) {
  ea = a.exposure_a ? expf(a.exposure_a[0]) : 1.f;
  eb = a.exposure_b ? a.exposure_b[0] : 0.f;
}

// Everything the fused forward+backward of one pixel produces (upstream gradient 1):
// gradient of the loss wrt the rendered pixel and the four reduction terms.
struct PixelLoss {
  float g[3], gd;        // dL/dimage (3 channels), dL/ddepth
  float rgb, dep;        // |.| sums of the value
  float ga, gb;          // exposure-gradient terms (ga still to be multiplied by exp(a))
};

__device__ __forceinline__ PixelLoss pixel_loss(const B2gsLossArgs& a, float ea, float eb, float w_rgb,
                                                float w_dep, int i, float c0, float c1, float c2,
                                                float dep, float op, float4 gt) {
  const PixelTerms t = pixel_terms_of(a, ea, eb, i, c0, c1, c2, dep, op, gt);
  PixelLoss r;
  const float img[3] = {c0, c1, c2};
  r.ga = 0.f; r.gb = 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // d|x|/dx * d(image_ab*m)/d(image_ab) = sign * m ; image_ab = ea*image + eb
    const float g_ab = w_rgb * t.o * sgnf(t.d[c]) * t.m;
    r.g[c] = g_ab * ea;
    r.ga += g_ab * img[c];
    r.gb += g_ab;
  }
  r.gd = w_dep * sgnf(t.dd) * t.md;
  r.rgb = t.o * (fabsf(t.d[0]) + fabsf(t.d[1]) + fabsf(t.d[2]));
  r.dep = fabsf(t.dd);
  return r;
}

}  // namespace b2gs
