// expand.cuh — warp-cooperative expansion of 32 tile rectangles into their instances.
//
// Each lane owns one Gaussian whose splat touches `touched` = w*h tiles starting at
// tile (x0, y0).  The warp walks the concatenation of the 32 rectangles with lanes
// striding over it, so a huge splat is spread over all lanes instead of serialising one
// thread (the reference's per-thread double loop, CR/rasterizer_impl.cu:97-109).
// Used twice per forward: to count instances per tile (preprocess_fwd.cu) and to
// scatter them into their tile segments (binning.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2gs {

struct WarpRects {
  uint32_t pre[32];  // inclusive prefix of `touched` inside the warp
  uint32_t org[32];  // x0 | y0 << 16
  uint32_t w[32];    // rect width in tiles
  uint32_t a[32];    // payload (depth bits)
  uint32_t b[32];    // payload (Gaussian id)
};

// f(tile, a, b) is called once per (Gaussian, tile) instance.  All 32 lanes must call.
template <class F>
__device__ __forceinline__ void warp_expand_rects(WarpRects& s, int lane, uint32_t touched,
                                                  uint32_t org, uint32_t w, uint32_t a, uint32_t b,
                                                  uint32_t gx, F f) {
  uint32_t pre = touched;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, pre, d);
    if (lane >= d) pre += n;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
  if (total == 0) return;  // warp-uniform
  s.pre[lane] = pre;
  s.org[lane] = org;
  s.w[lane] = w;
  s.a[lane] = a;
  s.b[lane] = b;
  __syncwarp();
  for (uint32_t k = lane; k < total; k += 32) {
    int lo = 0;  // smallest j with pre[j] > k
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      if (s.pre[lo + step - 1] <= k) lo += step;
    }
    const uint32_t local = k - (lo ? s.pre[lo - 1] : 0u);
    const uint32_t ww = s.w[lo];
    const uint32_t o = s.org[lo];
    const uint32_t dy = local / ww, dx = local - dy * ww;
    f(((o >> 16) + dy) * gx + (o & 0xffffu) + dx, s.a[lo], s.b[lo]);
  }
  __syncwarp();
}

}  // namespace b2gs
