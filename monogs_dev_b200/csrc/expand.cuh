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
};

// For every (Gaussian, tile) instance:  token = issue(tile);  ...;  commit(token, tile, owner)
// where `owner` is the lane (0..31) whose Gaussian the instance belongs to.
// Instances are handled U per lane per trip with all `issue` calls ahead of the `commit`s, so
// that U atomics (with return) are in flight per lane instead of one: a huge splat spread over
// the warp is a chain of total/32 dependent L2 round trips otherwise.  All 32 lanes must call.
template <int U, class Issue, class Commit>
__device__ __forceinline__ void warp_expand_rects(WarpRects& s, int lane, uint32_t touched,
                                                  uint32_t org, uint32_t w, uint32_t gx, Issue issue,
                                                  Commit commit) {
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
  __syncwarp();
  for (uint32_t k0 = lane; k0 < total; k0 += 32 * U) {
    int lo[U];
    uint32_t token[U], tile[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t k = k0 + 32 * u;
      lo[u] = -1;
      if (k < total) {
        int l = 0;  // smallest j with pre[j] > k
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
          if (s.pre[l + step - 1] <= k) l += step;
        }
        const uint32_t local = k - (l ? s.pre[l - 1] : 0u);
        const uint32_t ww = s.w[l];
        const uint32_t o = s.org[l];
        const uint32_t dy = local / ww, dx = local - dy * ww;
        lo[u] = l;
        tile[u] = ((o >> 16) + dy) * gx + (o & 0xffffu) + dx;
        token[u] = issue(tile[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (lo[u] >= 0) commit(token[u], tile[u], lo[u]);
    }
  }
  __syncwarp();
}

}  // namespace b2gs
