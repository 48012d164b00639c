// render_fwd.cu — forward alpha blending, one CTA per 16x16 tile.
// Replaces renderCUDA<3> forward (CR/forward.cu:263-392).
//
// Differences in structure (results identical up to fp32 rounding of identical
// expressions):
//  * a warp covers an 8x4 pixel block (not 16x2) so that its 32 lanes see
//    nearly the same splats -> warp-level early termination and coherent skips;
//  * the tile's batch of splats is staged in shared memory as packed 16 B
//    records (conic+opacity, rgb+depth): no per-pair global colour fetch
//    (CR/forward.cu:365 reads features[] from global for every contributing pair);
//  * n_touched (CR/forward.cu:369-371) is counted with one ballot per
//    (warp, splat) into shared memory and flushed with ONE global atomic per
//    (tile, splat) instead of one per (pixel, splat).
#include "common.cuh"

namespace b2gs {

constexpr int RF_BATCH = 256;

__global__ void __launch_bounds__(256)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W,
                  int H, int gx, const float2* __restrict__ xy, const float4* __restrict__ conic_o,
                  const float4* __restrict__ rgbd, const float* __restrict__ bg,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_depth,
                  float* __restrict__ out_opacity, int* __restrict__ n_touched) {
  __shared__ float2 s_xy[RF_BATCH];
  __shared__ float4 s_co[RF_BATCH];
  __shared__ float4 s_cd[RF_BATCH];
  __shared__ int s_id[RF_BATCH];
  __shared__ int s_cnt[RF_BATCH];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % gx, tile_y = tile / gx;
  // warp -> 8x4 sub-block (2 across, 4 down); lane -> pixel inside it
  const int px = tile_x * TILE + (warp & 1) * 8 + (lane & 7);
  const int py = tile_y * TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const int pix_id = W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);

  bool done = !inside;
  float T = 1.0f;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last_contributor = 0;

  for (int base = 0; base < n; base += RF_BATCH) {
    // whole tile finished?  (also the barrier that protects the staging buffers)
    if (__syncthreads_count(done) == TILE_PIX) break;
    const int cnt = min(RF_BATCH, n - base);
    if (tid < cnt) {
      const int id = (int)point_list[range.x + base + tid];
      s_id[tid] = id;
      s_xy[tid] = xy[id];
      s_co[tid] = conic_o[id];
      s_cd[tid] = rgbd[id];
    }
    s_cnt[tid] = 0;
    __syncthreads();

    if (!__all_sync(0xffffffffu, done)) {
      for (int j = 0; j < cnt; ++j) {
        const float2 p = s_xy[j];
        const float4 co = s_co[j];
        const float dx = p.x - pixfx, dy = p.y - pixfy;
        const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
        bool touch = false;
        if (!done && !(power > 0.0f)) {
          const float alpha = fminf(0.99f, co.w * expf(power));
          if (!(alpha < 1.0f / 255.0f)) {
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) {
              done = true;
            } else {
              const float4 cd = s_cd[j];
              C0 += cd.x * alpha * T;
              C1 += cd.y * alpha * T;
              C2 += cd.z * alpha * T;
              D += cd.w * alpha * T;
              touch = test_T > 0.5f;
              T = test_T;
              last_contributor = (uint32_t)(base + j + 1);
            }
          }
        }
        const unsigned m = __ballot_sync(0xffffffffu, touch);
        if (m != 0u && lane == 0) atomicAdd(&s_cnt[j], __popc(m));
        // warp-level early termination, checked every 8 splats
        if ((j & 7) == 7 && __all_sync(0xffffffffu, done)) break;
      }
    }
    __syncthreads();
    if (tid < cnt) {
      const int c = s_cnt[tid];
      if (c) atomicAdd(&n_touched[s_id[tid]], c);
    }
  }

  if (inside) {
    final_T[pix_id] = T;
    n_contrib[pix_id] = last_contributor;
    const size_t plane = (size_t)H * W;
    out_color[0 * plane + pix_id] = C0 + T * bg[0];
    out_color[1 * plane + pix_id] = C1 + T * bg[1];
    out_color[2 * plane + pix_id] = C2 + T * bg[2];
    out_depth[pix_id] = D;
    out_opacity[pix_id] = 1 - T;
  }
}

int launch_render_fwd(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                      const B2gsForwardOut& out, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  B2GS_PROF("render_fwd", st);
  render_fwd_kernel<<<gx * gy, 256, 0, st>>>(im.ranges, b.point_list, s.W, s.H, gx, g.xy, g.conic_o,
                                             g.rgbd, s.bg, im.final_T, im.n_contrib, out.color,
                                             out.depth, out.opacity, out.n_touched);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

}  // namespace b2gs

extern "C" int b2gs_debug_export_image(const B2gsScene* scene, const B2gsWorkspace* ws,
                                       float* final_T, uint32_t* n_contrib, void* stream) {
  using namespace b2gs;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ImageWS im = carve_image(ws->image, scene->W, scene->H, nullptr);
  const size_t N = (size_t)scene->W * scene->H;
  if (final_T)
    B2GS_CUDA(cudaMemcpyAsync(final_T, im.final_T, N * 4, cudaMemcpyDeviceToDevice, st));
  if (n_contrib)
    B2GS_CUDA(cudaMemcpyAsync(n_contrib, im.n_contrib, N * 4, cudaMemcpyDeviceToDevice, st));
  return B2GS_OK;
}
