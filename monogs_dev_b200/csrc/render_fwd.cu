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
//  * each warp first culls the staged batch against its own 8x4 block (cull.cuh, one
//    lane per splat) and only iterates over the survivors' bit mask;
//  * n_touched (CR/forward.cu:369-371) is kept as per-lane bit flags, counted with
//    ballots only for splats that were touched, into shared memory, and flushed with
//    ONE global atomic per (tile, splat) instead of one per (pixel, splat).
#include "common.cuh"
#include "cull.cuh"

namespace b2gs {

constexpr int RF_BATCH = 256;
constexpr unsigned RF_FULL = 0xffffffffu;

__global__ void __launch_bounds__(256)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W,
                  int H, int gx, const float2* __restrict__ xy, const float4* __restrict__ conic_o,
                  const float4* __restrict__ rgbd, const float* __restrict__ bg,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_depth,
                  float* __restrict__ out_opacity, int* __restrict__ n_touched) {
  __shared__ SplatRec s_rec[RF_BATCH];
  __shared__ int s_cnt[RF_BATCH];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % gx, tile_y = tile / gx;
  // warp -> 8x4 sub-block (2 across, 4 down); lane -> pixel inside it
  const int bx = tile_x * TILE + (warp & 1) * 8, by = tile_y * TILE + (warp >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const int pix_id = W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);

  const uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(s_cnt);
  int done = !inside;
  float T = 1.0f;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, D = 0.f;
  uint32_t last_contributor = 0;

  for (int base = 0; base < n; base += RF_BATCH) {
    // whole tile finished?  (also the barrier that protects the staging buffers)
    if (__syncthreads_count(done) == TILE_PIX) break;
    const int cnt = min(RF_BATCH, n - base);
    if (tid < cnt) {
      const uint32_t v = point_list[range.x + base + tid];   // id << 8 | block mask
      const int id = (int)(v >> 8);
      SplatRec& r = s_rec[tid];
      r.co = conic_o[id];
      r.cd = rgbd[id];
      r.xy = xy[id];
      r.id = id;
      r.bm = v & 0xffu;
    }
    s_cnt[tid] = 0;
    __syncthreads();

    if (!__all_sync(RF_FULL, done)) {
      for (int c0 = 0; c0 < cnt; c0 += 32) {
        // can the splat reach this warp's 8x4 block at all?  (bit computed once by the binning
        // stage, cull.cuh: splat_block_mask)
        const int jj = c0 + lane;
        const bool keep = jj < cnt && ((s_rec[jj].bm >> warp) & 1);
        unsigned mask = __ballot_sync(RF_FULL, keep);
        while (mask) {
          const int k = __ffs(mask) - 1;
          mask &= mask - 1;
          const int j = c0 + k;
          const SplatRec& r = s_rec[j];
          const float2 p = r.xy;
          const float4 co = r.co;
          const float dx = p.x - pixfx, dy = p.y - pixfy;
          const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
          bool touch = false;
          if (!done && !(power > 0.0f)) {
            const float alpha = fminf(0.99f, co.w * expf(power));
            if (!(alpha < 1.0f / 255.0f)) {
              const float test_T = T * (1 - alpha);
              if (test_T < 0.0001f) {
                done = 1;
              } else {
                const float4 cd = r.cd;
                const float w = alpha * T;
                C0 += cd.x * w;
                C1 += cd.y * w;
                C2 += cd.z * w;
                D += cd.w * w;
                touch = test_T > 0.5f;
                T = test_T;
                last_contributor = (uint32_t)(base + j + 1);
              }
            }
          }
          // n_touched: how many pixels saw this splat with T still above 0.5
          const unsigned tb = __ballot_sync(RF_FULL, touch);
          if (tb != 0u && lane == 0) {   // plain red.shared: one lane, nothing to aggregate
            asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(cnt_base + 4u * (uint32_t)j), "r"(__popc(tb)) : "memory");
          }
        }
        if (__all_sync(RF_FULL, done)) break;  // warp-level early termination
      }
    }
    __syncthreads();
    if (tid < cnt) {
      const int c = s_cnt[tid];
      if (c) atomicAdd(&n_touched[s_rec[tid].id], c);
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
