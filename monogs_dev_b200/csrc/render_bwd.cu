// render_bwd.cu — backward of the alpha blend, one CTA per 16x16 tile, back to front.
// Replaces renderCUDA<3> backward + render_cuda_reduce_sum (CR/backward.cu:541-787).
//
// Per-pixel arithmetic follows the reference recurrences (T recovered by division,
// accum_rec / last_alpha / last_color replay, CR/backward.cu:700-757).  What is
// redesigned is the reduction of the 10 per-splat partials:
//   reference: per splat per tile 3 block barriers + a 256-wide, 8-level shared
//              memory tree over 13 floats + 10 atomics from thread 0;
//   here:      no block barrier inside a batch.  Each warp (an 8x4 pixel block)
//              votes; warps where every lane skips the splat do nothing; active
//              warps fold the 10 partials with a 12-shuffle split butterfly
//              (each step halves the value set instead of reducing every value
//              over all 32 lanes: 5+3+2+1+1 shuffles instead of 50) and the 10
//              lanes that end up owning a total issue one red.global.add.f32 each
//              into a packed 48 B accumulator row per splat.
//   Before any per-pixel work each warp culls the staged batch against its own 8x4
//   block (cull.cuh): one lane tests one splat, the warp walks the survivors' bit mask.
//   Splats behind every pixel's last contributor are never staged: the block
//   starts at max(n_contrib) over the tile, each warp at its own max.
#include "common.cuh"
#include "cull.cuh"

namespace b2gs {

constexpr int RB_BATCH = 256;
constexpr unsigned FULL = 0xffffffffu;

// accumulator row layout (ACC_STRIDE floats):
// 0 dmean2D.x  1 dmean2D.y  2 dconic.x  3 dconic.y  4 dconic.w
// 5 dopacity   6 dcolor.r   7 dcolor.g  8 dcolor.b  9 ddepth   10,11 pad

// 1/x for x in [0.01, 1] (x = 1 - alpha, alpha <= 0.99): MUFU.RCP + one Newton step, <= 1 ulp,
// without the range / denormal handling of a full-range reciprocal.
__device__ __forceinline__ float rcp_unit_range(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(r, fmaf(-x, r, 1.0f), r);
}

__device__ __forceinline__ float xchg_add(bool upper, float lo_val, float hi_val, int offset) {
  // lanes with the bit clear keep lo_val and hand hi_val to their partner (and vice versa)
  const float send = upper ? lo_val : hi_val;
  const float keep = upper ? hi_val : lo_val;
  return keep + __shfl_xor_sync(FULL, send, offset);
}

// owner lanes of the fold below and the accumulator column each one ends up holding: lane
// constants, formed once per thread (and kept opaque so that they are not re-derived per splat)
__device__ __forceinline__ int fold10_slot(int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  const int local = (b3 ? 3 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0);
  // bit0 clear, and (b3,b2,b1) not a padding slot
  const bool owner = !(lane & 1) && !(b2 && (b3 || b1));
  int slot = owner ? (b4 ? 5 : 0) + local : -1;
  asm volatile("" : "+r"(slot));
  return slot;
}

// slot_row = acc + slot (owners) / nullptr; row_off = splat id * ACC_STRIDE
__device__ __forceinline__ void warp_fold10_and_red(const float v[10], int lane,
                                                    float* __restrict__ slot_row, size_t row_off) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
  float w0 = xchg_add(b4, v[0], v[5], 16);
  float w1 = xchg_add(b4, v[1], v[6], 16);
  float w2 = xchg_add(b4, v[2], v[7], 16);
  float w3 = xchg_add(b4, v[3], v[8], 16);
  float w4 = xchg_add(b4, v[4], v[9], 16);
  float u0 = xchg_add(b3, w0, w3, 8);
  float u1 = xchg_add(b3, w1, w4, 8);
  float u2 = xchg_add(b3, w2, 0.f, 8);
  float t0 = xchg_add(b2, u0, u2, 4);
  float t1 = xchg_add(b2, u1, 0.f, 4);
  float s = xchg_add(b1, t0, t1, 2);
  s += __shfl_xor_sync(FULL, s, 1);
  if (slot_row) atomicAdd(slot_row + row_off, s);
}

__global__ void __launch_bounds__(256)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W,
                  int H, int gx, const float* __restrict__ bg, const float2* __restrict__ xy,
                  const float4* __restrict__ conic_o, const float4* __restrict__ rgbd,
                  const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dpixels_depth,
                  float* __restrict__ acc) {
  __shared__ SplatRec s_rec[RB_BATCH];
  __shared__ int s_max;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % gx, tile_y = tile / gx;
  const int bx = tile_x * TILE + (warp & 1) * 8, by = tile_y * TILE + (warp >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < W && py < H;
  const int pix_id = W * py + px;
  const float pixfx = (float)px, pixfy = (float)py;

  const uint2 range = ranges[tile];

  const float T_final = inside ? final_Ts[pix_id] : 0.f;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;

  // accum_rec_c: blended colour behind the current splat; diff_c: (colour - accum_rec) of the last
  // contributor.  The reference's replay  accum = last_alpha * last_colour + (1 - last_alpha) * accum
  // (CR/backward.cu:716-727) is written as  accum += last_alpha * (last_colour - accum): the
  // difference is the one the previous contributor already formed for its dL_dalpha.
  float accum_rec0 = 0.f, accum_rec1 = 0.f, accum_rec2 = 0.f, accum_recd = 0.f;
  float diff0 = 0.f, diff1 = 0.f, diff2 = 0.f, diffd = 0.f;
  float dLp0 = 0.f, dLp1 = 0.f, dLp2 = 0.f, dLpd = 0.f;
  if (inside) {
    const size_t plane = (size_t)H * W;
    dLp0 = dL_dpixels[0 * plane + pix_id];
    dLp1 = dL_dpixels[1 * plane + pix_id];
    dLp2 = dL_dpixels[2 * plane + pix_id];
    dLpd = dL_dpixels_depth[pix_id];
  }
  float last_alpha = 0.f;
  float bg_dot_dpixel = 0.f;
  bg_dot_dpixel += bg[0] * dLp0;
  bg_dot_dpixel += bg[1] * dLp1;
  bg_dot_dpixel += bg[2] * dLp2;
  const float neg_Tf_bg = -T_final * bg_dot_dpixel;   // dL_dalpha += (-T_final / (1 - alpha)) * bg . dL_dpixel

  const float ddelx_dx = 0.5f * W;
  const float ddely_dy = 0.5f * H;
  const int slot = fold10_slot(lane);
  float* const slot_row = slot >= 0 ? acc + slot : nullptr;

  // deepest contributor over the warp / the tile
  const int warp_max = __reduce_max_sync(FULL, last_contributor);
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
  __syncthreads();
  const int tile_max = s_max;  // positions [0, tile_max) of the range matter

  for (int base = 0; base < tile_max; base += RB_BATCH) {
    const int cnt = min(RB_BATCH, tile_max - base);
    if (base) __syncthreads();
    if (tid < cnt) {
      const int pos = tile_max - 1 - (base + tid);
      const uint32_t v = point_list[range.x + pos];   // id << 8 | block mask
      const int id = (int)(v >> 8);
      SplatRec& r = s_rec[tid];
      r.co = conic_o[id];
      r.cd = rgbd[id];
      r.xy = xy[id];
      r.id = id;
      r.bm = v & 0xffu;
    }
    __syncthreads();

    // entry j is at position tile_max-1-(base+j); this warp needs pos < warp_max
    int j0 = tile_max - base - warp_max;
    if (j0 < 0) j0 = 0;
    for (int c0 = j0 & ~31; c0 < cnt; c0 += 32) {
      // one lane per splat: can it reach this warp's 8x4 block at all? (cull.cuh)
      const int jj = c0 + lane;
      const bool keep = jj >= j0 && jj < cnt && ((s_rec[jj].bm >> warp) & 1);
      unsigned mask = __ballot_sync(FULL, keep);
      while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1;
        const int j = c0 + k;
        const int pos = tile_max - 1 - (base + j);
        const SplatRec& r = s_rec[j];
        const float2 p = r.xy;
        const float4 co = r.co;
        const float dx = p.x - pixfx, dy = p.y - pixfy;
        const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
        bool skip = !(pos < last_contributor);  // also false-inside lanes: last_contributor == 0
        skip |= power > 0.0f;
        const float G = expf(power);
        const float alpha = fminf(0.99f, co.w * G);
        skip |= alpha < 1.0f / 255.0f;
        if (__all_sync(FULL, skip)) continue;

        // skipping lanes contribute exact zeros: every partial is a product with one of these three
        float dL_dalpha = 0.f, dchannel_dcolor = 0.f;
        const float Gs = skip ? 0.f : G;
        if (!skip) {
          const float4 cd = r.cd;
          const float inv1ma = rcp_unit_range(1.f - alpha);
          T = T * inv1ma;  // reference: T / (1 - alpha); differs by <= 1 ulp
          dchannel_dcolor = alpha * T;

          accum_rec0 = fmaf(last_alpha, diff0, accum_rec0);
          accum_rec1 = fmaf(last_alpha, diff1, accum_rec1);
          accum_rec2 = fmaf(last_alpha, diff2, accum_rec2);
          accum_recd = fmaf(last_alpha, diffd, accum_recd);
          diff0 = cd.x - accum_rec0;
          diff1 = cd.y - accum_rec1;
          diff2 = cd.z - accum_rec2;
          diffd = cd.w - accum_recd;
          float d = diff0 * dLp0;
          d = fmaf(diff1, dLp1, d);
          d = fmaf(diff2, dLp2, d);
          d = fmaf(diffd, dLpd, d);
          last_alpha = alpha;
          dL_dalpha = fmaf(inv1ma, neg_Tf_bg, d * T);
        }
        const float dL_dG = co.w * dL_dalpha;
        const float gdx = Gs * dx;
        const float gdy = Gs * dy;
        const float dG_ddelx = -gdx * co.x - gdy * co.y;
        const float dG_ddely = -gdy * co.z - gdx * co.y;
        const float hg = -0.5f * dL_dG;
        float v[10];
        v[0] = dL_dG * dG_ddelx * ddelx_dx;
        v[1] = dL_dG * dG_ddely * ddely_dy;
        v[2] = hg * gdx * dx;
        v[3] = hg * gdx * dy;
        v[4] = hg * gdy * dy;
        v[5] = Gs * dL_dalpha;
        v[6] = dchannel_dcolor * dLp0;
        v[7] = dchannel_dcolor * dLp1;
        v[8] = dchannel_dcolor * dLp2;
        v[9] = dchannel_dcolor * dLpd;
        warp_fold10_and_red(v, lane, slot_row, (size_t)r.id * ACC_STRIDE);
      }
    }
  }
}

int launch_render_bwd(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                      const B2gsBackwardIO& io, float* acc, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  B2GS_PROF("render_bwd", st);
  render_bwd_kernel<<<gx * gy, 256, 0, st>>>(im.ranges, b.point_list, s.W, s.H, gx, s.bg, g.xy,
                                             g.conic_o, g.rgbd, im.final_T, im.n_contrib,
                                             io.dL_dcolor, io.dL_ddepth, acc);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

}  // namespace b2gs
