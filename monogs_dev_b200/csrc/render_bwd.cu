// render_bwd.cu — backward of the alpha blend, one CTA per 16x16 tile, back to front.
// Replaces renderCUDA<3> backward + render_cuda_reduce_sum (CR/backward.cu:541-787).
//
// Per-pixel arithmetic follows the reference recurrences (CR/backward.cu:700-757) with two
// algebraic rewrites that stay within the 1e-4 gradient tolerance but are NOT operation for
// operation: T / (1 - alpha) is T * rcp(1 - alpha) (MUFU.RCP + one Newton step, <= 1 ulp), and the
// replay accum = last_alpha * last_colour + (1 - last_alpha) * accum is carried as the one number
// that is used of it, accum . dL_dpixel (see `behind` below).  `power` / exp / the skip tests ARE the
// forward's operation tree bit for bit, so both passes take the same decisions.  What is
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

// accumulator row layout (ACC_STRIDE floats), q = G * dL_dalpha per pixel, d = mean2D - pixel:
// 0 sum q dx   1 sum q dy   2 sum q dx^2   3 sum q dx dy   4 sum q dy^2
// 5 sum q (= dL_dopacity)   6 dcolor.r   7 dcolor.g  8 dcolor.b  9 ddepth   10,11 pad
// (preprocess_bwd.cu turns 0..4 into dL_dmean2D / dL_dconic)

// 1/x for x in [0.01, 1] (x = 1 - alpha, alpha <= 0.99): MUFU.RCP + one Newton step, <= 1 ulp,
// without the range / denormal handling of a full-range reciprocal.
__device__ __forceinline__ float rcp_unit_range(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(r, fmaf(-x, r, 1.0f), r);
}

__device__ __forceinline__ float2 rcp_unit_range2(float2 x) {   // both pixels, Newton step packed
  float2 r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.x) : "f"(x.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r.y) : "f"(x.y));
  const float2 e = __ffma2_rn(make_float2(-x.x, -x.y), r, make_float2(1.0f, 1.0f));
  return __ffma2_rn(r, e, r);
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
  // the adds of a level are issued in pairs where two run side by side (FADD2: keep.x + recv.x,
  // keep.y + recv.y) — 3 issue slots fewer per fold
#define B2GS_KEEP(up, lo, hi) ((up) ? (hi) : (lo))
#define B2GS_RECV(up, lo, hi, off) __shfl_xor_sync(FULL, (up) ? (lo) : (hi), off)
  const float2 w01 = __fadd2_rn(make_float2(B2GS_KEEP(b4, v[0], v[5]), B2GS_KEEP(b4, v[1], v[6])),
                                make_float2(B2GS_RECV(b4, v[0], v[5], 16), B2GS_RECV(b4, v[1], v[6], 16)));
  const float2 w23 = __fadd2_rn(make_float2(B2GS_KEEP(b4, v[2], v[7]), B2GS_KEEP(b4, v[3], v[8])),
                                make_float2(B2GS_RECV(b4, v[2], v[7], 16), B2GS_RECV(b4, v[3], v[8], 16)));
  const float w4 = xchg_add(b4, v[4], v[9], 16);
  const float2 u01 = __fadd2_rn(make_float2(B2GS_KEEP(b3, w01.x, w23.y), B2GS_KEEP(b3, w01.y, w4)),
                                make_float2(B2GS_RECV(b3, w01.x, w23.y, 8), B2GS_RECV(b3, w01.y, w4, 8)));
  const float u2 = xchg_add(b3, w23.x, 0.f, 8);
  const float2 t01 = __fadd2_rn(make_float2(B2GS_KEEP(b2, u01.x, u2), B2GS_KEEP(b2, u01.y, 0.f)),
                                make_float2(B2GS_RECV(b2, u01.x, u2, 4), B2GS_RECV(b2, u01.y, 0.f, 4)));
#undef B2GS_KEEP
#undef B2GS_RECV
  float s = xchg_add(b1, t01.x, t01.y, 2);
  s += __shfl_xor_sync(FULL, s, 1);
  if (slot_row) atomicAdd(slot_row + row_off, s);
}

// ---- packed fp32x2 helpers: .x = the lane's upper pixel, .y = its lower pixel (4 rows down) ----
// FFMA2 / FMUL2 / FADD2 (sm_100) issue one instruction for both pixels; each half is the IEEE
// single-rounding operation of the scalar instruction.
using f2 = float2;
__device__ __forceinline__ f2 bc(float s) { return make_float2(s, s); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float hsum(f2 a) { return a.x + a.y; }

constexpr int RB_THREADS = 128;   // 4 warps x (8x8 pixels, two per lane) = one 16x16 tile

// COUNT: debug instantiation that also counts the work it executes (b2gs_debug_set_counters):
// [4] warp visits, [5] warp visits that run the blend / gradient arithmetic, [6] contributing pairs
template <bool COUNT>
__global__ void __launch_bounds__(RB_THREADS)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ point_list, int W,
                  int H, int gx, const float* __restrict__ bg, const float2* __restrict__ xy,
                  const float4* __restrict__ conic_o, const float4* __restrict__ rgbd,
                  const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dpixels_depth,
                  float* __restrict__ acc, unsigned long long* __restrict__ counters, float expf_k,
                  float expf_252) {
  __shared__ SplatRec s_rec[RB_BATCH];
  __shared__ int s_max;
  unsigned c_visit = 0, c_full = 0, c_contrib = 0;
  pdl_wait();

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = (int)order[blockIdx.x];   // longest lists first (tile_scan_kernel)
  const int tile_x = tile % gx, tile_y = tile / gx;
  // warp -> 8x8 quadrant of the tile; lane -> column lane & 7, rows lane >> 3 and (lane >> 3) + 4:
  // the upper pixel lies in the 8x4 cull block 4 * (warp >> 1) + (warp & 1), the lower one in the
  // block two bits up (cull.cuh numbers the blocks 2 across, 4 down)
  const int bx = tile_x * TILE + (warp & 1) * 8, by = tile_y * TILE + (warp >> 1) * 8;
  const int px = bx + (lane & 7), pyA = by + (lane >> 3), pyB = pyA + 4;
  const bool insideA = px < W && pyA < H, insideB = px < W && pyB < H;
  const int pixA = W * pyA + px, pixB = W * pyB + px;
  const float pixfx = (float)px;
  const f2 npixfy = make_float2(-(float)pyA, -(float)pyB);
  const uint32_t wmask = 0x5u << (4 * (warp >> 1) + (warp & 1));

  const uint2 range = ranges[tile];

  const f2 T_final = make_float2(insideA ? final_Ts[pixA] : 0.f, insideB ? final_Ts[pixB] : 0.f);
  f2 T = T_final;
  const int lastA = insideA ? (int)n_contrib[pixA] : 0, lastB = insideB ? (int)n_contrib[pixB] : 0;

  // behind = accum . dL_dpixel, accum_c being the colour / depth blended behind the current splat.
  // The reference replays  accum_c = last_alpha * last_colour_c + (1 - last_alpha) * accum_c  per
  // channel when the NEXT contributor is visited and then forms  sum_c (colour_c - accum_c) *
  // dL_dpixel_c  (CR/backward.cu:716-732).  Only that dot product is ever used, and the replay is
  // linear, so the dot product itself is carried:  d = colour . dL_dpixel - behind,  behind +=
  // alpha * d  — one scalar per pixel instead of four channels, 6 instead of 12 packed
  // instructions per visit; the rounding error is of the same order (eps * |accum . dL_dpixel|
  // either way).  A pixel that skips a splat runs the same recurrences with alpha = 0 — a
  // transparent contributor: behind + 0 * d, T * 1/(1 - 0), both exact — so the two pixels of a
  // lane need no divergent state updates.
  f2 behind = bc(0.f);
  f2 dLp0 = bc(0.f), dLp1 = bc(0.f), dLp2 = bc(0.f), dLpd = bc(0.f);
  {
    const size_t plane = (size_t)H * W;
    if (insideA) {
      dLp0.x = dL_dpixels[0 * plane + pixA];
      dLp1.x = dL_dpixels[1 * plane + pixA];
      dLp2.x = dL_dpixels[2 * plane + pixA];
      dLpd.x = dL_dpixels_depth[pixA];
    }
    if (insideB) {
      dLp0.y = dL_dpixels[0 * plane + pixB];
      dLp1.y = dL_dpixels[1 * plane + pixB];
      dLp2.y = dL_dpixels[2 * plane + pixB];
      dLpd.y = dL_dpixels_depth[pixB];
    }
  }
  f2 bg_dot = mul2(bc(bg[0]), dLp0);
  bg_dot = fma2(bc(bg[1]), dLp1, bg_dot);
  bg_dot = fma2(bc(bg[2]), dLp2, bg_dot);
  // dL_dalpha += (-T_final / (1 - alpha)) * bg . dL_dpixel
  const f2 neg_Tf_bg = mul2(make_float2(-T_final.x, -T_final.y), bg_dot);

  const int slot = fold10_slot(lane);
  float* const slot_row = slot >= 0 ? acc + slot : nullptr;

  // deepest contributor over the warp / the tile
  const int warp_max = __reduce_max_sync(FULL, max(lastA, lastB));
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
  __syncthreads();
  const int tile_max = s_max;  // positions [0, tile_max) of the range matter

  for (int base = 0; base < tile_max; base += RB_BATCH) {
    const int cnt = min(RB_BATCH, tile_max - base);
    if (base) __syncthreads();
    for (int t = tid; t < cnt; t += RB_THREADS) {
      const int pos = tile_max - 1 - (base + t);
      const uint32_t v = point_list[range.x + pos];   // id << 8 | block mask
      const int id = (int)(v >> 8);
      SplatRec& r = s_rec[t];
      r.id = id;
      r.bm = v & 0xffu;
      if (v & 0xffu) {        // a splat that reaches no block of the tile is never visited
        r.co = conic_o[id];
        r.cd = rgbd[id];
        r.xy = xy[id];
      }
    }
    __syncthreads();

    // entry j is at position pos = tile_max-1-(base+j); a pixel is past its last contributor
    // unless pos < last  <=>  j > tile_max-1-base-last
    const int behindA = tile_max - 1 - base - lastA, behindB = tile_max - 1 - base - lastB;
    // this warp needs pos < warp_max
    int j0 = tile_max - base - warp_max;
    if (j0 < 0) j0 = 0;
    for (int c0 = j0 & ~31; c0 < cnt; c0 += 32) {
      // one lane per splat: can it reach this warp's 8x8 quadrant at all? (cull.cuh)
      const int jj = c0 + lane;
      const bool keep = jj >= j0 && jj < cnt && (s_rec[jj].bm & wmask);
      unsigned mask = __ballot_sync(FULL, keep);
      while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1;
        const int j = c0 + k;
        const SplatRec& r = s_rec[j];
        const float2 p = r.xy;
        const float4 co = r.co;
        // power: the forward's operation tree, operation for operation (render_fwd.cu), so that
        // both passes take the same skip decisions
        const float dx = p.x - pixfx;
        const f2 dy = add2(bc(p.y), npixfy);
        const float adx = co.x * dx, nbdx = -(co.y * dx);
        const f2 cdy2 = mul2(dy, mul2(bc(co.z), dy));
        const f2 quad = fma2(bc(dx), bc(adx), cdy2);
        const f2 power = fma2(quad, bc(-0.5f), mul2(dy, bc(nbdx)));
        const f2 G = expf_pair(power, expf_k, expf_252);
        const f2 og = mul2(bc(co.w), G);
        const float alphaA = fminf(0.99f, og.x), alphaB = fminf(0.99f, og.y);
        const bool skipA = j <= behindA || power.x > 0.0f || alphaA < 1.0f / 255.0f;
        const bool skipB = j <= behindB || power.y > 0.0f || alphaB < 1.0f / 255.0f;
        if (COUNT) ++c_visit;
        // no warp vote on "every pixel skips" (0.8 % of the visits of the benched view: the block
        // mask is tight); such a visit adds exact zeros to the splat's row
        if (COUNT) { ++c_full; c_contrib += (unsigned)!skipA + (unsigned)!skipB; }

        const f2 alpha = make_float2(skipA ? 0.f : alphaA, skipB ? 0.f : alphaB);
        const f2 Gs = make_float2(skipA ? 0.f : G.x, skipB ? 0.f : G.y);
        const float4 cd = r.cd;
        const f2 oma = add2(bc(1.f), make_float2(-alpha.x, -alpha.y));
        const f2 inv1ma = rcp_unit_range2(oma);
        T = mul2(T, inv1ma);  // reference: T / (1 - alpha); differs by <= 1 ulp
        const f2 dchannel_dcolor = mul2(alpha, T);

        // d = sum_c (colour_c - accum_c) * dL_dpixel_c = colour . dL_dpixel - behind
        f2 d = fma2(bc(cd.x), dLp0, make_float2(-behind.x, -behind.y));
        d = fma2(bc(cd.y), dLp1, d);
        d = fma2(bc(cd.z), dLp2, d);
        d = fma2(bc(cd.w), dLpd, d);
        behind = fma2(alpha, d, behind);   // what lies behind the NEXT entry (skipping pixel: + 0)
        const f2 dL_dalpha = fma2(inv1ma, neg_Tf_bg, mul2(d, T));

        // per-splat sums over the pixels.  The conic, the opacity and the viewport factors do not
        // depend on the pixel, so what is folded are the moments of q = G * dL_dalpha about the
        // splat centre (d = mean - pixel):  S0 = sum q,  Sx = sum q dx,  Sy = sum q dy,  Sxx, Sxy,
        // Syy; preprocess_bwd forms dL_dmean2D = o * (-a Sx - b Sy) * W/2, ...,  dL_dconic =
        // -o/2 * (Sxx, Sxy, Syy) from them (CR/backward.cu:735-757 distributes the same factors
        // over every pixel).  Both pixels of a lane share dx, so the x moments are dx times the
        // lane's sums.  A skipping pixel contributes exact zeros (q carries Gs = 0, the colour
        // terms alpha = 0).
        const f2 q = mul2(Gs, dL_dalpha);
        const f2 qy = mul2(q, dy);
        float v[10];
        v[5] = hsum(q);
        v[1] = hsum(qy);
        v[4] = hsum(mul2(qy, dy));
        v[0] = v[5] * dx;
        v[2] = v[0] * dx;
        v[3] = v[1] * dx;
        v[6] = hsum(mul2(dchannel_dcolor, dLp0));
        v[7] = hsum(mul2(dchannel_dcolor, dLp1));
        v[8] = hsum(mul2(dchannel_dcolor, dLp2));
        v[9] = hsum(mul2(dchannel_dcolor, dLpd));
        warp_fold10_and_red(v, lane, slot_row, (size_t)r.id * ACC_STRIDE);
      }
    }
  }
  if (COUNT) {
    c_contrib = __reduce_add_sync(FULL, c_contrib);
    if (lane == 0) {
      atomicAdd(&counters[4], (unsigned long long)c_visit);
      atomicAdd(&counters[5], (unsigned long long)c_full);
      atomicAdd(&counters[6], (unsigned long long)c_contrib);
    }
  }
}

int launch_render_bwd(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                      const B2gsBackwardIO& io, float* acc, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  B2GS_PROF("render_bwd", st);
  unsigned long long* cnt = debug_counters();
  if (cnt)
    B2GS_LAUNCH(render_bwd_kernel<true>, gx * gy, RB_THREADS, st, im.ranges, im.order, b.point_list, s.W, s.H, gx, s.bg, g.xy,
                                                            g.conic_o, g.rgbd, im.final_T, im.n_contrib,
                                                            io.dL_dcolor, io.dL_ddepth, acc, cnt, expf_k_host(), 252.0f);
  else
    B2GS_LAUNCH(render_bwd_kernel<false>, gx * gy, RB_THREADS, st, im.ranges, im.order, b.point_list, s.W, s.H, gx, s.bg, g.xy,
                                                             g.conic_o, g.rgbd, im.final_T, im.n_contrib,
                                                             io.dL_dcolor, io.dL_ddepth, acc, nullptr, expf_k_host(), 252.0f);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

}  // namespace b2gs
