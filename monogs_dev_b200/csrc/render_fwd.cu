// render_fwd.cu — forward alpha blending, one CTA (128 threads) per 16x16 tile.
// Replaces renderCUDA<3> forward (CR/forward.cu:263-392).
//
// Structure (per-pixel results identical to the reference: same operation tree for `power`,
// the library's expf bit for bit, the reference's skip / termination tests):
//  * a warp covers an 8x8 quadrant of the tile, each lane TWO pixels (rows r and r+4); the
//    per-pixel arithmetic of the pair is packed fp32x2 (FFMA2 / FMUL2 / FADD2: one issue slot
//    for both pixels, each half an IEEE single-rounding op);
//  * a batch of the tile's list is staged in shared memory as one 48 B SplatRec per splat
//    (conic+opacity, rgb+depth, mean, id, block mask): a visit is one address and three
//    broadcast loads, no per-pair global colour fetch (CR/forward.cu:365);
//  * the block mask (cull.cuh) was computed once by the instance scatter and arrives in the low
//    byte of the sorted list word; a warp only walks the splats whose mask reaches its quadrant;
//  * warp-level early termination; n_touched (CR/forward.cu:369-371) is two ballots per visit
//    while some pixel of the warp still has T > 0.5, one red.shared from lane 0, and ONE global
//    atomic per (tile, splat) instead of one per (pixel, splat).
#include "common.cuh"
#include "cull.cuh"
#include "loss_terms.cuh"
#include <string.h>

namespace b2gs {

constexpr int RF_BATCH = 256;
constexpr int RF_THREADS = 128;   // 4 warps x (8x8 pixels, two per lane) = one 16x16 tile
constexpr unsigned RF_FULL = 0xffffffffu;

// packed fp32x2: .x = the lane's upper pixel, .y = its lower pixel (4 rows down); each half is the
// IEEE single-rounding operation of the scalar instruction (FFMA2 / FMUL2 / FADD2, sm_100)
using f2 = float2;
__device__ __forceinline__ f2 bc(float s) { return make_float2(s, s); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }

// COUNT: debug instantiation that also counts the work it executes (b2gs_debug_set_counters):
// [0] warp visits, [1] warp visits that evaluate exp / blend, [2] contributing pixel-splat pairs
// LOSS: the view's mapping / tracking loss and its gradient wrt the pixel are formed in the epilogue
// from the blended values still in registers (B2gsFusedLoss); per-tile partial sums go to
// fl.partials for b2gs_window_finalize.
template <bool COUNT, bool LOSS>
__global__ void __launch_bounds__(RF_THREADS)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ point_list, int W,
                  int H, int gx, const float2* __restrict__ xy, const float4* __restrict__ conic_o,
                  const float4* __restrict__ rgbd, const float* __restrict__ bg,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_depth,
                  float* __restrict__ out_opacity, int* __restrict__ n_touched,
                  unsigned long long* __restrict__ counters, const __grid_constant__ B2gsFusedLoss fl,
                  float expf_k, float expf_252) {
  __shared__ SplatRec s_rec[RF_BATCH];
  unsigned c_visit = 0, c_full = 0, c_contrib = 0, c_blocks = 0;
  __shared__ int s_cnt[RF_BATCH];
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
  const int n = (int)(range.y - range.x);

  // fused loss: this lane's two ground-truth pixels are pulled into L2 now (prefetch: no register is
  // held across the blend loop — as register prefetches they cost the kernel a CTA per SM) and
  // loaded in the epilogue
  if (LOSS) {
    if (insideA) prefetch_gt(fl.args, pixA);
    if (insideB) prefetch_gt(fl.args, pixB);
  }
  uint32_t cnt_base = (uint32_t)__cvta_generic_to_shared(s_cnt);
  asm volatile("" : "+r"(cnt_base));   // keep it in its register (else re-derived from %cluster_ctaid per visit)
  bool doneA = !insideA, doneB = !insideB;
  f2 T = bc(1.0f);
  f2 C0 = bc(0.f), C1 = bc(0.f), C2 = bc(0.f), D = bc(0.f);
  uint32_t lastA = 0, lastB = 0;

  for (int base = 0; base < n; base += RF_BATCH) {
    // whole tile finished?  (also the barrier that protects the staging buffers)
    if (__syncthreads_count(doneA && doneB) == RF_THREADS) break;
    const int cnt = min(RF_BATCH, n - base);
    for (int t = tid; t < RF_BATCH; t += RF_THREADS) {
      if (t < cnt) {
        const uint32_t v = point_list[range.x + base + t];   // id << 8 | block mask
        const int id = (int)(v >> 8);
        SplatRec& r = s_rec[t];
        r.bm = v & 0xffu;
        r.id = id;
        if (v & 0xffu) {      // a splat that reaches no block of the tile is never visited
          r.co = conic_o[id];
          r.cd = rgbd[id];
          r.xy = xy[id];
        }
      }
      s_cnt[t] = 0;
    }
    __syncthreads();

    if (!__all_sync(RF_FULL, doneA && doneB)) {
      int lastjA = -1, lastjB = -1;
      for (int c0 = 0; c0 < cnt; c0 += 32) {
        // can the splat reach this warp's quadrant at all?  (bits computed once by the binning
        // stage, cull.cuh: splat_block_mask)
        const int jj = c0 + lane;
        const bool keep = jj < cnt && (s_rec[jj].bm & wmask);
        unsigned mask = __ballot_sync(RF_FULL, keep);
        // n_touched counts pixels whose transmittance is still above 0.5 after the splat; T only
        // falls, so once no pixel of the warp is above 0.5 the two ballots per visit are dead code
        const bool hot = __any_sync(RF_FULL, (!doneA && T.x > 0.5f) || (!doneB && T.y > 0.5f));
        while (mask) {
          const int k = __ffs(mask) - 1;
          mask &= mask - 1;
          const int j = c0 + k;
          const SplatRec& r = s_rec[j];
          const float2 p = r.xy;
          const float4 co = r.co;
          // power = -0.5f * (a dx dx + c dy dy) - b dx dy in the operation tree the reference's
          // expression compiles to (CR/forward.cu:346): (c*dy)*dy, fma(dx, a*dx, .), (b*dx)*dy,
          // fma(., -0.5, -.)  — render_bwd.cu evaluates the same tree
          const float dx = p.x - pixfx;
          const f2 dy = add2(bc(p.y), npixfy);
          const float adx = co.x * dx, nbdx = -(co.y * dx);
          const f2 cdy2 = mul2(dy, mul2(bc(co.z), dy));
          const f2 quad = fma2(bc(dx), bc(adx), cdy2);
          const f2 power = fma2(quad, bc(-0.5f), mul2(dy, bc(nbdx)));
          bool actA = !doneA && !(power.x > 0.0f), actB = !doneB && !(power.y > 0.0f);
          if (COUNT) { ++c_visit; c_blocks += __popc(r.bm & wmask); }
          // no warp vote on (actA || actB) here: the block mask already says that the splat reaches
          // this quadrant, so "no live pixel with power <= 0" never happened in 1.6 M visits of the
          // benched view (ncu source counters), and the vote + branch were 6 of 88 instructions per
          // visit.  A visit that no pixel takes blends alpha = 0 everywhere: exact no-ops.
          if (COUNT) ++c_full;
          const f2 og = mul2(bc(co.w), expf_pair(power, expf_k, expf_252));
          const float alphaA = fminf(0.99f, og.x), alphaB = fminf(0.99f, og.y);
          actA = actA && !(alphaA < 1.0f / 255.0f);
          actB = actB && !(alphaB < 1.0f / 255.0f);
          const f2 test_T = mul2(T, add2(bc(1.f), make_float2(-alphaA, -alphaB)));
          if (actA && test_T.x < 0.0001f) { doneA = true; actA = false; }
          if (actB && test_T.y < 0.0001f) { doneB = true; actB = false; }
          // a pixel that does not take the splat blends it with alpha = 0: exact no-ops
          const f2 w = mul2(make_float2(actA ? alphaA : 0.f, actB ? alphaB : 0.f), T);
          const float4 cd = r.cd;
          C0 = fma2(bc(cd.x), w, C0);
          C1 = fma2(bc(cd.y), w, C1);
          C2 = fma2(bc(cd.z), w, C2);
          D = fma2(bc(cd.w), w, D);
          T = make_float2(actA ? test_T.x : T.x, actB ? test_T.y : T.y);
          if (COUNT) c_contrib += (unsigned)actA + (unsigned)actB;
          // last contributor inside this batch (list position = base + j + 1, formed after the batch)
          lastjA = actA ? j : lastjA;
          lastjB = actB ? j : lastjB;
          // n_touched: how many pixels saw this splat with T still above 0.5
          if (hot) {
            // one warp-wide integer add (REDUX) instead of two ballots + two popcounts
            const unsigned mine = (unsigned)(actA && test_T.x > 0.5f) + (unsigned)(actB && test_T.y > 0.5f);
            const unsigned tot = __reduce_add_sync(RF_FULL, mine);
            if (tot != 0u && lane == 0) {   // plain red.shared: one lane, nothing to aggregate
              asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(cnt_base + 4u * (uint32_t)j), "r"(tot) : "memory");
            }
          }
        }
        if (__all_sync(RF_FULL, doneA && doneB)) break;  // warp-level early termination
      }
      if (lastjA >= 0) lastA = (uint32_t)(base + lastjA + 1);
      if (lastjB >= 0) lastB = (uint32_t)(base + lastjB + 1);
    }
    __syncthreads();
    for (int t = tid; t < cnt; t += RF_THREADS) {
      const int c = s_cnt[t];
      if (c) atomicAdd(&n_touched[s_rec[t].id], c);
    }
  }

  if (COUNT) {
    c_contrib = __reduce_add_sync(RF_FULL, c_contrib);
    if (lane == 0) {
      atomicAdd(&counters[0], (unsigned long long)c_visit);
      atomicAdd(&counters[1], (unsigned long long)c_full);
      atomicAdd(&counters[2], (unsigned long long)c_contrib);
      atomicAdd(&counters[3], (unsigned long long)c_blocks);   // (splat, 8x4 block) pairs among the visits
    }
  }
  const size_t plane = (size_t)H * W;
  if (insideA) {
    final_T[pixA] = T.x;
    n_contrib[pixA] = lastA;
    out_color[0 * plane + pixA] = C0.x + T.x * bg[0];
    out_color[1 * plane + pixA] = C1.x + T.x * bg[1];
    out_color[2 * plane + pixA] = C2.x + T.x * bg[2];
    out_depth[pixA] = D.x;
    out_opacity[pixA] = 1 - T.x;
  }
  if (insideB) {
    final_T[pixB] = T.y;
    n_contrib[pixB] = lastB;
    out_color[0 * plane + pixB] = C0.y + T.y * bg[0];
    out_color[1 * plane + pixB] = C1.y + T.y * bg[1];
    out_color[2 * plane + pixB] = C2.y + T.y * bg[2];
    out_depth[pixB] = D.y;
    out_opacity[pixB] = 1 - T.y;
  }
  if (LOSS) {
    // loss + its gradient for this tile's pixels (loss_terms.cuh), upstream gradient 1
    __shared__ float s_loss[RF_THREADS / 32][4];
    float ea, eb;
    exposure(fl.args, ea, eb);
    const bool mono = fl.args.mode & 2;
    const float w_rgb = (mono ? 1.f : fl.args.alpha) / (3.0f * (float)fl.args.N);
    const float w_dep = mono ? 0.f : (1.f - fl.args.alpha) / (float)fl.args.N;
    float red[4] = {0.f, 0.f, 0.f, 0.f};
    if (insideA) {
      const PixelLoss r = pixel_loss(fl.args, ea, eb, w_rgb, w_dep, pixA, C0.x + T.x * bg[0], C1.x + T.x * bg[1],
                                     C2.x + T.x * bg[2], D.x, 1 - T.x, load_gt(fl.args, pixA));
      fl.dL_dimage[0 * plane + pixA] = r.g[0]; fl.dL_dimage[1 * plane + pixA] = r.g[1];
      fl.dL_dimage[2 * plane + pixA] = r.g[2]; fl.dL_ddepth[pixA] = r.gd;
      red[0] += r.rgb; red[1] += r.dep; red[2] += r.ga; red[3] += r.gb;
    }
    if (insideB) {
      const PixelLoss r = pixel_loss(fl.args, ea, eb, w_rgb, w_dep, pixB, C0.y + T.y * bg[0], C1.y + T.y * bg[1],
                                     C2.y + T.y * bg[2], D.y, 1 - T.y, load_gt(fl.args, pixB));
      fl.dL_dimage[0 * plane + pixB] = r.g[0]; fl.dL_dimage[1 * plane + pixB] = r.g[1];
      fl.dL_dimage[2 * plane + pixB] = r.g[2]; fl.dL_ddepth[pixB] = r.gd;
      red[0] += r.rgb; red[1] += r.dep; red[2] += r.ga; red[3] += r.gb;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) red[k] += __shfl_xor_sync(RF_FULL, red[k], d);
      if (lane == 0) s_loss[warp][k] = red[k];
    }
    __syncthreads();
    if (tid < 4) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < RF_THREADS / 32; ++w) v += s_loss[w][tid];   // fixed order: deterministic
      fl.partials[4 * (size_t)tile + tid] = v;
    }
  }
}

int launch_render_fwd(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                      const B2gsForwardOut& out, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  B2GS_PROF("render_fwd", st);
  unsigned long long* cnt = debug_counters();
  B2gsFusedLoss fl;
  memset(&fl, 0, sizeof(fl));
  if (out.fused_loss) fl = *out.fused_loss;
#define B2GS_RF_LAUNCH(C_, L_)                                                                        \
  B2GS_LAUNCH((render_fwd_kernel<C_, L_>), gx * gy, RF_THREADS, st,                                  \
      im.ranges, im.order, b.point_list, s.W, s.H, gx, g.xy, g.conic_o, g.rgbd, s.bg, im.final_T,     \
      im.n_contrib,                                                                                  \
      out.color, out.depth, out.opacity, out.n_touched, cnt, fl, expf_k_host(), 252.0f)
  if (out.fused_loss) {
    if (cnt) B2GS_RF_LAUNCH(true, true); else B2GS_RF_LAUNCH(false, true);
  } else {
    if (cnt) B2GS_RF_LAUNCH(true, false); else B2GS_RF_LAUNCH(false, false);
  }
#undef B2GS_RF_LAUNCH
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

}  // namespace b2gs

namespace b2gs {
__global__ void expf_pair_check_kernel(uint32_t first, uint32_t last, unsigned long long* out) {
  const unsigned long long span = (unsigned long long)last - first + 1ull;
  unsigned long long bad = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < span;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t bits = first + (uint32_t)i;
    const float x = __uint_as_float(bits);
    const float y = __uint_as_float(first + (uint32_t)((i * 2654435761ull) % span));   // an unrelated partner
    const float2 e = expf_pair(make_float2(x, y));
    const float ex = expf(x), ey = expf(y);
    bad += (__float_as_uint(e.x) != __float_as_uint(ex)) || (__float_as_uint(e.y) != __float_as_uint(ey));
  }
  bad = __reduce_add_sync(0xffffffffu, (unsigned)bad);   // < 2^32 per warp: a lane visits < 2^27 patterns
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, bad);
}
}  // namespace b2gs

extern "C" int b2gs_debug_expf_pair_check(uint32_t first_bits, uint32_t last_bits,
                                          unsigned long long* mismatches, void* stream) {
  using namespace b2gs;
  if (!mismatches || last_bits < first_bits) {
    set_error("b2gs_debug_expf_pair_check: NULL output or empty bit range");
    return B2GS_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B2GS_CUDA(cudaMemsetAsync(mismatches, 0, sizeof(unsigned long long), st));
  expf_pair_check_kernel<<<NUM_SMS_HINT * 16, 256, 0, st>>>(first_bits, last_bits, mismatches);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

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
