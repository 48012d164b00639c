// loss.cu — the per-pixel side of one mapping / tracking iteration as two HBM-bound
// kernels, and the SE(3) pose update as one tiny kernel.
//
// Replaces the ~15 elementwise / reduction torch kernels per view of
//   get_loss_mapping(_rgb|_rgbd)    utils/slam_utils.py:91-128
//   get_loss_tracking(_rgb|_rgbd)   utils/slam_utils.py:56-88
// (exposure affine exp(a)*image+b, rgb boundary mask, depth mask, opacity weighting,
// masked L1, two means, alpha blend) and, in backward, autograd's replay of them.
// Forward reads 32-36 B/pixel and writes one scalar; backward reads the same and writes
// dL/dimage, dL/ddepth (16 B/pixel) plus the exposure gradients.
//
// Also update_pose (utils/pose_utils.py:76-93): T_w2c <- Exp(tau) T_w2c entirely on the
// device, refreshing the view / full-projection matrices and the camera centre in place,
// so that a whole optimisation step is CUDA-graph capturable (no .cpu() round trip).
#include "common.cuh"
#include "loss_terms.cuh"
#include "pose_math.cuh"

namespace b2gs {

constexpr int LOSS_THREADS = 256;
constexpr int LOSS_MAX_BLOCKS = 148 * 8;

__device__ __forceinline__ PixelTerms pixel_terms(const B2gsLossArgs& a, float ea, float eb, int i) {
  const size_t N = (size_t)a.N;
  return pixel_terms_of(a, ea, eb, i, a.image[i], a.image[N + i], a.image[2 * N + i],
                        (a.mode & 2) ? 0.f : a.depth[i], (a.mode & 1) ? a.opacity[i] : 1.f, load_gt(a, i));
}

__device__ __forceinline__ float block_sum(float v, float* s_red) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = lane < LOSS_THREADS / 32 ? s_red[lane] : 0.f;
#pragma unroll
    for (int d = 4; d > 0; d >>= 1) r += __shfl_xor_sync(0xffffffffu, r, d);
  }
  return r;  // valid in thread 0
}

// per-block partials: [0] sum |rgb term|, [1] sum |depth term|
__global__ void __launch_bounds__(LOSS_THREADS)
loss_fwd_kernel(const __grid_constant__ B2gsLossArgs a, float* __restrict__ partial) {
  __shared__ float s_red[LOSS_THREADS / 32];
  float ea, eb;
  exposure(a, ea, eb);
  float rgb = 0.f, dep = 0.f;
  for (int i = blockIdx.x * LOSS_THREADS + threadIdx.x; i < a.N; i += gridDim.x * LOSS_THREADS) {
    const PixelTerms t = pixel_terms(a, ea, eb, i);
    rgb += t.o * (fabsf(t.d[0]) + fabsf(t.d[1]) + fabsf(t.d[2]));
    dep += fabsf(t.dd);
  }
  rgb = block_sum(rgb, s_red);
  dep = block_sum(dep, s_red);
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = rgb;
    partial[2 * blockIdx.x + 1] = dep;
  }
}

// one block: fixed-order (deterministic) sum of the per-block partials in double
__global__ void __launch_bounds__(LOSS_THREADS)
loss_fwd_finalize_kernel(int blocks, int N, float alpha, int mono, const float* __restrict__ partial,
                         float* __restrict__ loss) {
  __shared__ double s_a[LOSS_THREADS], s_b[LOSS_THREADS];
  double ra = 0.0, rb = 0.0;
  for (int i = threadIdx.x; i < blocks; i += LOSS_THREADS) {
    ra += partial[2 * i];
    rb += partial[2 * i + 1];
  }
  s_a[threadIdx.x] = ra;
  s_b[threadIdx.x] = rb;
  __syncthreads();
  for (int d = LOSS_THREADS / 2; d > 0; d >>= 1) {
    if (threadIdx.x < d) {
      s_a[threadIdx.x] += s_a[threadIdx.x + d];
      s_b[threadIdx.x] += s_b[threadIdx.x + d];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double l_rgb = s_a[0] / (3.0 * (double)N), l_dep = s_b[0] / (double)N;
    // monocular: l1.mean() alone (utils/slam_utils.py:72,109); rgbd: alpha blend (:88,128)
    loss[0] = mono ? (float)l_rgb : (float)((double)alpha * l_rgb + (1.0 - (double)alpha) * l_dep);
  }
}

// per-block partials: [0] dL/d exposure_a, [1] dL/d exposure_b
__global__ void __launch_bounds__(LOSS_THREADS)
loss_bwd_kernel(const __grid_constant__ B2gsLossArgs a, const float* __restrict__ upstream,
                float* __restrict__ dL_dimage, float* __restrict__ dL_ddepth,
                float* __restrict__ dL_dopacity, float* __restrict__ partial) {
  __shared__ float s_red[LOSS_THREADS / 32];
  float ea, eb;
  exposure(a, ea, eb);
  const float up = upstream ? upstream[0] : 1.f;
  const bool mono = a.mode & 2;
  const float w_rgb = up * (mono ? 1.f : a.alpha) / (3.0f * (float)a.N);
  const float w_dep = mono ? 0.f : up * (1.f - a.alpha) / (float)a.N;
  const size_t N = (size_t)a.N;
  float ga = 0.f, gb = 0.f;
  for (int i = blockIdx.x * LOSS_THREADS + threadIdx.x; i < a.N; i += gridDim.x * LOSS_THREADS) {
    const PixelTerms t = pixel_terms(a, ea, eb, i);
    float absum = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      // d|x|/dx * d(image_ab*m)/d(image_ab) = sign * m ; image_ab = ea*image + eb
      const float g_ab = w_rgb * t.o * sgnf(t.d[c]) * t.m;
      const float img = a.image[c * N + i];
      dL_dimage[c * N + i] = g_ab * ea;
      ga += g_ab * img;
      gb += g_ab;
      absum += fabsf(t.d[c]);
    }
    dL_ddepth[i] = w_dep * sgnf(t.dd) * t.md;
    if (dL_dopacity) dL_dopacity[i] = (a.mode & 1) ? w_rgb * absum : 0.f;
  }
  ga = block_sum(ga, s_red);
  gb = block_sum(gb, s_red);
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = ga * ea;  // d exp(a)/da = exp(a)
    partial[2 * blockIdx.x + 1] = gb;
  }
}

__global__ void __launch_bounds__(LOSS_THREADS)
loss_bwd_finalize_kernel(int blocks, const float* __restrict__ partial, float* __restrict__ dL_dexp) {
  __shared__ double s_a[LOSS_THREADS], s_b[LOSS_THREADS];
  double ra = 0.0, rb = 0.0;
  for (int i = threadIdx.x; i < blocks; i += LOSS_THREADS) {
    ra += partial[2 * i];
    rb += partial[2 * i + 1];
  }
  s_a[threadIdx.x] = ra;
  s_b[threadIdx.x] = rb;
  __syncthreads();
  for (int d = LOSS_THREADS / 2; d > 0; d >>= 1) {
    if (threadIdx.x < d) {
      s_a[threadIdx.x] += s_a[threadIdx.x + d];
      s_b[threadIdx.x] += s_b[threadIdx.x + d];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    dL_dexp[0] = (float)s_a[0];
    dL_dexp[1] = (float)s_b[0];
  }
}

static inline int loss_blocks(int N) {
  int b = (N + LOSS_THREADS - 1) / LOSS_THREADS;
  return b < 1 ? 1 : (b > LOSS_MAX_BLOCKS ? LOSS_MAX_BLOCKS : b);
}

static int check_loss(const B2gsLossArgs* a) {
  if (!a) { set_error("loss: NULL args"); return B2GS_ERR_INVALID; }
  if (a->N <= 0) { set_error("loss: N must be > 0"); return B2GS_ERR_INVALID; }
  if (a->mode < 0 || a->mode > 3) { set_error("loss: bad mode %d", a->mode); return B2GS_ERR_INVALID; }
  if (!a->image || !a->gt_image) { set_error("loss: image / gt_image are NULL"); return B2GS_ERR_INVALID; }
  if (!(a->mode & 2) && (!a->depth || !a->gt_depth)) { set_error("loss: depth / gt_depth are NULL"); return B2GS_ERR_INVALID; }
  if ((a->mode & 1) && !a->opacity) { set_error("loss: tracking needs opacity"); return B2GS_ERR_INVALID; }
  return B2GS_OK;
}

}  // namespace b2gs

using namespace b2gs;

extern "C" size_t b2gs_loss_scratch_bytes(int32_t N) {
  (void)N;
  return align_up((size_t)LOSS_MAX_BLOCKS * 2 * sizeof(float), 256);
}

extern "C" int b2gs_loss_forward(const B2gsLossArgs* args, float* loss, void* scratch,
                                 size_t scratch_bytes, void* stream) {
  int rc = check_loss(args);
  if (rc) return rc;
  if (!loss || !scratch || scratch_bytes < b2gs_loss_scratch_bytes(args->N)) {
    set_error("loss: NULL output or scratch too small");
    return B2GS_ERR_WORKSPACE;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int blocks = loss_blocks(args->N);
  float* partial = reinterpret_cast<float*>(scratch);
  {
    B2GS_PROF("loss_fwd", st);
    loss_fwd_kernel<<<blocks, LOSS_THREADS, 0, st>>>(*args, partial);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  loss_fwd_finalize_kernel<<<1, LOSS_THREADS, 0, st>>>(blocks, args->N, args->alpha,
                                                       (args->mode & 2) ? 1 : 0, partial, loss);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

extern "C" int b2gs_loss_backward(const B2gsLossArgs* args, const float* dL_dloss, float* dL_dimage,
                                  float* dL_ddepth, float* dL_dopacity, float* dL_dexposure,
                                  void* scratch, size_t scratch_bytes, void* stream) {
  int rc = check_loss(args);
  if (rc) return rc;
  if (!dL_dimage || !dL_ddepth || !dL_dexposure || !scratch ||
      scratch_bytes < b2gs_loss_scratch_bytes(args->N)) {
    set_error("loss backward: NULL output or scratch too small");
    return B2GS_ERR_WORKSPACE;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int blocks = loss_blocks(args->N);
  float* partial = reinterpret_cast<float*>(scratch);
  {
    B2GS_PROF("loss_bwd", st);
    loss_bwd_kernel<<<blocks, LOSS_THREADS, 0, st>>>(*args, dL_dloss, dL_dimage, dL_ddepth,
                                                     dL_dopacity, partial);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  loss_bwd_finalize_kernel<<<1, LOSS_THREADS, 0, st>>>(blocks, partial, dL_dexposure);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

// ---- pose update ---------------------------------------------------------------------
namespace b2gs {

__global__ void pose_update_kernel(float* __restrict__ state, const float* __restrict__ proj,
                                   float* __restrict__ rot_delta, float* __restrict__ trans_delta,
                                   float* __restrict__ tau_norm, const int32_t* __restrict__ guard) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (guard && *guard) return;   // the step is being discarded (b2gs_set_update_guard)
  pose_update_device(state, proj, rot_delta, trans_delta, tau_norm);
}

}  // namespace b2gs

extern "C" int b2gs_pose_update(float* cam_state, const float* projmatrix_raw, float* rot_delta,
                                float* trans_delta, float* tau_norm, void* stream) {
  if (!cam_state || !projmatrix_raw || !rot_delta || !trans_delta) {
    set_error("b2gs_pose_update: NULL pointer");
    return B2GS_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  pose_update_kernel<<<1, 32, 0, st>>>(cam_state, projmatrix_raw, rot_delta, trans_delta, tau_norm,
                                       update_guard());
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

// ---- isotropic-scale regulariser -------------------------------------------------------
// utils/slam_backend.py:229-231:  scaling = exp(_scaling);
//   loss += 10 * |scaling - scaling.mean(dim=1, keepdim)|.mean()
// Forward value and the gradient wrt the RAW log-scales in one pass, accumulated into the
// optimiser's gradient buffer (replaces ~8 torch kernels fwd + their autograd replay).
namespace b2gs {
__global__ void __launch_bounds__(LOSS_THREADS)
iso_reg_kernel(int P, const float* __restrict__ raw_scaling, float weight,
               float* __restrict__ dL_draw, float* __restrict__ partial) {
  __shared__ float s_red[LOSS_THREADS / 32];
  const float w = weight / (3.0f * (float)P);
  float acc = 0.f;
  for (int i = blockIdx.x * LOSS_THREADS + threadIdx.x; i < P; i += gridDim.x * LOSS_THREADS) {
    const float s0 = expf(raw_scaling[3 * i]), s1 = expf(raw_scaling[3 * i + 1]), s2 = expf(raw_scaling[3 * i + 2]);
    const float m = (s0 + s1 + s2) / 3.0f;
    const float d0 = s0 - m, d1 = s1 - m, d2 = s2 - m;
    acc += fabsf(d0) + fabsf(d1) + fabsf(d2);
    if (dL_draw) {
      const float g0 = sgnf(d0), g1 = sgnf(d1), g2 = sgnf(d2);
      const float gm = (g0 + g1 + g2) / 3.0f;          // through the mean
      // d exp(x) = exp(x) dx; red.global.add: the views' backward kernels may be adding to the same
      // arena on other streams
      atomicAdd(&dL_draw[3 * i], w * (g0 - gm) * s0);
      atomicAdd(&dL_draw[3 * i + 1], w * (g1 - gm) * s1);
      atomicAdd(&dL_draw[3 * i + 2], w * (g2 - gm) * s2);
    }
  }
  acc = block_sum(acc, s_red);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(LOSS_THREADS)
iso_reg_finalize_kernel(int blocks, float scale, const float* __restrict__ partial, float* __restrict__ loss) {
  __shared__ double s_a[LOSS_THREADS];
  double ra = 0.0;
  for (int i = threadIdx.x; i < blocks; i += LOSS_THREADS) ra += partial[i];
  s_a[threadIdx.x] = ra;
  __syncthreads();
  for (int d = LOSS_THREADS / 2; d > 0; d >>= 1) {
    if (threadIdx.x < d) s_a[threadIdx.x] += s_a[threadIdx.x + d];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = (float)(s_a[0] * (double)scale);
}
}  // namespace b2gs

extern "C" int b2gs_scale_regulariser(int32_t P, const float* raw_scaling, float weight,
                                      float* dL_draw_scaling, float* loss, void* scratch,
                                      size_t scratch_bytes, void* stream) {
  if (P <= 0 || !raw_scaling || !loss || !scratch || scratch_bytes < b2gs_loss_scratch_bytes(P)) {
    set_error("b2gs_scale_regulariser: bad arguments");
    return B2GS_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int blocks = loss_blocks(P);
  float* partial = reinterpret_cast<float*>(scratch);
  B2GS_PROF("iso_reg", st);
  iso_reg_kernel<<<blocks, LOSS_THREADS, 0, st>>>(P, raw_scaling, weight, dL_draw_scaling, partial);
  B2GS_LAUNCH_CHECK(st, 0);
  iso_reg_finalize_kernel<<<1, LOSS_THREADS, 0, st>>>(blocks, weight / (3.0f * (float)P), partial, loss);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}
