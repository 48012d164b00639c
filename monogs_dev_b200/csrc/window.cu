// window.cu — the scalar tail of one mapping-window / tracking step as TWO small kernels, so that a
// captured step contains no framework (torch) kernel at all:
//
//   b2gs_window_finalize   after every local view's forward + backward: per view, the fixed-order
//       (deterministic) sums of the fused loss partials (render_fwd epilogue) and of the
//       preprocess-backward partial rows -> the view's loss value, dL/d exposure_a,b and dL_dtau;
//       then, by the last block, the step's total loss and the OR of the views' binning-overflow
//       flags (the guard word of the update kernels).
//       Replaces loss_fwd_finalize + loss_bwd_finalize + tau_finalize per view, torch's
//       `loss = loss + term` adds, autograd's AccumulateGrad adds into the camera parameters,
//       `torch.stack(status).max()`.
//   b2gs_pose_step         Adam over (cam_rot_delta, cam_trans_delta, exposure_a, exposure_b) of every
//       local keyframe (torch.optim.Adam(opt_params) semantics, utils/slam_backend.py:434-471,
//       utils/slam_frontend.py:132-161) followed by update_pose (utils/pose_utils.py:76-93) and the
//       step-counter ticks.  Replaces the keyframe optimizer's launches, one pose_update launch per
//       keyframe and the gradient zero fills.
#include "common.cuh"
#include "pose_math.cuh"
#include <math.h>
#include <string.h>

namespace b2gs {

constexpr int WF_THREADS = 1024;
constexpr int WF_MAX_VIEWS = 32;
constexpr int PS_MAX_CAMS = 24;

struct FinalizeTable {
  B2gsViewFinalize v[WF_MAX_VIEWS];
  int n;
};

__global__ void __launch_bounds__(WF_THREADS)
window_finalize_kernel(const __grid_constant__ FinalizeTable t, float* __restrict__ loss_total,
                       int32_t* __restrict__ guard, uint32_t* __restrict__ ticket) {
  __shared__ double s_red[WF_THREADS / 32][11];
  __shared__ bool s_last;
  pdl_wait();
  const B2gsViewFinalize& v = t.v[blockIdx.x];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // r[0..3]: loss partial sums, r[4..10]: tau[6] + regulariser.  Every thread's share is summed in a
  // fixed order and the cross-thread reduction order is fixed too: deterministic given the partials.
  // The kernel is pure latency (a few hundred KB, one CTA per view): every load of a thread is
  // issued before the first add (WF_ROWS rows per trip, all in flight together).
  double r[11] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  constexpr int WF_ROWS = 4;
  if (v.loss_partials) {
    const float4* lp = reinterpret_cast<const float4*>(v.loss_partials);
    for (int i0 = threadIdx.x; i0 < v.loss_rows; i0 += WF_ROWS * WF_THREADS) {
      float4 p[WF_ROWS];
#pragma unroll
      for (int u = 0; u < WF_ROWS; ++u) {
        const int i = i0 + u * WF_THREADS;
        p[u] = i < v.loss_rows ? __ldcg(lp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < WF_ROWS; ++u) { r[0] += p[u].x; r[1] += p[u].y; r[2] += p[u].z; r[3] += p[u].w; }
    }
  }
  if (v.tau_partials) {
    const float4* tp = reinterpret_cast<const float4*>(v.tau_partials);
    for (int i0 = threadIdx.x; i0 < v.tau_rows; i0 += WF_ROWS * WF_THREADS) {
      float4 a[WF_ROWS], b[WF_ROWS];
#pragma unroll
      for (int u = 0; u < WF_ROWS; ++u) {
        const int i = i0 + u * WF_THREADS;
        const bool ok = i < v.tau_rows;
        a[u] = ok ? __ldcg(tp + 2 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
        b[u] = ok ? __ldcg(tp + 2 * i + 1) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < WF_ROWS; ++u) {
        r[4] += a[u].x; r[5] += a[u].y; r[6] += a[u].z; r[7] += a[u].w;
        r[8] += b[u].x; r[9] += b[u].y; r[10] += b[u].z;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 11; ++k) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) r[k] += __shfl_xor_sync(0xffffffffu, r[k], d);
    if (lane == 0) s_red[warp][k] = r[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      double x = 0.0;
      for (int w = 0; w < WF_THREADS / 32; ++w) x += s_red[w][k];   // fixed order
      r[k] = x;
    }
    const bool mono = v.mode & 2;
    const double l_rgb = r[0] / (3.0 * (double)v.N), l_dep = r[1] / (double)v.N;
    // monocular: l1.mean() alone (utils/slam_utils.py:72,109); rgbd: alpha blend (:88,128)
    double loss = v.loss_partials ? (mono ? l_rgb : (double)v.alpha * l_rgb + (1.0 - (double)v.alpha) * l_dep) : 0.0;
    loss += r[10] * (double)v.reg_scale;
    if (v.loss) v.loss[0] = (float)loss;
    if (v.dL_dexposure) {
      const float ea = v.exposure_a ? expf(v.exposure_a[0]) : 1.f;
      v.dL_dexposure[0] = (float)r[2] * ea;    // d exp(a)/da = exp(a)
      v.dL_dexposure[1] = (float)r[3];
    }
    if (v.dL_dtau) {
#pragma unroll
      for (int k = 0; k < 6; ++k) v.dL_dtau[k] = (float)r[4 + k];
    }
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == (uint32_t)(t.n - 1);
    if (s_last) {
      __threadfence();
      double total = 0.0;
      uint32_t over = 0;
      for (int k = 0; k < t.n; ++k) {       // view order: deterministic
        if (t.v[k].loss) total += (double)*(volatile float*)t.v[k].loss;
        if (t.v[k].status) over |= *(volatile const uint32_t*)(t.v[k].status + 1);
      }
      if (loss_total) loss_total[0] = (float)total;
      if (guard) guard[0] = (int32_t)over;
      ticket[2] |= over;                     // sticky: "some launch since the caller cleared it overflowed"
      *ticket = 0u;                          // ready for the next launch
    }
  }
}

struct PoseTable {
  B2gsPoseCam c[PS_MAX_CAMS];
  int n;
};

// torch.optim.Adam on one scalar (foreach path: lerp_, mul_/addcmul_, sqrt / bias_correction2_sqrt + eps,
// addcdiv_ with -lr / bias_correction1)
__device__ __forceinline__ void adam_scalar(float& p, float g, float& m, float& v, float beta1, float beta2,
                                            float eps, float lr, double step) {
  const double bc1 = 1.0 - pow((double)beta1, step);
  const float bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, step));
  const float w1 = (float)(1.0 - (double)beta1), w2 = (float)(1.0 - (double)beta2);
  m = __fmaf_rn(w1, __fsub_rn(g, m), m);
  v = __fmaf_rn(__fmul_rn(w2, g), g, __fmul_rn(v, beta2));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
  p = __fmaf_rn((float)(-((double)lr / bc1)), __fdiv_rn(m, denom), p);
}

__global__ void __launch_bounds__(32)
pose_step_kernel(const __grid_constant__ PoseTable t, long long* __restrict__ step_counter,
                 long long* __restrict__ extra_tick, float beta1, float beta2, float eps, int do_tick,
                 const int32_t* __restrict__ guard, uint32_t* __restrict__ ticket) {
  // every load of the kernel is issued up front (this kernel is pure latency): lane k < 8 its
  // parameter / gradient / moments, lanes 8..19 the camera block's R | T for the pose update
  __shared__ float s_rt[12];
  pdl_wait();
  const int k = threadIdx.x;
  const bool have = (int)blockIdx.x < t.n;
  const B2gsPoseCam& c = t.c[have ? blockIdx.x : 0];
  const int32_t g_word = guard ? *guard : 0;
  const long long step_ll = *step_counter;
  if (have && k >= 8 && k < 20 && c.optimize_pose) s_rt[k - 8] = c.cam_state[k - 8];
  if (g_word) return;            // the step is being discarded: no update, no tick
  const double step = (double)(step_ll + 1);
  if (have) {
    if (k < 8) {
      // 0-2 rot (theta <- dL_dtau[3:6]), 3-5 trans (rho <- dL_dtau[0:3]), 6 exposure_a, 7 exposure_b
      const bool pose = k < 6;
      if (pose ? c.optimize_pose != 0 : c.optimize_exposure != 0) {
        float* p = k < 3 ? c.rot_delta + k : k < 6 ? c.trans_delta + (k - 3) : k == 6 ? c.exposure_a : c.exposure_b;
        const float g = k < 3 ? c.dL_dtau[3 + k] : k < 6 ? c.dL_dtau[k - 3] : c.dL_dexposure[k - 6];
        const float lr = k < 3 ? c.lr_rot : k < 6 ? c.lr_trans : k == 6 ? c.lr_exposure_a : c.lr_exposure_b;
        float pv = *p, m = c.moments[k], v = c.moments[8 + k];
        adam_scalar(pv, g, m, v, beta1, beta2, eps, lr, step);
        *p = pv; c.moments[k] = m; c.moments[8 + k] = v;
      }
    }
    __syncwarp();
    if (k == 0 && c.optimize_pose)
      pose_update_device(c.cam_state, c.projmatrix_raw, c.rot_delta, c.trans_delta, c.tau_norm);
  }
  if (threadIdx.x == 0 && do_tick) {
    __threadfence();
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) {   // every block has read the counters
      *step_counter += 1;
      if (extra_tick) *extra_tick += 1;
      *ticket = 0u;
    }
  }
}

}  // namespace b2gs

using namespace b2gs;

extern "C" int b2gs_window_finalize(const B2gsViewFinalize* views, int32_t n_views, float* loss_total,
                                    int32_t* guard, uint32_t* ticket, void* stream) {
  if (n_views < 1 || n_views > WF_MAX_VIEWS || !views || !ticket) {
    set_error("b2gs_window_finalize: 1..%d views and a ticket word are needed", WF_MAX_VIEWS);
    return B2GS_ERR_INVALID;
  }
  FinalizeTable t;
  memset(&t, 0, sizeof(t));
  for (int k = 0; k < n_views; ++k) {
    t.v[k] = views[k];
    if ((views[k].loss_partials && (views[k].loss_rows < 0 || views[k].N <= 0)) ||
        (views[k].tau_partials && views[k].tau_rows < 0) ||
        (((uintptr_t)views[k].loss_partials | (uintptr_t)views[k].tau_partials) & 15)) {
      set_error("b2gs_window_finalize: view %d has bad partial tables", k);
      return B2GS_ERR_INVALID;
    }
  }
  t.n = n_views;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B2GS_PROF("window_finalize", st);
  B2GS_LAUNCH(window_finalize_kernel, n_views, WF_THREADS, st, t, loss_total, guard, ticket);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

extern "C" int b2gs_pose_step(const B2gsPoseCam* cams, int32_t n_cams, int64_t* step_counter,
                              int64_t* extra_tick, float beta1, float beta2, float eps, uint32_t* ticket,
                              void* stream) {
  if (n_cams < 0 || (n_cams > 0 && !cams) || !step_counter || !ticket) {
    set_error("b2gs_pose_step: bad arguments");
    return B2GS_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int done = 0;
  do {
    PoseTable t;
    memset(&t, 0, sizeof(t));
    int k = 0;
    for (; k < PS_MAX_CAMS && done + k < n_cams; ++k) {
      const B2gsPoseCam& c = cams[done + k];
      if (!c.moments || (c.optimize_pose && (!c.cam_state || !c.projmatrix_raw || !c.rot_delta ||
                                              !c.trans_delta || !c.dL_dtau)) ||
          (c.optimize_exposure && (!c.exposure_a || !c.exposure_b || !c.dL_dexposure))) {
        set_error("b2gs_pose_step: camera %d has NULL pointers", done + k);
        return B2GS_ERR_INVALID;
      }
      t.c[k] = c;
    }
    t.n = k;
    done += k;
    const int last = done >= n_cams;
    B2GS_PROF("pose_step", st);
    B2GS_LAUNCH(pose_step_kernel, k > 0 ? k : 1, 32, st, t, reinterpret_cast<long long*>(step_counter),
                reinterpret_cast<long long*>(extra_tick), beta1, beta2, eps, last, update_guard(), ticket);
    B2GS_LAUNCH_CHECK(st, 0);
  } while (done < n_cams);
  return B2GS_OK;
}
