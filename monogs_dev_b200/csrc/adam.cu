// adam.cu — fused Adam step over all parameter groups of GaussianModel in ONE launch.
// Replaces torch.optim.Adam(l, lr=0.0, eps=1e-15).step() over 6 groups
// (GS/scene/gaussian_model.py:245-283, called at utils/slam_backend.py:307) and the
// per-keyframe pose/exposure optimizer (utils/slam_backend.py:310).
//
// HBM-bound: 28 B per parameter (read p,g,m,v; write p,m,v).  Rounding follows
// torch's op sequence (lerp_, mul_, addcmul_, sqrt, div, add, addcdiv_) with
// explicit _rn intrinsics so that nvcc cannot fuse across what are separate
// kernels in torch.
#include "common.cuh"
#include <math.h>
#include <string.h>

namespace b2gs {

constexpr int ADAM_MAX_SEG = 16;
constexpr int ADAM_CHUNK = 4096;  // floats per block

struct AdamTable {
  float* p[ADAM_MAX_SEG];
  const float* g[ADAM_MAX_SEG];
  float* m[ADAM_MAX_SEG];
  float* v[ADAM_MAX_SEG];
  long long n[ADAM_MAX_SEG];
  long long first_chunk[ADAM_MAX_SEG + 1];
  float neg_step_size[ADAM_MAX_SEG];
  float lr[ADAM_MAX_SEG];
  const float* lr_dev[ADAM_MAX_SEG];
  int num;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float w1, float beta2,
                                         float w2, float bc2_sqrt, float eps, float neg_step) {
  m = __fmaf_rn(w1, __fsub_rn(g, m), m);                       // exp_avg.lerp_(grad, 1-beta1)
  v = __fmaf_rn(__fmul_rn(w2, g), g, __fmul_rn(v, beta2));     // mul_(beta2).addcmul_(g, g, 1-beta2)
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), eps);
  p = __fmaf_rn(neg_step, __fdiv_rn(m, denom), p);             // addcdiv_(m, denom, -step_size)
}

__global__ void __launch_bounds__(256)
adam_kernel(const __grid_constant__ AdamTable t, float w1, float beta2, float w2, float bc2_sqrt,
            float eps, float beta1, const long long* __restrict__ step_dev) {
  // which segment does this block's chunk belong to?
  int s = 0;
  const long long chunk = blockIdx.x;
#pragma unroll 1
  while (s + 1 < t.num && chunk >= t.first_chunk[s + 1]) ++s;
  const long long begin = (chunk - t.first_chunk[s]) * ADAM_CHUNK;
  const long long n = t.n[s];
  long long end = begin + ADAM_CHUNK;
  if (end > n) end = n;
  float* __restrict__ p = t.p[s];
  const float* __restrict__ g = t.g[s];
  float* __restrict__ m = t.m[s];
  float* __restrict__ v = t.v[s];
  float neg_step = t.neg_step_size[s];
  if (step_dev != nullptr) {
    // CUDA-graph capturable form: the step count lives on the device (incremented by
    // adam_tick_kernel just before); same double-precision scalar prep as the host path
    const double step = (double)*step_dev;
    const double bc1 = 1.0 - pow((double)beta1, step);
    bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, step));
    const float lr = t.lr_dev[s] ? t.lr_dev[s][0] : t.lr[s];
    neg_step = (float)(-((double)lr / bc1));
  }

  const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
  if (vec) {
    const long long end4 = begin + ((end - begin) & ~3LL);
    for (long long i = begin + 4LL * threadIdx.x; i < end4; i += 4LL * blockDim.x) {
      float4 pp = *reinterpret_cast<float4*>(p + i);
      const float4 gg = *reinterpret_cast<const float4*>(g + i);
      float4 mm = *reinterpret_cast<float4*>(m + i);
      float4 vv = *reinterpret_cast<float4*>(v + i);
      adam_one(pp.x, gg.x, mm.x, vv.x, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      adam_one(pp.y, gg.y, mm.y, vv.y, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      adam_one(pp.z, gg.z, mm.z, vv.z, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      adam_one(pp.w, gg.w, mm.w, vv.w, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      *reinterpret_cast<float4*>(p + i) = pp;
      *reinterpret_cast<float4*>(m + i) = mm;
      *reinterpret_cast<float4*>(v + i) = vv;
    }
    for (long long i = end4 + threadIdx.x; i < end; i += blockDim.x) {
      float pp = p[i], mm = m[i], vv = v[i];
      adam_one(pp, g[i], mm, vv, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      p[i] = pp; m[i] = mm; v[i] = vv;
    }
  } else {
    for (long long i = begin + threadIdx.x; i < end; i += blockDim.x) {
      float pp = p[i], mm = m[i], vv = v[i];
      adam_one(pp, g[i], mm, vv, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      p[i] = pp; m[i] = mm; v[i] = vv;
    }
  }
}

__global__ void adam_tick_kernel(long long* step) { *step += 1; }

}  // namespace b2gs

using namespace b2gs;

static int adam_impl(const B2gsAdamSegment* segments, int32_t num_segments, int64_t step,
                     long long* step_dev, float beta1, float beta2, float eps, void* stream) {
  if (num_segments < 0 || (num_segments > 0 && !segments)) {
    set_error("b2gs_adam_step: bad segments");
    return B2GS_ERR_INVALID;
  }
  if (step < 1) { set_error("b2gs_adam_step: step must be >= 1"); return B2GS_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (step_dev) {
    adam_tick_kernel<<<1, 1, 0, st>>>(step_dev);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  // scalar prep in double, as torch does in Python (torch/optim/adam.py)
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float bc2_sqrt = (float)sqrt(bc2);
  const float w1 = (float)(1.0 - (double)beta1);
  const float w2 = (float)(1.0 - (double)beta2);
  int done = 0;
  while (done < num_segments) {
    AdamTable t;
    memset(&t, 0, sizeof(t));
    long long chunks = 0;
    int k = 0;
    for (; k < ADAM_MAX_SEG && done + k < num_segments; ++k) {
      const B2gsAdamSegment& sgm = segments[done + k];
      if (sgm.n < 0 || (sgm.n > 0 && (!sgm.param || !sgm.grad || !sgm.exp_avg || !sgm.exp_avg_sq))) {
        set_error("b2gs_adam_step: segment %d has NULL pointers", done + k);
        return B2GS_ERR_INVALID;
      }
      t.p[k] = sgm.param; t.g[k] = sgm.grad; t.m[k] = sgm.exp_avg; t.v[k] = sgm.exp_avg_sq;
      t.n[k] = sgm.n;
      t.first_chunk[k] = chunks;
      chunks += (sgm.n + ADAM_CHUNK - 1) / ADAM_CHUNK;
      t.neg_step_size[k] = (float)(-((double)sgm.lr / bc1));
      t.lr[k] = sgm.lr;
      t.lr_dev[k] = step_dev ? sgm.lr_dev : nullptr;
    }
    t.first_chunk[k] = chunks;
    t.num = k;
    if (chunks > 0) {
      if (chunks > 0x7fffffffLL) { set_error("b2gs_adam_step: too many elements"); return B2GS_ERR_INVALID; }
      B2GS_PROF("adam", st);
      adam_kernel<<<(unsigned)chunks, 256, 0, st>>>(t, w1, beta2, w2, bc2_sqrt, eps, beta1, step_dev);
      B2GS_LAUNCH_CHECK(st, 0);
    }
    done += k;
  }
  return B2GS_OK;
}

extern "C" int b2gs_adam_step(const B2gsAdamSegment* segments, int32_t num_segments, int64_t step,
                              float beta1, float beta2, float eps, void* stream) {
  return adam_impl(segments, num_segments, step, nullptr, beta1, beta2, eps, stream);
}

extern "C" int b2gs_adam_step_dev(const B2gsAdamSegment* segments, int32_t num_segments,
                                  int64_t* step_counter, float beta1, float beta2, float eps,
                                  void* stream) {
  if (!step_counter) { set_error("b2gs_adam_step_dev: NULL step counter"); return B2GS_ERR_INVALID; }
  return adam_impl(segments, num_segments, 1, reinterpret_cast<long long*>(step_counter), beta1,
                   beta2, eps, stream);
}
