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
            float eps, float beta1, const long long* __restrict__ step_dev,
            const int32_t* __restrict__ guard, int flags) {
  pdl_wait();
  if (guard && *guard) return;   // the step is being discarded (b2gs_set_update_guard)
  const bool zero = flags & B2GS_ADAM_ZERO_GRAD;
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
  float* __restrict__ g = const_cast<float*>(t.g[s]);
  float* __restrict__ m = t.m[s];
  float* __restrict__ v = t.v[s];
  float neg_step = t.neg_step_size[s];
  if (step_dev != nullptr) {
    // CUDA-graph capturable form: the step count lives on the device (incremented by
    // adam_tick_kernel just before); same double-precision scalar prep as the host path.
    // Once per block, not once per thread: the two double-precision pow() are ~500 instructions,
    // more than a thread's whole share of the update (4 float4 trips).
    __shared__ float s_bc[2];
    if (threadIdx.x == 0) {
      const double step = (double)(*step_dev + ((flags & B2GS_ADAM_NO_TICK) ? 1 : 0));
      const double bc1 = 1.0 - pow((double)beta1, step);
      const float lr = t.lr_dev[s] ? t.lr_dev[s][0] : t.lr[s];
      s_bc[0] = (float)sqrt(1.0 - pow((double)beta2, step));
      s_bc[1] = (float)(-((double)lr / bc1));
    }
    __syncthreads();
    bc2_sqrt = s_bc[0];
    neg_step = s_bc[1];
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
      if (zero) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (long long i = end4 + threadIdx.x; i < end; i += blockDim.x) {
      float pp = p[i], mm = m[i], vv = v[i];
      adam_one(pp, g[i], mm, vv, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      p[i] = pp; m[i] = mm; v[i] = vv;
      if (zero) g[i] = 0.f;
    }
  } else {
    for (long long i = begin + threadIdx.x; i < end; i += blockDim.x) {
      float pp = p[i], mm = m[i], vv = v[i];
      adam_one(pp, g[i], mm, vv, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      p[i] = pp; m[i] = mm; v[i] = vv;
      if (zero) g[i] = 0.f;
    }
  }
}

__global__ void adam_tick_kernel(long long* step, const int32_t* __restrict__ guard) {
  pdl_wait();
  if (guard && *guard) return;
  *step += 1;
}

// ---- sharded step over NVLink peer memory --------------------------------------------------
// Data-parallel ranks hold replicas of the parameter arena and their own gradient arena, both
// in memory every peer has mapped.  One kernel replaces  all_reduce(grad) -> Adam -> zero_grad:
// the rank that owns a 4 KB chunk (chunk index mod world) loads that chunk of EVERY rank's
// gradients over NVLink (summed in rank order: deterministic), clears them, updates its private
// moments, and stores the new parameters into every replica.  Per rank: 1/world of the Adam
// traffic plus (world-1)/world x 12 B per parameter over the links, instead of a 2(world-1)/world
// x 4 B ring all-reduce followed by a full 28 B/parameter Adam.  Replicas stay bit-identical
// because each element is computed exactly once.
constexpr int PEER_MAX = 16;

struct PeerTable {
  float* param[PEER_MAX];
  float* grad[PEER_MAX];
  uint32_t* signal[PEER_MAX];
  float* param_mc;   // multicast views of the arenas (NVLS), or nullptr
  float* grad_mc;
  int world, rank;
};

// NVLink-switch multicast accesses (sm_90+): one load returns the SUM over every rank's copy
// (the switch reduces in flight), one store lands in every rank's copy.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc_addr) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc_addr) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(float* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// All-to-all flag barrier between the ranks' streams (one thread per peer).  Slot [src] on the
// destination's pad: 0 -> 1 by the arriving rank, 1 -> 0 by the owner; CAS on both sides makes
// back-to-back barriers safe.  Bounded spin (about 4 s): a dead peer must not hang the GPU.
// With a guard word: every rank also posts its guard value on all pads (slots 16 + src) before
// arriving, and leaves the OR over all ranks in its own guard word — the ranks then agree on
// whether this step's update is discarded.
// With status words (b2gs_peer_barrier_status): this rank's guard value is not read from *guard but
// formed here as the OR of status[2 i + 1], the overflow flags the forward passes of the rank's
// n_status views left (B2gsForwardOut.status) — the caller need not wait for b2gs_window_finalize.
__global__ void peer_barrier_kernel(const __grid_constant__ PeerTable pt, uint32_t* __restrict__ fail,
                                    int32_t* __restrict__ guard, long long* __restrict__ tick,
                                    const int32_t* __restrict__ tick_guard,
                                    const uint32_t* __restrict__ status, int n_status) {
  pdl_wait();
  const int t = threadIdx.x;
  const bool peer = t < pt.world && t != pt.rank;
  if (guard && t < pt.world) {
    uint32_t mine = 0;
    if (status) {
      for (int i = 0; i < n_status; ++i) mine |= status[2 * i + 1];
    } else {
      mine = (uint32_t)*guard;
    }
    volatile uint32_t* slot = pt.signal[t] + PEER_MAX + pt.rank;
    *slot = (uint32_t)(mine != 0);
  }
  if (peer) {
    __threadfence_system();
    const long long t0 = clock64();
    const long long limit = 8000000000LL;
    uint32_t* dst = pt.signal[t] + pt.rank;
    bool ok = true;
    while (ok && atomicCAS_system(dst, 0u, 1u) != 0u)
      if (clock64() - t0 > limit) { *fail = 1u; ok = false; }
    uint32_t* src = pt.signal[pt.rank] + t;
    while (ok && atomicCAS_system(src, 1u, 0u) != 1u)
      if (clock64() - t0 > limit) { *fail = 1u; ok = false; }
    __threadfence_system();
  }
  __syncthreads();
  if (guard && t == 0) {
    uint32_t any = 0;
    for (int r = 0; r < pt.world; ++r) any |= *(volatile uint32_t*)(pt.signal[pt.rank] + PEER_MAX + r);
    // a timed-out barrier means some peer's gradients may not be final: discard the update (the
    // host reads fail_flag with the overflow flags and raises, mapping.MappingWindow._settle)
    any |= *(volatile uint32_t*)fail;
    *guard = (int32_t)any;
  }
  // B2GS_ADAM_TICK_LATE: the step counter advances here, after every block of the update has read it
  if (tick && t == 0 && !(tick_guard && *tick_guard) && !*(volatile uint32_t*)fail) *tick += 1;
}

constexpr int PEER_CHUNK = 1024;   // floats per ownership chunk = one float4 per thread of a block

__global__ void __launch_bounds__(256)
adam_peer_kernel(const __grid_constant__ AdamTable t, const __grid_constant__ PeerTable pt, float w1,
                 float beta2, float w2, float eps, float beta1, const long long* __restrict__ step_dev,
                 const int32_t* __restrict__ guard, const uint32_t* __restrict__ fail, int flags) {
  pdl_wait();
  if (guard && *guard) return;   // some rank overflowed: every rank discards this step
  if (*fail) return;             // a barrier timed out: peers' arenas may not be final
  const long long chunk = (long long)blockIdx.x * pt.world + pt.rank;   // chunks this rank owns
  if (chunk >= t.first_chunk[t.num]) return;
  int s = 0;
#pragma unroll 1
  while (s + 1 < t.num && chunk >= t.first_chunk[s + 1]) ++s;
  const long long begin = (chunk - t.first_chunk[s]) * PEER_CHUNK;
  const long long n = t.n[s];
  long long end = begin + PEER_CHUNK;
  if (end > n) end = n;
  float* __restrict__ m = t.m[s];
  float* __restrict__ v = t.v[s];
  const long long poff = t.p[s] - pt.param[pt.rank], goff = t.g[s] - pt.grad[pt.rank];
  // bias corrections: double-precision pow once per block, not once per thread
  __shared__ float s_bc[2];
  if (threadIdx.x == 0) {
    const double step = (double)(*step_dev + ((flags & B2GS_ADAM_NO_TICK) ? 1 : 0));
    const double bc1 = 1.0 - pow((double)beta1, step);
    const float lr = t.lr_dev[s] ? t.lr_dev[s][0] : t.lr[s];
    s_bc[0] = (float)sqrt(1.0 - pow((double)beta2, step));
    s_bc[1] = (float)(-((double)lr / bc1));
  }
  __syncthreads();
  const float bc2_sqrt = s_bc[0], neg_step = s_bc[1];
  const int W = pt.world;

  const bool vec = (((poff | goff) & 3) == 0) && ((((uintptr_t)m | (uintptr_t)v) & 15) == 0);
  const long long i = begin + 4LL * threadIdx.x;
  if (vec && i + 4 <= end && pt.param_mc != nullptr) {
    // NVLS: the switch sums the ranks' gradient copies and replicates the stores
    float4 gg = multimem_ld_reduce_add(pt.grad_mc + goff + i);
    float4 pp = *reinterpret_cast<const float4*>(pt.param[pt.rank] + poff + i);
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    adam_one(pp.x, gg.x, mm.x, vv.x, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.y, gg.y, mm.y, vv.y, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.z, gg.z, mm.z, vv.z, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.w, gg.w, mm.w, vv.w, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
    multimem_st(pt.param_mc + poff + i, pp);
  } else if (vec && i + 4 <= end) {
    // all peers' loads in flight together (one NVLink round trip), then the sum in rank order
    float4 x[PEER_MAX];
#pragma unroll
    for (int r = 0; r < PEER_MAX; ++r)
      if (r < W) x[r] = __ldcg(reinterpret_cast<const float4*>(pt.grad[r] + goff + i));
    float4 pp = *reinterpret_cast<const float4*>(pt.param[pt.rank] + poff + i);
    float4 mm = *reinterpret_cast<float4*>(m + i);
    float4 vv = *reinterpret_cast<float4*>(v + i);
    float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < PEER_MAX; ++r)
      if (r < W) {
        gg.x = __fadd_rn(gg.x, x[r].x); gg.y = __fadd_rn(gg.y, x[r].y);
        gg.z = __fadd_rn(gg.z, x[r].z); gg.w = __fadd_rn(gg.w, x[r].w);
      }
    adam_one(pp.x, gg.x, mm.x, vv.x, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.y, gg.y, mm.y, vv.y, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.z, gg.z, mm.z, vv.z, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    adam_one(pp.w, gg.w, mm.w, vv.w, w1, beta2, w2, bc2_sqrt, eps, neg_step);
    *reinterpret_cast<float4*>(m + i) = mm;
    *reinterpret_cast<float4*>(v + i) = vv;
#pragma unroll
    for (int r = 0; r < PEER_MAX; ++r)
      if (r < W) __stcg(reinterpret_cast<float4*>(pt.param[r] + poff + i), pp);
  } else {
    for (long long j = i; j < end && j < i + 4; ++j) {
      float g = 0.f;
      for (int r = 0; r < W; ++r) g = __fadd_rn(g, __ldcg(pt.grad[r] + goff + j));
      float pp = pt.param[pt.rank][poff + j], mm = m[j], vv = v[j];
      adam_one(pp, g, mm, vv, w1, beta2, w2, bc2_sqrt, eps, neg_step);
      m[j] = mm; v[j] = vv;
      for (int r = 0; r < W; ++r) __stcg(pt.param[r] + poff + j, pp);
    }
  }
}

}  // namespace b2gs

using namespace b2gs;

static int adam_impl(const B2gsAdamSegment* segments, int32_t num_segments, int64_t step,
                     long long* step_dev, float beta1, float beta2, float eps, void* stream, int flags = 0) {
  if (num_segments < 0 || (num_segments > 0 && !segments)) {
    set_error("b2gs_adam_step: bad segments");
    return B2GS_ERR_INVALID;
  }
  if (step < 1) { set_error("b2gs_adam_step: step must be >= 1"); return B2GS_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (step_dev && !(flags & B2GS_ADAM_NO_TICK)) {
    B2GS_LAUNCH(adam_tick_kernel, 1, 1, st, step_dev, update_guard());
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
      // the camera / exposure optimizer is a handful of floats: latency, not bandwidth — its own name
      long long total = 0;
      for (int q = 0; q < k; ++q) total += t.n[q];
      B2GS_PROF(total < 65536 ? "adam_small" : "adam", st);
      B2GS_LAUNCH(adam_kernel, (unsigned)chunks, 256, st, t, w1, beta2, w2, bc2_sqrt, eps, beta1,
                  (const long long*)step_dev, update_guard(), flags);
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

extern "C" int b2gs_adam_step_fused(const B2gsAdamSegment* segments, int32_t num_segments,
                                    int64_t* step_counter, float beta1, float beta2, float eps,
                                    int32_t flags, void* stream) {
  if (!step_counter) { set_error("b2gs_adam_step_fused: NULL step counter"); return B2GS_ERR_INVALID; }
  return adam_impl(segments, num_segments, 1, reinterpret_cast<long long*>(step_counter), beta1, beta2,
                   eps, stream, flags);
}

extern "C" int b2gs_adam_step_peer(const B2gsAdamSegment* segments, int32_t num_segments,
                                   int64_t* step_counter, float beta1, float beta2, float eps,
                                   const B2gsPeerArena* peers, void* stream) {
  return b2gs_adam_step_peer_fused(segments, num_segments, step_counter, beta1, beta2, eps, peers, 0, stream);
}

extern "C" int b2gs_adam_step_peer_fused(const B2gsAdamSegment* segments, int32_t num_segments,
                                         int64_t* step_counter, float beta1, float beta2, float eps,
                                         const B2gsPeerArena* peers, int32_t flags, void* stream) {
  if (!step_counter || !peers) { set_error("b2gs_adam_step_peer: NULL step counter / peers"); return B2GS_ERR_INVALID; }
  if (peers->world < 1 || peers->world > PEER_MAX || peers->rank < 0 || peers->rank >= peers->world ||
      !peers->param_base || !peers->grad_base || !peers->signal || !peers->fail_flag) {
    set_error("b2gs_adam_step_peer: bad peer table (world %d, rank %d, at most %d ranks)", peers->world,
              peers->rank, PEER_MAX);
    return B2GS_ERR_INVALID;
  }
  if (num_segments < 0 || num_segments > ADAM_MAX_SEG || (num_segments > 0 && !segments)) {
    set_error("b2gs_adam_step_peer: 0..%d segments", ADAM_MAX_SEG);
    return B2GS_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int32_t* guard = const_cast<int32_t*>(update_guard());
  PeerTable pt;
  memset(&pt, 0, sizeof(pt));
  pt.world = peers->world; pt.rank = peers->rank;
  if (peers->param_mc && peers->grad_mc && pt.world > 1) { pt.param_mc = peers->param_mc; pt.grad_mc = peers->grad_mc; }
  for (int r = 0; r < pt.world; ++r) {
    pt.param[r] = peers->param_base[r]; pt.grad[r] = peers->grad_base[r]; pt.signal[r] = peers->signal[r];
    if (!pt.param[r] || !pt.grad[r] || !pt.signal[r]) { set_error("b2gs_adam_step_peer: NULL peer pointer"); return B2GS_ERR_INVALID; }
  }
  AdamTable t;
  memset(&t, 0, sizeof(t));
  long long chunks = 0;
  for (int k = 0; k < num_segments; ++k) {
    const B2gsAdamSegment& sgm = segments[k];
    if (sgm.n < 0 || (sgm.n > 0 && (!sgm.param || !sgm.grad || !sgm.exp_avg || !sgm.exp_avg_sq))) {
      set_error("b2gs_adam_step_peer: segment %d has NULL pointers", k);
      return B2GS_ERR_INVALID;
    }
    t.p[k] = sgm.param; t.g[k] = sgm.grad; t.m[k] = sgm.exp_avg; t.v[k] = sgm.exp_avg_sq;
    t.n[k] = sgm.n;
    t.first_chunk[k] = chunks;
    chunks += (sgm.n + PEER_CHUNK - 1) / PEER_CHUNK;
    t.lr[k] = sgm.lr;
    t.lr_dev[k] = sgm.lr_dev;
  }
  t.first_chunk[num_segments] = chunks;
  t.num = num_segments;
  // every rank's gradients are final ...  (the barrier also leaves the OR of all ranks' guard words
  // in *guard, and sets it when a peer timed out: the update below is then skipped on every rank)
  if (!(flags & B2GS_ADAM_NO_FIRST_BARRIER)) {
    B2GS_LAUNCH(peer_barrier_kernel, 1, 32, st, pt, peers->fail_flag, guard, (long long*)nullptr,
                (const int32_t*)nullptr, (const uint32_t*)nullptr, 0);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  // ... and only now may the step counter advance: with the rank-local guard value a rank whose own
  // forward did not overflow would tick while an overflowed rank does not, and the replicas' bias
  // corrections would diverge for good
  if (!(flags & B2GS_ADAM_NO_TICK)) {
    B2GS_LAUNCH(adam_tick_kernel, 1, 1, st, reinterpret_cast<long long*>(step_counter), (const int32_t*)guard);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  if (chunks > 0) {
    const long long blocks = (chunks + pt.world - 1) / pt.world;
    if (blocks > 0x7fffffffLL) { set_error("b2gs_adam_step_peer: too many elements"); return B2GS_ERR_INVALID; }
    B2GS_PROF("adam_peer", st);
    B2GS_LAUNCH(adam_peer_kernel, (unsigned)blocks, 256, st, t, pt, (float)(1.0 - (double)beta1), beta2,
                (float)(1.0 - (double)beta2), eps, beta1, reinterpret_cast<const long long*>(step_counter),
                (const int32_t*)guard, (const uint32_t*)peers->fail_flag, flags);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  // ... and every replica has been written, every gradient arena read, by all of its chunk owners
  B2GS_LAUNCH(peer_barrier_kernel, 1, 32, st, pt, peers->fail_flag, (int32_t*)nullptr,
              (flags & B2GS_ADAM_TICK_LATE) ? reinterpret_cast<long long*>(step_counter) : (long long*)nullptr,
              (const int32_t*)guard, (const uint32_t*)nullptr, 0);
  B2GS_LAUNCH_CHECK(st, 0);
  if (flags & B2GS_ADAM_NO_ZERO) return B2GS_OK;
  // zero_grad: every rank clears ITS OWN gradient arena (local HBM writes) now that all owners have
  // read it — clearing it through the owners' remote stores would put a second arena's worth of
  // zeros on every rank's NVLink ingress next to the parameter broadcast
  {
    const float* lo = nullptr; const float* hi = nullptr;
    for (int k = 0; k < num_segments; ++k) {
      if (segments[k].n <= 0) continue;
      const float* a = segments[k].grad; const float* b = a + segments[k].n;
      if (!lo || a < lo) lo = a;
      if (!hi || b > hi) hi = b;
    }
    const float* base = pt.grad[pt.rank];
    // the segments must form one arena (pads of < 4 floats between them): then it is one memset
    long long covered = 0;
    for (int k = 0; k < num_segments; ++k) covered += (segments[k].n + 3) / 4 * 4;
    if (lo && (hi - lo) <= covered && lo >= base) {
      B2GS_CUDA(cudaMemsetAsync(const_cast<float*>(lo), 0, (size_t)(hi - lo) * sizeof(float), st));
    } else {
      for (int k = 0; k < num_segments; ++k)
        if (segments[k].n > 0)
          B2GS_CUDA(cudaMemsetAsync(const_cast<float*>(segments[k].grad), 0, (size_t)segments[k].n * sizeof(float), st));
    }
  }
  return B2GS_OK;
}

static int peer_table(const B2gsPeerArena* peers, b2gs::PeerTable* pt) {
  if (!peers || peers->world < 1 || peers->world > b2gs::PEER_MAX || peers->rank < 0 || peers->rank >= peers->world ||
      !peers->signal || !peers->fail_flag) {
    b2gs::set_error("bad peer table");
    return B2GS_ERR_INVALID;
  }
  memset(pt, 0, sizeof(*pt));
  pt->world = peers->world; pt->rank = peers->rank;
  for (int r = 0; r < pt->world; ++r) {
    pt->signal[r] = peers->signal[r];
    if (!pt->signal[r]) { b2gs::set_error("NULL signal pad"); return B2GS_ERR_INVALID; }
  }
  return B2GS_OK;
}

extern "C" int b2gs_peer_barrier(const B2gsPeerArena* peers, void* stream) {
  b2gs::PeerTable pt;
  int rc = peer_table(peers, &pt);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B2GS_LAUNCH(b2gs::peer_barrier_kernel, 1, 32, st, pt, peers->fail_flag, const_cast<int32_t*>(b2gs::update_guard()),
              (long long*)nullptr, (const int32_t*)nullptr, (const uint32_t*)nullptr, 0);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

extern "C" int b2gs_peer_barrier_status(const B2gsPeerArena* peers, const uint32_t* status, int32_t n_views,
                                        void* stream) {
  if (!status || n_views < 1) { b2gs::set_error("b2gs_peer_barrier_status: status words of >= 1 view needed"); return B2GS_ERR_INVALID; }
  if (!b2gs::update_guard()) { b2gs::set_error("b2gs_peer_barrier_status: no update guard set (b2gs_set_update_guard)"); return B2GS_ERR_INVALID; }
  b2gs::PeerTable pt;
  int rc = peer_table(peers, &pt);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B2GS_LAUNCH(b2gs::peer_barrier_kernel, 1, 32, st, pt, peers->fail_flag, const_cast<int32_t*>(b2gs::update_guard()),
              (long long*)nullptr, (const int32_t*)nullptr, status, (int)n_views);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}

extern "C" int b2gs_memset_async(void* device_ptr, int value, size_t bytes, void* stream) {
  if (!device_ptr && bytes) { b2gs::set_error("b2gs_memset_async: NULL pointer"); return B2GS_ERR_INVALID; }
  B2GS_CUDA(cudaMemsetAsync(device_ptr, value, bytes, reinterpret_cast<cudaStream_t>(stream)));
  return B2GS_OK;
}
