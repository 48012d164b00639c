// densify.cu — densify_and_prune on the packed arenas as three kernels.
// Replaces the tensor-op sequence of GaussianModel.densify_and_prune / densify_and_clone /
// densify_and_split / prune_points / cat_tensors_to_optimizer / _prune_optimizer
// (GS/scene/gaussian_model.py:468-555, 593-689): there, the whole model (6 parameter tensors + 12
// Adam moment tensors) is re-materialised four times (clone cat, split cat, split prune, final
// prune) with boolean masks bounced through the host.  Here:
//   1. densify_plan_kernel    one pass over the Gaussians: clone / split / prune decisions of the
//                             WHOLE sequence per Gaussian (a clone or a child is pruned by the final
//                             mask iff its own opacity / size says so, and both are functions of the
//                             parent's row), a 3-bit code per Gaussian and per-block counts
//   2. densify_scan_kernel    one CTA: exclusive scan of the per-block counts -> block offsets, totals
//   3. densify_apply_kernel   each Gaussian writes its surviving rows — original, clone, N children —
//                             of every group of the three arenas (parameters, exp_avg, exp_avg_sq)
//                             straight to their final position in the NEW arenas, in the reference's
//                             final order [kept originals | kept clones | kept children, copy-major];
//                             children get their sampled position and shrunk scale here too.
// One host read (the totals, to size the new arenas) sits between 2 and 3 — the reference has the
// same need at every one of its four re-materialisations.
#include "common.cuh"
#include <string.h>

namespace b2gs {

constexpr int DN_THREADS = 256;
constexpr int DN_MAX_GROUPS = 8;

enum : uint32_t { DN_ORIG = 1u, DN_CLONE = 2u, DN_CHILD = 4u, DN_IS_CLONE = 8u, DN_IS_SPLIT = 16u };

struct DensifyWS {
  uint8_t* code;        // [P]
  uint32_t* block_cnt;  // [5][nb]: kept originals, kept clones, kept split parents (children / N), clones, splits
  uint32_t* block_off;  // [5][nb] exclusive
  unsigned long long* totals;  // [8]
  int nb;
};

static DensifyWS carve_densify(void* base, int32_t P, size_t* total) {
  Carver c(base);
  DensifyWS w;
  const size_t n = (size_t)(P > 0 ? P : 1);
  w.nb = (int)((n + DN_THREADS - 1) / DN_THREADS);
  w.code = c.take<uint8_t>(n);
  w.block_cnt = c.take<uint32_t>((size_t)5 * w.nb);
  w.block_off = c.take<uint32_t>((size_t)5 * w.nb);
  w.totals = c.take<unsigned long long>(8);
  if (total) *total = c.total();
  return w;
}

__device__ __forceinline__ float sigmoidf_torch(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(DN_THREADS)
densify_plan_kernel(int P, const float* __restrict__ scaling, const float* __restrict__ opacity,
                    const float* __restrict__ accum, const float* __restrict__ denom, float max_grad,
                    float min_opacity, float dense_thr, float ws_thr, int mss_on, float k,
                    uint8_t* __restrict__ code, uint32_t* __restrict__ block_cnt, int nb) {
  __shared__ uint32_t s_cnt[DN_THREADS / 32][5];
  const int i = blockIdx.x * DN_THREADS + threadIdx.x;
  uint32_t c = 0;
  if (i < P) {
    // grads = xyz_gradient_accum / denom; grads[isnan] = 0  (gaussian_model.py:666-667)
    float g = accum[i] / denom[i];
    if (g != g) g = 0.f;
    const float s0 = expf(scaling[3 * i]), s1 = expf(scaling[3 * i + 1]), s2 = expf(scaling[3 * i + 2]);
    const float smax = fmaxf(s0, fmaxf(s1, s2));
    const bool sel = g >= max_grad;                       // torch.norm of a 1-vector = |g| = g
    const bool big = smax > dense_thr;                    // percent_dense * scene_extent
    const bool is_clone = sel && !big, is_split = sel && big;
    // final prune (gaussian_model.py:676-689); max_radii2D was zeroed by the postfix, so big_vs is False
    const bool low = sigmoidf_torch(opacity[i]) < min_opacity;
    const bool wide = mss_on && smax > ws_thr;            // 0.1 * extent
    // children: scaling' = log(scaling / (0.8 N)); the prune test exponentiates it again
    const float cmax = fmaxf(expf(logf(s0 / k)), fmaxf(expf(logf(s1 / k)), expf(logf(s2 / k))));
    const bool cwide = mss_on && cmax > ws_thr;
    if (!is_split && !low && !wide) c |= DN_ORIG;
    if (is_clone && !low && !wide) c |= DN_CLONE;
    if (is_split && !low && !cwide) c |= DN_CHILD;
    if (is_clone) c |= DN_IS_CLONE;
    if (is_split) c |= DN_IS_SPLIT;
    code[i] = (uint8_t)c;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int b = 0; b < 5; ++b) {
    const uint32_t n = __popc(__ballot_sync(0xffffffffu, (c >> b) & 1u));
    if (lane == 0) s_cnt[warp][b] = n;
  }
  __syncthreads();
  if (threadIdx.x < 5) {
    uint32_t n = 0;
    for (int w = 0; w < DN_THREADS / 32; ++w) n += s_cnt[w][threadIdx.x];
    block_cnt[(size_t)threadIdx.x * nb + blockIdx.x] = n;
  }
}

// one CTA of 1024 threads, one warp-row per quantity would waste lanes: thread t scans a
// contiguous slice of each of the five rows, slices are combined by a block scan
__global__ void __launch_bounds__(1024)
densify_scan_kernel(int nb, const uint32_t* __restrict__ cnt, uint32_t* __restrict__ off,
                    unsigned long long* __restrict__ totals) {
  __shared__ unsigned long long s_w[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int per = (nb + 1023) / 1024;
  const int lo = min(nb, tid * per), hi = min(nb, lo + per);
  for (int q = 0; q < 5; ++q) {
    unsigned long long mine = 0;
    for (int b = lo; b < hi; ++b) mine += cnt[(size_t)q * nb + b];
    unsigned long long incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long n = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += n;
    }
    __syncthreads();
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const unsigned long long w = s_w[lane];
      unsigned long long wi = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long n = __shfl_up_sync(0xffffffffu, wi, d);
        if (lane >= d) wi += n;
      }
      s_w[lane] = wi - w;
      if (lane == 31) totals[q] = wi;
    }
    __syncthreads();
    unsigned long long run = s_w[warp] + incl - mine;
    for (int b = lo; b < hi; ++b) {
      off[(size_t)q * nb + b] = (uint32_t)run;
      run += cnt[(size_t)q * nb + b];
    }
  }
}

struct DensifyTable {
  const float* p[DN_MAX_GROUPS]; const float* m[DN_MAX_GROUPS]; const float* v[DN_MAX_GROUPS];
  float* np[DN_MAX_GROUPS]; float* nm[DN_MAX_GROUPS]; float* nv[DN_MAX_GROUPS];
  int w[DN_MAX_GROUPS];
  int n, xyz, scaling, rotation;
  const int32_t* aux_in[2]; int32_t* aux_out[2];
};

__device__ __forceinline__ void copy_row(const DensifyTable& t, int grp, long long src, long long dst,
                                         bool fresh) {
  const int w = t.w[grp];
  const float* p = t.p[grp] + src * w;
  const float* m = t.m[grp] + src * w;
  const float* v = t.v[grp] + src * w;
  float* np = t.np[grp] + dst * w;
  float* nm = t.nm[grp] + dst * w;
  float* nv = t.nv[grp] + dst * w;
  for (int k = 0; k < w; ++k) {
    np[k] = p[k];
    nm[k] = fresh ? 0.f : m[k];     // cat_tensors_to_optimizer: new rows start with zero moments
    nv[k] = fresh ? 0.f : v[k];
  }
}

__global__ void __launch_bounds__(DN_THREADS)
densify_apply_kernel(int P, int N, const __grid_constant__ DensifyTable t, const uint8_t* __restrict__ code,
                     const uint32_t* __restrict__ off, int nb, long long n_orig, long long n_clone,
                     long long n_child, long long n_split, const float* __restrict__ samples, float kk) {
  __shared__ uint32_t s_warp[DN_THREADS / 32][4];
  const int i = blockIdx.x * DN_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t c = i < P ? code[i] : 0u;
  // block-local exclusive ranks of: kept original, kept clone, kept split parent, split parent
  const uint32_t bits[4] = {c & DN_ORIG, c & DN_CLONE, c & DN_CHILD, c & DN_IS_SPLIT};
  uint32_t rank[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t b = __ballot_sync(0xffffffffu, bits[q] != 0u);
    rank[q] = __popc(b & ((1u << lane) - 1u));
    if (lane == 0) s_warp[warp][q] = __popc(b);
  }
  __syncthreads();
  const int rows[4] = {0, 1, 2, 4};    // rows of the count table the four ranks belong to
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint32_t base = off[(size_t)rows[q] * nb + blockIdx.x];
    for (int w = 0; w < warp; ++w) base += s_warp[w][q];
    rank[q] += base;
  }
  if (i >= P || (c & (DN_ORIG | DN_CLONE | DN_CHILD)) == 0u) return;
  if (c & DN_ORIG) {
    for (int g = 0; g < t.n; ++g) copy_row(t, g, i, rank[0], false);
    for (int a = 0; a < 2; ++a)
      if (t.aux_in[a]) t.aux_out[a][rank[0]] = t.aux_in[a][i];
  }
  if (c & DN_CLONE) {
    const long long dst = n_orig + rank[1];
    for (int g = 0; g < t.n; ++g) copy_row(t, g, i, dst, true);
    for (int a = 0; a < 2; ++a)
      if (t.aux_in[a]) t.aux_out[a][dst] = t.aux_in[a][i];
  }
  if (c & DN_CHILD) {
    // densify_and_split (gaussian_model.py:593-626): samples ~ N(0, scaling), rotated by R(q / |q|)
    // (build_rotation, GS/utils/general_utils.py:113-135), added to the parent's mean;
    // scaling' = log(scaling / (0.8 N)).  Sample of copy k of split parent r: row k * n_split + r.
    const float* q_ = t.p[t.rotation] + 4ll * i;
    const float qn = sqrtf(q_[0] * q_[0] + q_[1] * q_[1] + q_[2] * q_[2] + q_[3] * q_[3]);
    const float r = q_[0] / qn, x = q_[1] / qn, y = q_[2] / qn, z = q_[3] / qn;
    const float R[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    const float* sc = t.p[t.scaling] + 3ll * i;
    const float s[3] = {expf(sc[0]), expf(sc[1]), expf(sc[2])};
    const float* mu = t.p[t.xyz] + 3ll * i;
    for (int k = 0; k < N; ++k) {
      const long long dst = n_orig + n_clone + (long long)k * n_child + rank[2];
      for (int g = 0; g < t.n; ++g) copy_row(t, g, i, dst, true);
      const float* u = samples + 3ll * ((long long)k * n_split + rank[3]);
      const float v0 = u[0] * s[0], v1 = u[1] * s[1], v2 = u[2] * s[2];
      float* nx = t.np[t.xyz] + 3ll * dst;
      nx[0] = (R[0][0] * v0 + R[0][1] * v1 + R[0][2] * v2) + mu[0];
      nx[1] = (R[1][0] * v0 + R[1][1] * v1 + R[1][2] * v2) + mu[1];
      nx[2] = (R[2][0] * v0 + R[2][1] * v1 + R[2][2] * v2) + mu[2];
      float* ns = t.np[t.scaling] + 3ll * dst;
      ns[0] = logf(s[0] / kk); ns[1] = logf(s[1] / kk); ns[2] = logf(s[2] / kk);
      for (int a = 0; a < 2; ++a)
        if (t.aux_in[a]) t.aux_out[a][dst] = t.aux_in[a][i];
    }
  }
}

static int check_args(const B2gsDensifyArgs* a) {
  if (!a || a->P < 0 || a->n_groups < 1 || a->n_groups > DN_MAX_GROUPS || a->N < 1) {
    set_error("densify: bad arguments (1..%d groups, N >= 1)", DN_MAX_GROUPS);
    return B2GS_ERR_INVALID;
  }
  const int idx[4] = {a->xyz, a->opacity, a->scaling, a->rotation};
  const int want[4] = {3, 1, 3, 4};
  for (int k = 0; k < 4; ++k)
    if (idx[k] < 0 || idx[k] >= a->n_groups || a->width[idx[k]] != want[k]) {
      set_error("densify: xyz / opacity / scaling / rotation groups must have widths 3 / 1 / 3 / 4");
      return B2GS_ERR_INVALID;
    }
  if (a->P > 0 && (!a->grad_accum || !a->denom)) { set_error("densify: NULL statistics"); return B2GS_ERR_INVALID; }
  return B2GS_OK;
}

}  // namespace b2gs

using namespace b2gs;

extern "C" size_t b2gs_densify_bytes(int32_t P) {
  size_t t = 0;
  carve_densify(nullptr, P, &t);
  return t;
}

extern "C" int b2gs_densify_plan(const B2gsDensifyArgs* a, void* workspace, size_t workspace_bytes,
                                 uint64_t* totals, void* stream) {
  int rc = check_args(a);
  if (rc) return rc;
  size_t need = 0;
  DensifyWS w = carve_densify(workspace, a->P, &need);
  if (!workspace || workspace_bytes < need || !totals) {
    set_error("densify workspace too small: %zu < %zu (or NULL totals)", workspace_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (a->P == 0) {
    B2GS_CUDA(cudaMemsetAsync(totals, 0, 8 * sizeof(uint64_t), st));
    return B2GS_OK;
  }
  {
    B2GS_PROF("densify_plan", st);
    densify_plan_kernel<<<w.nb, DN_THREADS, 0, st>>>(a->P, a->param[a->scaling], a->param[a->opacity],
                                                     a->grad_accum, a->denom, a->max_grad, a->min_opacity,
                                                     a->dense_threshold, a->world_size_threshold,
                                                     a->max_screen_size_on, a->split_divisor, w.code,
                                                     w.block_cnt, w.nb);
    B2GS_LAUNCH_CHECK(st, 0);
  }
  densify_scan_kernel<<<1, 1024, 0, st>>>(w.nb, w.block_cnt, w.block_off, w.totals);
  B2GS_LAUNCH_CHECK(st, 0);
  B2GS_CUDA(cudaMemcpyAsync(totals, w.totals, 8 * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
  return B2GS_OK;
}

extern "C" int b2gs_densify_apply(const B2gsDensifyArgs* a, const void* workspace, const int64_t* totals_host,
                                  const float* samples, float* const* new_param, float* const* new_exp_avg,
                                  float* const* new_exp_avg_sq, int32_t* const* new_aux, void* stream) {
  int rc = check_args(a);
  if (rc) return rc;
  if (!workspace || !totals_host || !new_param || !new_exp_avg || !new_exp_avg_sq) {
    set_error("densify apply: NULL argument");
    return B2GS_ERR_INVALID;
  }
  if (a->P == 0) return B2GS_OK;
  DensifyWS w = carve_densify(const_cast<void*>(workspace), a->P, nullptr);
  const long long n_orig = totals_host[0], n_clone = totals_host[1], n_child = totals_host[2], n_split = totals_host[4];
  if (n_split > 0 && !samples) { set_error("densify apply: %lld split parents need samples", n_split); return B2GS_ERR_INVALID; }
  DensifyTable t;
  memset(&t, 0, sizeof(t));
  t.n = a->n_groups; t.xyz = a->xyz; t.scaling = a->scaling; t.rotation = a->rotation;
  for (int g = 0; g < a->n_groups; ++g) {
    t.p[g] = a->param[g]; t.m[g] = a->exp_avg[g]; t.v[g] = a->exp_avg_sq[g]; t.w[g] = a->width[g];
    t.np[g] = new_param[g]; t.nm[g] = new_exp_avg[g]; t.nv[g] = new_exp_avg_sq[g];
    if (a->width[g] > 0 && (!t.p[g] || !t.m[g] || !t.v[g] || !t.np[g] || !t.nm[g] || !t.nv[g])) {
      set_error("densify apply: group %d has NULL arenas", g);
      return B2GS_ERR_INVALID;
    }
  }
  for (int k = 0; k < 2; ++k) {
    t.aux_in[k] = a->aux[k];
    t.aux_out[k] = (a->aux[k] && new_aux) ? new_aux[k] : nullptr;
    if (t.aux_in[k] && !t.aux_out[k]) t.aux_in[k] = nullptr;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  B2GS_PROF("densify_apply", st);
  densify_apply_kernel<<<w.nb, DN_THREADS, 0, st>>>(a->P, a->N, t, w.code, w.block_off, w.nb, n_orig, n_clone,
                                                    n_child, n_split, samples, a->split_divisor);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}
