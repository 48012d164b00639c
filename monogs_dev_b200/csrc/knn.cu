// knn.cu — exact mean squared distance to the 3 nearest neighbours (distCUDA2).
// Replaces SimpleKNN::knn and its kernels (KNN/simple_knn.cu:63-221).
//
// Same contract: out[i] = (d1 + d2 + d3) / 3 with d1<=d2<=d3 the three smallest
// squared distances from point i to any OTHER point (excluded by index, so
// coincident points give 0).  Distances use the reference's expression shape, so
// results are bit-identical to the reference whatever the traversal order.
//
// B200-first redesign:
//  * no cudaMalloc / thrust vectors / blocking D2H (KNN/simple_knn.cu:188-208):
//    the bounding box stays on the device, all storage is caller workspace;
//  * points are gathered once into Morton order as float4 (coalesced afterwards);
//  * a 3-level box hierarchy (32 / 1024 / 32768 points) replaces the flat list of
//    1024-point boxes that every thread scans in full (KNN/simple_knn.cu:165-179);
//  * one warp owns one 32-point leaf: pruning decisions are warp votes (no
//    divergence), candidate leaves are loaded coalesced and broadcast from shared
//    memory, the warp's own leaf / 1024-box / 32768-box are visited first so the
//    pruning radius is tight before the global sweep;
//  * the Morton sort (thrust::sort_by_key there, KNN/simple_knn.cu:210-213) is an own
//    stable LSD radix sort of the 30-bit codes: 4 passes of 8 bits, per pass a per-tile
//    digit histogram, one scan, and a stable scatter whose in-tile ranks come from
//    warp match_any votes (no library call anywhere in the library).
#include "common.cuh"
#include <float.h>
#include <mutex>

namespace b2gs {

struct Box { float lo[3]; float hi[3]; };  // 24 B

struct KnnWS {
  uint32_t* bbox;      // 6 ordered-uint encodings: min xyz, max xyz
  uint32_t* morton;    // [P]
  uint32_t* iota;      // [P]
  uint32_t* morton_s;  // [P]
  uint32_t* idx_s;     // [P]
  float4* pts;         // [P] sorted points, w = original index bits
  Box* l0; Box* l1; Box* l2;
  uint32_t* hist;      // [256][tiles] digit counts of one radix pass, then their per-digit exclusive scan
  uint32_t* totals;    // [256] keys per digit
  int n0, n1, n2, tiles;
};

constexpr int RS_THREADS = 256;                       // 8 warps
constexpr int RS_PER_WARP = 512;                      // consecutive keys owned by one warp
constexpr int RS_TILE = (RS_THREADS / 32) * RS_PER_WARP;   // 4096 keys per block

static KnnWS carve_knn(void* base, int32_t P, size_t* total) {
  Carver c(base);
  KnnWS k;
  size_t n = (size_t)(P > 0 ? P : 1);
  k.n0 = (int)((n + 31) / 32);
  k.n1 = (k.n0 + 31) / 32;
  k.n2 = (k.n1 + 31) / 32;
  k.bbox = c.take<uint32_t>(8);
  k.morton = c.take<uint32_t>(n);
  k.iota = c.take<uint32_t>(n);
  k.morton_s = c.take<uint32_t>(n);
  k.idx_s = c.take<uint32_t>(n);
  k.pts = c.take<float4>((size_t)k.n0 * 32);
  k.l0 = c.take<Box>(k.n0);
  k.l1 = c.take<Box>(k.n1);
  k.l2 = c.take<Box>(k.n2);
  k.tiles = (int)((n + RS_TILE - 1) / RS_TILE);
  k.hist = c.take<uint32_t>((size_t)256 * k.tiles);
  k.totals = c.take<uint32_t>(256);
  if (total) *total = c.total();
  return k;
}

// order-preserving float <-> uint encoding for atomicMin/Max
__device__ __forceinline__ uint32_t f2ord(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void knn_init_kernel(uint32_t* bbox) {
  if (threadIdx.x < 3) bbox[threadIdx.x] = 0xffffffffu;
  else if (threadIdx.x < 6) bbox[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) knn_bbox_kernel(int P, const float* __restrict__ pts,
                                                       uint32_t* __restrict__ bbox) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float v = pts[3 * i + a];
      lo[a] = fminf(lo[a], v);
      hi[a] = fmaxf(hi[a], v);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(&bbox[a], f2ord(lo[a]));
      atomicMax(&bbox[3 + a], f2ord(hi[a]));
    }
  }
}

__device__ __forceinline__ uint32_t spread3(uint32_t x) {  // 10 bits -> every third bit
  x = (x | (x << 16)) & 0x030000FFu;
  x = (x | (x << 8)) & 0x0300F00Fu;
  x = (x | (x << 4)) & 0x030C30C3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ pts,
                                                         const uint32_t* __restrict__ bbox,
                                                         uint32_t* __restrict__ codes,
                                                         uint32_t* __restrict__ iota) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t code = 0;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float lo = ord2f(bbox[a]), hi = ord2f(bbox[3 + a]);
    float ext = hi - lo;
    float u = ext > 0.f ? (pts[3 * i + a] - lo) / ext : 0.f;
    u = fminf(fmaxf(u, 0.f), 1.f);
    code |= spread3((uint32_t)(u * 1023.f)) << a;
  }
  codes[i] = code;
  iota[i] = (uint32_t)i;
}

// ---- stable LSD radix sort of (code, index) pairs, 8 bits per pass ------------------------------
// A block owns a tile of 4096 consecutive pairs, a warp 512 consecutive ones, so "tile order" is
// warp-major and a pair's rank among equal digits is  (pairs of earlier tiles) + (earlier warps of the
// tile) + (earlier rounds of the warp) + (lower lanes of the round) — every term a count, no sort
// inside.  Lanes with the same digit find each other with __match_any_sync.
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(int n, int shift, const uint32_t* __restrict__ keys, uint32_t* __restrict__ hist, int tiles) {
  __shared__ uint32_t s_cnt[256];
  s_cnt[threadIdx.x] = 0u;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long base = (long long)blockIdx.x * RS_TILE + (long long)warp * RS_PER_WARP;
#pragma unroll 4
  for (int r = 0; r < RS_PER_WARP / 32; ++r) {
    const long long i = base + r * 32 + lane;
    const bool ok = i < n;
    const uint32_t d = ok ? (keys[i] >> shift) & 255u : 256u;     // 256: padding lanes group together
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    if (ok && lane == __ffs(m) - 1) atomicAdd(&s_cnt[d], (uint32_t)__popc(m));
  }
  __syncthreads();
  hist[(size_t)threadIdx.x * tiles + blockIdx.x] = s_cnt[threadIdx.x];
}

// one CTA per digit: exclusive scan of that digit's row of per-tile counts (coalesced), row total to
// totals[digit]; the scatter blocks turn the 256 totals into digit bases themselves
constexpr int RS_SCAN_THREADS = 512;
__global__ void __launch_bounds__(RS_SCAN_THREADS)
radix_scan_kernel(int tiles, uint32_t* __restrict__ hist, uint32_t* __restrict__ totals) {
  __shared__ uint32_t s_w[RS_SCAN_THREADS / 32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t* row = hist + (size_t)blockIdx.x * tiles;
  if (tid == 0) s_carry = 0u;
  __syncthreads();
  for (int t0 = 0; t0 < tiles; t0 += RS_SCAN_THREADS) {
    const int t = t0 + tid;
    const uint32_t c = t < tiles ? row[t] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += v;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t before = s_carry;
    for (int w = 0; w < warp; ++w) before += s_w[w];
    if (t < tiles) row[t] = before + incl - c;
    __syncthreads();
    if (tid == RS_SCAN_THREADS - 1) s_carry = before + incl;
    __syncthreads();
  }
  if (tid == 0) totals[blockIdx.x] = s_carry;
}

__global__ void __launch_bounds__(RS_THREADS)
radix_scatter_kernel(int n, int shift, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                     const uint32_t* __restrict__ hist, const uint32_t* __restrict__ totals, int tiles,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
  __shared__ uint32_t s_cnt[RS_THREADS / 32][256];
  __shared__ uint32_t s_wsum[RS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int d = lane; d < 256; d += 32) s_cnt[warp][d] = 0u;
  __syncwarp();
  const long long base = (long long)blockIdx.x * RS_TILE + (long long)warp * RS_PER_WARP;
  constexpr int ROUNDS = RS_PER_WARP / 32;
  uint32_t key[ROUNDS], val[ROUNDS];
#pragma unroll
  for (int r = 0; r < ROUNDS; ++r) {
    const long long i = base + r * 32 + lane;
    const bool ok = i < n;
    key[r] = ok ? keys[i] : 0xffffffffu;
    val[r] = ok ? vals[i] : 0u;
    const uint32_t d = ok ? (key[r] >> shift) & 255u : 256u;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    if (ok && lane == __ffs(m) - 1) s_cnt[warp][d] += (uint32_t)__popc(m);   // one lane per digit group
    __syncwarp();
  }
  __syncthreads();
  {  // per digit: digit base (exclusive scan of the 256 totals) + this tile's offset in the digit's
     // row, then the warps of the tile in order
    const int d = threadIdx.x;
    const uint32_t tot = totals[d];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t digit_base = incl - tot;
    for (int w = 0; w < warp; ++w) digit_base += s_wsum[w];
    uint32_t run = digit_base + hist[(size_t)d * tiles + blockIdx.x];
#pragma unroll
    for (int w = 0; w < RS_THREADS / 32; ++w) {
      const uint32_t c = s_cnt[w][d];
      s_cnt[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < ROUNDS; ++r) {
    const long long i = base + r * 32 + lane;
    const bool ok = i < n;
    const uint32_t d = ok ? (key[r] >> shift) & 255u : 256u;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    uint32_t dst = 0;
    if (ok) dst = s_cnt[warp][d] + (uint32_t)__popc(m & ((1u << lane) - 1u));
    __syncwarp();
    if (ok && lane == __ffs(m) - 1) s_cnt[warp][d] += (uint32_t)__popc(m);
    __syncwarp();
    if (ok) {
      keys_out[dst] = key[r];
      vals_out[dst] = val[r];
    }
  }
}

// gather into Morton order + leaf boxes (one warp per 32-point leaf)
__global__ void __launch_bounds__(256) knn_leaf_kernel(int P, const float* __restrict__ pts,
                                                       const uint32_t* __restrict__ idx_s,
                                                       float4* __restrict__ sorted, Box* __restrict__ l0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  float4 p = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, __uint_as_float(0xffffffffu));
  if (i < P) {
    const uint32_t id = idx_s[i];
    p = make_float4(pts[3 * id], pts[3 * id + 1], pts[3 * id + 2], __uint_as_float(id));
    lo[0] = hi[0] = p.x; lo[1] = hi[1] = p.y; lo[2] = hi[2] = p.z;
  }
  if (i < ((P + 31) / 32) * 32) sorted[i] = p;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  }
  if (lane == 0 && (i / 32) * 32 < P) {
    Box b;
#pragma unroll
    for (int a = 0; a < 3; ++a) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
    l0[i / 32] = b;
  }
}

// parent boxes: one warp per parent, lane k reads child k
__global__ void __launch_bounds__(256) knn_parent_kernel(int n_child, const Box* __restrict__ child,
                                                         Box* __restrict__ parent) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int n_parent = (n_child + 31) / 32;
  if (w >= n_parent) return;
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  const int c = w * 32 + lane;
  if (c < n_child) {
    Box b = child[c];
#pragma unroll
    for (int a = 0; a < 3; ++a) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  }
  if (lane == 0) {
    Box b;
#pragma unroll
    for (int a = 0; a < 3; ++a) { b.lo[a] = lo[a]; b.hi[a] = hi[a]; }
    parent[w] = b;
  }
}

__device__ __forceinline__ float sq3(float dx, float dy, float dz) {
  return dx * dx + dy * dy + dz * dz;
}

// squared distance from p to the box (0 inside) — KNN/simple_knn.cu:119-129
__device__ __forceinline__ float box_dist2(const Box& b, float px, float py, float pz) {
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (px < b.lo[0] || px > b.hi[0]) dx = fminf(fabsf(px - b.lo[0]), fabsf(px - b.hi[0]));
  if (py < b.lo[1] || py > b.hi[1]) dy = fminf(fabsf(py - b.lo[1]), fabsf(py - b.hi[1]));
  if (pz < b.lo[2] || pz > b.hi[2]) dz = fminf(fabsf(pz - b.lo[2]), fabsf(pz - b.hi[2]));
  return sq3(dx, dy, dz);
}

__device__ __forceinline__ void kbest3(float dist, float* best) {  // KNN/simple_knn.cu:131-145
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (best[j] > dist) {
      float t = best[j];
      best[j] = dist;
      dist = t;
    }
  }
}

struct LeafVisitor {
  const float4* __restrict__ pts;
  float4* s_c;  // this warp's 32-entry staging slot
  int lane;
  float px, py, pz;
  float best[3];
  unsigned visits;   // leaves scanned by this warp (debug work counter)

  __device__ __forceinline__ void leaf(int b0, int count, bool own) {
    visits += (unsigned)count;
    __syncwarp();
    s_c[lane] = pts[(size_t)b0 * 32 + lane];
    __syncwarp();
    for (int k = 0; k < count; ++k) {
      if (own && k == lane) continue;
      const float4 c = s_c[k];
      kbest3(sq3(c.x - px, c.y - py, c.z - pz), best);
    }
  }
};

__global__ void __launch_bounds__(256)
knn_search_kernel(int P, int n0, int n1, int n2, const float4* __restrict__ pts,
                  const Box* __restrict__ l0, const Box* __restrict__ l1, const Box* __restrict__ l2,
                  float* __restrict__ out, unsigned long long* __restrict__ counters) {
  __shared__ float4 s_stage[8][32];
  const int warp_in_block = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int my0 = blockIdx.x * 8 + warp_in_block;  // leaf owned by this warp
  if (my0 >= n0) return;
  const int my1 = my0 >> 5, my2 = my0 >> 10;
  const int i = my0 * 32 + lane;
  const float4 me = pts[i];
  const bool valid = i < P;

  LeafVisitor v;
  v.pts = pts; v.s_c = s_stage[warp_in_block]; v.lane = lane;
  v.px = me.x; v.py = me.y; v.pz = me.z;
  v.best[0] = v.best[1] = v.best[2] = FLT_MAX;
  v.visits = 0;
  // padding lanes must never keep a box alive
  const float NEVER = -1.f;

  auto leaf_count = [&](int b0) { return min(32, P - b0 * 32); };
  auto wants = [&](const Box& b) {
    const float d = box_dist2(b, v.px, v.py, v.pz);
    const float lim = valid ? v.best[2] : NEVER;
    return __any_sync(0xffffffffu, !(d > lim));
  };
  auto visit_l1 = [&](int b1, int skip0) {
    const int first = b1 * 32, last = min(n0, first + 32);
    for (int b0 = first; b0 < last; ++b0) {
      if (b0 == skip0) continue;
      if (!wants(l0[b0])) continue;
      v.leaf(b0, leaf_count(b0), false);
    }
  };
  auto visit_l2 = [&](int b2, int skip1) {
    const int first = b2 * 32, last = min(n1, first + 32);
    for (int b1 = first; b1 < last; ++b1) {
      if (b1 == skip1) continue;
      if (!wants(l1[b1])) continue;
      visit_l1(b1, -1);
    }
  };

  v.leaf(my0, leaf_count(my0), true);  // own leaf (self excluded by index)
  visit_l1(my1, my0);                  // own 1024-box
  visit_l2(my2, my1);                  // own 32768-box
  for (int b2 = 0; b2 < n2; ++b2) {    // everything else
    if (b2 == my2) continue;
    if (!wants(l2[b2])) continue;
    visit_l2(b2, -1);
  }
  if (valid) out[__float_as_uint(me.w)] = (v.best[0] + v.best[1] + v.best[2]) / 3.0f;
  // every lane evaluates one distance per candidate point of a visited leaf
  if (counters && lane == 0) atomicAdd(&counters[7], 32ull * v.visits);
}

}  // namespace b2gs

using namespace b2gs;

extern "C" size_t b2gs_knn_bytes(int32_t P) {
  size_t t = 0;
  carve_knn(nullptr, P, &t);
  return t;
}

extern "C" int b2gs_knn3_dist2(const float* points, int32_t P, float* out, void* workspace,
                               size_t workspace_bytes, void* stream) {
  if (P < 0) { set_error("b2gs_knn3_dist2: negative P"); return B2GS_ERR_INVALID; }
  if (P == 0) return B2GS_OK;
  if (!points || !out) { set_error("b2gs_knn3_dist2: NULL points/out"); return B2GS_ERR_INVALID; }
  size_t need = 0;
  KnnWS k = carve_knn(workspace, P, &need);
  if (!workspace || workspace_bytes < need) {
    set_error("knn workspace too small: %zu < %zu", workspace_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  knn_init_kernel<<<1, 32, 0, st>>>(k.bbox);
  int blocks = (P + 255) / 256;
  knn_bbox_kernel<<<blocks < 1184 ? blocks : 1184, 256, 0, st>>>(P, points, k.bbox);
  knn_morton_kernel<<<blocks, 256, 0, st>>>(P, points, k.bbox, k.morton, k.iota);
  B2GS_LAUNCH_CHECK(st, 0);
  note_launches(2);
  {
    // 30-bit codes: 4 passes of 8 bits, ping-ponging (morton, iota) <-> (morton_s, idx_s); an even
    // number of passes leaves the sorted pairs in (morton, iota)
    B2GS_PROF("knn_morton_sort", st);
    uint32_t *ka = k.morton, *va = k.iota, *kb = k.morton_s, *vb = k.idx_s;
    for (int shift = 0; shift < 32; shift += 8) {
      radix_hist_kernel<<<k.tiles, RS_THREADS, 0, st>>>(P, shift, ka, k.hist, k.tiles);
      radix_scan_kernel<<<256, RS_SCAN_THREADS, 0, st>>>(k.tiles, k.hist, k.totals);
      radix_scatter_kernel<<<k.tiles, RS_THREADS, 0, st>>>(P, shift, ka, va, k.hist, k.totals, k.tiles, kb, vb);
      uint32_t* t = ka; ka = kb; kb = t;
      t = va; va = vb; vb = t;
    }
    B2GS_LAUNCH_CHECK(st, 0);
    note_launches(11);
  }
  knn_leaf_kernel<<<(k.n0 * 32 + 255) / 256, 256, 0, st>>>(P, points, k.iota, k.pts, k.l0);
  knn_parent_kernel<<<(k.n1 * 32 + 255) / 256, 256, 0, st>>>(k.n0, k.l0, k.l1);
  knn_parent_kernel<<<(k.n2 * 32 + 255) / 256, 256, 0, st>>>(k.n1, k.l1, k.l2);
  B2GS_PROF("knn_search", st);
  knn_search_kernel<<<(k.n0 + 7) / 8, 256, 0, st>>>(P, k.n0, k.n1, k.n2, k.pts, k.l0, k.l1, k.l2, out,
                                                    debug_counters());
  B2GS_LAUNCH_CHECK(st, 0);
  note_launches(3);
  return B2GS_OK;
}

// ---- no-allocation variant --------------------------------------------------------------------
// b2gs_knn3_dist2 with its workspace in a file-scope device array, for callers that cannot hand
// in device memory of their own (under CUDA-graph capture, serving loops, harnesses that forbid
// cudaMalloc).  Every device that loads the library holds these bytes until the process exits.
// All calls on a device share the array: two calls that overlap on the GPU (different streams)
// race.  Nothing in it has to start at zero — knn_init_kernel resets the bounding-box words and
// every other region is written in full before it is read — so the last call's contents are
// simply overwritten.
constexpr int32_t KNN_STATIC_MAX_POINTS = 262144;
constexpr size_t KNN_STATIC_BYTES = 9u << 20;             // carve_knn(262144) needs 8 658 432
static __device__ __align__(256) unsigned char g_knn_static[KNN_STATIC_BYTES];

namespace {
constexpr int KNN_STATIC_MAX_DEVICES = 64;
std::mutex g_knn_static_mu;
void* g_knn_static_ptr[KNN_STATIC_MAX_DEVICES] = {};      // per device: the array's DEVICE address
}  // namespace

extern "C" int32_t b2gs_knn3_dist2_static_max_points(void) { return KNN_STATIC_MAX_POINTS; }

extern "C" int b2gs_knn3_dist2_static(const float* points, int32_t P, float* out, void* stream) {
  if (P < 0) { set_error("b2gs_knn3_dist2_static: negative P"); return B2GS_ERR_INVALID; }
  if (P > KNN_STATIC_MAX_POINTS) {
    set_error("b2gs_knn3_dist2_static: %d points exceed the built-in scratch (at most %d); use "
              "b2gs_knn3_dist2 with a caller workspace", P, KNN_STATIC_MAX_POINTS);
    return B2GS_ERR_INVALID;
  }
  if (P == 0) return B2GS_OK;
  size_t need = 0;
  carve_knn(nullptr, P, &need);
  if (need > KNN_STATIC_BYTES) {      // only if carve_knn grows without this file's constants
    set_error("b2gs_knn3_dist2_static: scratch layout needs %zu bytes, built-in array has %zu", need,
              (size_t)KNN_STATIC_BYTES);
    return B2GS_ERR_WORKSPACE;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= KNN_STATIC_MAX_DEVICES) {
    set_error("b2gs_knn3_dist2_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return B2GS_ERR_CUDA;
  }
  void* ws = nullptr;
  {
    // the array's name is a host-side handle, not a device address: resolve it once per device,
    // with that device current
    std::lock_guard<std::mutex> lock(g_knn_static_mu);
    if (!g_knn_static_ptr[dev]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_knn_static);
      if (e != cudaSuccess || !p) {
        set_error("b2gs_knn3_dist2_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        return B2GS_ERR_CUDA;
      }
      g_knn_static_ptr[dev] = p;
    }
    ws = g_knn_static_ptr[dev];
  }
  return b2gs_knn3_dist2(points, P, out, ws, KNN_STATIC_BYTES, stream);
}
