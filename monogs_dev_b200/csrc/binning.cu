// binning.cu — tile ranges, (tile, Gaussian) instance scatter and per-tile depth sort.
//
// Replaces: cub InclusiveSum + D2H (CR/rasterizer_impl.cu:280-284), duplicateWithKeys
// (:70-111), the 64-bit cub SortPairs over [0, 32+bit) (:306-311) and identifyTileRanges
// (:116-138).
//
// B200-first restructuring (DESIGN.md "Binning without a global sort"): the reference
// sorts R 12-byte (tile<<32|depth, id) pairs with 6 global radix passes after a host
// sync.  The same final order — (tile, depth bits, Gaussian id); the id tie-break is
// what radix-sort stability plus id-order emission gives the reference — is produced
// here with no global sort, no scan over the Gaussians, no library call and no sync:
//   1. preprocess_fwd counts instances per tile          (T counters, red.add)
//   2. tile_scan_kernel: counters -> ranges, cursors, num_rendered, overflow flag
//   3. scatter_instances_kernel: each instance's 8 B key (depth bits<<32 | id<<8 | block mask) goes
//      to its tile's segment through an atomic cursor (arbitrary order inside a segment)
//   4. tile_sort_kernel: one CTA per tile sorts its segment (bitonic network on the
//      unique 64-bit keys, shared memory; oversized segments run it in global/L2) and
//      writes the sorted id list.
// Everything is L2-resident at the BASELINE sizes (R = 1.3 M -> 10 MB of keys).
#include "common.cuh"
#include "expand.cuh"
#include "cull.cuh"
#include <cooperative_groups.h>

namespace b2gs {

GeomWS carve_geom(void* base, int32_t P, size_t* total) {
  Carver c(base);
  GeomWS g;
  size_t n = (size_t)(P > 0 ? P : 1);
  g.xy = c.take<float2>(n);
  g.conic_o = c.take<float4>(n);
  g.rgbd = c.take<float4>(n);
  g.rect = c.take<ushort4>(n);
  g.tiles_touched = c.take<uint32_t>(n);
  g.clamped = c.take<uint8_t>(n);
  if (total) *total = c.total();
  return g;
}

ImageWS carve_image(void* base, int32_t W, int32_t H, size_t* total) {
  Carver c(base);
  ImageWS im;
  size_t N = (size_t)W * H;
  size_t tiles = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  im.final_T = c.take<float>(N);
  im.n_contrib = c.take<uint32_t>(N);
  im.ranges = c.take<uint2>(tiles);
  im.cursor = c.take<uint32_t>(tiles * CURSOR_STRIDE);
  im.status = c.take<uint32_t>(8);
  im.order = c.take<uint32_t>(tiles);
  if (total) *total = c.total();
  return im;
}

BinWS carve_bin(void* base, int64_t cap, size_t* total) {
  Carver c(base);
  BinWS b;
  size_t n = (size_t)(cap > 0 ? cap : 1);
  b.point_list = c.take<uint32_t>(n);
  b.keys = c.take<uint64_t>(n);
  if (total) *total = c.total();
  return b;
}

// ---- counters -> ranges ------------------------------------------------------------
// Exclusive scan of the T per-tile instance counts.  Writes ranges (the reference leaves empty
// tiles at (0,0), CR/rasterizer_impl.cu:313-321), turns the counters into scatter cursors,
// publishes num_rendered and the overflow flag, and lists the tiles longest list first (`order`,
// half-octave length classes) for the render kernels.
// One thread-block CLUSTER of 8 CTAs: as a single 1024-thread CTA the kernel was issue-bound on
// one SM (32 warps x ~1400 instructions = 8 us in the middle of every view's dependent chain).
// Each CTA scans an eighth of the tiles; chunk totals and class censuses cross the cluster
// through distributed shared memory between two cluster barriers.
constexpr int SCAN_CTAS = 8;
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_CLASSES = 64;
// length class of a tile: half octaves of its instance count, LONGEST = class 0
__device__ __forceinline__ uint32_t length_class(uint32_t c) {
  if (c == 0) return SCAN_CLASSES - 1;
  const int msb = 31 - __clz(c);
  const uint32_t half = msb > 0 ? (c >> (msb - 1)) & 1u : 0u;
  const int k = 2 * msb + (int)half;           // 0 .. 63, growing with c
  return (uint32_t)(SCAN_CLASSES - 2 - min(k, SCAN_CLASSES - 2));
}

// Position of a tile inside its length class (phase 2), or just the class census (phase 1), with ONE
// shared atomic per (warp, class present in the warp): neighbouring tiles mostly share a class.
// Inactive lanes pass class SCAN_CLASSES (a dump slot).  Inside a warp the order is by lane.
__device__ __forceinline__ uint32_t class_take(uint32_t* s_class, uint32_t cls, int lane) {
  const unsigned peers = __match_any_sync(0xffffffffu, cls);
  const int leader = __ffs(peers) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(&s_class[cls], (uint32_t)__popc(peers));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
}

__global__ void __cluster_dims__(SCAN_CTAS, 1, 1) __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(uint32_t tiles, long long cap, uint32_t* __restrict__ cursor,
                 uint2* __restrict__ ranges, uint32_t* __restrict__ status,
                 uint32_t* __restrict__ status_out, uint32_t* __restrict__ order) {
  namespace cg = cooperative_groups;
  __shared__ uint32_t s_class[SCAN_CLASSES + 1];     // census of this CTA's tiles (+ dump slot), then the class cursors
  __shared__ uint32_t s_pub[SCAN_CLASSES];           // the census as the peers read it
  __shared__ unsigned long long s_total;             // instances of this CTA's tiles (read by the peers)
  __shared__ unsigned long long s_w64[SCAN_THREADS / 32];
  __shared__ uint32_t s_all[SCAN_CTAS][SCAN_CLASSES];
  __shared__ unsigned long long s_tot[SCAN_CTAS];
  pdl_wait();
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t rank = cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid <= SCAN_CLASSES) s_class[tid] = 0u;
  __syncthreads();
  const uint32_t chunk = (tiles + SCAN_CTAS - 1) / SCAN_CTAS;
  const uint32_t c_lo = min(tiles, rank * chunk), c_hi = min(tiles, c_lo + chunk);
  const uint32_t per = (chunk + SCAN_THREADS - 1) / SCAN_THREADS;      // the same in every CTA
  const uint32_t lo = min(c_hi, c_lo + (uint32_t)tid * per), hi = min(c_hi, lo + per);
  // up to SCAN_REG counters per thread stay in registers (2 x 512 x 8 tiles covers 1920x1080);
  // beyond that the second pass reads them again
  constexpr uint32_t SCAN_REG = 2;
  const bool in_regs = per <= SCAN_REG;
  uint32_t cnt[SCAN_REG], cls[SCAN_REG];
  unsigned long long mine = 0;
  if (in_regs) {
#pragma unroll
    for (uint32_t k = 0; k < SCAN_REG; ++k) cnt[k] = (lo + k < hi) ? cursor[(lo + k) * CURSOR_STRIDE] : 0u;
#pragma unroll
    for (uint32_t k = 0; k < SCAN_REG; ++k) {
      mine += cnt[k];
      cls[k] = lo + k < hi ? length_class(cnt[k]) : (uint32_t)SCAN_CLASSES;
      class_take(s_class, cls[k], lane);
    }
  } else {
    for (uint32_t k = 0; k < per; ++k) {       // uniform trip count: class_take is a warp operation
      const uint32_t t = lo + k;
      const uint32_t c = t < hi ? cursor[t * CURSOR_STRIDE] : 0u;
      mine += c;
      class_take(s_class, t < hi ? length_class(c) : (uint32_t)SCAN_CLASSES, lane);
    }
  }
  // CTA-wide exclusive scan of `mine` (64-bit: the total may exceed 2^32)
  unsigned long long incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long n = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += n;
  }
  if (lane == 31) s_w64[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const unsigned long long w = lane < SCAN_THREADS / 32 ? s_w64[lane] : 0ull;
    unsigned long long wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long n = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += n;
    }
    if (lane < SCAN_THREADS / 32) s_w64[lane] = wi - w;  // exclusive prefix of the warp
    if (lane == 31) s_total = wi;
  } else if (warp == 1 || warp == 2) {
    const int k = (warp - 1) * 32 + lane;
    s_pub[k] = s_class[k];
  }
  cluster.sync();        // every CTA's s_total / s_pub is in place (also a CTA barrier)

  // gather the cluster's totals and censuses (distributed shared memory), then a second cluster
  // barrier: nobody reads a peer's shared memory after it, so CTAs may finish independently
  {
    const int c = tid >> 6, k = tid & 63;                 // SCAN_THREADS == SCAN_CTAS * SCAN_CLASSES
    s_all[c][k] = *cluster.map_shared_rank(&s_pub[k], c);
    if (tid < SCAN_CTAS) s_tot[tid] = *cluster.map_shared_rank(&s_total, tid);
  }
  cluster.sync();

  unsigned long long before = 0, grand = 0;
#pragma unroll
  for (int c = 0; c < SCAN_CTAS; ++c) {
    const unsigned long long v = s_tot[c];
    if ((uint32_t)c < rank) before += v;
    grand += v;
  }
  const bool over = grand >= (1ull << 31) || (cap >= 0 && grand > (unsigned long long)cap);
  if (rank == 0 && tid == 0) {
    status[0] = (uint32_t)min(grand, 0xffffffffull);
    status[1] = over ? 1u : 0u;
    if (status_out) {
      status_out[0] = (uint32_t)min(grand, 0xffffffffull);
      status_out[1] = over ? 1u : 0u;
    }
  }
  if (warp == 0) {
    // class start offsets over the whole image (two classes per lane) + the share of the CTAs
    // before this one inside each class -> this CTA's class cursors
    uint32_t tot[2] = {0u, 0u}, mineoff[2] = {0u, 0u};
#pragma unroll
    for (int c = 0; c < SCAN_CTAS; ++c) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t v = s_all[c][2 * lane + h];
        tot[h] += v;
        if ((uint32_t)c < rank) mineoff[h] += v;
      }
    }
    uint32_t incl2 = tot[0] + tot[1];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, incl2, d);
      if (lane >= d) incl2 += n;
    }
    s_class[2 * lane] = incl2 - tot[0] - tot[1] + mineoff[0];
    s_class[2 * lane + 1] = incl2 - tot[1] + mineoff[1];
  }
  __syncthreads();
  uint32_t run = (uint32_t)(before + s_w64[warp] + incl - mine);
  if (in_regs) {
#pragma unroll
    for (uint32_t k = 0; k < SCAN_REG; ++k) {
      const uint32_t t = lo + k, c = cnt[k];
      const uint32_t at = class_take(s_class, cls[k], lane);
      if (t < hi) {
        ranges[t] = (c && !over) ? make_uint2(run, run + c) : make_uint2(0u, 0u);
        cursor[t * CURSOR_STRIDE] = run;
        order[at] = t;
      }
      run += c;
    }
    return;
  }
  for (uint32_t k = 0; k < per; ++k) {
    const uint32_t t = lo + k;
    const uint32_t c = t < hi ? cursor[t * CURSOR_STRIDE] : 0u;
    const uint32_t at = class_take(s_class, t < hi ? length_class(c) : (uint32_t)SCAN_CLASSES, lane);
    if (t < hi) {
      // on overflow every tile is rendered empty; the caller sees status[1] and retries
      ranges[t] = (c && !over) ? make_uint2(run, run + c) : make_uint2(0u, 0u);
      cursor[t * CURSOR_STRIDE] = run;
      order[at] = t;
    }
    run += c;
  }
}

int launch_tile_scan(const B2gsScene& s, const ImageWS& im, int64_t cap, uint32_t* status_out,
                     cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  B2GS_PROF("tile_scan", st);
  B2GS_LAUNCH(tile_scan_kernel, SCAN_CTAS, SCAN_THREADS, st, (uint32_t)(gx * gy), (long long)cap, im.cursor,
              im.ranges, im.status, status_out, im.order);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

// ---- instance scatter ----------------------------------------------------------------
// key = depth bits << 32 | Gaussian id << 8 | block mask.  The low byte says which of the
// tile's eight 8x4 warp blocks the splat can touch (cull.cuh); it is computed here, where the
// splat's conic is already in the owner lane's hands and the atomic's latency hides the
// arithmetic, rides through the per-tile sort for free (ids are unique, so the order is still
// (depth, id)), and reaches both render kernels inside the sorted list word.
__global__ void __launch_bounds__(256)
scatter_instances_kernel(int P, uint32_t gx, const uint32_t* __restrict__ tiles_touched,
                         const ushort4* __restrict__ rect, const float2* __restrict__ xy,
                         const float4* __restrict__ conic_o, const float4* __restrict__ rgbd,
                         uint32_t* __restrict__ cursor, const uint32_t* __restrict__ status,
                         uint64_t* __restrict__ keys) {
  __shared__ WarpRects s_rects[8];
  __shared__ SplatCullBox s_cull[8][32];
  __shared__ uint32_t s_depth[8][32];
  pdl_wait();
  if (status[1]) return;  // overflow: nothing may be written
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t touched = 0, org = 0, w = 1;
  if (i < P) {
    touched = tiles_touched[i];
    if (touched) {
      const ushort4 r = rect[i];
      org = (uint32_t)r.x | ((uint32_t)r.y << 16);
      w = (uint32_t)(r.z - r.x);
      s_depth[warp][lane] = __float_as_uint(rgbd[i].w);
      s_cull[warp][lane] = splat_cull_setup(xy[i], conic_o[i]);
    }
  }
  const uint32_t id0 = (uint32_t)(blockIdx.x * blockDim.x + warp * 32);   // id of lane 0's Gaussian
  warp_expand_rects<4>(
      s_rects[warp], lane, touched, org, w, gx,
      [&](uint32_t tile) { return atomicAdd(&cursor[tile * CURSOR_STRIDE], 1u); },
      [&](uint32_t slot, uint32_t tile, int owner) {
        const uint32_t ty = tile / gx, tx = tile - ty * gx;
        const uint32_t bm = splat_block_mask(s_cull[warp][owner], (float)(tx * TILE), (float)(ty * TILE));
        keys[slot] = ((uint64_t)s_depth[warp][owner] << 32) | ((id0 + (uint32_t)owner) << 8) | bm;
      });
}

// ---- per-tile sort -------------------------------------------------------------------
// Bitonic network in its all-ascending form (first step of every merge stage compares
// mirrored positions, the rest are half-cleaners).  Because every compare-exchange puts
// the minimum at the lower index, positions >= n behave as +infinity without being
// stored: a pair whose upper index is >= n is skipped.  So the segment needs no padding
// and the same code runs on shared memory or directly on the global segment.
constexpr int SORT_THREADS = 256;
constexpr int SORT_SMEM_KEYS = 4096;  // 32 KB

__device__ __forceinline__ void bitonic_sort_block(uint64_t* __restrict__ a, uint32_t n) {
  uint32_t N = 2;
  while (N < n) N <<= 1;
  const uint32_t half = N >> 1;
  for (uint32_t k = 2; k <= N; k <<= 1) {
    const uint32_t hk = k >> 1;
    for (uint32_t t = threadIdx.x; t < half; t += SORT_THREADS) {
      const uint32_t off = t & (hk - 1);  // hk, j are powers of two
      const uint32_t i = 2 * t - off, l = 2 * (t - off) + k - 1 - off;
      if (l < n) {
        const uint64_t x = a[i], y = a[l];
        if (y < x) { a[i] = y; a[l] = x; }
      }
    }
    __syncthreads();
    for (uint32_t j = k >> 2; j > 0; j >>= 1) {
      for (uint32_t t = threadIdx.x; t < half; t += SORT_THREADS) {
        const uint32_t i = 2 * t - (t & (j - 1)), l = i + j;
        if (l < n) {
          const uint64_t x = a[i], y = a[l];
          if (y < x) { a[i] = y; a[l] = x; }
        }
      }
      __syncthreads();
    }
  }
}

// 32 keys, one per lane, ascending: bitonic network on shuffles (no shared memory, no barrier)
__device__ __forceinline__ uint64_t warp_sort32(uint64_t key, int lane) {
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const uint64_t other = __shfl_xor_sync(0xffffffffu, key, j);
      const bool take_min = ((lane & j) == 0) == ((lane & k) == 0);
      key = ((key < other) == take_min) ? key : other;   // equal keys (padding): either one
    }
  }
  return key;
}

// Segments up to MERGE_KEYS: every warp sorts 32-key chunks in registers, then log2(n/32) merge
// levels in which each key finds its output slot by a binary search of the sibling run (keys are
// unique, so rank = own position + number of smaller sibling keys): 5 barriers for a 512-key
// tile instead of the 45 of a full bitonic network.  Keys of a right run count sibling keys <=
// themselves, written as v < key + 1; the padding is 2^64 - 2 so that this cannot wrap.  A warp's
// 32 keys share run and sibling, so the full-sibling fast path (fixed trip count, no bounds
// test) is warp-uniform; only the last run pair of a level takes the bounded search.
// Larger segments: the bitonic network in shared memory (<= SORT_SMEM_KEYS) or in global / L2.
constexpr uint32_t MERGE_KEYS = SORT_SMEM_KEYS / 2;
constexpr uint64_t SORT_PAD = ~0ull - 1;

template <uint32_t L>
__device__ __forceinline__ void merge_level(const uint64_t* __restrict__ in, uint64_t* __restrict__ out,
                                            uint32_t nn) {
  for (uint32_t i = threadIdx.x; i < nn; i += SORT_THREADS) {
    const uint64_t key = in[i];
    const uint32_t pos = i & (L - 1), run = i - pos;   // run start
    const uint32_t sib = run ^ L;                       // sibling run start
    const uint64_t kk = key + ((run & L) ? 1ull : 0ull);
    const uint64_t* sp = in + sib;
    uint32_t lo = 0;                                    // number of sibling keys < kk
    if (sib + L <= nn) {
      // count is in [lo, lo + 2 st] before each probe; lo <= L - 1 throughout, so no bounds test
#pragma unroll
      for (uint32_t st = L >> 1; st > 0; st >>= 1)
        if (sp[lo + st - 1] < kk) lo += st;
      if (sp[lo] < kk) lo += 1;
    } else if (sib < nn) {
      const uint32_t len = nn - sib;
      for (uint32_t st = L; st > 0; st >>= 1) {
        const uint32_t m = lo + st;
        if (m <= len && sp[m - 1] < kk) lo = m;
      }
    }
    out[(run & ~L) + pos + lo] = key;
  }
}

__global__ void __launch_bounds__(SORT_THREADS)
tile_sort_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ order,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                 uint32_t* __restrict__ cursor) {
  __shared__ uint64_t s_keys[SORT_SMEM_KEYS];
  pdl_wait();
  const uint32_t tile = order[blockIdx.x];
  // the scatter is done with this tile's cursor: leave it zero for the next forward on this workspace
  // (B2gsForwardOut::tile_counters_clean)
  if (threadIdx.x == 0) cursor[tile * CURSOR_STRIDE] = 0u;
  const uint2 r = ranges[tile];
  const uint32_t n = r.y - r.x;
  if (n == 0) return;
  uint64_t* seg = keys + r.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (n <= MERGE_KEYS) {
    const uint32_t nn = (n + 31u) & ~31u;               // padded to whole warps
    for (uint32_t c = warp * 32; c < nn; c += SORT_THREADS) {
      const uint32_t i = c + lane;
      s_keys[i] = warp_sort32(i < n ? seg[i] : SORT_PAD, lane);
    }
    __syncthreads();
    uint64_t* a = s_keys;
    uint64_t* b = s_keys + MERGE_KEYS;
#define B2GS_MERGE(LV)                                                        \
    if (LV < nn) {                                                            \
      merge_level<LV>(a, b, nn);                                              \
      __syncthreads();                                                        \
      uint64_t* t = a; a = b; b = t;                                          \
    }
    B2GS_MERGE(32u) B2GS_MERGE(64u) B2GS_MERGE(128u) B2GS_MERGE(256u) B2GS_MERGE(512u) B2GS_MERGE(1024u)
#undef B2GS_MERGE
    for (uint32_t i = threadIdx.x; i < n; i += SORT_THREADS) point_list[r.x + i] = (uint32_t)a[i];
  } else if (n <= SORT_SMEM_KEYS) {
    for (uint32_t i = threadIdx.x; i < n; i += SORT_THREADS) s_keys[i] = seg[i];
    __syncthreads();
    bitonic_sort_block(s_keys, n);
    for (uint32_t i = threadIdx.x; i < n; i += SORT_THREADS) point_list[r.x + i] = (uint32_t)s_keys[i];
  } else {
    __syncthreads();
    bitonic_sort_block(seg, n);
    for (uint32_t i = threadIdx.x; i < n; i += SORT_THREADS) point_list[r.x + i] = (uint32_t)seg[i];
  }
}

int launch_binning(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                   int64_t cap, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  if (cap <= 0 || s.P == 0) {   // no sort will run: clear the tile counters here
    B2GS_CUDA(cudaMemsetAsync(im.cursor, 0, sizeof(uint32_t) * (size_t)gx * gy * CURSOR_STRIDE, st));
    return B2GS_OK;
  }
  {
    B2GS_PROF("scatter_instances", st);
    B2GS_LAUNCH(scatter_instances_kernel, (s.P + 255) / 256, 256, st, s.P, (uint32_t)gx, (const uint32_t*)g.tiles_touched,
                                                               g.rect, g.xy, g.conic_o, g.rgbd, im.cursor,
                                                               im.status,
                                                               b.keys);
    B2GS_LAUNCH_CHECK(st, s.debug);
  }
  B2GS_PROF("tile_sort", st);
  B2GS_LAUNCH(tile_sort_kernel, gx * gy, SORT_THREADS, st, (const uint2*)im.ranges, (const uint32_t*)im.order, b.keys,
              b.point_list, im.cursor);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

// ---- debug export (parity tests) ---------------------------------------------------
// Materialises the reference's sorted key list: (tile << 32) | depth bits, and the ids.
__global__ void export_keys_kernel(const uint2* __restrict__ ranges,
                                   const uint32_t* __restrict__ point_list,
                                   const float4* __restrict__ rgbd, uint64_t* __restrict__ keys,
                                   uint32_t* __restrict__ values) {
  const uint2 r = ranges[blockIdx.x];
  for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x) {
    const uint32_t id = point_list[i] >> 8;   // low byte: block mask
    if (keys) keys[i] = ((uint64_t)blockIdx.x << 32) | __float_as_uint(rgbd[id].w);
    if (values) values[i] = id;
  }
}

}  // namespace b2gs

extern "C" int b2gs_debug_export_binning(const B2gsScene* scene, const B2gsWorkspace* ws,
                                         int64_t cap, uint64_t* keys, uint32_t* values,
                                         uint32_t* ranges, void* stream) {
  using namespace b2gs;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  GeomWS g = carve_geom(ws->geom, scene->P, nullptr);
  ImageWS im = carve_image(ws->image, scene->W, scene->H, nullptr);
  BinWS b = carve_bin(ws->binning, cap, nullptr);
  const size_t tiles = (size_t)((scene->W + TILE - 1) / TILE) * ((scene->H + TILE - 1) / TILE);
  if (cap > 0 && (keys || values)) {
    export_keys_kernel<<<(uint32_t)tiles, 256, 0, st>>>(im.ranges, b.point_list, g.rgbd, keys, values);
    B2GS_LAUNCH_CHECK(st, 1);
  }
  if (ranges) {
    B2GS_CUDA(cudaMemcpyAsync(ranges, im.ranges, tiles * sizeof(uint2), cudaMemcpyDeviceToDevice, st));
  }
  return B2GS_OK;
}
