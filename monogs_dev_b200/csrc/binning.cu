// binning.cu — depth ordering, instance offsets, (tile, Gaussian) instance emission,
// tile binning and tile ranges.
//
// Replaces: cub InclusiveSum + D2H (CR/rasterizer_impl.cu:280-284), duplicateWithKeys
// (:70-111), the 64-bit SortPairs over [0, 32+bit) (:306-311) and identifyTileRanges
// (:116-138).
//
// B200-first restructuring (DESIGN.md "binning"): the reference sorts R 12-byte
// (tile<<32|depth, id) pairs with 6 radix passes.  The same final order
// (tile, depth bits, Gaussian id — the id tie-break comes from radix-sort
// stability and emission in id order) is produced here by
//   1. a stable 32-bit sort of the P depth keys (id payload)       ~16 B x P x 4 passes
//   2. emitting instances in that (depth, id) order, warp-cooperatively
//   3. a stable sort of the R (tile, id) pairs on tile bits only   ~16 B x R x 2 passes
// Stable sort by tile of a (depth,id)-ordered list is ordered by (tile, depth, id):
// exactly the reference's sorted list, bit for bit (tests/test_binning_gpu.py).
#include "common.cuh"
#include <cub/cub.cuh>

namespace b2gs {

struct TouchedInSortedOrder {
  const uint32_t* touched;
  const uint32_t* id_sorted;
  __host__ __device__ __forceinline__ uint32_t operator()(const uint32_t& i) const {
    return touched[id_sorted[i]];
  }
};
using CountIt = cub::CountingInputIterator<uint32_t>;
using TouchedIt = cub::TransformInputIterator<uint32_t, TouchedInSortedOrder, CountIt>;

static size_t depth_temp_bytes(int32_t P) {
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, P);
  TouchedInSortedOrder op{nullptr, nullptr};
  TouchedIt it(CountIt(0), op);
  cub::DeviceScan::InclusiveSum(nullptr, b, it, (uint32_t*)nullptr, P);
  return a > b ? a : b;
}

static size_t tile_temp_bytes(int64_t R) {
  size_t a = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)R);
  return a;
}

GeomWS carve_geom(void* base, int32_t P, size_t* total) {
  Carver c(base);
  GeomWS g;
  size_t n = (size_t)(P > 0 ? P : 1);
  g.xy = c.take<float2>(n);
  g.conic_o = c.take<float4>(n);
  g.rgbd = c.take<float4>(n);
  g.depth_key = c.take<uint32_t>(n);
  g.rect = c.take<ushort4>(n);
  g.tiles_touched = c.take<uint32_t>(n);
  g.clamped = c.take<uint8_t>(n);
  g.iota = c.take<uint32_t>(n);
  g.key_sorted = c.take<uint32_t>(n);
  g.id_sorted = c.take<uint32_t>(n);
  g.offsets = c.take<uint32_t>(n);
  g.misc = c.take<uint32_t>(64);
  g.temp_bytes = depth_temp_bytes((int32_t)n);
  g.temp = c.take<char>(g.temp_bytes);
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
  if (total) *total = c.total();
  return im;
}

BinWS carve_bin(void* base, int64_t R, size_t* total) {
  Carver c(base);
  BinWS b;
  size_t n = (size_t)(R > 0 ? R : 1);
  b.point_list = c.take<uint32_t>(n);
  b.tile_sorted = c.take<uint32_t>(n);
  b.tile_unsorted = c.take<uint32_t>(n);
  b.id_unsorted = c.take<uint32_t>(n);
  b.temp_bytes = tile_temp_bytes((int64_t)n);
  b.temp = c.take<char>(b.temp_bytes);
  if (total) *total = c.total();
  return b;
}

// ---- depth order + offsets -------------------------------------------------------
int launch_depth_order(const B2gsScene& s, const GeomWS& g, cudaStream_t st) {
  size_t tb = g.temp_bytes;
  {
    B2GS_PROF("depth_sort(cub)", st);
    B2GS_CUDA(cub::DeviceRadixSort::SortPairs(g.temp, tb, (const uint32_t*)g.depth_key, g.key_sorted,
                                              (const uint32_t*)g.iota, g.id_sorted, s.P, 0, 32, st));
  }
  if (s.debug) B2GS_CUDA(cudaStreamSynchronize(st));
  TouchedInSortedOrder op{g.tiles_touched, g.id_sorted};
  TouchedIt it(CountIt(0), op);
  tb = g.temp_bytes;
  {
    B2GS_PROF("offset_scan(cub)", st);
    B2GS_CUDA(cub::DeviceScan::InclusiveSum(g.temp, tb, it, g.offsets, s.P, st));
  }
  if (s.debug) B2GS_CUDA(cudaStreamSynchronize(st));
  return B2GS_OK;
}

// ---- instance emission -------------------------------------------------------------
// One warp expands 32 consecutive (depth-ordered) Gaussians into their contiguous
// span of instances; lanes stride over the span so stores are coalesced no matter
// how unevenly the 32 splats are sized (the reference's per-thread double loop,
// CR/rasterizer_impl.cu:97-109, serialises big splats on one thread).
__global__ void __launch_bounds__(256)
emit_instances_kernel(int P, int gx, const uint32_t* __restrict__ id_sorted,
                      const uint32_t* __restrict__ offsets,
                      const uint32_t* __restrict__ tiles_touched,
                      const ushort4* __restrict__ rect, uint32_t* __restrict__ tile_out,
                      uint32_t* __restrict__ id_out) {
  __shared__ uint32_t s_pre[8][32];
  __shared__ uint32_t s_id[8][32];
  __shared__ uint32_t s_org[8][32];  // x0 | y0 << 16
  __shared__ uint32_t s_w[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;

  uint32_t id = 0, touched = 0, incl = 0;
  ushort4 r = make_ushort4(0, 0, 0, 0);
  if (i < P) {
    id = id_sorted[i];
    touched = tiles_touched[id];
    incl = offsets[i];
    if (touched) r = rect[id];
  } else {
    incl = offsets[P - 1];  // empty tail lanes sit at the end of the span
  }
  // span start of this warp in the instance arrays
  const uint32_t warp_base = __shfl_sync(0xffffffffu, incl - touched, 0);
  const uint32_t pre = incl - warp_base;  // inclusive prefix inside the warp
  const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
  if (total == 0) return;
  s_pre[warp][lane] = pre;
  s_id[warp][lane] = id;
  s_org[warp][lane] = (uint32_t)r.x | ((uint32_t)r.y << 16);
  s_w[warp][lane] = (uint32_t)(r.z - r.x);
  __syncwarp();
  for (uint32_t k = lane; k < total; k += 32) {
    // smallest j with pre[j] > k
    int lo = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      if (s_pre[warp][lo + step - 1] <= k) lo += step;
    }
    const uint32_t pj = s_pre[warp][lo];
    const uint32_t prev = lo ? s_pre[warp][lo - 1] : 0u;
    (void)pj;
    const uint32_t local = k - prev;
    const uint32_t w = s_w[warp][lo];
    const uint32_t org = s_org[warp][lo];
    const uint32_t dy = local / w, dx = local - dy * w;
    const uint32_t tile = ((org >> 16) + dy) * (uint32_t)gx + (org & 0xffffu) + dx;
    tile_out[warp_base + k] = tile;
    id_out[warp_base + k] = s_id[warp][lo];
  }
}

__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t R, const uint32_t* __restrict__ tile_sorted, uint2* __restrict__ ranges) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R) return;
  uint32_t cur = tile_sorted[i];
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    uint32_t prev = tile_sorted[i - 1];
    if (cur != prev) {
      ranges[prev].y = i;
      ranges[cur].x = i;
    }
  }
  if (i == R - 1) ranges[cur].y = R;
}

int launch_binning(const B2gsScene& s, const GeomWS& g, const ImageWS& im, const BinWS& b,
                   int64_t R, cudaStream_t st) {
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  const uint32_t tiles = (uint32_t)gx * gy;
  B2GS_CUDA(cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * tiles, st));
  if (R <= 0) return B2GS_OK;
  {
    B2GS_PROF("emit_instances", st);
    emit_instances_kernel<<<(s.P + 255) / 256, 256, 0, st>>>(s.P, gx, g.id_sorted, g.offsets,
                                                            g.tiles_touched, g.rect,
                                                            b.tile_unsorted, b.id_unsorted);
    B2GS_LAUNCH_CHECK(st, s.debug);
  }
  size_t tb = b.temp_bytes;
  {
    B2GS_PROF("tile_sort(cub)", st);
    B2GS_CUDA(cub::DeviceRadixSort::SortPairs(b.temp, tb, (const uint32_t*)b.tile_unsorted,
                                              b.tile_sorted, (const uint32_t*)b.id_unsorted,
                                              b.point_list, (int)R, 0, tile_bits(tiles), st));
  }
  if (s.debug) B2GS_CUDA(cudaStreamSynchronize(st));
  B2GS_PROF("tile_ranges", st);
  tile_ranges_kernel<<<(uint32_t)((R + 255) / 256), 256, 0, st>>>((uint32_t)R, b.tile_sorted,
                                                                 im.ranges);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

// ---- debug export (parity tests) ---------------------------------------------------
__global__ void export_keys_kernel(uint32_t R, const uint32_t* __restrict__ tile_sorted,
                                   const uint32_t* __restrict__ point_list,
                                   const uint32_t* __restrict__ depth_key,
                                   uint64_t* __restrict__ keys, uint32_t* __restrict__ values) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R) return;
  uint32_t id = point_list[i];
  if (keys) keys[i] = ((uint64_t)tile_sorted[i] << 32) | depth_key[id];
  if (values) values[i] = id;
}

}  // namespace b2gs

extern "C" int b2gs_debug_export_binning(const B2gsScene* scene, const B2gsWorkspace* ws,
                                         int64_t R, uint64_t* keys, uint32_t* values,
                                         uint32_t* ranges, void* stream) {
  using namespace b2gs;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  GeomWS g = carve_geom(ws->geom, scene->P, nullptr);
  ImageWS im = carve_image(ws->image, scene->W, scene->H, nullptr);
  BinWS b = carve_bin(ws->binning, R, nullptr);
  if (R > 0 && (keys || values)) {
    export_keys_kernel<<<(uint32_t)((R + 255) / 256), 256, 0, st>>>(
        (uint32_t)R, b.tile_sorted, b.point_list, g.depth_key, keys, values);
    B2GS_LAUNCH_CHECK(st, 1);
  }
  if (ranges) {
    const size_t tiles = (size_t)((scene->W + TILE - 1) / TILE) * ((scene->H + TILE - 1) / TILE);
    B2GS_CUDA(cudaMemcpyAsync(ranges, im.ranges, tiles * sizeof(uint2), cudaMemcpyDeviceToDevice, st));
  }
  return B2GS_OK;
}
