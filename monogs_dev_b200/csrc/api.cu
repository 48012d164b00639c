// api.cu — the extern "C" surface of libb2gs.so (see include/b2gs.h).
// Host orchestration replacing CudaRasterizer::Rasterizer::forward / ::backward
// (CR/rasterizer_impl.cu:198-344, :348-455).  Stream-aware, re-entrant, no
// allocations: all buffers come from the caller.
#include "common.cuh"
#include <atomic>
#include <mutex>
#include <vector>
#include <string>
#include <map>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

namespace b2gs {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

static thread_local const int32_t* g_guard = nullptr;
const int32_t* update_guard() { return g_guard; }
// process-wide, not thread-local: the backward runs on autograd's device thread
static std::atomic<unsigned long long*> g_counters{nullptr};
unsigned long long* debug_counters() { return g_counters.load(std::memory_order_relaxed); }

static int pdl_from_env() {
  const char* e = getenv("B2GS_PDL");
  return (e && e[0] == '0') ? 0 : 1;
}
static std::atomic<int> g_pdl{-1};
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) { v = pdl_from_env(); g_pdl.store(v, std::memory_order_relaxed); }
  return v != 0;
}

void note_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// ---- per-kernel event timing ------------------------------------------------------
struct ProfRec { const char* name; cudaEvent_t a, b; };
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;

ProfScope::ProfScope(const char* name, cudaStream_t stream) : slot(-1), st(stream) {
  if (!g_prof_on) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (g_prof.size() >= 200000) return;
  ProfRec r;
  r.name = name;
  if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
  cudaEventRecord(r.a, stream);
  g_prof.push_back(r);
  slot = (int)g_prof.size() - 1;
}
ProfScope::~ProfScope() {
  if (slot < 0) return;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (slot < (int)g_prof.size()) cudaEventRecord(g_prof[slot].b, st);
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error '%s' (%d) at %s:%d in %s", cudaGetErrorString(e), (int)e, file, line, what);
  return B2GS_ERR_CUDA;
}

static int check_scene(const B2gsScene* s, bool backward) {
  if (!s) { set_error("scene is NULL"); return B2GS_ERR_INVALID; }
  if (s->P < 0 || s->W <= 0 || s->H <= 0) {
    set_error("bad sizes P=%d W=%d H=%d", s->P, s->W, s->H);
    return B2GS_ERR_INVALID;
  }
  if (s->P == 0) return B2GS_OK;
  if (!s->means3D || !s->opacities || !s->viewmatrix || !s->projmatrix || !s->bg) {
    set_error("means3D/opacities/viewmatrix/projmatrix/bg must be non-NULL");
    return B2GS_ERR_INVALID;
  }
  if ((s->shs == nullptr) == (s->colors_precomp == nullptr)) {
    // message of DGR/diff_gaussian_rasterization/__init__.py:209
    set_error("Please provide excatly one of either SHs or precomputed colors!");
    return B2GS_ERR_INVALID;
  }
  const bool has_sr = s->scales != nullptr && s->rotations != nullptr;
  if (has_sr == (s->cov3D_precomp != nullptr) || ((s->scales != nullptr) != (s->rotations != nullptr))) {
    // message of DGR/diff_gaussian_rasterization/__init__.py:212
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return B2GS_ERR_INVALID;
  }
  if (s->shs) {
    if (!s->campos) { set_error("campos must be non-NULL when SHs are used"); return B2GS_ERR_INVALID; }
    if (s->D < 0 || s->D > 3 || (s->D + 1) * (s->D + 1) > s->M) {
      set_error("SH degree %d needs %d coefficients, shs has M=%d", s->D, (s->D + 1) * (s->D + 1), s->M);
      return B2GS_ERR_INVALID;
    }
  }
  if (s->raw_params) {
    if (!s->shs || !s->scales || !s->rotations) {
      set_error("raw_params needs shs (DC part), scales and rotations");
      return B2GS_ERR_INVALID;
    }
    if (s->M > 1 && !s->shs_rest) {
      set_error("raw_params with M=%d needs shs_rest", s->M);
      return B2GS_ERR_INVALID;
    }
  }
  if (backward && !s->projmatrix_raw) {
    set_error("projmatrix_raw must be non-NULL in backward");
    return B2GS_ERR_INVALID;
  }
  return B2GS_OK;
}

}  // namespace b2gs

using namespace b2gs;

extern "C" int b2gs_version(void) { return 100; }
extern "C" const char* b2gs_last_error(void) { return g_err; }
extern "C" void b2gs_set_update_guard(const int32_t* device_flag) { b2gs::g_guard = device_flag; }
extern "C" void b2gs_debug_set_counters(uint64_t* device_counters) {
  b2gs::g_counters.store(reinterpret_cast<unsigned long long*>(device_counters), std::memory_order_relaxed);
}

extern "C" void b2gs_set_pdl(int on) { b2gs::g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }

extern "C" uint64_t b2gs_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" void b2gs_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
}

// Synchronises the device, folds all recorded scopes into per-name totals, writes
// "name count total_ms\n" lines into buf, clears the records.  Returns bytes written.
extern "C" int b2gs_profile_read(char* buf, int cap) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  cudaDeviceSynchronize();
  std::map<std::string, std::pair<long long, double>> acc;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      auto& e = acc[r.name];
      e.first += 1;
      e.second += ms;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof.clear();
  cudaGetLastError();
  int off = 0;
  for (auto& kv : acc) {
    int n = snprintf(buf + off, cap > off ? cap - off : 0, "%s %lld %.6f\n", kv.first.c_str(),
                     kv.second.first, kv.second.second);
    if (n < 0 || off + n >= cap) break;
    off += n;
  }
  return off;
}

extern "C" size_t b2gs_geom_bytes(int32_t P) {
  size_t t = 0;
  carve_geom(nullptr, P, &t);
  return t;
}
extern "C" size_t b2gs_image_bytes(int32_t W, int32_t H) {
  size_t t = 0;
  carve_image(nullptr, W, H, &t);
  return t;
}
extern "C" size_t b2gs_binning_bytes(int64_t R, int32_t W, int32_t H) {
  (void)W; (void)H;
  size_t t = 0;
  carve_bin(nullptr, R, &t);
  return t;
}
extern "C" size_t b2gs_backward_scratch_bytes(int32_t P) {
  const size_t blocks = ((size_t)(P > 0 ? P : 1) + 127) / 128;
  return align_up(((size_t)(P > 0 ? P : 1) * ACC_STRIDE + blocks * 8) * sizeof(float), 256);
}
extern "C" const float* b2gs_backward_tau_partials(const void* scratch, int32_t P, int32_t* blocks) {
  if (blocks) *blocks = (int32_t)(((size_t)(P > 0 ? P : 1) + 127) / 128);
  return reinterpret_cast<const float*>(scratch) + (size_t)(P > 0 ? P : 0) * ACC_STRIDE;
}

// shared front half: preprocess (+ per-tile instance counts) and the tile scan
static int forward_front(const B2gsScene& s, const B2gsWorkspace* ws, const B2gsForwardOut* out,
                         int64_t cap, cudaStream_t st, GeomWS* g_out, ImageWS* im_out) {
  if (!out->radii || !out->n_touched) { set_error("radii/n_touched outputs must be non-NULL"); return B2GS_ERR_INVALID; }
  // Gaussian ids travel in 24 bits of the per-tile sort key (the low 8 carry the block mask)
  if (s.P >= (1 << 24)) { set_error("P too large: at most 16,777,215 Gaussians"); return B2GS_ERR_INVALID; }
  size_t need = 0;
  GeomWS g = carve_geom(ws->geom, s.P, &need);
  if (!ws->geom || ws->geom_bytes < need) {
    set_error("geometry workspace too small: %zu < %zu", ws->geom_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  ImageWS im = carve_image(ws->image, s.W, s.H, &need);
  if (!ws->image || ws->image_bytes < need) {
    set_error("image workspace too small: %zu < %zu", ws->image_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  const size_t tiles = (size_t)((s.W + TILE - 1) / TILE) * ((s.H + TILE - 1) / TILE);
  if (!(cap >= 0 && out->tile_counters_clean))
    B2GS_CUDA(cudaMemsetAsync(im.cursor, 0, sizeof(uint32_t) * tiles * CURSOR_STRIDE, st));
  float* acc_zero = nullptr;
  if (out->bwd_scratch) {
    if (out->bwd_scratch_bytes < b2gs_backward_scratch_bytes(s.P)) {
      set_error("bwd_scratch too small: %zu < %zu", out->bwd_scratch_bytes, b2gs_backward_scratch_bytes(s.P));
      return B2GS_ERR_WORKSPACE;
    }
    acc_zero = reinterpret_cast<float*>(out->bwd_scratch);
  }
  int rc = launch_preprocess_fwd(s, g, im, out->radii, out->n_touched, acc_zero, st);
  if (rc) return rc;
  rc = launch_tile_scan(s, im, cap, cap >= 0 ? out->status : nullptr, st);
  if (rc) return rc;
  *g_out = g;
  *im_out = im;
  return B2GS_OK;
}

static int forward_empty(const B2gsScene& s, const B2gsWorkspace* ws, cudaStream_t st) {
  // P == 0: every tile empty; render still writes the background image
  size_t need = 0;
  ImageWS im = carve_image(ws->image, s.W, s.H, &need);
  if (!ws->image || ws->image_bytes < need) {
    set_error("image workspace too small: %zu < %zu", ws->image_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  const size_t tiles = (size_t)((s.W + TILE - 1) / TILE) * ((s.H + TILE - 1) / TILE);
  B2GS_CUDA(cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * tiles, st));
  B2GS_CUDA(cudaMemsetAsync(im.status, 0, sizeof(uint32_t) * 8, st));
  // tile order of the per-tile kernels: run the scan on all-zero counters (identity-like order)
  B2GS_CUDA(cudaMemsetAsync(im.cursor, 0, sizeof(uint32_t) * tiles * CURSOR_STRIDE, st));
  return launch_tile_scan(s, im, 0, nullptr, st);
}

extern "C" int b2gs_forward_plan(const B2gsScene* scene, const B2gsWorkspace* ws,
                                 const B2gsForwardOut* out, void* stream, int64_t* num_rendered) {
  int rc = check_scene(scene, false);
  if (rc) return rc;
  if (!ws || !out || !num_rendered) { set_error("NULL ws/out/num_rendered"); return B2GS_ERR_INVALID; }
  *num_rendered = 0;
  const B2gsScene& s = *scene;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (s.P == 0) return B2GS_OK;  // nothing to plan; b2gs_forward_render clears the ranges
  GeomWS g;
  ImageWS im;
  rc = forward_front(s, ws, out, -1, st, &g, &im);
  if (rc) return rc;
  uint32_t status[2] = {0, 0};
  B2GS_CUDA(cudaMemcpyAsync(status, im.status, sizeof(status), cudaMemcpyDeviceToHost, st));
  B2GS_CUDA(cudaStreamSynchronize(st));
  if (status[1]) { set_error("too many instances (>= 2^31)"); return B2GS_ERR_INVALID; }
  *num_rendered = (int64_t)status[0];
  return B2GS_OK;
}

static int forward_back(const B2gsScene& s, const B2gsWorkspace* ws, const B2gsForwardOut* out,
                        const GeomWS& g, const ImageWS& im, int64_t cap, cudaStream_t st) {
  if (!out->color || !out->depth || !out->opacity) { set_error("image outputs must be non-NULL"); return B2GS_ERR_INVALID; }
  if (out->fused_loss) {
    const B2gsFusedLoss& f = *out->fused_loss;
    if (f.args.N != s.W * s.H || f.args.mode < 0 || f.args.mode > 3 || !f.args.gt_image || !f.dL_dimage ||
        !f.dL_ddepth || !f.partials || (!(f.args.mode & 2) && !f.args.gt_depth)) {
      set_error("fused_loss: N must be W*H, mode 0..3, gt / gradient / partial pointers non-NULL");
      return B2GS_ERR_INVALID;
    }
  }
  BinWS b = carve_bin(ws->binning, cap, nullptr);
  int rc = launch_binning(s, g, im, b, cap, st);
  if (rc) return rc;
  if (out->event_after_binning)
    B2GS_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(out->event_after_binning), st));
  return launch_render_fwd(s, g, im, b, *out, st);
}

extern "C" int b2gs_forward_render(const B2gsScene* scene, const B2gsWorkspace* ws,
                                   const B2gsForwardOut* out, int64_t R, void* stream) {
  int rc = check_scene(scene, false);
  if (rc) return rc;
  if (!ws || !out) { set_error("NULL ws/out"); return B2GS_ERR_INVALID; }
  const B2gsScene& s = *scene;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (s.P == 0) R = 0;
  if (R < 0 || R >= ((int64_t)1 << 31)) { set_error("bad instance count: %lld", (long long)R); return B2GS_ERR_INVALID; }
  size_t need = 0;
  carve_bin(nullptr, R, &need);
  if (R > 0 && (!ws->binning || ws->binning_bytes < need)) {
    set_error("binning workspace too small: %zu < %zu", ws->binning_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  if (s.P == 0) {
    rc = forward_empty(s, ws, st);
    if (rc) return rc;
  }
  GeomWS g = carve_geom(ws->geom, s.P, nullptr);
  ImageWS im = carve_image(ws->image, s.W, s.H, nullptr);
  return forward_back(s, ws, out, g, im, R, st);
}

extern "C" int64_t b2gs_binning_capacity(size_t binning_bytes) {
  // inverse of b2gs_binning_bytes: point_list (4 B) + keys (8 B) per instance
  int64_t n = (int64_t)(binning_bytes / 12);
  while (n > 0) {
    size_t need = 0;
    carve_bin(nullptr, n, &need);
    if (need <= binning_bytes) break;
    n -= (int64_t)((need - binning_bytes + 11) / 12);
  }
  for (;;) {   // the step above may undershoot by the alignment slack of the three segments
    size_t need = 0;
    carve_bin(nullptr, n + 1, &need);
    if (need > binning_bytes) break;
    ++n;
  }
  return n > 0 ? n : 0;
}

extern "C" int b2gs_forward(const B2gsScene* scene, const B2gsWorkspace* ws,
                            const B2gsForwardOut* out, void* stream) {
  int rc = check_scene(scene, false);
  if (rc) return rc;
  if (!ws || !out) { set_error("NULL ws/out"); return B2GS_ERR_INVALID; }
  const B2gsScene& s = *scene;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t cap = b2gs_binning_capacity(ws->binning ? ws->binning_bytes : 0);
  if (cap <= 0) { set_error("b2gs_forward needs a binning workspace (capacity > 0)"); return B2GS_ERR_WORKSPACE; }
  GeomWS g = carve_geom(ws->geom, s.P, nullptr);
  ImageWS im;
  if (s.P == 0) {
    rc = forward_empty(s, ws, st);
    if (rc) return rc;
    im = carve_image(ws->image, s.W, s.H, nullptr);
  } else {
    rc = forward_front(s, ws, out, cap, st, &g, &im);
    if (rc) return rc;
  }
  if (s.P == 0 && out->status)   // (otherwise the tile scan has written it)
    B2GS_CUDA(cudaMemsetAsync(out->status, 0, 2 * sizeof(uint32_t), st));
  return forward_back(s, ws, out, g, im, cap, st);
}

extern "C" int b2gs_backward(const B2gsScene* scene, const B2gsWorkspace* ws, int64_t R,
                             const B2gsBackwardIO* io, void* stream) {
  int rc = check_scene(scene, true);
  if (rc) return rc;
  if (!ws || !io) { set_error("NULL ws/io"); return B2GS_ERR_INVALID; }
  const B2gsScene& s = *scene;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!io->dL_dtau && !io->defer_tau) { set_error("dL_dtau must be non-NULL"); return B2GS_ERR_INVALID; }
  if (io->iso_reg_weight != 0.f && !(s.raw_params && io->accumulate && io->dL_dscales)) {
    set_error("iso_reg_weight needs raw_params, accumulate and dL_dscales");
    return B2GS_ERR_INVALID;
  }
  if (s.P == 0) {
    if (io->dL_dtau) B2GS_CUDA(cudaMemsetAsync(io->dL_dtau, 0, 6 * sizeof(float), st));
    if (io->defer_tau && io->scratch) B2GS_CUDA(cudaMemsetAsync(io->scratch, 0, io->scratch_bytes, st));
    return B2GS_OK;
  }
  if (!io->dL_dcolor || !io->dL_ddepth || !io->radii) {
    set_error("backward: required pointer is NULL");
    return B2GS_ERR_INVALID;
  }
  if (s.raw_params && io->dL_dsh && s.M > 1 && !io->dL_dsh_rest) {
    set_error("backward: raw_params needs dL_dsh_rest when dL_dsh is requested");
    return B2GS_ERR_INVALID;
  }
  const size_t need = b2gs_backward_scratch_bytes(s.P);
  if (!io->scratch || io->scratch_bytes < need) {
    set_error("backward scratch too small: %zu < %zu", io->scratch_bytes, need);
    return B2GS_ERR_WORKSPACE;
  }
  GeomWS g = carve_geom(ws->geom, s.P, nullptr);
  ImageWS im = carve_image(ws->image, s.W, s.H, nullptr);
  BinWS b = carve_bin(ws->binning, R, nullptr);
  float* acc = reinterpret_cast<float*>(io->scratch);
  if (!io->scratch_zeroed)
    B2GS_CUDA(cudaMemsetAsync(acc, 0, (size_t)s.P * ACC_STRIDE * sizeof(float), st));
  if (R > 0) {
    rc = launch_render_bwd(s, g, im, b, *io, acc, st);
    if (rc) return rc;
  }
  return launch_preprocess_bwd(s, g, *io, acc, st);
}

namespace b2gs {
__global__ void export_geom_kernel(int P, GeomWS g, float* depths, float* means2D, float* conic_o,
                                   float* rgb, uint32_t* touched) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const bool vis = g.tiles_touched[i] != 0;
  if (touched) touched[i] = g.tiles_touched[i];
  float4 cd = vis ? g.rgbd[i] : make_float4(0, 0, 0, 0);
  float4 co = vis ? g.conic_o[i] : make_float4(0, 0, 0, 0);
  float2 p = vis ? g.xy[i] : make_float2(0, 0);
  if (depths) depths[i] = cd.w;
  if (means2D) { means2D[2 * i] = p.x; means2D[2 * i + 1] = p.y; }
  if (conic_o) { conic_o[4 * i] = co.x; conic_o[4 * i + 1] = co.y; conic_o[4 * i + 2] = co.z; conic_o[4 * i + 3] = co.w; }
  if (rgb) { rgb[3 * i] = cd.x; rgb[3 * i + 1] = cd.y; rgb[3 * i + 2] = cd.z; }
}
}  // namespace b2gs

extern "C" int b2gs_debug_export_geom(const B2gsScene* scene, const B2gsWorkspace* ws, float* depths,
                                      float* means2D, float* conic_opacity, float* rgb,
                                      uint32_t* tiles_touched, void* stream) {
  if (!scene || !ws) { set_error("NULL scene/ws"); return B2GS_ERR_INVALID; }
  if (scene->P == 0) return B2GS_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  GeomWS g = carve_geom(ws->geom, scene->P, nullptr);
  export_geom_kernel<<<(scene->P + 255) / 256, 256, 0, st>>>(scene->P, g, depths, means2D,
                                                             conic_opacity, rgb, tiles_touched);
  B2GS_LAUNCH_CHECK(st, 1);
  return B2GS_OK;
}
