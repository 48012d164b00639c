/*
 * b2gs.h — C ABI of libb2gs.so, the B200 (sm_100a) differentiable Gaussian-splatting
 * hot path.  Plain C: raw device pointers, sizes, an opaque stream handle
 * (cudaStream_t passed as void*), int status codes.  No torch types.
 *
 * Every entry point replaces one piece of the reference's native interface
 * (paths relative to the reference tree; DGR = submodules/diff-gaussian-rasterization,
 * CR = DGR/cuda_rasterizer, KNN = submodules/simple-knn, GS = gaussian_splatting):
 *
 *   b2gs_forward_plan + b2gs_forward_render, or b2gs_forward
 *        <- _C.rasterize_gaussians            DGR/rasterize_points.h:17-38, DGR/ext.cpp:16
 *           CudaRasterizer::Rasterizer::forward   CR/rasterizer.h:35-61, CR/rasterizer_impl.cu:198-344
 *   b2gs_backward
 *        <- _C.rasterize_gaussians_backward   DGR/rasterize_points.h:40-64, DGR/ext.cpp:17
 *           CudaRasterizer::Rasterizer::backward  CR/rasterizer.h:63-92, CR/rasterizer_impl.cu:348-455
 *           + the [P,6] -> [6] tau reduction of DGR/diff_gaussian_rasterization/__init__.py:152
 *   b2gs_mark_visible
 *        <- _C.mark_visible                   DGR/rasterize_points.h:66-69, CR/rasterizer_impl.cu:141-153
 *   b2gs_knn3_dist2
 *        <- simple_knn._C.distCUDA2           KNN/spatial.cu:15-26, KNN/simple_knn.cu:185-221
 *   b2gs_adam_step
 *        <- torch.optim.Adam(eps=1e-15).step  GS/scene/gaussian_model.py:283, utils/slam_backend.py:307
 *   b2gs_*_bytes
 *        <- required<T>() / fromChunk         CR/rasterizer_impl.h:21-73, CR/rasterizer_impl.cu:155-194
 *
 * Ownership: the caller allocates every buffer (outputs, gradients, workspaces);
 * the library never allocates or frees device memory and keeps no global state
 * apart from a thread-local last-error string.  All work is enqueued on the
 * given stream.  The only host synchronisation is in b2gs_forward_plan (it must
 * return the instance count so the caller can size the binning workspace — the
 * reference has the same sync at CR/rasterizer_impl.cu:284; b2gs_forward is the
 * sync-free alternative) and after every kernel when args->debug != 0
 * (CR/auxiliary.h:166-173).
 *
 * Matrix convention (identical to the reference): all 4x4 inputs
 * are the transpose of the math matrix stored row-major, i.e. column-major math
 * matrices: element (row r, col c) is m[4*c + r].
 *
 * Absent optional inputs are NULL pointers (reference: empty tensors whose
 * data_ptr() is null, DGR/diff_gaussian_rasterization/__init__.py:214-228).
 */
#ifndef B2GS_H_INCLUDED
#define B2GS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B2GS_API __attribute__((visibility("default")))
#else
#define B2GS_API
#endif

#define B2GS_OK 0
#define B2GS_ERR_INVALID 1   /* bad argument (message in b2gs_last_error) */
#define B2GS_ERR_CUDA 2      /* a CUDA call / kernel failed */
#define B2GS_ERR_WORKSPACE 3 /* a workspace is too small */

#define B2GS_TILE 16         /* tile edge in pixels (CR/config.h:16-17) */

/* Arguments shared by forward and backward: the 13 fields of
 * GaussianRasterizationSettings plus the per-Gaussian inputs. */
typedef struct B2gsScene {
  int32_t P;               /* number of Gaussians */
  int32_t D;               /* active SH degree 0..3 */
  int32_t M;               /* SH coefficients per Gaussian in `shs` (0 when colors_precomp is used) */
  int32_t W, H;            /* image size in pixels */
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int32_t prefiltered;     /* CR/auxiliary.h:155-160: trap if a culled point shows up */
  int32_t debug;           /* sync + check after every launch */
  const float* bg;             /* device [3] */
  const float* viewmatrix;     /* device [16] */
  const float* projmatrix;     /* device [16] */
  const float* projmatrix_raw; /* device [16]; backward only (may be NULL in forward) */
  const float* campos;         /* device [3] */
  const float* means3D;        /* device [P,3] */
  const float* shs;            /* device [P,M,3] or NULL */
  const float* colors_precomp; /* device [P,3] or NULL */
  const float* opacities;      /* device [P] */
  const float* scales;         /* device [P,3] or NULL */
  const float* rotations;      /* device [P,4] (r,x,y,z) or NULL */
  const float* cov3D_precomp;  /* device [P,6] or NULL */
  /* Fused parameter activations (GS/scene/gaussian_model.py:76-101,
   * GS/gaussian_renderer/__init__.py:76-93): when raw_params != 0 the per-Gaussian inputs are the
   * RAW optimiser parameters and the kernels apply the activations themselves —
   *   scales = exp(scales), rotations = rotations / max(|rotations|, 1e-12),
   *   opacities = sigmoid(opacities), SH = cat(shs [P,1,3], shs_rest [P,M-1,3]) —
   * and the backward returns gradients with respect to the raw parameters. */
  int32_t raw_params;
  const float* shs_rest;       /* device [P,M-1,3]; only with raw_params (NULL if M == 1) */
} B2gsScene;

/* Workspaces.  Layout is private to the library; the same three buffers must be
 * handed back, unmodified, to b2gs_backward. */
typedef struct B2gsWorkspace {
  void* geom;   size_t geom_bytes;    /* >= b2gs_geom_bytes(P) */
  void* image;  size_t image_bytes;   /* >= b2gs_image_bytes(W,H) */
  void* binning; size_t binning_bytes;/* >= b2gs_binning_bytes(R,W,H); may be NULL in forward_plan */
} B2gsWorkspace;

struct B2gsFusedLoss;    /* defined with the loss entry points below */

typedef struct B2gsForwardOut {
  float* color;      /* device [3,H,W] */
  float* depth;      /* device [1,H,W] */
  float* opacity;    /* device [1,H,W] */
  int32_t* radii;    /* device [P] */
  int32_t* n_touched;/* device [P]; zero-filled by the library */
  uint32_t* status;  /* device [2], optional (b2gs_forward only): [0] = number of (Gaussian, tile)
                      * instances, [1] = 1 if that exceeded the binning capacity (then the image
                      * is background only and the caller must retry with a larger workspace) */
  /* optional, loop-level fusion; all may be NULL / 0:
   * fused_loss: the mapping / tracking loss of this view and its gradient are computed in the
   *   render kernel's epilogue from the pixel values it holds in registers (no separate pass over
   *   the image, no torch autograd node): see B2gsFusedLoss.
   * bwd_scratch: the backward's scratch buffer; the forward then clears the accumulator rows of
   *   the visible Gaussians itself and b2gs_backward(scratch_zeroed = 1) skips its memset. */
  const struct B2gsFusedLoss* fused_loss;
  void* bwd_scratch; size_t bwd_scratch_bytes;
  /* optional cudaEvent_t recorded on the stream after the binning stage (preprocess, scan, scatter,
   * sort) and before the render kernel: lets a caller that issues several views on several streams
   * start the NEXT view's HBM-bound front end when this view's issue-bound render begins, instead
   * of all front ends first and all renders after (window.py: the views form a software pipeline). */
  void* event_after_binning;
  /* b2gs_forward only: the caller promises that the tile counters inside ws->image are zero — they
   * are after a zero-filled allocation and after every completed b2gs_forward on that workspace
   * (the per-tile sort clears its tile's counter once the scatter is done with it) — so the call
   * skips its own fill of them (one node less at the head of every view's chain). */
  int32_t tile_counters_clean;
} B2gsForwardOut;

typedef struct B2gsBackwardIO {
  const float* dL_dcolor;  /* device [3,H,W] */
  const float* dL_ddepth;  /* device [1,H,W] */
  const int32_t* radii;    /* device [P], from the forward */
  /* outputs: every row is written by the library (zeros for culled Gaussians);
   * pointers marked optional may be NULL to skip that output. */
  float* dL_dmeans3D;   /* [P,3]; may be NULL (pose-only backward) */
  float* dL_dmeans2D;   /* [P,3] optional (z always 0; x,y scaled by W/2, H/2 — CR/backward.cu:643-644) */
  float* dL_dopacity;   /* [P]; may be NULL (pose-only backward) */
  float* dL_dcolors;    /* [P,3] optional: gradient wrt colors_precomp */
  float* dL_dsh;        /* [P,M,3] optional: gradient wrt shs */
  float* dL_dscales;    /* [P,3] optional */
  float* dL_drotations; /* [P,4] optional */
  float* dL_dcov3D;     /* [P,6] optional */
  float* dL_dtau;       /* [6] = (rho(3), theta(3)), already summed over Gaussians */
  void* scratch; size_t scratch_bytes; /* >= b2gs_backward_scratch_bytes(P) */
  /* accumulate != 0: per-Gaussian gradient outputs are ADDED to (rows of culled Gaussians left
   * untouched) instead of written — lets every view of a mapping window sum straight into the
   * optimiser's gradient arena (replaces autograd's AccumulateGrad adds and the zero fills). */
  int32_t accumulate;
  float* dL_dsh_rest;          /* [P,M-1,3]; with raw_params: dL_dsh is the [P,1,3] DC part */
  /* optional densification statistics of this view, fused (GS/scene/gaussian_model.py:691-695,
   * utils/slam_backend.py:277-284): for radii > 0:  xyz_gradient_accum += |dL_dmeans2D.xy|,
   * denom += 1, max_radii2D = max(max_radii2D, radii).  Any may be NULL. */
  float* densify_grad_accum;   /* [P] */
  float* densify_denom;        /* [P] */
  float* max_radii2D;          /* [P] */
  int32_t scratch_zeroed;      /* the forward was given this scratch as out->bwd_scratch */
  /* defer_tau != 0: dL_dtau may be NULL; the per-block partial rows (8 floats: tau[6], the
   * regulariser partial, pad) stay in the scratch for b2gs_window_finalize
   * (b2gs_backward_tau_partials tells where). */
  int32_t defer_tau;
  /* iso_reg_weight != 0 (needs raw_params + accumulate): this call also adds the gradient of the
   * isotropic-scale regulariser  weight * mean|exp(s) - mean_row exp(s)|  (utils/slam_backend.py:
   * 229-231) to dL_dscales for EVERY Gaussian, visible or not, and leaves per-block partial sums of
   * its value in slot 6 of the tau partial rows.  Give it to exactly one view of a step — or, with
   * data-parallel ranks, to one view per rank with iso_reg_stride = world and iso_reg_offset = rank:
   * the call then covers only the Gaussians i with i % stride == offset (stride 0 means all). */
  float iso_reg_weight;
  int32_t iso_reg_stride, iso_reg_offset;
  /* optional device words read when the call RUNS (graph replays included): the densification
   * statistics above are left untouched if any non-NULL word is non-zero.  The sync-free mapping
   * window points them at: the view's own overflow flag (B2gsForwardOut.status + 1: a
   * background-only pass must not count as "seen"), the update guard (still holding the
   * previous step's verdict: the replays after a discarded step repeat it), and a per-view host
   * mask set for the re-run of a discarded step (the views that did fit were counted already).
   * Keeps xyz_gradient_accum / denom exact under GS/scene/gaussian_model.py:691-695 when a
   * step is discarded and repeated. */
  const int32_t* stats_gate[3];
} B2gsBackwardIO;

B2GS_API int b2gs_version(void);
B2GS_API const char* b2gs_last_error(void);
/* number of the library's kernels launched so far in this process (every kernel in the library is
 * hand-written: no CUB / Thrust / cuBLAS call anywhere). */
B2GS_API uint64_t b2gs_kernel_launches(void);

/* Per-kernel timing: while enabled, every kernel of the library is
 * bracketed by CUDA events on its launching stream.  b2gs_profile_read synchronises,
 * writes "name count total_ms" lines into buf and clears the records. */
B2GS_API void b2gs_profile_enable(int on);
B2GS_API int b2gs_profile_read(char* buf, int cap);

/* Programmatic dependent launch between the consecutive kernels of a call / a step (default on;
 * environment B2GS_PDL=0 or b2gs_set_pdl(0) turns it off): hides the launch latency between
 * dependent kernels, in eager streams and as programmatic edges of a captured CUDA graph. */
B2GS_API void b2gs_set_pdl(int on);

B2GS_API size_t b2gs_geom_bytes(int32_t P);
B2GS_API size_t b2gs_image_bytes(int32_t W, int32_t H);
B2GS_API size_t b2gs_binning_bytes(int64_t R, int32_t W, int32_t H);
B2GS_API size_t b2gs_backward_scratch_bytes(int32_t P);
/* where b2gs_backward(defer_tau) leaves its per-block partial rows inside `scratch`, and how many */
B2GS_API const float* b2gs_backward_tau_partials(const void* scratch, int32_t P, int32_t* blocks);

/* Two-call form (exact sizing, like the reference: one small D2H read + sync).
 * Phase 1: per-Gaussian preprocessing, per-tile instance counts, tile ranges.
 * Writes out->radii, zero-fills out->n_touched, returns the number of
 * (Gaussian,tile) instances in *num_rendered (synchronises the stream). */
B2GS_API int b2gs_forward_plan(const B2gsScene* scene, const B2gsWorkspace* ws,
                      const B2gsForwardOut* out, void* stream, int64_t* num_rendered);

/* Phase 2: instance scatter, per-tile depth sort, alpha blending.
 * ws->binning must hold b2gs_binning_bytes(num_rendered, W, H). */
B2GS_API int b2gs_forward_render(const B2gsScene* scene, const B2gsWorkspace* ws,
                        const B2gsForwardOut* out, int64_t num_rendered, void* stream);

/* One-call form WITHOUT any host synchronisation (CUDA-graph capturable): the caller
 * supplies a binning workspace of any size; its capacity in instances is
 * b2gs_binning_capacity(ws->binning_bytes).  If the scene needs more, out->status[1] is
 * set, the images are background only, and the caller retries with a larger workspace
 * (status[0] holds the required count).  Replaces the reference's sync at
 * CR/rasterizer_impl.cu:284. */
B2GS_API int64_t b2gs_binning_capacity(size_t binning_bytes);
B2GS_API int b2gs_forward(const B2gsScene* scene, const B2gsWorkspace* ws,
                 const B2gsForwardOut* out, void* stream);

/* binning_capacity: num_rendered as given to b2gs_forward_render, or
 * b2gs_binning_capacity(ws->binning_bytes) after b2gs_forward. */
B2GS_API int b2gs_backward(const B2gsScene* scene, const B2gsWorkspace* ws, int64_t binning_capacity,
                  const B2gsBackwardIO* io, void* stream);

/* present[i] = 1 if view-space z > 0.2 (CR/auxiliary.h:139-164). */
B2GS_API int b2gs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present, void* stream);

/* Parity/debug helper: materialise the reference's sorted 64-bit key list
 * (tile<<32 | depth bits), the sorted Gaussian ids and the tile ranges from the
 * library's private workspaces.  keys[R], values[R], ranges[2*tiles]. */
B2GS_API int b2gs_debug_export_binning(const B2gsScene* scene, const B2gsWorkspace* ws, int64_t num_rendered,
                              uint64_t* keys, uint32_t* values, uint32_t* ranges, void* stream);
/* Parity/debug helper: copy per-Gaussian forward state out of the geometry workspace.
 * Any pointer may be NULL.  depths[P], means2D[P,2], conic_opacity[P,4], rgb[P,3], tiles_touched[P]. */
B2GS_API int b2gs_debug_export_geom(const B2gsScene* scene, const B2gsWorkspace* ws, float* depths,
                           float* means2D, float* conic_opacity, float* rgb,
                           uint32_t* tiles_touched, void* stream);
/* Parity/debug helper: final_T[N], n_contrib[N] from the image workspace. */
B2GS_API int b2gs_debug_export_image(const B2gsScene* scene, const B2gsWorkspace* ws, float* final_T,
                            uint32_t* n_contrib, void* stream);

/* Parity/debug helper: the render kernels evaluate exp() of a lane's two pixels with one packed
 * instruction sequence (csrc/common.cuh: expf_pair) that must return expf()'s bits, since the
 * reference's skip decisions (CR/forward.cu:346-362) hang on them.  Evaluates both for EVERY float
 * whose bit pattern lies in [first_bits, last_bits] and writes the number of patterns whose results
 * differ (either half of the pair) to mismatches[0] (device, uint64). */
B2GS_API int b2gs_debug_expf_pair_check(uint32_t first_bits, uint32_t last_bits,
                                        unsigned long long* mismatches, void* stream);

/* mean squared distance to the 3 nearest neighbours (exact). points [P,3] -> out [P]. */
B2GS_API size_t b2gs_knn_bytes(int32_t P);
B2GS_API int b2gs_knn3_dist2(const float* points, int32_t P, float* out, void* workspace,
                    size_t workspace_bytes, void* stream);

/* b2gs_knn3_dist2 WITHOUT a caller workspace and without any allocation: the scratch is a static
 * device array inside the library (9 MiB, held by every device that loads the library until the
 * process exits), enough for b2gs_knn3_dist2_static_max_points() = 262144 points; more points are
 * rejected with B2GS_ERR_INVALID.  For callers that cannot call cudaMalloc (under CUDA-graph
 * capture, serving loops, harnesses that forbid it).  No cudaMalloc / cudaFree / allocating
 * library call is made.  All calls on a device share the one array: calls that overlap on the GPU
 * (different streams) race; calls on one stream are safe.  Same result as b2gs_knn3_dist2.
 * No counterpart in the reference. */
B2GS_API int32_t b2gs_knn3_dist2_static_max_points(void);
B2GS_API int b2gs_knn3_dist2_static(const float* points, int32_t P, float* out, void* stream);

/* Fused Adam over a set of parameter segments (torch.optim.Adam semantics:
 * no weight decay, no amsgrad, bias correction from `step`, 1-based).
 * Segment s covers n[s] floats at param[s]/grad[s]/exp_avg[s]/exp_avg_sq[s]
 * with learning rate lr[s].  All pointer arrays are HOST arrays of device pointers. */
typedef struct B2gsAdamSegment {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t n; float lr;
  const float* lr_dev; /* optional DEVICE scalar that overrides lr (b2gs_adam_step_dev only):
                        * lets a captured graph follow a learning-rate schedule */
} B2gsAdamSegment;
B2GS_API int b2gs_adam_step(const B2gsAdamSegment* segments, int32_t num_segments, int64_t step,
                   float beta1, float beta2, float eps, void* stream);
/* Same, CUDA-graph capturable: the step count is a DEVICE int64 that the call increments
 * (on the stream) before using it, so a captured graph stays correct across replays. */
B2GS_API int b2gs_adam_step_dev(const B2gsAdamSegment* segments, int32_t num_segments,
                       int64_t* step_counter, float beta1, float beta2, float eps, void* stream);

/* Sharded optimiser step over NVLink peer memory — replaces, for data-parallel ranks that hold
 * replicas of the parameter arena,  dist.all_reduce(grad) + Adam.step() + zero_grad()
 * (the reference has no multi-GPU path).  Every rank calls it with the same
 * segments (pointers into ITS arenas).  Chunks of 1024 floats are owned round-robin by rank; the
 * owner sums that chunk of every rank's gradients in rank order, clears them, updates its own
 * Adam moments (moments of chunks a rank does not own stay untouched) and writes the new
 * parameters into every replica.  Two inter-rank flag barriers bracket the kernel on the stream.
 * All arrays are HOST arrays of DEVICE pointers valid in this process (peer mappings). */
typedef struct B2gsPeerArena {
  int32_t world, rank;
  float* const* param_base;   /* [world] base of every rank's parameter arena */
  float* const* grad_base;    /* [world] base of every rank's gradient arena */
  uint32_t* const* signal;    /* [world] every rank's signal pad: >= 32 zero-initialised uint32 */
  uint32_t* fail_flag;        /* local device word, set to 1 if a barrier timed out (about 4 s) */
  /* optional NVLink-switch multicast mappings of the SAME two arenas (CUDA multicast objects bound on
   * every rank, e.g. torch symmetric memory's handle.multicast_ptr): when both are non-NULL the
   * owner of a chunk reads the gradient sum with one multimem.ld_reduce (the switch adds the ranks'
   * copies) and broadcasts the new parameters / the cleared gradients with multimem.st — per-rank
   * link traffic drops from 3 (world-1)/world to 3/world of the arena.  NULL: plain peer loads/stores. */
  float* param_mc;
  float* grad_mc;
} B2gsPeerArena;
B2GS_API int b2gs_adam_step_peer(const B2gsAdamSegment* segments, int32_t num_segments,
                        int64_t* step_counter, float beta1, float beta2, float eps,
                        const B2gsPeerArena* peers, void* stream);

/* ---- the scalar tail of a window / tracking step (csrc/window.cu) ----------------------------
 * One entry per local view: where its fused-loss partials (B2gsFusedLoss.partials, `loss_rows` =
 * tiles) and its backward partial rows (b2gs_backward_tau_partials) are, and where the results go. */
typedef struct B2gsViewFinalize {
  const float* loss_partials; int32_t loss_rows;   /* NULL: this view has no loss term */
  int32_t N; int32_t mode; float alpha;            /* as in B2gsLossArgs */
  const float* exposure_a;                         /* device scalar or NULL */
  const float* tau_partials; int32_t tau_rows;     /* NULL: no pose gradient wanted */
  float reg_scale;    /* iso_reg_weight / (3P) if this view's backward carried the regulariser, else 0 */
  const uint32_t* status;   /* [2] of this view's b2gs_forward, or NULL */
  float* loss;              /* [1] out: loss of the view (+ the regulariser value if reg_scale != 0) */
  float* dL_dtau;           /* [6] out (rho, theta), may be NULL */
  float* dL_dexposure;      /* [2] out (a, b), may be NULL */
} B2gsViewFinalize;
/* loss_total[0] = sum of the views' losses (view order), guard[0] = OR of the views' overflow flags
 * (either may be NULL).  `ticket`: FOUR zero-initialised device words: [0] is the launch's ticket
 * (left zero), [2] accumulates the OR of every launch's overflow flag until the caller clears it (a
 * tracking loop replays many iterations and looks once per frame).  <= 32 views. */
B2GS_API int b2gs_window_finalize(const B2gsViewFinalize* views, int32_t n_views, float* loss_total,
                         int32_t* guard, uint32_t* ticket, void* stream);

/* Adam (torch.optim.Adam semantics, one shared 1-based step count) over the camera parameters of
 * every local keyframe + update_pose + zeroed deltas, ONE launch; honours the update guard. */
typedef struct B2gsPoseCam {
  float* cam_state; const float* projmatrix_raw;   /* as in b2gs_pose_update */
  float* rot_delta; float* trans_delta; float* exposure_a; float* exposure_b;
  const float* dL_dtau;        /* [6] (rho, theta) */
  const float* dL_dexposure;   /* [2] */
  float* moments;              /* [16] exp_avg[8], exp_avg_sq[8] in the order rot, trans, a, b */
  float lr_rot, lr_trans, lr_exposure_a, lr_exposure_b;
  int32_t optimize_pose, optimize_exposure;
  float* tau_norm;             /* optional [1]: |(rho, theta)| of this update (update_pose's convergence test) */
} B2gsPoseCam;
/* step_counter: device int64, incremented by the call (after use: the kernels read count + 1);
 * extra_tick: optional second device counter incremented alongside (the Gaussian optimizer's, when
 * b2gs_adam_step_fused ran with B2GS_ADAM_NO_TICK).  `ticket` as above. */
B2GS_API int b2gs_pose_step(const B2gsPoseCam* cams, int32_t n_cams, int64_t* step_counter,
                   int64_t* extra_tick, float beta1, float beta2, float eps, uint32_t* ticket,
                   void* stream);

/* b2gs_adam_step_dev / b2gs_adam_step_peer with options:
 *   B2GS_ADAM_ZERO_GRAD  clear each gradient element after reading it (optimizer.zero_grad fused);
 *   B2GS_ADAM_NO_TICK    use step_counter + 1 and leave the counter alone (the caller ticks it after
 *                        the update, e.g. b2gs_pose_step(extra_tick)) — saves the 1-thread tick launch. */
#define B2GS_ADAM_ZERO_GRAD 1
#define B2GS_ADAM_NO_TICK 2
/* b2gs_adam_step_peer_fused only: */
#define B2GS_ADAM_TICK_LATE 4         /* with NO_TICK: the SECOND barrier advances the counter (nothing else
                                       * on the stream has to: the camera step can run beside the exchange) */
#define B2GS_ADAM_NO_ZERO 8           /* leave this rank's gradient arena as it is: the caller clears it
                                       * (b2gs_memset_async) off the critical path, before the next backward */
#define B2GS_ADAM_NO_FIRST_BARRIER 16 /* the caller has already run b2gs_peer_barrier on this stream */
B2GS_API int b2gs_adam_step_fused(const B2gsAdamSegment* segments, int32_t num_segments,
                         int64_t* step_counter, float beta1, float beta2, float eps, int32_t flags,
                         void* stream);
/* The first of the two barriers of b2gs_adam_step_peer on its own: all ranks' streams meet, and the
 * update guard of the calling thread (b2gs_set_update_guard) becomes the OR over all ranks. */
B2GS_API int b2gs_peer_barrier(const B2gsPeerArena* peers, void* stream);
/* The same, but this rank's contribution to the guard is the OR of status[2 i + 1], i < n_views — the
 * overflow flags its forward passes left in B2gsForwardOut.status ((instances, overflow) pairs, one
 * per view, contiguous) — instead of the guard word's current value: the barrier then depends on the
 * views' backward passes only, and b2gs_window_finalize (called with guard = NULL) can run beside it.
 * No counterpart in the reference (its loop is single-GPU). */
B2GS_API int b2gs_peer_barrier_status(const B2gsPeerArena* peers, const uint32_t* status, int32_t n_views,
                             void* stream);
/* cudaMemsetAsync through the C ABI (a step built from libb2gs calls only needs no framework kernel). */
B2GS_API int b2gs_memset_async(void* device_ptr, int value, size_t bytes, void* stream);
B2GS_API int b2gs_adam_step_peer_fused(const B2gsAdamSegment* segments, int32_t num_segments,
                              int64_t* step_counter, float beta1, float beta2, float eps,
                              const B2gsPeerArena* peers, int32_t flags, void* stream);

/* ---- densify_and_prune on the packed arenas (csrc/densify.cu) --------------------------------
 * Replaces GaussianModel.densify_and_prune and everything under it (GS/scene/gaussian_model.py:
 * 468-555, 593-689) with plan -> (one host read of the totals) -> apply.  Groups: the parameter
 * tensors of the model, `width[g]` floats per Gaussian, with their Adam moments.  The new arenas
 * receive [kept originals | kept clones | kept children (copy-major)], the reference's final order;
 * new rows get zero moments, kept rows keep theirs. */
typedef struct B2gsDensifyArgs {
  int32_t P, n_groups;                       /* <= 8 groups */
  const float* param[8]; const float* exp_avg[8]; const float* exp_avg_sq[8];
  int32_t width[8];
  int32_t xyz, opacity, scaling, rotation;   /* which groups they are (widths 3 / 1 / 3 / 4) */
  const float* grad_accum; const float* denom;   /* [P] densification statistics */
  float max_grad, min_opacity;
  float dense_threshold;                     /* percent_dense * scene_extent, rounded once by the caller */
  float world_size_threshold;                /* 0.1 * extent */
  int32_t max_screen_size_on;                /* the reference's `if max_screen_size:` */
  int32_t N; float split_divisor;            /* children per split parent, 0.8 * N */
  const int32_t* aux[2];                     /* optional int32 [P] arrays that ride along (unique_kfIDs, n_obs) */
} B2gsDensifyArgs;
B2GS_API size_t b2gs_densify_bytes(int32_t P);
/* totals (device, 8 x uint64): [0] kept originals, [1] kept clones, [2] kept split parents (each has
 * N kept children), [3] clones selected, [4] split parents selected.  New size = [0] + [1] + N * [2]. */
B2GS_API int b2gs_densify_plan(const B2gsDensifyArgs* args, void* workspace, size_t workspace_bytes,
                      uint64_t* totals, void* stream);
/* totals_host: the five counts read back.  samples: device [N * totals[4], 3] unit normals, row
 * k * totals[4] + r belongs to copy k of the r-th split parent (torch.normal's draw order,
 * gaussian_model.py:606-611).  new_*: HOST arrays of n_groups device pointers (the new arenas),
 * new_aux: 2 device pointers or NULL. */
B2GS_API int b2gs_densify_apply(const B2gsDensifyArgs* args, const void* workspace, const int64_t* totals_host,
                       const float* samples, float* const* new_param, float* const* new_exp_avg,
                       float* const* new_exp_avg_sq, int32_t* const* new_aux, void* stream);

/* RGB-D keyframe back-projection (the seed point cloud of new Gaussians; feeds b2gs_knn3_dist2).
 * Replaces GaussianModel.create_pcd_from_image_and_depth's host round trip through open3d 0.17
 * (GS/scene/gaussian_model.py:107-156: RGBDImage.create_from_color_and_depth(depth_scale=1,
 * depth_trunc, convert_rgb_to_intensity=False) + PointCloud.create_from_rgbd_image(intrinsic,
 * extrinsic, project_valid_depth_only=True)).
 *   image [3,H,W] float32 (cam.original_image), exposure_a/b device scalars or NULL,
 *   depth [H,W] float32, cam2world: HOST, 12 doubles = first three rows of inverse(extrinsic).
 * Dense outputs in pixel (row-major) order: xyz [H*W,3], rgb [H*W,3] (uint8-quantised / 255),
 * valid [H*W] (depth > 0 after truncation); rows of invalid pixels hold zeros. */
B2GS_API int b2gs_backproject_rgbd(int32_t W, int32_t H, const float* image, const float* exposure_a,
                          const float* exposure_b, const float* depth, double fx, double fy,
                          double cx, double cy, const double* cam2world, float depth_trunc,
                          float* xyz, float* rgb, uint8_t* valid, void* stream);

/* Update guard: while `device_flag` (a device int32, or NULL to clear) is set for the calling host
 * thread, the state-updating launches enqueued by b2gs_adam_step_dev / b2gs_adam_step_peer /
 * b2gs_pose_update read it ON THE DEVICE and do nothing when it is non-zero.  A CUDA-graph step
 * can thus discard its own update after a sync-free forward overflowed (b2gs_forward status) with
 * no host round trip inside the step.  b2gs_adam_step_peer additionally ORs the flag over all
 * ranks (the result is left in the flag), so data-parallel replicas agree. */
B2GS_API void b2gs_set_update_guard(const int32_t* device_flag);

/* Debug work counters (measurement only, bench.py roofline): while `device_counters` (8 device
 * uint64, or NULL to clear) is set (process-wide), the render kernels launch a counting
 * instantiation and ADD to it:  [0] forward warp visits (a warp = 64 pixels looking at one splat of
 * its tile's list), [1] forward visits that evaluate exp + blend, [2] forward contributing
 * pixel-splat pairs (alpha >= 1/255, not terminated: the pairs CR/forward.cu:359-371 blends),
 * [4] backward warp visits, [5] backward visits that run the gradient arithmetic, [6] backward
 * contributing pairs (CR/backward.cu:700-757), [7] kNN search: point-to-point distance evaluations. */
B2GS_API void b2gs_debug_set_counters(uint64_t* device_counters);

/* ---- fused per-pixel losses (utils/slam_utils.py:56-128) ---------------------------------
 * mode: 0 mapping rgbd (get_loss_mapping_rgbd :113-128), 1 tracking rgbd (:74-88),
 *       2 mapping rgb / monocular (:101-110), 3 tracking rgb / monocular (:64-72).
 * image_ab = exp(exposure_a) * image + exposure_b (:57,95); NULL exposure pointers mean the
 * `initialization` branch (image_ab = image, :92-93).  All pointers are device pointers. */
typedef struct B2gsLossArgs {
  int32_t N;                      /* pixels (H*W) */
  int32_t mode;
  float alpha;                    /* config Training.alpha, default 0.95 */
  float rgb_boundary_threshold;   /* config Training.rgb_boundary_threshold */
  const float* image;             /* [3,N] rendered colour */
  const float* depth;             /* [N] rendered depth (rgbd modes) */
  const float* opacity;           /* [N] rendered opacity (tracking modes) */
  const float* gt_image;          /* [3,N] */
  const float* gt_depth;          /* [N] (rgbd modes) */
  const float* exposure_a;        /* [1] or NULL */
  const float* exposure_b;        /* [1] or NULL */
  const float* grad_mask;         /* [N] or NULL: viewpoint.grad_mask (tracking modes) */
} B2gsLossArgs;
/* The same loss fused into the forward render (B2gsForwardOut.fused_loss), upstream gradient 1:
 * args.image / depth / opacity are ignored (the kernel uses the pixel it has just blended).
 * Writes dL/dimage, dL/ddepth for the backward and, per 16x16 tile, the four partial sums
 * (sum |rgb term|, sum |depth term|, dL/d exposure_a / exp(a), dL/d exposure_b) that
 * b2gs_window_finalize turns into the loss value and the exposure gradients. */
typedef struct B2gsFusedLoss {
  B2gsLossArgs args;
  float* dL_dimage;   /* [3,N] */
  float* dL_ddepth;   /* [N] */
  float* partials;    /* [tiles,4] */
} B2gsFusedLoss;
B2GS_API size_t b2gs_loss_scratch_bytes(int32_t N);
B2GS_API int b2gs_loss_forward(const B2gsLossArgs* args, float* loss, void* scratch,
                      size_t scratch_bytes, void* stream);
/* dL_dloss: device scalar or NULL (= 1).  dL_dopacity optional.  dL_dexposure[2] = (a, b). */
B2GS_API int b2gs_loss_backward(const B2gsLossArgs* args, const float* dL_dloss, float* dL_dimage,
                       float* dL_ddepth, float* dL_dopacity, float* dL_dexposure, void* scratch,
                       size_t scratch_bytes, void* stream);

/* Isotropic-scale regulariser of the mapping loss (utils/slam_backend.py:229-231):
 * loss[0] = weight * mean(|exp(raw) - mean_row(exp(raw))|) over the [P,3] log-scales, and its
 * gradient wrt the RAW log-scales ADDED to dL_draw_scaling (may be NULL: value only). */
B2GS_API int b2gs_scale_regulariser(int32_t P, const float* raw_scaling, float weight,
                           float* dL_draw_scaling, float* loss, void* scratch, size_t scratch_bytes,
                           void* stream);

/* update_pose (utils/pose_utils.py:76-93) on the device: T_w2c <- Exp((trans_delta, rot_delta))
 * T_w2c; refreshes viewmatrix, full projection and camera centre in place, zeroes the deltas.
 * cam_state: float[64] = R[9] row-major | T[3] | pad[4] | viewmatrix[16] | projmatrix[16] |
 * campos[3] | pad.  tau_norm (optional, [1]) receives |tau| for the caller's convergence test. */
B2GS_API int b2gs_pose_update(float* cam_state, const float* projmatrix_raw, float* rot_delta,
                     float* trans_delta, float* tau_norm, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2GS_H_INCLUDED */
