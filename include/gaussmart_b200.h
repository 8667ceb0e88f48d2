/*
 * gaussmart_b200 -- C ABI of the B200-native 2D Gaussian surfel rasterizer and kNN scale
 * initialiser.  This is the drop-in boundary: exactly what the reference's pybind module
 * (submodules/diff-surfel-rasterization/ext.cpp:15-19, submodules/simple-knn/ext.cpp:15-17)
 * binds, restated with plain pointers and sizes.  No torch types cross this boundary.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - every entry point is asynchronous on `stream` (a cudaStream_t passed as void*) except
 *     where noted; none of them allocates device memory -- the caller owns all buffers;
 *   - return value: 0 on success, a cudaError_t (1..9999) on a CUDA failure, or a GSM_ERR_*
 *     code; gsm_last_error() returns a human-readable description for the calling thread;
 *   - nothing throws across the boundary.
 */
#ifndef GAUSSMART_B200_H
#define GAUSSMART_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSM_ABI_VERSION 9

#define GSM_OK 0
#define GSM_ERR_INVALID_ARGUMENT 10001
#define GSM_ERR_BINNING_OVERFLOW 10002 /* binning_capacity < num_rendered; *num_rendered_host is valid */
#define GSM_ERR_WORKSPACE_TOO_SMALL 10003

/* Bytes of one packed per-surfel record written by the forward preprocess and read by the
 * compositing kernels (5 x float4: Tu|x, Tv|y, Tw|opacity, normal|depth, rgb|clamp-bits). */
#define GSM_SURFEL_RECORD_BYTES 80
/* Floats of one per-surfel gradient accumulation record written by the backward compositing
 * kernel (dTu3 dTv3 dTw3 dmean2D2 dcolor3 dnormal3 dopacity1 = 18, padded to 20). */
#define GSM_GRAD_RECORD_FLOATS 20
/* Ranks of one NVLink / NVSwitch domain the view-parallel exchange can address directly. */
#define GSM_MAX_PEERS 8

const char* gsm_last_error(void);
int gsm_abi_version(void);

/* ---- optional per-stage device timing (bench / profiling aid; off by default) -----------
 * When enabled, every stage of gsm_forward / gsm_backward / gsm_knn_dist2 is bracketed by a
 * cudaEvent pair on the launching stream.  gsm_stage_timing_read synchronises the outstanding
 * events and returns, per stage, the summed device milliseconds and the number of launches.
 * The reference has no equivalent (its only hook is debug's sync-after-stage, auxiliary.h:296-303). */
#define GSM_NUM_STAGES 14
int gsm_stage_count(void);
const char* gsm_stage_name(int stage);
int gsm_stage_timing_enable(int on); /* returns the previous setting */
int gsm_stage_timing_read(double* sum_ms /*[GSM_NUM_STAGES]*/, long long* calls /*[GSM_NUM_STAGES]*/, int reset);

/* Bench aid: the SIMT FP32 FMA rate this GPU sustains (the denominator of the compositing kernels' roofline): launches
 * 4 CTAs x 512 threads per SM, each thread running `iters` rounds of 8 independent register-resident FMAs; time it with
 * CUDA events.  out: device scratch of >= 4 * SMs * 512 floats; *flops_out (host) receives the flops of the launch. */
int gsm_fp32_peak_probe(int iters, float* out, size_t out_floats, double* flops_out, void* stream);

/* ---- workspace sizing (host-only arithmetic, no CUDA calls) ---------------------------- */

/* geometry workspace (must survive until backward): records + tile counts/offsets + scan temp.
 * Replaces GeometryState::fromChunk / required<GeometryState> (rasterizer_impl.cu:155-170). */
size_t gsm_geometry_bytes(int P);
/* image workspace (must survive until backward): final_T[3,H,W], n_contrib[2,H,W], ranges[Nt].
 * Replaces ImageState::fromChunk (rasterizer_impl.cu:172-179). */
size_t gsm_image_bytes(int width, int height);
/* binning workspace for `capacity` (surfel,tile) instances.  Its first
 * gsm_binning_persistent_bytes(capacity) bytes must survive until backward -- a copy of that prefix is as good
 * as the original: the depth-sorted surfel list u32[capacity], the per-instance block masks u8[capacity]
 * (written by the tile sort, REFINED by the forward compositing: bits of (block, entry) pairs without a blended
 * pixel are cleared) and the per-entry pixel sets u32[8][capacity] (which pixels of block b blended list entry i,
 * written by the forward compositing; gsm_backward walks them).  The rest (keys, sort scratch) is transient.
 * Replaces BinningState::fromChunk (rasterizer_impl.cu:181-194). */
size_t gsm_binning_bytes(size_t capacity, int width, int height);
size_t gsm_binning_persistent_bytes(size_t capacity);
/* byte offsets of the sorted keys (u64[R]) inside the binning workspace, for parity tests. */
size_t gsm_binning_sorted_keys_offset(size_t capacity);
/* byte offset of ranges (uint2[Nt]) / n_contrib (u32[2,H,W]) inside the image workspace. */
size_t gsm_image_ranges_offset(int width, int height);
size_t gsm_image_ncontrib_offset(int width, int height);

/* ---- forward: replaces CudaRasterizer::Rasterizer::forward (rasterizer.h:33-58,
 *      rasterizer_impl.cu:198-345) as called by RasterizeGaussiansCUDA (rasterize_points.cu:39-134)
 */
typedef struct gsm_forward_args {
  int P;             /* surfels */
  int D;             /* active SH degree 0..3 */
  int M;             /* allocated SH coefficients per surfel (0 if shs == NULL) */
  int width, height;
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int prefiltered;   /* kept for API parity; a culled surfel with prefiltered set is an error flag */
  int debug;         /* bit 0: synchronise + check after every stage (auxiliary.h:296-303);
                        bit 1: disable the conservative per-warp culling in the compositing kernels
                        (every (pixel, list entry) pair is evaluated, as the reference does);
                        bit 2: inference render (render.py, torch.no_grad): no backward will follow, so the
                        compositing kernel skips what it would hand to gsm_backward; calling gsm_backward on
                        the buffers of such a forward is invalid (undefined gradients);
                        bit 3: test aid: take the binning path of P > 2^24 surfels (block masks gathered by the tile
                        sort instead of travelling in the top byte of the pair values) at any P */
  const float* background;       /* [3] */
  const float* means3D;          /* [P,3] */
  const float* shs;              /* [P,M,3] or NULL */
  const float* colors_precomp;   /* [P,3] or NULL */
  const float* opacities;        /* [P] */
  const float* scales;           /* [P,2] or NULL (then transMat_precomp) */
  const float* rotations;        /* [P,4] (w,x,y,z) or NULL */
  const float* transMat_precomp; /* [P,9] rows Tu,Tv,Tw or NULL */
  const float* viewmatrix;       /* [16] row-major, row-vector convention */
  const float* projmatrix;       /* [16] */
  const float* campos;           /* [3] */
  float* out_color;              /* [3,H,W] */
  float* out_others;             /* [7,H,W]: depth, alpha, normal xyz, median depth, distortion */
  int* radii;                    /* [P] */
  void* geometry; size_t geometry_bytes;
  void* image;    size_t image_bytes;
  void* binning;  size_t binning_bytes; size_t binning_capacity;
  int* num_rendered_host;        /* HOST pointer: receives R (valid on GSM_OK and on overflow) */
  uint64_t* work_counters;       /* optional DEVICE u64[4], caller-zeroed, ADDED to: [0] (pixel,entry)
                                    pairs the reference's loop would evaluate, [1] contributing pairs,
                                    [2] records staged, [3] pairs evaluated after culling.
                                    NULL in production; used by bench.py for the roofline model. */
  /* raw-parameter entry, both optional, zero / NULL for the reference's behaviour:
     the getters of scene/gaussian_model.py:104-123 fused into the preprocess kernels */
  const float* shs_rest;         /* when given: shs = _features_dc [P,1,3] and shs_rest = _features_rest [P,M-1,3]
                                    (get_features without the torch.cat copy, gaussian_model.py:115-118) */
  int raw_params;                /* 1: scales = _scaling, rotations = _rotation, opacities = _opacity as STORED;
                                    exp / F.normalize / sigmoid (gaussian_model.py:104-113) are applied in-kernel,
                                    evaluated as torch's CUDA kernels evaluate them */
  void* status;                  /* optional gsm_status handle (gsm_status_create): selects the ASYNCHRONOUS forward */
} gsm_forward_args;

/* SYNCHRONOUS mode (status == NULL; what the reference does, rasterizer_impl.cu:282): runs preprocess + scan, waits
 * for R (one stream synchronisation), then binning, sort and compositing.  On GSM_ERR_BINNING_OVERFLOW nothing
 * after the scan has run; grow the binning workspace to *num_rendered_host and call again.
 *
 * ASYNCHRONOUS mode (status != NULL): the whole forward is enqueued and the call returns WITHOUT waiting for the
 * device -- no launch shape depends on R (the scatter is P-sized, sort and compositing are tile-sized).  The binning
 * workspace must be given (sized for binning_capacity); *num_rendered_host is set to -1.  After the scan the device
 * writes (R, overflow, prefiltered-violation) into the status slot's pinned host record and the slot's event is
 * recorded; gsm_status_wait / gsm_status_query read it.  If R > binning_capacity the frame is rendered EMPTY
 * (scatter, sort, both compositing passes see no instances: background image, zero allmap, zero gradients) and
 * `overflow` is set: the caller grows the workspace and renders the frame again. */
int gsm_forward(const gsm_forward_args* args, void* stream);

/* Status slot of an asynchronous forward: 16 bytes of pinned host memory + one cudaEvent.  A slot serves one forward
 * at a time (create a few and recycle them).  Host-side calls; gsm_status_wait blocks the calling thread until the
 * forward's SCAN has finished (not the whole forward) and returns R, the overflow flag and the prefiltered-violation
 * flag ("Point is filtered although prefiltered is set", auxiliary.h:202-206); gsm_status_query returns 1 when that
 * point has been reached, 0 when not yet, and never blocks. */
int gsm_status_create(void** status_out);
int gsm_status_destroy(void* status);
int gsm_status_query(void* status);
int gsm_status_wait(void* status, long long* num_rendered, int* overflow, int* prefiltered_violation);

/* ---- backward: replaces CudaRasterizer::Rasterizer::backward (rasterizer.h:60-87,
 *      rasterizer_impl.cu:346-458) as called by RasterizeGaussiansBackwardCUDA
 *      (rasterize_points.cu:136-233).  Every output is fully written (zeros for culled
 *      surfels), so the caller may pass uninitialised memory. */
typedef struct gsm_backward_args {
  int P, D, M;
  int width, height;
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int debug;            /* bits as gsm_forward_args.debug */
  const float* background;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* transMat_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int* radii;
  const void* geometry;  /* as left by gsm_forward */
  const void* image;
  const void* binning;   /* the persistent prefix of the forward's binning workspace (or a copy) */
  size_t binning_capacity; /* the capacity the forward was called with (locates the strip masks) */
  const float* dL_dout_color;  /* [3,H,W] */
  const float* dL_dout_others; /* [7,H,W] */
  void* grad_workspace; size_t grad_workspace_bytes; /* >= gsm_grad_workspace_bytes(P), transient */
  float* dL_dmeans2D;   /* [P,3] */
  float* dL_dcolors;    /* [P,3] */
  float* dL_dopacity;   /* [P] */
  float* dL_dmeans3D;   /* [P,3] */
  float* dL_dtransMat;  /* [P,9] */
  float* dL_dsh;        /* [P,M,3] (may be NULL when M == 0) */
  float* dL_dscales;    /* [P,2] */
  float* dL_drotations; /* [P,4] */
  int accumulate;       /* 0: outputs are overwritten (reference semantics: fresh zero tensors,
                           rasterize_points.cu:187-195).  1: every output except dL_dmeans2D is
                           ADDED to (view-parallel accumulation over a rank's local views) */
  float* dL_drgb;       /* optional [P,3] (NULL: reference behaviour).  Factored SH gradient for the view-parallel
                           exchange: when given (with shs), the clamp-masked dL/dRGB of every surfel (zeros for
                           culled ones) is written here INSTEAD of the [P,M,3] dL_dsh rows (dL_dsh may be NULL);
                           gsm_sh_grad_expand rebuilds dL_dsh from it on any rank.  Never accumulated. */
  /* optional densification statistics of this view, fused into the backward preprocess (train.py:199-203,
     gaussian_model.py:551-553), all [P], all NULL for reference behaviour: for every visible surfel
     densify_grad_norm += ||dL_dmeans2D.xy||, densify_visible += 1, max_radii = max(max_radii, radii) */
  float* densify_grad_norm;
  float* densify_visible;
  int* max_radii;
  /* raw-parameter entry, as in gsm_forward_args (must match the forward's call): with shs_rest, dL_dsh receives
     d/d_features_dc [P,1,3] and dL_dsh_rest d/d_features_rest [P,M-1,3]; with raw_params, dL_dscales / dL_drotations /
     dL_dopacity are gradients w.r.t. the RAW parameters (chained through exp / normalize / sigmoid in the backward
     preprocess) and `opacities` (the raw ones) must be given */
  const float* shs_rest;
  float* dL_dsh_rest;
  const float* opacities;
  int raw_params;
  /* view-parallel exchange FUSED into the backward preprocess (new: the reference is single-GPU): with n_rgb_peers > 0
     (and shs given) the kernel pushes the clamp-masked dL/dRGB row of every surfel (zeros for culled ones) straight into
     dL_drgb_peers[k] for every rank k of the NVLink domain -- the [P,3] slot of this (rank, local view) in rank k's
     receive buffer, peer memory mapped into this process (CUDA IPC / symmetric memory), 16-byte aligned -- as coalesced
     128-bit stores over NVLink while the kernel runs: the all-gather of the factored exchange without a collective
     launch.  dL_drgb is then ignored. */
  float* dL_drgb_peers[GSM_MAX_PEERS];
  int n_rgb_peers;
} gsm_backward_args;

size_t gsm_grad_workspace_bytes(int P);
int gsm_backward(const gsm_backward_args* args, void* stream);

/* ---- view-parallel exchange, receiving side (new: the reference is single-GPU).  dL_dsh of one view is the
 *      rank-1 product basis(normalize(mean - campos_view)) (x) dL_dRGB (backward.cu:20-139), so ranks exchange
 *      dL_dRGB (12 B/surfel/view) instead of the 12M-byte row.  dL_dsh[p] (= or +=, `accumulate`) the sum over
 *      the n_views views of basis (x) dL_drgb_all[v][p]; D = active SH degree, M = allocated coefficients. */
int gsm_sh_grad_expand(int P, int D, int M, int n_views, const float* means3D /*[P,3]*/,
                       const float* campos_all /*[n_views,3]*/, const float* dL_drgb_all /*[n_views,P,3]*/,
                       float* dL_dsh /*[P,M,3]*/, int accumulate, void* stream);
/* the same with the views `view_stride` floats apart (>= 3 P; receive buffers pad every [P,3] slot to 16 bytes) */
int gsm_sh_grad_expand_strided(int P, int D, int M, int n_views, const float* means3D, const float* campos_all,
                               const float* dL_drgb_all, size_t view_stride, float* dL_dsh, int accumulate, void* stream);

/* ---- the all-reduce of the view-parallel exchange as ONE hand-written kernel over NVLink peer memory (two-shot:
 *      rank r sums shard r of every rank's buffer and writes the sums back to every rank).  bufs[k]: rank k's copy of
 *      the SAME buffer (symmetric / IPC-mapped), laid out as n_sum fp32 to be SUMMED followed by n_max int32 to be
 *      MAXED (the surfel gradients + densification sums, then max_radii); both counts multiples of 4, buffers 16-byte
 *      aligned.  The caller brackets the call with cross-rank barriers on the stream: every rank's buffer complete before,
 *      every shard written after.  Summation order is rank 0..n-1 on every rank: the result is bitwise identical
 *      everywhere. */
int gsm_peer_allreduce(void* const* bufs /*HOST array [n_ranks] of device pointers*/, int n_ranks, int rank,
                       size_t n_sum, size_t n_max, void* stream);
/* The same all-reduce with only WRITES on the wire (the production form: peer loads run at ~1/3 of the rate of posted peer
 * stores on NVLink 5 and starve concurrent kernels): gsm_peer_scatter pushes shard s of the caller's buffer into rank
 * s's staging area (slot [rank]; staging areas of gsm_peer_staging_floats(n, n_sum, n_max) floats, symmetric);
 * after a cross-rank barrier, gsm_peer_sum_broadcast sums the n staged copies of the caller's own shard with LOCAL
 * loads in rank order and pushes the result into every rank's buffer; a second barrier completes the all-reduce. */
size_t gsm_peer_staging_floats(int n_ranks, size_t n_sum, size_t n_max);
int gsm_peer_scatter(const void* local_buf, void* const* staging /*HOST [n_ranks]*/, int n_ranks, int rank, size_t n_sum,
                     size_t n_max, void* stream);
int gsm_peer_sum_broadcast(const void* my_staging, void* const* bufs /*HOST [n_ranks]*/, int n_ranks, int rank, size_t n_sum,
                           size_t n_max, void* stream);
/* the same through the NVSwitch multicast mapping of the buffer (NVLS: multimem.ld_reduce / multimem.st -- the switch
 * sums the ranks' copies and writes the result back to all of them); multicast_buf = the buffer's multicast address */
int gsm_peer_allreduce_multimem(void* multicast_buf, int n_ranks, int rank, size_t n_sum, size_t n_max, void* stream);

/* ---- photometric loss of the training iteration (the rasterizer's caller).  Replaces, fused,
 *      utils/loss_utils.py:16-17 (l1_loss) + :32-62 (ssim, _ssim: 11x11 Gaussian window, sigma 1.5, zero padding,
 *      C1 = 0.01^2, C2 = 0.03^2, mean) as combined at train.py:113-114 AND their autograd backward:
 *        loss = (1 - lambda_dssim) * mean|image - gt| + lambda_dssim * (1 - ssim(image, gt))
 *      image, gt: [C,H,W] fp32 on the device.  loss_out (device, may be NULL): [3] = loss, l1, ssim.
 *      dL_dimage [C,H,W] = grad_scale * dloss/dimage (overwritten).  workspace: caller-owned scratch of
 *      gsm_photometric_loss_workspace_bytes(C,H,W) bytes. */
size_t gsm_photometric_loss_workspace_bytes(int C, int H, int W);
int gsm_photometric_loss(const float* image, const float* gt, int C, int H, int W, float lambda_dssim,
                         float grad_scale, float* loss_out, float* dL_dimage, void* workspace,
                         size_t workspace_bytes, void* stream);
/* ---- allocation-free variant of gsm_photometric_loss.  NO counterpart in the reference: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 *      harnesses that forbid it).  Same inputs, outputs and numerics as gsm_photometric_loss, but the scratch (the SSIM
 *      sums and the three partial maps) lives in a file-scope `static __device__` array inside the library, sized for
 *      C*H*W <= GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS: 48 MiB + 256 B of device memory on every GPU that loads the
 *      library, held until exit.  Larger images are refused with GSM_ERR_WORKSPACE_TOO_SMALL (gsm_last_error() names
 *      the limit).  The call performs no cudaMalloc / cudaFree and calls no allocating library routine; its one-time
 *      set-up per device (cudaGetSymbolAddress of the array, upload of the 11 window weights) happens on the first
 *      call on that device, so make that call outside a graph capture.  The array is shared by every call on a device
 *      and keeps the last call's contents (what must start at zero is cleared in stream order): calls that overlap on
 *      one GPU -- different streams or host threads -- RACE on it and must be serialised by the caller. */
#define GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS 4194304 /* = 4 Mi pixel-channels, e.g. 3 x 1365 x 1024 */
size_t gsm_photometric_loss_static_max_elements(void);
int gsm_photometric_loss_static(const float* image, const float* gt, int C, int H, int W, float lambda_dssim,
                                float grad_scale, float* loss_out, float* dL_dimage, void* stream);

/* ---- optimizer step of the training iteration.  Replaces, fused into ONE launch over all parameter
 *      tensors, `gaussians.optimizer.step()` (train.py:212) of torch.optim.Adam(l, lr=0.0, eps=1e-15)
 *      (scene/gaussian_model.py:283-292): amsgrad off, weight_decay 0, every element updated.  Per tensor (in place):
 *        exp_avg = lerp(exp_avg, grad, 1 - beta1);  exp_avg_sq = exp_avg_sq * beta2 + (1 - beta2) * grad^2;
 *        param  -= lr / (1 - beta1^step) * exp_avg / (sqrt(exp_avg_sq) / sqrt(1 - beta2^step) + eps)
 *      `step` is the step count AFTER this update (>= 1), as torch keeps it.  `tensors` is a HOST array; all device
 *      pointers fp32, 16-byte aligned. */
typedef struct gsm_adam_tensor {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  size_t n;
  double lr, beta1, beta2, eps;
  int step;
} gsm_adam_tensor;
int gsm_adam_step(const gsm_adam_tensor* tensors, int n_tensors, void* stream);

/* ---- densification bookkeeping of the training iteration (second half).  Replaces, in ONE pass,
 *      train.py:199-203  `gaussians.max_radii2D[vf] = torch.max(gaussians.max_radii2D[vf], radii[vf])`  and
 *      GaussianModel.add_densification_stats (scene/gaussian_model.py:551-553)
 *        xyz_gradient_accum[uf] += torch.norm(viewspace_point_tensor.grad[uf], dim=-1, keepdim=True);  denom[uf] += 1
 *      radii int32 [P] (the rasterizer's output), viewspace_grad fp32 [P,3] (the gradient the backward leaves on
 *      means2D), update_filter bool/uint8 [P] or NULL (= radii > 0, render()'s visibility_filter);
 *      max_radii2D [P], xyz_gradient_accum [P,1], denom [P,1] fp32, updated in place.  16-byte aligned arrays. */
int gsm_densification_stats(const int* radii, const float* viewspace_grad, const unsigned char* update_filter,
                            float* max_radii2D, float* xyz_gradient_accum, float* denom, size_t P, void* stream);

/* ---- densify_and_prune of the training loop (last part).  Replaces GaussianModel.densify_and_prune
 *      (scene/gaussian_model.py:535-549) with everything it calls: densify_and_clone (:519-533), densify_and_split
 *      (:492-517), densification_postfix (:468-490), cat_tensors_to_optimizer (:446-466), prune_points /
 *      _prune_optimizer (:413-444); called from train.py:205-208.  Two calls with one host read between them (the new
 *      surfel count sizes the new tensors):
 *        gsm_densify_plan   flags every surfel (clone / split / prune), scans, leaves the four totals
 *                           (kept originals nA, kept clones nB, kept split parents nS, all split parents n_split)
 *                           in the workspace and, if totals_host != NULL, copies them there on the stream;
 *                           P_new = nA + nB + 2 nS; the caller draws normal_draw = N(0,1)[2 n_split, 3] (the same
 *                           draw torch.normal(mean, std) makes at gaussian_model.py:507, so the random stream matches)
 *        gsm_densify_apply  builds the source map and gathers EVERY tensor in one launch: rows of surfel order
 *                           [kept originals | clones | first children | second children].
 *      Roles: COPY for plain per-surfel rows (also int64 segments as 2 words), MOMENT for Adam's exp_avg / exp_avg_sq
 *      (kept for originals, zero for clones and children; in == NULL means all zero), XYZ and SCALING for the two
 *      tensors whose children rows are computed (xyz + R(q) (z * (exp(scaling), 0)); log(exp(scaling) / (0.8 N))).
 *      All tests are evaluated as the reference's torch statements evaluate them (fp32, sigmoid / exp as torch's
 *      kernels, Python scalars rounded to fp32).  src_map / zrow: caller-owned u32[P_new] scratch. */
#define GSM_DENSIFY_ROLE_COPY 0
#define GSM_DENSIFY_ROLE_MOMENT 1
#define GSM_DENSIFY_ROLE_XYZ 2
#define GSM_DENSIFY_ROLE_SCALING 3
typedef struct gsm_densify_args {
  size_t P;
  const float* xyz;                 /* [P,3] */
  const float* opacity;             /* [P]   raw (pre-sigmoid) */
  const float* scaling;             /* [P,2] raw (log) */
  const float* rotation;            /* [P,4] raw */
  const float* xyz_gradient_accum;  /* [P] */
  const float* denom;               /* [P] */
  float grad_threshold;             /* max_grad */
  float dense_limit;                /* percent_dense * extent */
  float min_opacity;
  float world_limit;                /* 0.1 * extent */
  float split_inv;                  /* fp32(1) / fp32(0.8 * N) */
  int has_size_limit;               /* max_screen_size is not None */
  int split_count;                  /* N = 2 */
  unsigned long long* totals_host;  /* optional HOST u64[4] */
} gsm_densify_args;
typedef struct gsm_densify_tensor {
  const void* in;   /* [P, words] 4-byte words */
  void* out;        /* [P_new, words], 16-byte aligned */
  int words;
  int role;
} gsm_densify_tensor;
size_t gsm_densify_workspace_bytes(size_t P);
int gsm_densify_plan(const gsm_densify_args* args, void* workspace, size_t workspace_bytes, void* stream);
int gsm_densify_apply(const gsm_densify_args* args, const void* workspace, const gsm_densify_tensor* tensors, int n_tensors,
                      size_t P_new, uint32_t* src_map, uint32_t* zrow, const float* normal_draw, void* stream);

/* ---- surface regularisers of the training iteration.  Replaces, fused, the allmap
 *      post-processing of render() (gaussian_renderer/__init__.py:117-147: world-space normal, nan_to_num'd median and
 *      expected depth, surf_depth, surf_normal = depth_to_normal(view, surf_depth) * alpha.detach()), depth_to_normal /
 *      depths_to_points (utils/point_utils.py:9-37) and the two loss terms of train.py:131-140
 *        normal_loss = lambda_normal * mean(1 - sum_c rend_normal_c * surf_normal_c),  dist_loss = lambda_dist * mean(rend_dist)
 *      AND their autograd backward: dL_dallmap[7,H,W] = grad_scale * d(normal_loss + dist_loss)/d(allmap), overwritten.
 *      Per-camera constants (host floats): ray_matrix (row-major 3x3, ray_dir(x, y) = [x y 1] * ray_matrix, i.e.
 *      intrins.inverse().T @ c2w[:3,:3].T of point_utils.py:21), ray_origin (c2w[:3,3]), view_rotation
 *      (world_view_transform[:3,:3] as stored).  Optional outputs may be NULL.  One deviation: where alpha == 0 torch's
 *      backward of D/alpha yields 0/0 = NaN (pixels nothing was blended into, which no surfel reads); here 0 is written. */
typedef struct gsm_regularizer_args {
  const float* allmap; /* [7,H,W] device */
  int H, W;
  float ray_matrix[9];
  float ray_origin[3];
  float view_rotation[9];
  float depth_ratio, lambda_normal, lambda_dist, grad_scale;
  float* loss_out;     /* device [3]: normal_loss + dist_loss, normal_loss, dist_loss (may be NULL) */
  float* dL_dallmap;   /* device [7,H,W] */
  float* rend_normal;  /* optional device [3,H,W]: render_pkg["rend_normal"] */
  float* surf_normal;  /* optional device [3,H,W]: render_pkg["surf_normal"] */
  float* surf_depth;   /* optional device [H,W]:   render_pkg["surf_depth"] */
  void* workspace;     /* gsm_surface_regularizer_workspace_bytes(H, W) */
  size_t workspace_bytes;
} gsm_regularizer_args;
size_t gsm_surface_regularizer_workspace_bytes(int H, int W);
int gsm_surface_regularizer(const gsm_regularizer_args* args, void* stream);

/* ---- markVisible: replaces CudaRasterizer::Rasterizer::markVisible (rasterizer.h:24-31,
 *      rasterizer_impl.cu:141-153; rasterize_points.cu:235-254).  present is uint8[P]. */
int gsm_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* ---- distCUDA2: replaces SimpleKNN::knn (simple_knn.h:18, simple_knn.cu:186-222) as called by
 *      distCUDA2 (spatial.cu:15-26): mean squared distance to the 3 nearest other points. */
size_t gsm_knn_workspace_bytes(int P);
int gsm_knn_dist2(int P, const float* points /*[P,3]*/, float* mean_dist2 /*[P]*/,
                  void* workspace, size_t workspace_bytes, void* stream);

/* ---- 32-bit key/value radix sort (stable, ascending on key bits [0, end_bit)): the hand-written
 *      replacement of the cub::DeviceRadixSort::SortPairs call of the Morton ordering
 *      (simple_knn.cu:213-214).  Exposed so that parity tests can hold it to a stable reference
 *      sort directly; gsm_knn_dist2 uses it internally.  in/out buffers must not overlap. */
size_t gsm_sort_pairs32_temp_bytes(size_t n);
int gsm_sort_pairs32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* values_in,
                     uint32_t* values_out, size_t n, int end_bit, void* temp, size_t temp_bytes,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GAUSSMART_B200_H */
