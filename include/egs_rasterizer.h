/*
 * egs_rasterizer.h -- C-ABI of the B200-native differentiable Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for the hot path named by BASELINE.json `north_star`:
 * the functions below replace, one for one, the inner C++ API of the reference submodule
 * `submodules/diff-gaussian-rasterization` (abbreviated DGR below):
 *
 *   egs_forward        <- CudaRasterizer::Rasterizer::forward      DGR/cuda_rasterizer/rasterizer.h:31-55
 *                         (impl DGR/cuda_rasterizer/rasterizer_impl.cu:197-339)
 *   egs_backward       <- CudaRasterizer::Rasterizer::backward     rasterizer.h:57-89 (impl :343-447)
 *   egs_mark_visible   <- CudaRasterizer::Rasterizer::markVisible  rasterizer.h:24-29 (impl :141-153)
 *   egs_adam_step      <- torch.optim.Adam.step as configured in scene/gaussian_model.py:180-202
 *   egs_se3_apply      <- GaussianModel.apply_trans_rot_new        scene/gaussian_model.py:939-986
 *   egs_adam_se3_step  <- reverse_trans_rot_new + Adam.step        scene/gaussian_model.py:1037-1060
 *   egs_exchange_sum / egs_exchange_adam_step <- the gradient sum of view-data-parallel training
 *                         (no reference counterpart; loss.backward() + optimizer.step(), train.py:124,177)
 *   egs_ply_pack / egs_ply_unpack <- GaussianModel.save_ply / load_ply  scene/gaussian_model.py:375-481
 *   egs_model_forward / egs_model_backward
 *                      <- get_scaling/get_opacity/get_features     scene/gaussian_model.py:126-153
 *                         get_covariance/get_rotated_covariance    scene/gaussian_model.py:29-63,167-171
 *                         build_rotation/build_scaling_rotation    utils/general_utils.py:110-158
 *   egs_image_loss     <- l1_loss + ssim (+ gradient mask hook)    utils/loss_utils.py:57-106
 *   egs_densify_stats  <- add_densification_stats + max_radii2D    scene/gaussian_model.py:735-740
 *   egs_knn_mean_dist2 <- distCUDA2                                submodules/simple-knn/spatial.cu:15-27
 *   egs_compact_plan / egs_compact_rows
 *                      <- prune_points / _prune_optimizer          scene/gaussian_model.py:506-557
 *   egs_densify_select / egs_append_rows / egs_split_samples
 *                      <- densify_and_clone / densify_and_split / cat_tensors_to_optimizer
 *                                                                  scene/gaussian_model.py:243-272,559-677
 *
 * Conventions
 *   - plain C, no torch types, no exceptions; every pointer named `d_*` or documented as
 *     "device" is a CUDA device pointer on the current device; `stream` is a cudaStream_t
 *     passed as void* (NULL = legacy default stream).
 *   - all floating point is fp32, contiguous row-major, exactly the reference layouts:
 *     viewmatrix/projmatrix are the transposed (row-vector) 4x4 matrices, sh is [P,M,3],
 *     cov3D is [xx,xy,xz,yy,yz,zz], quaternions are (w,x,y,z) and are NOT normalised.
 *   - an absent optional input is a NULL pointer (the reference's empty tensor).
 *   - return value: EGS_OK (0) or a negative EgsStatus; egs_last_error() returns a
 *     thread-local human-readable message for the last failure.
 *   - there is NO CPU fallback: if no CUDA device is usable the calls fail with
 *     EGS_ERR_CUDA.
 */
#ifndef EGS_RASTERIZER_H_
#define EGS_RASTERIZER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EGS_ABI_VERSION 2
#define EGS_TILE_X 16          /* DGR/cuda_rasterizer/config.h:16 */
#define EGS_TILE_Y 16          /* DGR/cuda_rasterizer/config.h:17 */
#define EGS_NUM_CHANNELS 3     /* DGR/cuda_rasterizer/config.h:15 */
#define EGS_MAX_PEERS 8        /* GPUs of one NVLink domain; also the largest batch of views of one exchange */

typedef enum EgsStatus {
  EGS_OK = 0,
  EGS_ERR_INVALID_ARGUMENT = -1, /* bad shape / both-or-neither of an exclusive input pair */
  EGS_ERR_CUDA = -2,             /* a CUDA runtime call or kernel launch failed */
  EGS_ERR_ALLOC = -3,            /* a workspace callback returned NULL */
  EGS_ERR_TOO_LARGE = -4,        /* num_rendered or P beyond the supported range */
  EGS_ERR_PREFILTERED = -5       /* args->prefiltered was set and a Gaussian was culled all the same. The
                                    reference prints "Point is filtered although prefiltered is set" and
                                    traps the kernel (auxiliary.h:154-161); here the forward returns this
                                    code with the same message, read back with the instance count */
} EgsStatus;

/* Workspace callback: must return a device pointer to at least `bytes` bytes, 256-byte
 * aligned, that stays valid until the matching egs_backward has run. Replaces the three
 * std::function<char*(size_t)> resizers of rasterize_points.cu:27-33. */
typedef void* (*EgsAllocFn)(void* ctx, size_t bytes);

typedef struct EgsForwardArgs {
  int P;                /* number of Gaussians */
  int sh_degree;        /* active degree D, 0..3 */
  int sh_coeffs;        /* M = coefficients per channel present in `shs` (0 when shs == NULL) */
  int width, height;
  const float* background;     /* device [3] */
  const float* means3D;        /* device [P,3] */
  const float* shs;            /* device [P,M,3] or NULL */
  const float* colors_precomp; /* device [P,3]   or NULL (exactly one of shs / colors_precomp) */
  const float* opacities;      /* device [P] */
  const float* scales;         /* device [P,3] or NULL */
  const float* rotations;      /* device [P,4] or NULL */
  const float* cov3D_precomp;  /* device [P,6] or NULL (exactly one of scales+rotations / cov3D) */
  float scale_modifier;
  const float* viewmatrix;     /* device [16] */
  const float* projmatrix;     /* device [16] */
  const float* campos;         /* device [3] */
  float tan_fovx, tan_fovy;
  int prefiltered;
  int debug;                   /* non-zero: synchronise and check after every stage */
  /* outputs (device, caller-allocated, need not be zeroed) */
  float* out_color;            /* [3,H,W] */
  float* out_depth;            /* [1,H,W]  un-normalised sum depth*alpha*T (forward.cu:360) */
  float* out_alpha;            /* [1,H,W]  sum alpha*T                     (forward.cu:361) */
  int* radii;                  /* [P] */
  /* workspaces: three opaque blobs that must be handed unchanged to egs_backward */
  EgsAllocFn geom_alloc;   void* geom_ctx;    /* size = egs_geom_bytes(P) */
  EgsAllocFn binning_alloc; void* binning_ctx; /* size = egs_binning_bytes(num_rendered) */
  EgsAllocFn image_alloc;  void* image_ctx;   /* size = egs_image_bytes(width,height) */
  /* Optional: the caller's guess of num_rendered (0 = none), e.g. 1.2 x the value of the previous
   * iteration. With a guess the binning workspace is requested for that many instances BEFORE the
   * count is known and every kernel is enqueued without waiting for it; the one device->host
   * read is then waited for after the last launch (it completes early in the forward), so the
   * GPU never idles on the host and the call returns about one blend earlier. If the real count
   * exceeds the guess, the enqueued work has skipped itself, binning_alloc is called a second
   * time with the exact size and binning + blend are enqueued again: results never depend on the
   * guess. */
  int num_rendered_hint;
} EgsForwardArgs;

/* Sizes of the three workspaces; pure functions of their arguments so that backward can
 * re-derive the internal layout (rasterizer_impl.h:67-73 `required<T>`). The binning workspace is
 * laid out for a capacity = num_rendered rounded up to a multiple of 4096, and its size is
 * strictly increasing in that capacity, so egs_binning_capacity() recovers it from the size. */
size_t egs_geom_bytes(int P);
size_t egs_binning_bytes(int num_rendered, int width, int height);
int egs_binning_capacity(size_t binning_bytes);
size_t egs_image_bytes(int width, int height);

/* Forward render. On success *num_rendered is the number of (Gaussian, tile) instances.
 * Performs exactly one device->host read (num_rendered), like rasterizer_impl.cu:281-282. */
int egs_forward(const EgsForwardArgs* args, void* stream, int* num_rendered);

typedef struct EgsBackwardArgs {
  int P, sh_degree, sh_coeffs, num_rendered;
  int width, height;
  const float* background;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  float scale_modifier;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  float tan_fovx, tan_fovy;
  const int* radii;            /* [P] from forward */
  const float* out_alpha;      /* [1,H,W] from forward (T_final = 1 - alpha, backward.cu:463) */
  const void* geom_buffer;     /* the three blobs filled by egs_forward */
  const void* binning_buffer;
  const void* image_buffer;
  const float* dL_dout_color;  /* [3,H,W] */
  const float* dL_dout_depth;  /* [1,H,W], or NULL: no gradient flows into the depth image */
  const float* dL_dout_alpha;  /* [1,H,W], or NULL likewise; with both NULL (a loss on the colour image only)
                                  the tile kernel leaves their recurrences out */
  /* scratch: egs_backward_scratch_bytes(P) bytes, device, 256-byte aligned; contents are
   * overwritten (zeroed by the call, then used as atomic accumulators). */
  void* scratch;
  /* outputs (device, caller-allocated, need not be zeroed; every element is written) */
  float* dL_dmeans2D;          /* [P,3]  z component is always 0 */
  float* dL_dcolors;           /* [P,3]  written only when colors_precomp != NULL, else may be NULL */
  float* dL_dopacity;          /* [P] */
  float* dL_dmeans3D;          /* [P,3] */
  float* dL_dcov3D;            /* [P,6]  written only when cov3D_precomp != NULL, else may be NULL */
  float* dL_dsh;               /* [P,M,3] written only when shs != NULL, else may be NULL */
  float* dL_dscales;           /* [P,3]  written only when scales != NULL, else may be NULL */
  float* dL_drotations;        /* [P,4]  written only when rotations != NULL, else may be NULL */
  int debug;
  /* Size in bytes of the binning blob forward was given (0 = egs_binning_bytes(num_rendered)):
   * when forward ran with a num_rendered_hint the blob is laid out for the hinted capacity. */
  size_t binning_bytes;
  /* Non-zero: only dL_dcolors is wanted (requires colors_precomp). Everything that feeds the other
   * gradients is skipped in both backward kernels and no other output is written (they may be
   * NULL). This is the backward of the trainers' per-iteration label render
   * (gaussian_renderer/render_helper.py:40-66: geometry detached, the label as colour). */
  int colors_only;
  /* Optional, with shs (or, in egs_backward_raw, the model's features): device, EGS_DRGB_SLOT_FLOATS(P)
   * floats, 16-byte aligned. When set, the SH gradient is returned in factored form: rows [P,3] with
   * dL_drgb[i] = dL/dRGB of Gaussian i, clamped channels zeroed, followed (at float offset 3 P rounded up
   * to a multiple of 4) by the three floats of `campos`; dL_dsh (d_features_dc / d_features_rest) is NOT
   * written and may be NULL. dL/dsh[i][k][c] = basis_k(direction from campos to means3D[i]) * dL_drgb[i][c]:
   * 3 floats per Gaussian carry what 3 M floats would, which is what the view-data-parallel exchange
   * moves (egs_exchange_sum_factored expands the sum over the views). */
  float* dL_drgb;
  /* View-data-parallel: the same slot (rows + campos) is also written, by the same kernel, to
   * drgb_remote[0 .. drgb_n_remote) -- this view's slot in the OTHER ranks' gather buffers, peer-mapped --
   * or, when drgb_multicast is set, with one multicast store per 16 bytes that reaches every rank's
   * buffer through the NVLink switch. The stores ride along with the kernel's HBM traffic, so after
   * the backward every rank already holds every view's factors locally. */
  int drgb_n_remote;
  float* drgb_remote[EGS_MAX_PEERS];
  float* drgb_multicast;
} EgsBackwardArgs;

size_t egs_backward_scratch_bytes(int P);
/* floats in one dL_drgb slot: the rows, padded to a multiple of 4, then campos (3) and one spare */
#define EGS_DRGB_SLOT_FLOATS(P) ((((long long)(P) * 3 + 3) / 4) * 4 + 4)

int egs_backward(const EgsBackwardArgs* args, void* stream);

/* Backward pass for hosts that must not allocate device memory. No counterpart in the reference
 * (its binding allocates ten gradient tensors and its workspaces on every call).
 *
 * Who needs it: a process whose GPU memory is owned by somebody else's pool -- an inference or
 * render server that reserved the whole device at start-up, a plugin running inside such a
 * host, a loop with a fixed memory budget -- cannot hand egs_backward a fresh scratch buffer,
 * and cudaMalloc/cudaFree would also stall every stream of the device while they run.
 *
 * What it does: exactly egs_backward, except that `args->scratch` is ignored (may be NULL) and
 * the 48-byte-per-Gaussian accumulator records live in an array that is part of the library
 * image itself: EGS_STATIC_SCRATCH_MAX_GAUSSIANS records = 12 MiB, resident on every GPU that
 * loads the library until the process exits. The call performs no cudaMalloc, cudaFree or any
 * other allocating routine; inputs, the forward blobs and all gradient outputs are still the
 * caller's buffers.
 *
 * Limits and rules:
 *   - args->P above EGS_STATIC_SCRATCH_MAX_GAUSSIANS is refused with EGS_ERR_TOO_LARGE;
 *   - the first call on each device looks up the array's device address (a one-time set-up
 *     with that device current); make that first call before starting any stream capture;
 *   - all calls on a device share the one array. The accumulators are cleared at the start of
 *     every call, so stale contents never leak into a result, but two calls that overlap on
 *     the GPU (different streams, or different host threads) race on it: serialise them. */
#define EGS_STATIC_SCRATCH_MAX_GAUSSIANS 262144
int egs_backward_static_scratch(const EgsBackwardArgs* args, void* stream);

/* present[i] = !(view-space z of means3D[i] <= 0.2), 1 byte per Gaussian (rasterizer_impl.cu:54-66,
 * auxiliary.h:150: a point whose depth is NaN is reported visible, as the reference does). */
int egs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, unsigned char* present, void* stream);

/* ---- optimizer step (north_star item 5) ------------------------------------------------ */

/* One Adam parameter group: param/grad/exp_avg/exp_avg_sq are device [n] fp32. Follows
 * torch.optim.Adam(lr, betas, eps) with bias correction, no weight decay, no amsgrad, as
 * configured at scene/gaussian_model.py:198 (eps = 1e-15). `step` is the 1-based step count
 * AFTER this update (torch increments before use). lr, the betas and eps are doubles because
 * torch derives 1-beta and the bias corrections in double before rounding to fp32. */
typedef struct EgsAdamGroup {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  long long n;
  double lr;
  int step;
} EgsAdamGroup;

/* Fused multi-group Adam: one launch for all groups (host array of up to EGS_ADAM_MAX_GROUPS). */
#define EGS_ADAM_MAX_GROUPS 16
int egs_adam_step(const EgsAdamGroup* groups, int n_groups, double beta1, double beta2, double eps,
                  void* stream);

/* Rigid SE(3) transform of object Gaussians (scene/gaussian_model.py:939-986,
 * utils/geometry_utils.py:188-193): for rows with is_object[i] == which_object,
 * out[i] = R * xyz[i] + t, else out[i] = xyz[i]. `Rt` is device [12]: row-major R (9) then t (3).
 * When cov_in != NULL, also rotates the packed covariance: cov_out = R * cov * R^T. */
int egs_se3_apply(int P, const float* xyz, const int* is_object, int which_object,
                  const float* Rt, float* xyz_out, const float* cov_in, float* cov_out,
                  void* stream);

/* Fused "pull gradient back through the rigid map + Adam on canonical xyz" (SURVEY.md 8a A-24):
 * g_canonical = R^T * g_moved for object rows, then the Adam update of egs_adam_step. */
int egs_adam_se3_step(int P, float* xyz, const float* grad_moved, float* exp_avg, float* exp_avg_sq,
                      const int* is_object, int which_object, const float* Rt, double lr, int step,
                      double beta1, double beta2, double eps, void* stream);

/* ---- gradient exchange of view-data-parallel training --------------------- */

/* Every rank renders another camera view of the same scene; the one exchange of the path is the
 * sum of the parameter gradients (the reference has no multi-GPU mode: this replaces what a
 * DistributedDataParallel wrapper around train.py's loss.backward(), train.py:124, would do with
 * NCCL). All ranks keep their gradients in ONE flat fp32 buffer of the same layout, mapped into every
 * peer's address space (CUDA IPC / symmetric memory over NVLink), and one kernel per rank does the
 * whole exchange: rank r sums slice r of all the buffers in rank order and writes the sums back into
 * every buffer, between two flag barriers that live in the signal words. No host synchronisation, no
 * staging copies; every rank ends with bit-identical sums. */
#define EGS_EXCHANGE_SIGNAL_WORDS 64
typedef struct EgsPeerGroup {
  int world, rank;
  float* grads[EGS_MAX_PEERS];      /* flat gradient buffer of every rank, as mapped on THIS rank; 16-byte aligned */
  uint32_t* signals[EGS_MAX_PEERS]; /* EGS_EXCHANGE_SIGNAL_WORDS u32 of every rank, zero before the first call */
  float* grads_multicast;           /* NVLink-switch multicast mapping of the same gradient buffers, or NULL:
                                       with it the sums are formed inside the switch (multimem.ld_reduce) and
                                       stored to all ranks by one store; the order of the additions is then
                                       the switch's, still one result for all ranks */
} EgsPeerGroup;

/* grads[p][0..n) <- scale * sum over q of grads[q][0..n), on all ranks p. n_floats % 4 == 0.
 * `epoch` is the number of this exchange: the same on all ranks, 1 at the first call and one higher at
 * every call on the same signal words (both entry points count). `max_ctas` bounds the grid (0: as many
 * CTAs as the device can hold at once, from its SM count and the occupancy calculator; more than that
 * is refused, because the flag barriers need the whole grid resident). A peer that does not show up
 * within the timeout makes the kernel skip the exchange, leave every buffer as it was and set
 * signals[rank][EGS_EXCHANGE_SIGNAL_WORDS - 1] to 1. */
int egs_exchange_sum(const EgsPeerGroup* group, long long n_floats, float scale, uint32_t epoch, int max_ctas,
                     void* stream);
/* The same exchange with the SH gradient in factored form. Per view dL/dsh has rank one,
 * dL/dsh[i][k][c] = basis_k(direction from the view's camera centre to means3D[i]) * dL/dRGB[i][c], so a
 * view-data-parallel step lets egs_backward write the three floats of dL/dRGB (EgsBackwardArgs.dL_drgb,
 * pushed to every rank's gather buffer by the backward kernel itself: drgb_remote / drgb_multicast)
 * instead of the 3 M floats of dL/dsh, and this call
 *   - sums the flat buffer of the OTHER gradients (group->grads, n_floats) over the ranks like
 *     egs_exchange_sum, and
 *   - reads every view's slot (rows of dL/dRGB + camera centre) from THIS rank's memory, forms each
 *     view's products with the basis evaluated at this rank's own (replicated) means3D, and adds them
 *     in view order into dL_dsh, a plain local tensor: sum over views of the dL/dsh each would have
 *     written (the same products, the same order of additions, hence the same bits as exchanging the
 *     expanded rows), times `scale`.
 * Per GPU (views - 1) * 12 bytes per Gaussian cross NVLink for the SH part instead of about 2 * 12 M.
 * The two barriers of the exchange also fence the slots: a view's slot is complete when its rank has
 * entered, and nobody overwrites a slot before every rank has left.
 * Same epoch / max_ctas / timeout rules as egs_exchange_sum (one epoch counter for all three calls).
 *
 * A batch of views (BASELINE configs[3]: 8 views on 1, 2, 4 or 8 GPUs, rank r rendering views r, r + world,
 * ...): set n_views to the size of the batch and give, for every view in its global order, its slot in
 * this rank's gather buffer and the flat buffer its backward wrote (on whichever rank, as mapped here).
 * The sums then run over the views in that order, so the result does not depend on how the batch was
 * spread over the GPUs; the summed flat buffer lands in group->grads[r] on every rank r, which may be one
 * of that rank's view buffers. world = 1 is the single-GPU accumulation of the batch. */
typedef struct EgsShFactors {
  int P, sh_degree, sh_coeffs;
  const float* means3D;                  /* device [P,3], this rank's copy */
  const float* drgb[EGS_MAX_PEERS];      /* per rank, or per view when n_views > 0: EGS_DRGB_SLOT_FLOATS(P)
                                            floats each, in this rank's memory */
  float* dL_dsh;                         /* device [P, sh_coeffs, 3], 16-byte aligned, every element written */
  int n_views;                           /* 0: one view per rank */
  const float* view_flat[EGS_MAX_PEERS]; /* n_views > 0: the flat buffer of every view */
} EgsShFactors;
int egs_exchange_sum_factored(const EgsPeerGroup* group, long long n_floats, const EgsShFactors* sh, float scale,
                              uint32_t epoch, int max_ctas, void* stream);

/* How long an exchange kernel waits for its peers before it gives up (default 20 000 ms; process-wide,
 * read at launch). A kernel that gave up has skipped its data phase -- gradients and parameters are
 * left as they were -- and set the error word; the caller must check it before trusting the step. */
int egs_exchange_set_timeout_ms(unsigned long long ms);

/* One parameter group inside the flat layout: floats [begin, end), both multiples of 4. Elements whose
 * offset from `begin`, modulo `period`, is below `split` use lr_a, the others lr_b (features_dc and
 * features_rest share the rows of the SH gradient but not the learning rate,
 * scene/gaussian_model.py:191-192); period == 0: lr_a throughout. */
typedef struct EgsExchangeSegment {
  long long begin, end;
  double lr_a, lr_b;
  int period, split;
} EgsExchangeSegment;
#define EGS_EXCHANGE_MAX_SEGMENTS 8

typedef struct EgsExchangeAdamArgs {
  float* params[EGS_MAX_PEERS];  /* flat parameter buffer of every rank, same layout as the gradients */
  float* params_multicast;       /* multicast mapping of the parameter buffers, or NULL (then peer stores) */
  float* exp_avg;                /* THIS rank's shard of the moments: egs_exchange_shard_floats() floats each, */
  float* exp_avg_sq;             /* element k belongs to flat element rank * shard + k */
  const EgsExchangeSegment* segments;   /* host array, ascending, disjoint */
  int n_segments;
  int step;                      /* Adam step number of this update (>= 1) */
  double beta1, beta2, eps;
} EgsExchangeAdamArgs;
long long egs_exchange_shard_floats(long long n_floats, int world);

/* The exchange with the optimiser inside it: rank r sums slice r of the gradients as above, applies
 * the Adam update of egs_adam_step to slice r of the parameters with its shard of the moments, and
 * writes the new parameters into every rank's parameter buffer (reduce-scatter, update, all-gather in
 * one kernel; the moments exist once per node instead of once per rank). Floats outside every
 * segment (the screen-space gradient that rides along for the densification statistic) are summed
 * and written back like egs_exchange_sum does; gradients inside segments are left as they were. */
int egs_exchange_adam_step(const EgsPeerGroup* group, const EgsExchangeAdamArgs* adam, long long n_floats,
                           float scale, uint32_t epoch, int max_ctas, void* stream);

/* ---- GaussianModel activations --------------------- */

/* The optimiser's raw parameters and what render() derives from them every iteration
 * (gaussian_renderer/__init__.py:57-86 with compute_cov3D_python forced on, train.py:49):
 *   opacity = sigmoid(opacity_raw)                       get_opacity
 *   shs     = cat(features_dc, features_rest) on dim 1   get_features
 *   cov3D   = strip_symmetric(L' L'^T), L = R(q/|q|) diag(scale_modifier * exp(scaling_raw)),
 *             L' = train_R * (accum_R * L) on the selected rows, L' = L elsewhere
 *                                                        get_covariance / get_rotated_covariance
 * Selected rows: accum_R == NULL and train_R == NULL -> none (plain get_covariance);
 * otherwise is_object == NULL -> all (which_object=None), else rows with is_object[i] ==
 * which_object. accum_R NULL with train_R set means accum_R = identity (gaussian_model.py:55).
 *
 * row0_copies reproduces an indexing quirk of gaussian_model.py:49-59: the mask is a [P,1] tensor,
 * so nonzero(mask).squeeze() is a list of (row, 0) pairs, and indexing L with it reads and
 * overwrites row 0 once per selected row too. Whenever at least one row is selected, row 0 is
 * rotated as well, and in backward it receives its gradient once per index entry naming it.
 * Pass row0_copies = n_selected + (row 0 selected ? 1 : 0) when n_selected > 0, else 0; pass 0
 * always to get the "intended" per-row behaviour instead. */
typedef struct EgsModelArgs {
  int P;
  int sh_coeffs;               /* M of the concatenated feature tensor (1 + rest) */
  const float* scaling_raw;    /* device [P,3]   GaussianModel._scaling (log scale) */
  const float* rotation_raw;   /* device [P,4]   _rotation, un-normalised (w,x,y,z); 16-B aligned */
  const float* opacity_raw;    /* device [P,1]   _opacity logits */
  const float* features_dc;    /* device [P,1,3] */
  const float* features_rest;  /* device [P,M-1,3] (NULL when M == 1) */
  float scale_modifier;
  const int* is_object;        /* device [P] int32 (the reference's [P,1] tensor) or NULL */
  int which_object;
  const float* accum_R;        /* device [9] row-major 3x3, or NULL */
  const float* train_R;        /* device [9] row-major rot6d_to_matrix(obj_rotation_6d)
                                  (during_training, geometry_utils.py:26-28), or NULL */
  int row0_copies;
} EgsModelArgs;

/* Each output is optional (NULL = not wanted): cov3D [P,6], opacity [P,1], shs [P,M,3]. Only the
 * raw inputs an output needs have to be set. One launch for cov3D+opacity, one for shs. */
int egs_model_forward(const EgsModelArgs* model, float* cov3D, float* opacity, float* shs, void* stream);

/* Backward of the above. Each upstream gradient is optional and selects its outputs:
 * dL_dcov3D -> d_scaling_raw [P,3], d_rotation_raw [P,4] (and d_train_R, device [9], overwritten
 * with dL/d train_R when both it and model->train_R are set); dL_dopacity -> d_opacity_raw [P,1];
 * dL_dshs [P,M,3] -> d_features_dc [P,1,3], d_features_rest [P,M-1,3]. Every element of a
 * selected output is written. */
int egs_model_backward(const EgsModelArgs* model, const float* dL_dcov3D, const float* dL_dopacity,
                       const float* dL_dshs, float* d_scaling_raw, float* d_rotation_raw,
                       float* d_opacity_raw, float* d_features_dc, float* d_features_rest,
                       float* d_train_R, void* stream);

/* The same activations fused INTO the rasterizer (SURVEY.md 8f rank 1 as stated: "fuse activations
 * + covariance build + SH concat into preprocess"): the per-Gaussian forward kernel reads the raw
 * parameters directly and the per-Gaussian backward kernel writes the raw-parameter gradients, so
 * cov3D, the activated opacity and the concatenated SH tensor (and their gradients) never touch
 * HBM. Arithmetic is shared with egs_model_forward/backward (csrc/model_math.cuh), so
 * egs_forward_raw(args, model) renders bit-identically to egs_model_forward + egs_forward with
 * cov3D_precomp / opacities / shs set to its outputs.
 *
 * egs_forward_raw: args->opacities, shs, scales, rotations and cov3D_precomp must be NULL (they come
 * from `model`); args->sh_coeffs is ignored (model->sh_coeffs is used); args->colors_precomp may be
 * set (override_color), in which case the feature tensors are not read. Everything else as
 * egs_forward. */
int egs_forward_raw(const EgsForwardArgs* args, const EgsModelArgs* model, void* stream, int* num_rendered);

typedef struct EgsModelGrads {   /* device, caller-allocated, every element written */
  float* d_scaling_raw;    /* [P,3] */
  float* d_rotation_raw;   /* [P,4], 16-byte aligned */
  float* d_opacity_raw;    /* [P,1] */
  float* d_features_dc;    /* [P,1,3]   (NULL when colors_precomp was used) */
  float* d_features_rest;  /* [P,M-1,3] (NULL when M == 1 or colors_precomp was used) */
  float* d_train_R;        /* device [9], overwritten; NULL unless model->train_R is set and wanted */
} EgsModelGrads;

/* egs_backward_raw: the inputs of egs_backward (same NULL rules as egs_forward_raw); of its outputs
 * only dL_dmeans2D, dL_dmeans3D and (with colors_precomp) dL_dcolors are used, the rest is ignored
 * and the raw-parameter gradients go to `grads`. */
int egs_backward_raw(const EgsBackwardArgs* args, const EgsModelArgs* model, const EgsModelGrads* grads,
                     void* stream);

/* Per-iteration densification bookkeeping (the per-iteration half:
 * GaussianModel.add_densification_stats, scene/gaussian_model.py:735-740, and the max_radii2D
 * line of trainers/train_static_bg.py:128). For rows with radii[i] > 0:
 *   max_radii2D[i] = max(max_radii2D[i], radii[i]);
 *   xyz_gradient_accum[i] += sqrt(g.x^2 + g.y^2) with g = dL_dmeans2D[i] (viewspace_points.grad);
 *   denom[i] += 1.
 * All arrays are device, length P ([P,3] for dL_dmeans2D); no host sync (the reference's boolean
 * mask indexing costs several per iteration). */
int egs_densify_stats(int P, const int* radii, const float* dL_dmeans2D, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, void* stream);

/* Stable row compaction of up to EGS_COMPACT_MAX_TENSORS tensors at once (the
 * periodic half: GaussianModel.prune_points / _prune_optimizer, scene/gaussian_model.py:506-557,
 * i.e. `tensor[mask]` on every parameter, both Adam moments and the bookkeeping tensors).
 * egs_compact_plan scans keep[P] (device, one byte per row, non-zero = keep) into `scratch`
 * (device, egs_compact_scratch_bytes(P)) and returns the number of kept rows in *count (host; the
 * one synchronisation, needed to size the outputs). egs_compact_rows then copies, for every kept
 * row in order, row i of each src to row k of its dst (k = kept rows before i): one launch for all
 * tensors. row_bytes must be a multiple of 4; src and dst must not overlap. */
#define EGS_COMPACT_MAX_TENSORS 32
typedef struct EgsRowTensor {
  const void* src;   /* device [P, row_bytes] */
  void* dst;         /* device [count, row_bytes] */
  int row_bytes;
} EgsRowTensor;
size_t egs_compact_scratch_bytes(int P);
int egs_compact_plan(int P, const unsigned char* keep, void* scratch, size_t scratch_bytes, int* count, void* stream);
int egs_compact_rows(int P, const void* scratch, const EgsRowTensor* tensors, int n_tensors, void* stream);

/* The growing half of densification (same SURVEY.md row): GaussianModel.densify_and_clone /
 * densify_and_split with densification_postfix / cat_tensors_to_optimizer, scene/gaussian_model.py:243-272,
 * 559-677. Three steps around egs_compact_plan:
 *
 * egs_densify_select marks the rows to clone or split: grads = xyz_gradient_accum / denom with NaN -> 0
 * (:689-690; rows from n_graded on count as 0, the zero padding of :604-605), selected = grads >=
 * grad_threshold and max(exp(scaling_raw)) <= size_threshold (clone, :650-651) or > size_threshold
 * (split, :607-608), size_threshold = percent_dense * scene_extent; optionally only rows of one object
 * (:614-616, :653-655) and of one generation (:609-611). */
typedef struct EgsDensifySelect {
  int P;
  int n_graded;                  /* rows that have a gradient statistic (P before any append) */
  const float* grad_accum;       /* device [n_graded]  xyz_gradient_accum */
  const float* denom;            /* device [n_graded] */
  const float* scaling_raw;      /* device [P,3]  _scaling (log scale) */
  float grad_threshold, size_threshold;
  int split;                     /* 0: the clone side of the size test, 1: the split side */
  const int* is_object;          /* device [P] or NULL (which_object=None) */
  int which_object;
  const int* generation;         /* device [P] or NULL (split_prev_gen=True) */
  int only_generation;
} EgsDensifySelect;
int egs_densify_select(const EgsDensifySelect* args, unsigned char* selected /* device [P] */, void* stream);

/* After egs_compact_plan(P, selected, scratch, ..., &count): for every tensor at once,
 *   dst [P + copies * count, row] = cat(src, src[selected].repeat(copies, 1))          fill 0
 *                                   cat(src, zeros)     (Adam moments, :257-258)        fill 1
 *                                   cat(src, `value`)   (the generation stamp, :630)    fill 2
 * row_bytes a multiple of 4; src and dst must not overlap. */
typedef struct EgsAppendTensor {
  const void* src;   /* device [P, row_bytes] */
  void* dst;         /* device [P + copies * count, row_bytes] */
  int row_bytes;
  int fill;
  int value;         /* fill 2: the 4-byte pattern of the new rows */
} EgsAppendTensor;
int egs_append_rows(int P, const void* plan_scratch, int count, int copies, const EgsAppendTensor* tensors, int n_tensors,
                    void* stream);

/* Positions and scales of the rows a split adds (:619-622): for new row j (copy j / count of selected
 * row k = j % count), with s = exp(scaling_raw[k]) and R = build_rotation(rotation_raw[k]) (normalised),
 *   xyz_new[j] = R (noise[j] * s) + xyz[k],     scaling_new[j] = log(s / (0.8 * copies)).
 * `noise` is device [copies * count, 3] standard normal: torch.normal(0, s) of :619 is s * noise, so the
 * caller owns the random stream. xyz_new / scaling_new: device [copies * count, 3], e.g. the tail of the
 * tensors egs_append_rows produced. */
int egs_split_samples(int P, const void* plan_scratch, int count, int copies, const float* xyz, const float* scaling_raw,
                      const float* rotation_raw, const float* noise, float* xyz_new, float* scaling_new, void* stream);

/* Mean squared distance of every point to its three nearest neighbours, the point itself excluded
 * (distCUDA2 of submodules/simple-knn, used once at initialisation to set the
 * scales, scene/gaussian_model.py:301,312). points: device [P,3]; mean_dist2: device [P]. With fewer
 * than four points the missing neighbours count as FLT_MAX, like the reference.
 * workspace: egs_knn_workspace_bytes(P) bytes of device memory, 256-byte aligned (0 bytes for small
 * clouds, which are swept exhaustively). With it, large clouds are searched through a uniform grid
 * built on the device; with workspace == NULL the exhaustive sweep runs at any size (same result,
 * quadratic time). */
size_t egs_knn_workspace_bytes(int P);
int egs_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* workspace, size_t workspace_bytes,
                       void* stream);

/* ---- checkpoint records (save_ply / load_ply, scene/gaussian_model.py:340-481) ---- */

/* One record of 4-byte floats per Gaussian, in the reference's property order:
 *   x y z  nx ny nz (zeros)  f_dc_0..2  f_rest_0..  opacity  scale_0..2  rot_0..3  [label generation is_object]
 * f_rest_j with j = c * (sh_coeffs - 1) + k is features_rest[p][k][c] (gaussian_model.py:381). All device
 * pointers; features_dc [P,1,3], features_rest [P,sh_coeffs-1,3], label [P,1] float, generation and
 * is_object [P,1] int32 (stored as floats in the file, like the reference does). */
typedef struct EgsPlyColumns {
  const float *xyz, *features_dc, *features_rest, *opacity, *scaling, *rotation, *label;
  const int *generation, *is_object;
} EgsPlyColumns;
typedef struct EgsPlyColumnsOut {
  float *xyz, *features_dc, *features_rest, *opacity, *scaling, *rotation, *label;
  int *generation, *is_object;
} EgsPlyColumnsOut;
#define EGS_PLY_MAX_COLUMNS 128

/* 14 + 3 * sh_coeffs (+ 3 with the extras): 62 / 65 at SH degree 3. */
int egs_ply_record_floats(int sh_coeffs, int with_extras);
/* rows: device [P, egs_ply_record_floats()] <- the columns (what save_ply writes after the header). */
int egs_ply_pack(int P, int sh_coeffs, const EgsPlyColumns* cols, int with_extras, float* rows, void* stream);
/* The columns <- rows: device [P, record_floats] as read from a file whose properties may come in any
 * order: column_of (host, egs_ply_record_floats(sh_coeffs, 1) ints) gives, for every canonical column, its
 * position in the file's record, or -1 when the file lacks it (normals and the three extras only; a
 * missing label becomes label_default, generation 0, is_object is_object_default: load_ply :444-461). */
int egs_ply_unpack(int P, int sh_coeffs, const float* rows, int record_floats, const int* column_of,
                   const EgsPlyColumnsOut* cols, float label_default, int is_object_default, void* stream);

/* ---- image loss of the trainers ------------------------------------ */

/* loss = (1 - lambda_dssim) * mean|image - target| + lambda_dssim * (1 - SSIM(image, target))
 * with the reference's SSIM: 11x11 Gaussian window, sigma 1.5, zero padding, per channel, mean over
 * all elements (utils/loss_utils.py:57-106; call sites trainers/train_static_bg.py:100-102,
 * fine_all.py:99-100, fine_obj.py:142-143, coarse_obj_pose.py:252-253). Both terms are symmetric
 * in their two arguments, so the argument order of a call site does not matter; the gradient is
 * taken w.r.t. `image`. */
typedef struct EgsLossArgs {
  int channels, height, width;
  const float* image;      /* device [C,H,W]  the rendered image */
  const float* target;     /* device [C,H,W]  ground truth */
  const int* grad_mask;    /* device [H,W] int32 or NULL: the gradient is multiplied by (1 - mask),
                              the register_hook of trainers/train_static_bg.py:98 */
  float lambda_dssim;
} EgsLossArgs;

size_t egs_image_loss_workspace_bytes(int channels, int height, int width);

/* out_terms: device [3] = {L1 mean, SSIM mean, loss}. dL_dimage: device [C,H,W] = d loss / d image
 * (every element written), or NULL for the value only. Deterministic (fixed summation order).
 * workspace: device, 256-byte aligned, >= egs_image_loss_workspace_bytes(). */
int egs_image_loss(const EgsLossArgs* args, float* out_terms, float* dL_dimage, void* workspace,
                   size_t workspace_bytes, void* stream);

/* ---- introspection (tests / bench) ----------------------------------------------------- */

/* Views into the opaque blobs, for parity tests against the reference's intermediates
 * (SURVEY.md App. B). Pointers alias the blobs; nothing is copied. */
typedef struct EgsGeomView {
  const float* depths;         /* [P] */
  const float* means2D;        /* [P,2] */
  const float* conic_opacity;  /* [P,4] */
  const float* rgb_depth;      /* [P,4] rgb + depth packed for the blend kernels */
  const unsigned char* clamped;/* [P] bit c set = channel c clamped */
  const uint32_t* tiles_touched; /* [P] */
  const uint32_t* point_offsets; /* [P] inclusive scan */
} EgsGeomView;
typedef struct EgsBinningView {
  const uint64_t* keys_sorted;   /* [R]  valid only after egs_binning_keys() */
  const uint32_t* values_sorted; /* [R] */
} EgsBinningView;
typedef struct EgsImageView {
  const uint32_t* n_contrib;     /* [H*W] */
  const uint32_t* ranges;        /* [tiles,2] */
} EgsImageView;
int egs_geom_view(const void* geom_buffer, int P, EgsGeomView* out);
/* `num_rendered` here only selects the layout: pass egs_binning_capacity(size of the blob). */
int egs_binning_view(const void* binning_buffer, int num_rendered, int width, int height,
                     EgsBinningView* out);
int egs_image_view(const void* image_buffer, int width, int height, EgsImageView* out);

/* The sorted 64-bit key list of the reference (tile << 32 | depth bits, the sorted output of
 * cub::DeviceRadixSort::SortPairs, rasterizer_impl.cu:301-309), written into the binning blob where
 * egs_binning_view()'s keys_sorted points. No kernel of the path reads that list (the blends walk the
 * index list and the tile ranges), so forward does not write it; this call rebuilds it from the tile
 * ranges, the sorted index list and the depths for the parity tests. `num_rendered` selects the layout
 * like in egs_binning_view. */
int egs_binning_keys(const void* geom_buffer, int P, void* binning_buffer, int num_rendered,
                     const void* image_buffer, int width, int height, void* stream);

/* Process-wide switches for tests and A/B timing; the library never reads the environment. value NULL
 * or "" restores the default. Results never change (the parity tests run every setting):
 *   "sort"             "packed" | "pairs" | "pairs8"   force a binning path
 *   "presort"          "0" | "1"                       never / always sort the Gaussians front to back first
 *   "presort_scramble" "1"                             presort by keys unrelated to depth and raise its veto
 *   "blend_forward"    "v2"                            the two-pixels-per-thread forward blend
 *   "blend_backward"   "2" | "4"                       pixels per thread in the backward blend (default: by image size)
 *   "exchange_phases"  "a" | "b"                       timing only: run one phase of egs_exchange_sum_factored */
int egs_debug_option(const char* name, const char* value);

/* Built-in stage timing (the reference has no profiling hooks). While
 * enabled, every stage of egs_forward / egs_backward / egs_adam_step is bracketed by CUDA events
 * on the caller's stream. egs_profile_read waits for the recorded events, adds each stage's
 * elapsed milliseconds and launch count into ms[] / count[] (arrays of EGS_PROFILE_STAGES), and
 * forgets them. Stage i is named egs_profile_stage_name(i). */
#define EGS_PROFILE_STAGES 16
int egs_profile_enable(int on);
int egs_profile_read(double* ms, unsigned long long* count);
const char* egs_profile_stage_name(int stage);

/* Number of kernel launches issued by this library in this process since load (bench `gpu_launches`). */
unsigned long long egs_launch_count(void);
const char* egs_last_error(void);
int egs_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* EGS_RASTERIZER_H_ */
