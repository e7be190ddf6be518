/*
 * wheatsplat.h -- C ABI of libwheatsplat.so, the B200 (sm_100a) Gaussian-splatting hot path.
 *
 * This is the drop-in boundary for the three CUDA submodules Wheat3DGS calls
 * (paths below are into the reference tree, /root/reference/submodules/):
 *
 *   DGR = diff-gaussian-rasterization   pybind _C.rasterize_gaussians / _backward / mark_visible
 *                                       (DGR/ext.cpp:15-19, DGR/rasterize_points.h:18-70)
 *   FSR = flashsplat-rasterization      same names, + gt_mask / unique_label / num_obj
 *                                       (FSR/rasterize_points.h:18-46)
 *   KNN = simple-knn                    _C.distCUDA2 (KNN/ext.cpp:15-17, KNN/spatial.cu:15-26)
 *
 * Conventions
 *   - plain C: raw device pointers, ints, floats; no C++ types, no exceptions cross the ABI.
 *   - NULL pointer = "not provided" (the reference passes empty tensors whose data_ptr is
 *     nullptr, DGR/diff_gaussian_rasterization/__init__.py:200-210; tested at
 *     DGR/cuda_rasterizer/forward.cu:205,241).
 *   - the CALLER owns all memory (outputs and scratch).  The library never allocates device
 *     memory on the hot path.  Scratch sizes come from ws_*_bytes().  The reference instead
 *     grows three torch byte tensors through std::function callbacks
 *     (DGR/rasterize_points.cu:27-33,78-80); the forward is therefore split in two calls at
 *     exactly the point where the reference does its D2H read of num_rendered
 *     (DGR/cuda_rasterizer/rasterizer_impl.cu:282): ws_forward_prepare returns num_rendered,
 *     the caller sizes the binning scratch, ws_forward_render finishes.
 *   - every call takes the CUDA stream to run on (the reference uses the legacy default
 *     stream everywhere); calls are re-entrant per (device, stream) as long as scratch
 *     buffers are not shared.
 *   - return value: >= 0 success, < 0 error (WS_E_*); ws_last_error() gives the text.
 *   - matrices are the reference's: viewmatrix/projmatrix are float[16] stored transposed
 *     (row-vector convention), indexed m[0],m[4],m[8],m[12] for the first output component
 *     (DGR/cuda_rasterizer/auxiliary.h:58-66).
 */
#ifndef WHEATSPLAT_H_INCLUDED
#define WHEATSPLAT_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st *ws_stream_t;

#define WS_OK 0
#define WS_E_INVALID (-1)   /* bad argument (shape, NULL where required, both/neither of an either-or pair) */
#define WS_E_SCRATCH (-2)   /* scratch buffer too small */
#define WS_E_CUDA (-3)      /* a CUDA runtime call or kernel launch failed */
#define WS_E_UNSUPPORTED (-4)

#define WS_ABI_VERSION 3

/* Per-view constants = GaussianRasterizationSettings
 * (DGR/diff_gaussian_rasterization/__init__.py:160-172; FSR adds mask_grad, num_obj:
 * FSR/flashsplat_rasterization/__init__.py:201-215 -- num_obj travels in ws_flash). */
typedef struct ws_view {
    int32_t image_height;
    int32_t image_width;
    float tanfovx;
    float tanfovy;
    float scale_modifier;
    int32_t sh_degree;        /* active degree D, 0..3 */
    int32_t prefiltered;      /* callers always pass 0 (gaussian_renderer/__init__.py:51,143) */
    int32_t debug;            /* !=0: synchronise and check after every stage (auxiliary.h:166-173) */
    const float *bg;          /* device float[3] */
    const float *viewmatrix;  /* device float[16] */
    const float *projmatrix;  /* device float[16] */
    const float *campos;      /* device float[3] */
} ws_view;

/* The Gaussians = the tensor arguments of rasterize_gaussians (DGR/rasterize_points.h:18-38). */
typedef struct ws_gaussians {
    int32_t P;                   /* number of Gaussians, means3D is [P,3] */
    int32_t M;                   /* SH coefficients per Gaussian (sh is [P,M,3]); 0 if no SH */
    const float *means3D;        /* [P,3] */
    const float *shs;            /* [P,M,3] coefficient-major, RGB innermost, or NULL */
    const float *colors_precomp; /* [P,3] or NULL; exactly one of shs / colors_precomp */
    const float *opacities;      /* [P,1] post-sigmoid */
    const float *scales;         /* [P,3] post-exp, or NULL */
    const float *rotations;      /* [P,4] (w,x,y,z), NOT normalised in-kernel (forward.cu:127), or NULL */
    const float *cov3D_precomp;  /* [P,6] or NULL; exactly one of (scales,rotations) / cov3D_precomp */
    /* GaussianModel's getters folded into the kernels (scene/gaussian_model.py:101-121): */
    const float *shs_rest;       /* NULL, or features_rest [P,M-1,3]: then shs is features_dc [P,1,3] and the
                                    torch.cat of get_features (:113-116) is never materialised */
    int32_t raw_params;          /* != 0: opacities are logits (sigmoid in-kernel, :109-110), scales are
                                    log-scales (exp, :101-102), rotations unnormalised (normalize, :105-106);
                                    the backward returns gradients w.r.t. these raw tensors */
} ws_gaussians;

/* flashsplat additions (FSR/rasterize_points.cu:36-131).  Pass NULL for the plain DGR op. */
typedef struct ws_flash {
    int32_t num_obj;            /* used_count has num_obj+1 rows */
    const float *gt_mask;       /* [H,W] float object id per pixel (truncated to int); required */
    const int32_t *unique_label;/* [P] or NULL (FSR forward.cu:357-362) */
    int32_t *out_contrib_num;   /* [1,H,W], written for every pixel */
    float *used_count;          /* [num_obj+1,P]; ACCUMULATED into (+=): zero it for reference
                                   semantics, or keep one buffer across views to sum votes */
    float *image_xys;           /* [2,P]; caller pre-fills -1.0 (FSR/rasterize_points.cu:77) */
    float *gs_depth;            /* [P];   caller pre-fills -0.1 (FSR/rasterize_points.cu:78) */
} ws_flash;

/* Gradients = the 8 tensors rasterize_gaussians_backward returns
 * (DGR/rasterize_points.cu:207).  Every element of every non-NULL tensor is written
 * (zeros included), so the caller may allocate them uninitialised -- the reference's
 * torch::zeros fills (rasterize_points.cu:153-162) are not needed. */
typedef struct ws_grads {
    float *dL_dmeans2D;   /* [P,3]  (z = 0) */
    float *dL_dcolors;    /* [P,3]  grad of colors_precomp / of the SH-evaluated RGB */
    float *dL_dopacity;   /* [P,1] */
    float *dL_dmeans3D;   /* [P,3] */
    float *dL_dcov3D;     /* [P,6] */
    float *dL_dsh;        /* [P,M,3] or NULL when M == 0 */
    float *dL_dscales;    /* [P,3] */
    float *dL_drotations; /* [P,4] */
    float *dL_dsh_rest;   /* [P,M-1,3] when ws_gaussians.shs_rest is given (dL_dsh is then [P,1,3]), else NULL */
} ws_grads;

int ws_abi_version(void);
const char *ws_last_error(void);

/* Scratch sizes in bytes.  geom depends on P, image on W*H, binning on num_rendered. */
size_t ws_geom_bytes(int32_t P);
size_t ws_image_bytes(int32_t width, int32_t height);
size_t ws_binning_bytes(int64_t num_rendered, int32_t width, int32_t height);

/* Forward, phase A: preprocessCUDA + prefix sum of tiles_touched
 * (rasterizer_impl.cu:248-282).  Writes radii[P] (int32, every element) and, if fl != NULL,
 * image_xys / gs_depth.  Blocks the calling thread until the prefix sum has delivered num_rendered (the reference's
 * cudaMemcpy at :282): the count arrives through a mailbox in mapped pinned memory written from a side stream, so
 * `stream` itself is NOT drained -- the depth sort this call queued behind the prefix sum keeps running while the
 * caller sizes the binning scratch.  Not capturable into a CUDA graph (neither is the reference's copy).
 * Returns num_rendered >= 0. */
int64_t ws_forward_prepare(const ws_view *view, const ws_gaussians *g, int32_t *radii, void *geom,
                           size_t geom_bytes, const ws_flash *fl, ws_stream_t stream);

/* Forward, phase B: duplicateWithKeys, radix sort, identifyTileRanges, render
 * (rasterizer_impl.cu:284-339).  out_color [3,H,W], out_depth [1,H,W], out_alpha [1,H,W];
 * every pixel is written. */
int ws_forward_render(const ws_view *view, const ws_gaussians *g, const int32_t *radii, void *geom,
                      void *binning, size_t binning_bytes, void *image, size_t image_bytes,
                      int64_t num_rendered, float *out_color, float *out_depth, float *out_alpha,
                      const ws_flash *fl, ws_stream_t stream);

/* Backward (rasterizer_impl.cu:343-446).  geom/binning/image are the buffers of the matching
 * forward; out_alpha is the forward's alpha output (T_final = 1 - alpha, backward.cu:463). */
int ws_backward(const ws_view *view, const ws_gaussians *g, const int32_t *radii, void *geom,
                void *binning, void *image, int64_t num_rendered, const float *out_alpha,
                const float *dL_dcolor, const float *dL_ddepth, const float *dL_dalpha,
                const ws_grads *grads, const ws_flash *fl, ws_stream_t stream);

/* markVisible (DGR/rasterize_points.cu:210-229): present[i] = (view-space z > 0.2). */
int ws_mark_visible(int32_t P, const float *means3D, const float *viewmatrix, const float *projmatrix,
                    uint8_t *present, ws_stream_t stream);

/* simple-knn distCUDA2 (KNN/simple_knn.cu:186-222): mean squared distance to the 3 nearest
 * neighbours.  Synchronises `stream` once (the reference syncs twice for its bbox). */
size_t ws_knn_bytes(int32_t P);
int ws_knn_dist2(int32_t P, const float *points, float *mean_dists, void *scratch, size_t scratch_bytes,
                 ws_stream_t stream);

/* ws_knn_dist2 for callers that cannot provide scratch.  No counterpart in the reference (its distCUDA2 allocates
 * through thrust, KNN/simple_knn.cu:195-217); this entry point exists because ws_knn_dist2 asks the caller for
 * ws_knn_bytes(P) bytes of device memory and some callers have no means of getting them:
 *   - a binding that only speaks this C ABI (ctypes / cgo / JNI without a CUDA runtime of its own) and is handed the
 *     `points` and `mean_dists` pointers by another component: it can pass pointers on but has no cudaMalloc to call;
 *   - a plug-in inside a process whose device memory is owned by one pool (torch's caching allocator configured to
 *     take the whole card, a renderer with a fixed budget): a second allocator's cudaMalloc fails or fragments it;
 *   - code on a path where allocating is forbidden because it synchronises the device (cudaMalloc / cudaFree do).
 * The scratch is a fixed array inside the library image instead: 8 MiB of device memory on EVERY GPU that loads
 * libwheatsplat.so, held until the process exits whether or not this function is ever called.  It serves at most
 * WS_KNN_STATIC_MAX_POINTS points; a larger P returns WS_E_SCRATCH (ws_last_error says to use ws_knn_dist2) and
 * launches nothing.  The function calls no cudaMalloc, cudaFree or other allocating routine.  Results are bit-identical
 * to ws_knn_dist2 on the same input.
 * First call on a device: looks up the array's device address (cudaGetSymbolAddress, with that device current) and
 * remembers it per device ordinal; make this first call outside any stream capture.  Later calls only launch.
 * Sharing: every call on a device uses the same array and finds the previous call's contents in it; the function
 * clears the words that must start at zero itself, on `stream`.  Two calls whose kernels can overlap on the GPU (two
 * streams, two host threads, two processes' worth of streams in one context) race on the array and give wrong
 * distances: order them on one stream or with events.  The current device must be the one `points` lives on. */
#define WS_KNN_STATIC_MAX_POINTS 300000
int ws_knn_dist2_static(int32_t P, const float *points, float *mean_dists, ws_stream_t stream);

/* multi_instance_opt (run_3d_seg.py:54-72): labels[i,j] = argmax over normalize([sum_j - A_ij, A_ij])
 * with gamma added to row 0.  all_contrib float[num_rows,P] -> labels uint8[num_rows,P]. */
int ws_multi_instance_opt(int32_t num_rows, int32_t P, const float *all_contrib, float gamma,
                          uint8_t *labels, ws_stream_t stream);

/* Training loss either side of the rasterizer (replaces utils/loss_utils.py:17-62
 * as used by train_vanilla_3dgs.py:77-80): per image plane (= batch x channel, img/gt float
 * [planes,H,W]) the sum of the SSIM map (11x11 Gaussian window, sigma 1.5, zero padding) and the
 * sum of |img - gt|.  sums: float[planes][2] = {sum ssim, sum |img-gt|}.  The forward leaves the
 * three partial-derivative maps the backward convolves in `scratch` (ws_loss_bytes), which the
 * caller keeps until the backward.
 * Backward: dL_dimg = a * d(sum ssim)/dimg + b * d(sum |img-gt|)/dimg with {a, b} read from the
 * device array coef (float[2], or float[planes][2] when per_plane != 0). */
size_t ws_loss_bytes(int32_t planes, int32_t height, int32_t width);
int ws_l1_ssim_forward(int32_t planes, int32_t height, int32_t width, const float *img, const float *gt,
                       float *sums, void *scratch, size_t scratch_bytes, ws_stream_t stream);
int ws_l1_ssim_backward(int32_t planes, int32_t height, int32_t width, const float *img, const float *gt,
                        const float *coef, int32_t per_plane, const void *scratch, size_t scratch_bytes,
                        float *dL_dimg, ws_stream_t stream);

/* Batched eval_obj_labels (replaces utils/wheatgs_helper.py:245-271 = eval_wheatgs.py:29-64):
 * the segmentation image of one view from the per-object Gaussian sets, with ONE preprocess, ONE binning pass
 * keyed by (object, tile), ONE blend launch and ONE merge kernel instead of one flashsplat_render + two
 * image copies to the CPU per object.  Object sets are CSR lists by Gaussian: Gaussian g belongs to the objects
 * member_obj[member_start[g] .. member_start[g+1]) (ascending, each in 1..num_rows-1; row 0 = background is
 * never rendered, wheatgs_helper.py:254).  Three calls, split at the two counts the caller sizes memory by:
 *   ws_objeval_prepare -> number of (object, tile, Gaussian) instances R'  (geom as ws_forward_prepare)
 *   ws_objeval_bin     -> number of non-empty (object, tile) jobs J         (binning: ws_binning_bytes(R');
 *                                                                            scratch: ws_objeval_bytes(...))
 *   ws_objeval_render  -> pred_mask float[H*W]: object id per pixel, 0 = none (job_out: J*256*2 floats)
 * Per-object alpha and depth are bit-identical to the subset renders; the front/behind decision compares the
 * two means accumulated in double and rounded to float. */
typedef struct ws_objsets {
    int32_t num_rows;             /* rows of all_obj_labels (objects 1..num_rows-1) */
    const int32_t *member_start;  /* device int32[P+1] */
    const int32_t *member_obj;    /* device int32[member_start[P]] */
} ws_objsets;
size_t ws_objeval_bytes(int32_t num_rows, int32_t width, int32_t height, int64_t num_instances);
int64_t ws_objeval_prepare(const ws_view *view, const ws_gaussians *g, const ws_objsets *sets, int32_t *radii,
                           void *geom, size_t geom_bytes, ws_stream_t stream);
int64_t ws_objeval_bin(const ws_view *view, const ws_gaussians *g, const ws_objsets *sets, const int32_t *radii,
                       void *geom, void *binning, size_t binning_bytes, void *scratch, size_t scratch_bytes,
                       int64_t num_instances, ws_stream_t stream);
int ws_objeval_render(const ws_view *view, const ws_objsets *sets, int32_t P, void *geom, void *binning, void *scratch,
                      size_t scratch_bytes, int64_t num_instances, int64_t num_jobs, float *job_out,
                      size_t job_out_bytes, float *pred_mask, ws_stream_t stream);

/* Optimiser step and densification bookkeeping either side of the backward.
 * ws_adam_step = torch.optim.Adam(...).step() as GaussianModel.training_setup configures it
 * (scene/gaussian_model.py:167-186; called at train_vanilla_3dgs.py:110-112): betas/eps shared, one
 * learning rate and step count per parameter group, no weight decay, no amsgrad.  All groups (at most 8)
 * are updated by ONE kernel launch, in place.  Tensors are contiguous float; param / exp_avg / exp_avg_sq
 * 16-byte aligned, grad 4-byte aligned (16-byte aligned gradients are loaded as vectors). */
typedef struct ws_adam_group {
    float *param;        /* [numel] updated in place */
    const float *grad;   /* [numel] */
    float *exp_avg;      /* [numel] optimizer.state[p]["exp_avg"] */
    float *exp_avg_sq;   /* [numel] optimizer.state[p]["exp_avg_sq"] */
    int64_t numel;
    double lr;           /* param_group["lr"] */
    int64_t step;        /* state["step"] AFTER the increment of this call (>= 1) */
} ws_adam_group;
int ws_adam_step(int32_t n_groups, const ws_adam_group *groups, double beta1, double beta2, double eps,
                 ws_stream_t stream);
/* train_vanilla_3dgs.py:100-103 + GaussianModel.add_densification_stats (scene/gaussian_model.py:462-464):
 * for every Gaussian with radii > 0: max_radii2D = max(max_radii2D, radii),
 * xyz_gradient_accum += |dL_dmeans2D[:, :2]|, denom += 1.  radii int32[P], dL_dmeans2D float[P,3],
 * the three accumulators float[P]. */
int ws_densify_stats(int32_t P, const int32_t *radii, const float *dL_dmeans2D, float *max_radii2D,
                     float *xyz_gradient_accum, float *denom, ws_stream_t stream);

/* GaussianModel.densify_and_prune (scene/gaussian_model.py:440-458: densify_and_clone :420-438, densify_and_split
 * :392-418, prune_points :327-352, with cat_tensors_to_optimizer :354-374 and _prune_optimizer :310-325) as ONE plan
 * and ONE gather instead of three rounds of masked indexing + torch.cat over every parameter and Adam moment.
 * The surviving rows come out in the reference's order:
 *   [originals that are neither split nor pruned | their clones | first split children | second split children]
 * ws_densify_plan decides, per Gaussian i of the P present: grads = xyz_gradient_accum/denom (NaN -> 0);
 *   clone  iff |grads| >= max_grad and max(exp(scaling)) <= dense_scale   (= percent_dense * extent)
 *   split  iff  grads  >= max_grad and max(exp(scaling)) >  dense_scale
 *   pruned iff sigmoid(opacity) < min_opacity, or (use_screen != 0 and (0 > max_screen_size or
 *          max(scale) > big_scale (= 0.1 * extent)))    -- max_radii2D is all zeros when the reference tests it
 *          (densification_postfix :389 re-creates it before prune_points reads it), so it is not an input;
 *   a split child has its parent's opacity and scale*(1/1.6) (new_scaling :405), both children share the verdict.
 * flags uint8[P]: bit0 original kept, bit1 clone kept, bit2 children kept, bit3 split-selected.
 * counts_host[4] = {kept originals, kept clones, kept children per copy, split-selected S}; rows out =
 * counts[0] + counts[1] + 2*counts[2].  The call synchronises the stream (the caller sizes the outputs and draws
 * the 2S x 3 normal samples from it, as the reference's boolean indexing does many times over).
 * scratch: ws_densify_plan_bytes(P) bytes, kept for ws_densify_gather. */
size_t ws_densify_plan_bytes(int32_t P);
int ws_densify_plan(int32_t P, const float *xyz_gradient_accum, const float *denom, const float *opacity,
                    const float *scaling, float max_grad, float min_opacity, float dense_scale, float big_scale,
                    int32_t use_screen, float max_screen_size, uint8_t *flags, void *scratch, size_t scratch_bytes,
                    uint32_t *counts_host, ws_stream_t stream);
/* One array of the model that travels with the Gaussians: rows of `width` 32-bit words. */
#define WS_ROW_COPY 0      /* survivors, clones and children copy the source row (f_dc, f_rest, opacity, rotation, _which_object) */
#define WS_ROW_MOMENT 1    /* survivors copy, clones and children get 0 (exp_avg, exp_avg_sq: :361-362) */
#define WS_ROW_XYZ 2       /* children: build_rotation(rotation[i]) @ samples[row] + xyz[i] (:401-404), width 3 */
#define WS_ROW_SCALING 3   /* children: log(exp(scaling[i]) * (1/1.6)) (:405), width 3 */
typedef struct {
    const void *src;   /* [P, width] */
    void *dst;         /* [rows out, width] */
    int32_t width;
    int32_t role;      /* WS_ROW_* */
} ws_row_array;
/* Writes every array of the model at its new size.  samples float[2S,3]: the reference's
 * torch.normal(mean=0, std=get_scaling[selected].repeat(2,1)) (:399-400), row r < S for the first child of the
 * r-th split-selected Gaussian, S + r for the second; rotation float[P,4] raw.  src_row_out int32[rows out] (or NULL):
 * the source Gaussian of every output row. */
int ws_densify_gather(int32_t P, const uint8_t *flags, const void *scratch, size_t scratch_bytes,
                      const uint32_t *counts_host, const float *rotation, const float *samples, int32_t n_arrays,
                      const ws_row_array *arrays, int32_t *src_row_out, ws_stream_t stream);
/* GaussianModel.reset_opacity (scene/gaussian_model.py:234-237 + replace_tensor_to_optimizer :295-308):
 * opacity = inverse_sigmoid(min(sigmoid(opacity), 0.01)) in place, exp_avg = exp_avg_sq = 0 (either may be NULL). */
int ws_reset_opacity(int32_t P, float *opacity, float *exp_avg, float *exp_avg_sq, ws_stream_t stream);

/* Exchange step of data-parallel training: in-place SUM over ranks of a float buffer that lives in
 * symmetric memory (every rank has mapped every peer's copy), by one kernel on the caller's stream.  Rank r
 * reduces the r-th slice: with multicast_ptr != NULL (NVSwitch multicast address aliasing all copies) through
 * multimem.ld_reduce / multimem.st, else with loads from / stores to peer_ptrs[0..world) in rank order.  All
 * ranks call it with the same numel (a multiple of 4), after a barrier that orders every rank's writes to the
 * buffer before it, and follow it with a barrier before reading the sums.  num_ctas <= 0: one CTA per SM. */
int ws_allreduce_sum_f32(int32_t rank, int32_t world, int64_t numel, void *multicast_ptr, void *const *peer_ptrs,
                         int32_t num_ctas, ws_stream_t stream);
/* Tuning knobs of the multimem kernel (process-wide): independent 16-byte ld_reduce in flight per thread (2, 4 or 8;
 * default 4) and threads per CTA (default 512). */
int ws_allreduce_tune(int32_t unroll, int32_t threads);
/* EXPERIMENTAL, off by default and not part of the stable ABI (measured slower than the switch-only kernel at every
 * setting on 8 B200s, profiles/r02_exchange_split_sweep_8gpu.jsonl; kept for re-measurement on other fabrics).
 * With both a multicast address and peer pointers: the first `multicast_permille` thousandths of every rank's slice are
 * reduced through the switch (multimem), the rest at the same time by `peer_ctas` extra CTAs with peer loads and
 * stores (the in-switch reduction runs at about half the link rate; the peer path fills the links).  1000 = switch
 * only (default).  Process-wide. */
int ws_allreduce_split(int32_t multicast_permille, int32_t peer_ctas);

/* Exchange + optimiser step of data-parallel training fused into ONE kernel (reduce-scatter, Adam, all-gather):
 * rank r owns the r-th slice of the flat parameter space [0, numel).  It reads the slice's gradient SUM over ranks
 * (multimem.ld_reduce from grad_multicast, else peer loads from grad_peers[0..world) in rank order), applies
 * torch.optim.Adam's step (the element arithmetic of ws_adam_step) with exp_avg / exp_avg_sq indexed like the
 * parameters (only the slice is touched), and writes the new parameters into every rank's replica
 * (multimem.st to param_multicast, else stores to param_peers[0..world)).  param_peers is always required (the
 * local replica is read through param_peers[rank]).  A segment = one param group: elements [begin, end) of the
 * parameter space, their gradient at grad_offset in the gradient buffer, and the group's lr; segments tile
 * [0, numel) in order, everything in multiples of 4 floats.  `step` = state["step"] after the increment.
 * Bracket with the symmetric-memory barrier like ws_allreduce_sum_f32. */
typedef struct {
    int64_t begin, end;
    int64_t grad_offset;
    double lr;
} ws_shard_segment;
int ws_sharded_adam_step(int32_t rank, int32_t world, int64_t numel, int32_t n_segments, const ws_shard_segment *segments,
                         void *grad_multicast, void *const *grad_peers, void *param_multicast, void *const *param_peers,
                         float *exp_avg, float *exp_avg_sq, int64_t step, double beta1, double beta2, double eps,
                         int32_t num_ctas, ws_stream_t stream);

/* The stable LSD radix sort the rasterizer uses (replaces cub::DeviceRadixSort::SortPairs,
 * rasterizer_impl.cu:304-309), exposed for parity tests: sorts (u64 key, u32 value) pairs on
 * key bits [0, end_bit).  Results land in keys_out / vals_out; *_in are clobbered. */
size_t ws_sort_bytes(int64_t n);
int ws_sort_pairs_u64(int64_t n, uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in,
                      uint32_t *vals_out, int32_t end_bit, void *scratch, size_t scratch_bytes,
                      ws_stream_t stream);
/* Same sort on (u32 key, u32 value) pairs: the form depth-first binning uses (Gaussians by depth
 * bits, then instances by tile id). */
int ws_sort_pairs_u32(int64_t n, uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in,
                      uint32_t *vals_out, int32_t end_bit, void *scratch, size_t scratch_bytes,
                      ws_stream_t stream);

/* How the instance list is ordered.  0 (default): depth-first binning -- the P Gaussians are sorted
 * by depth bits once, instances are emitted in that order with their tile id as the only key and
 * stably sorted on it.  1: the reference's scheme (duplicateWithKeys' 64-bit tile|depth keys, one
 * sort over all instances, rasterizer_impl.cu:70-111,304-309).  -1: take it from the environment
 * (WS_BINNING=keys64).  Both produce the same point_list, ranges and images; a forward, its
 * backward and its ws_debug_fetch calls must run under the same mode. */
int ws_set_binning_mode(int32_t mode);

/* Parity/debug access to the internal arrays of a finished forward (device -> device copy
 * into caller memory).  name is one of: "depths" f32[P], "means2D" f32[P,2],
 * "conic_opacity" f32[P,4], "rgb" f32[P,3], "cov3D" f32[P,6], "tiles_touched" u32[P],
 * "point_list_keys" u64[R], "point_list" u32[R], "ranges" u32[T,2], "n_contrib" u32[H,W].
 * Returns bytes written or < 0. */
int64_t ws_debug_fetch(const char *name, int32_t P, int32_t width, int32_t height, int64_t num_rendered,
                       void *geom, void *binning, void *image, void *dst, size_t dst_bytes,
                       ws_stream_t stream);

/* Number of kernels this library launched since load (bench.py's gpu_launches counter). */
int64_t ws_launch_count(void);

/* Stage timing for bench.py's roofline figures.  While enabled, ws_forward_prepare /
 * ws_forward_render / ws_backward bracket every stage with CUDA events recorded on the caller's
 * stream.  ws_prof_read synchronises on the recorded events, adds the elapsed milliseconds and the
 * number of brackets per stage into ms[WS_PROF_NSTAGES] / calls[WS_PROF_NSTAGES] and forgets them.
 * Stages: 0 preprocess, 1 scan, 2 emit_keys, 3 sort (histogram + onesweep passes), 4 ranges_pack,
 * 5 render_fwd, 6 render_bwd, 7 preprocess_bwd. */
#define WS_PROF_NSTAGES 8
void ws_prof_enable(int32_t on);
int ws_prof_read(double *ms, int64_t *calls);
const char *ws_prof_stage_name(int32_t stage);

/* Work counters of a finished forward, for the blend kernels' flop figures:
 * counts[0] = N_eval = sum over pixels of n_contrib (pairs traversed up to the last contributor),
 * counts[1] = N_contrib = pairs among those with power <= 0 and alpha >= 1/255.
 * counts[2] = (warp, instance) pairs the 8x8 quadrant masks let through (4 x instances without the cull),
 * counts[3] = instances whose mask is empty (never fetched by the blend kernels).
 * counts is device uint64[4], zeroed by the call. */
int ws_debug_count_pairs(int32_t P, int32_t width, int32_t height, void *geom, void *binning, void *image,
                         int64_t num_rendered, uint64_t *counts, ws_stream_t stream);

/* FP32 issue-rate probe (bench.py's measured roofline denominator for the blend kernels): `ctas` CTAs of 256
 * threads each run iters x 32 dependent-chain FMAs (8 independent chains, three register operands); packed != 0
 * issues fma.rn.f32x2 (FFMA2) as the blend kernels do, else scalar FFMA.  out: device float[ctas * 256].
 * Returns the number of floating-point operations issued (or a negative status). */
int64_t ws_debug_fma_peak(int32_t packed, int32_t iters, int32_t ctas, float *out, ws_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* WHEATSPLAT_H_INCLUDED */
