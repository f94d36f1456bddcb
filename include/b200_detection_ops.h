/*
 * b200_detection_ops.h -- C-ABI of the B200-native detection-head hot path.
 *
 * This is the drop-in boundary beneath the PyTorch dispatcher shim
 * (cpu_vision_b200/csrc/torch_shim.cpp).  Every entry point takes plain
 * pointers, sizes and a cudaStream_t passed as void*; no torch types appear.
 * Nothing here allocates device memory unless the name ends in `_host`
 * (those own a private staging arena, see below), with one exception: the
 * first *_backward call on a device creates a 2 MB ring of task plans for the
 * persistent backward kernel; b200det_host_release() frees it.
 *
 * Each function names the reference interface it replaces; paths are relative
 * to the reference tree (kaisun24/CPU-Vision, a torchvision 0.20 fork).
 *
 * Return value: 0 on success, a positive cudaError_t from the CUDA runtime,
 * or a negative B200DET_E* code for argument errors.  b200det_last_error()
 * returns a thread-local message for the last non-zero return.
 */
#ifndef B200_DETECTION_OPS_H
#define B200_DETECTION_OPS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200DET_VERSION 100 /* 0.1.0 */
#define B200DET_MAX_LEVELS 8

/* element type of feature maps / rois / boxes.  Arithmetic is fp32 for F32,
 * F16 and BF16 I/O and fp64 for F64 (the reference dispatches double, float
 * and half: torchvision/csrc/ops/cuda/roi_align_kernel.cu:376). */
typedef enum {
  B200DET_F32 = 0,
  B200DET_F16 = 1,
  B200DET_BF16 = 2,
  B200DET_F64 = 3
} b200det_dtype;

/* memory order of a logically [N, C, H, W] tensor */
typedef enum {
  B200DET_NCHW = 0, /* torch.contiguous_format   */
  B200DET_NHWC = 1  /* torch.channels_last (the B200-native fast path) */
} b200det_layout;

typedef enum {
  B200DET_EINVAL = -1,   /* bad argument (null pointer, negative size, ...) */
  B200DET_ENOSUP = -2,   /* unsupported dtype/layout combination           */
  B200DET_EWORKSPACE = -3 /* workspace too small                           */
} b200det_status;

int b200det_version(void);
const char* b200det_last_error(void);

/* ------------------------------------------------------------------------
 * RoIAlign
 * --------------------------------------------------------------------- */

/* Forward.  Replaces roi_align_forward_kernel
 * (torchvision/csrc/ops/cuda/roi_align_kernel.cu:334-394; CPU twin
 * torchvision/csrc/ops/cpu/roi_align_kernel.cpp:282-328).
 *   input  : device, logical [N, C, H, W] in `in_layout`
 *   rois   : device, [K, 5] rows (batch_index, x1, y1, x2, y2) in the
 *            ARITHMETIC type of `dtype`: float for F32/F16/BF16 features,
 *            double for F64.  (The reference demands rois.dtype ==
 *            input.dtype, :350, and its autocast wrapper widens both to fp32,
 *            autocast/roi_align_kernel.cpp:12-31; the shim widens 16-bit rois
 *            the same way, so bf16/fp16 features keep full-precision boxes.)
 *   output : device, logical [K, C, PH, PW] in `out_layout`; every element is
 *            written (no pre-zeroing needed, unlike the reference's at::zeros)
 * spatial_scale is converted to the arithmetic type before use, like the
 * reference kernel argument (`const T spatial_scale`, :72). */
int b200det_roi_align_forward(
    const void* input, int in_layout, const void* rois, void* output,
    int out_layout, int dtype, int64_t N, int64_t C, int64_t H, int64_t W,
    int64_t K, int64_t PH, int64_t PW, double spatial_scale,
    int sampling_ratio, int aligned, void* stream);

/* Backward.  Replaces roi_align_backward_kernel
 * (torchvision/csrc/ops/cuda/roi_align_kernel.cu:396-466; CPU twin
 * torchvision/csrc/ops/cpu/roi_align_kernel.cpp:330-386).
 *   grad         : device, logical [K, C, PH, PW] with ARBITRARY element
 *                  strides grad_strides[4] = (k, c, ph, pw), as the reference
 *                  honours (:253-256, :435-438)
 *   grad_input   : device, logical [N, C, H, W] in `gi_layout`
 *   flags        : B200DET_BWD_ZERO_INIT  zero-fill grad_input in this call (fused
 *                    with the scatter; the reference's at::zeros, :427) -- without
 *                    it the call accumulates into what is there;
 *                  B200DET_BWD_DETERMINISTIC  use the tile-accumulating kernel
 *                    (no atomics, fixed summation order: bitwise reproducible;
 *                    channels_last, C % 128 == 0, sampling_ratio 1 or 2 only,
 *                    B200DET_ENOSUP otherwise).  Without it gradients are
 *                    accumulated with red.global.add in non-deterministic order,
 *                    like the reference (:440). */
#define B200DET_BWD_ZERO_INIT 1
#define B200DET_BWD_DETERMINISTIC 2
int b200det_roi_align_backward(
    const void* grad, const int64_t* grad_strides, const void* rois,
    void* grad_input, int gi_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, int sampling_ratio, int aligned, int flags,
    void* stream);

/* FPN level mapper.  Replaces LevelMapper.__call__
 * (torchvision/ops/poolers.py:73-84):
 *   lvl = floor(canonical_level + log2(sqrt(area) / canonical_scale) + eps)
 *   clamp to [k_min, k_max], minus k_min.
 *   rois   : device [K, 5] (column 0 ignored), float (arithmetic type rule as
 *            above; F64 is not supported here)
 *   levels : device int32 [K] */
int b200det_fpn_level_map(
    const void* rois, int dtype, int64_t K, int k_min, int k_max,
    int canonical_scale, int canonical_level, double eps, int32_t* levels,
    void* stream);

/* MultiScaleRoIAlign forward in ONE launch sequence.  Replaces the per-level
 * loop of _multiscale_roi_align (torchvision/ops/poolers.py:146-227): level
 * mapping, per-level where/gather, roi_align, and the index_put scatter back
 * into original RoI order.
 *   feats[l]   : device, logical [N, C, heights[l], widths[l]] in `layout`
 *   scales[l]  : spatial scale of level l (poolers.py:98-134)
 *   levels     : device int32 [K] scratch/out; filled by this call through
 *                b200det_fpn_level_map unless levels_given != 0
 *   output     : device, logical [K, C, PH, PW] in `out_layout`, RoI order kept
 * The reference calls roi_align with aligned=False here (poolers.py:204-210);
 * `aligned` is still a parameter so both are reachable. */
int b200det_msroi_align_forward(
    const void* const* feats, const int64_t* heights, const int64_t* widths,
    const double* scales, int num_levels, int layout, const void* rois,
    int32_t* levels, int levels_given, void* output, int out_layout, int dtype,
    int64_t N, int64_t C, int64_t K, int64_t PH, int64_t PW,
    int sampling_ratio, int aligned, int k_min, int k_max,
    int canonical_scale, int canonical_level, double eps, void* stream);

/* MultiScaleRoIAlign forward FUSED with the all-gather of the pooled rows (a "push" gather over NVLink peer
 * memory).  Replaces the pair the reference's multi-GPU recipes run after torchvision.ops on every rank,
 *   out = MultiScaleRoIAlign(...)                       (torchvision/ops/poolers.py:289-327)
 *   all_gather(out)                                     (references/detection/utils.py:82-106, dist.all_gather),
 * with ONE kernel: every pooled bin is stored to `output` and, in the same instruction stream, to the same row of
 * `peer_outputs[i]` -- the slice of THIS rank inside the gather buffer of peer GPU i, mapped into this process
 * (b200det_peer_open below, or any peer-accessible pointer) -- so the rows travel while the next ones are computed
 * and no separate collective pass reads them again.
 *   peer_outputs : host array of num_peers (<= 7) device pointers, each 16-byte aligned, each [K, C, PH, PW] in
 *                  `out_layout` like `output`
 * Only the channels_last fast path pushes (C % 128 == 0, sampling_ratio 1 or 2, pooled_width * sampling_ratio <= 32,
 * fp32 / fp16 / bf16, NHWC features and outputs); anything else returns B200DET_ENOSUP.  The caller orders the
 * ranks around the call (a barrier before the peers' buffers are overwritten and one after the kernel, before anybody
 * reads a gathered row): cpu_vision_b200/sharding.py PushGather does it with two stream-ordered 4-byte all-reduces. */
int b200det_msroi_align_forward_push(
    const void* const* feats, const int64_t* heights, const int64_t* widths,
    const double* scales, int num_levels, int layout, const void* rois,
    int32_t* levels, int levels_given, void* output, void* const* peer_outputs,
    int num_peers, int out_layout, int dtype, int64_t N, int64_t C, int64_t K,
    int64_t PH, int64_t PW, int sampling_ratio, int aligned, int k_min,
    int k_max, int canonical_scale, int canonical_level, double eps,
    void* stream);

/* Peer-visible device memory for the push gather (CUDA IPC; the reference has no counterpart, NCCL owns its
 * buffers there).  alloc / free: a cudaMalloc of its own, so that the handle covers exactly the buffer.  export:
 * writes the B200DET_PEER_HANDLE_BYTES opaque bytes another process of this node passes to open, which maps the
 * buffer into the caller's address space on its current device (peer access over NVLink is enabled lazily).
 * close before the owner frees.  These are set-up calls: they allocate and synchronise. */
#define B200DET_PEER_HANDLE_BYTES 64
int b200det_peer_alloc(size_t bytes, void** ptr);
int b200det_peer_free(void* ptr);
int b200det_peer_export(const void* ptr, void* handle);
int b200det_peer_open(const void* handle, void** ptr);
int b200det_peer_close(void* ptr);

/* MultiScaleRoIAlign backward: the autograd transpose of the call above.
 *   grad_feats[l] : device, logical [N, C, heights[l], widths[l]] in `layout`,
 *                   (flags as in b200det_roi_align_backward)
 *   levels        : device int32 [K], as produced by the forward */
int b200det_msroi_align_backward(
    const void* grad, const int64_t* grad_strides, const void* rois,
    const int32_t* levels, void* const* grad_feats, const int64_t* heights,
    const int64_t* widths, const double* scales, int num_levels, int layout,
    int dtype, int64_t N, int64_t C, int64_t K, int64_t PH, int64_t PW,
    int sampling_ratio, int aligned, int flags, void* stream);

/* ------------------------------------------------------------------------
 * NMS
 * --------------------------------------------------------------------- */

typedef enum {
  B200DET_BNMS_AUTO = -1,   /* reference rule for CUDA tensors: vanilla when
                               4*N > 20000 (torchvision/ops/boxes.py:72)     */
  B200DET_BNMS_COORD_TRICK = 0, /* _batched_nms_coordinate_trick, boxes.py:78-95 */
  B200DET_BNMS_VANILLA = 1      /* _batched_nms_vanilla, boxes.py:98-112        */
} b200det_bnms_mode;

/* Bytes of device scratch b200det_nms / b200det_batched_nms need for N boxes. */
size_t b200det_nms_workspace_bytes(int64_t N);

/* Greedy NMS, everything on the device (no D2H round trip, unlike the
 * reference's mask.to(kCPU) + host scan).  Replaces nms_kernel
 * (torchvision/csrc/ops/cuda/nms_kernel.cu:80-165; CPU twin
 * torchvision/csrc/ops/cpu/nms_kernel.cpp:9-108).
 *   boxes    : device [N, 4] (x1, y1, x2, y2); scores: device [N]; same dtype
 *   keep     : device int64 [N] capacity; the first *num_keep entries are the
 *              kept ORIGINAL indices in descending-score order (stable ties)
 *   num_keep : device int64 [1]
 * IoU is evaluated in fp32 as (inter / ((Sa + Sb) - inter)) > (float)thr,
 * like devIoU (nms_kernel.cu:16-29). */
int b200det_nms(
    const void* boxes, const void* scores, int dtype, int64_t N,
    double iou_threshold, int64_t* keep, int64_t* num_keep, void* workspace,
    size_t workspace_bytes, void* stream);

/* b200det_nms without ANY allocation and without a caller workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, result and launch sequence as b200det_nms;
 * the scratch is a fixed array in the library's device image (B200DET_NMS_STATIC_BYTES per GPU that loads the
 * library, held until exit), which limits N to B200DET_NMS_STATIC_MAX_BOXES: a larger N returns B200DET_EINVAL
 * with a message.  The call performs no cudaMalloc / cudaFree and calls no allocating library routine.  All
 * calls on a device share the array (it keeps the last call's contents; nothing in it has to start at zero):
 * calls that overlap on the GPU race -- keep them on one stream or order them with events. */
#define B200DET_NMS_STATIC_MAX_BOXES 8192
#define B200DET_NMS_STATIC_BYTES (10 * 1024 * 1024)
int b200det_nms_static(
    const void* boxes, const void* scores, int dtype, int64_t N,
    double iou_threshold, int64_t* keep, int64_t* num_keep, void* stream);

/* Class-aware NMS in one pass.  Replaces batched_nms
 * (torchvision/ops/boxes.py:44-112).
 *   idxs : device int64 [N] category of each box (the per-class compare of the VANILLA strategy and the
 *          class grouping of b200det_nms_segmented use the low 32 bits: ids must be distinct modulo 2^32)
 *   mode : b200det_bnms_mode.  COORD_TRICK reproduces the fp32 rounding of
 *          `boxes + idxs.to(boxes) * (boxes.max() + 1)`; VANILLA compares the
 *          unshifted boxes class by class. */
int b200det_batched_nms(
    const void* boxes, const void* scores, const int64_t* idxs, int dtype,
    int64_t N, double iou_threshold, int mode, int64_t* keep,
    int64_t* num_keep, void* workspace, size_t workspace_bytes, void* stream);

/* Segmented (per-image) class-aware NMS: S independent box sets in ONE launch
 * sequence, every set scanned by its own CTA.  Replaces the per-image Python
 * loops around batched_nms in RegionProposalNetwork.filter_proposals
 * (torchvision/models/detection/rpn.py:268-297) and
 * RoIHeads.postprocess_detections (roi_heads.py:693-723); per segment the
 * result is exactly batched_nms(boxes[valid], scores[valid], idxs[valid], thr)
 * mapped back to original indices and cut to the first top_n entries.
 *   seg_offsets : HOST int64 [S + 1], boxes of set s are [off[s], off[s+1])
 *   idxs        : device int64 [N] categories, or NULL (plain nms per set)
 *   valid       : device uint8 [N] or NULL; 0 = box was filtered out before
 *                 NMS (small-box / score filters keep static shapes this way)
 *   mode        : b200det_bnms_mode; AUTO applies the reference's rule
 *                 (boxes.py:72) to every set on its own box count
 *   top_n       : <= 0: no limit
 *   keep        : device int64 [N]; set s owns keep[off[s] .. off[s] + counts[s])
 *   counts      : device int64 [S] */
size_t b200det_nms_segmented_workspace_bytes(int64_t N, const int64_t* seg_offsets, int64_t S);
int b200det_nms_segmented(
    const void* boxes, const void* scores, const int64_t* idxs,
    const uint8_t* valid, int dtype, int64_t N, const int64_t* seg_offsets,
    int64_t S, double iou_threshold, int mode, int64_t top_n, int64_t* keep,
    int64_t* counts, void* workspace, size_t workspace_bytes, void* stream);

/* clip_boxes_to_image + remove_small_boxes + score threshold for all images in
 * one pass, shapes kept static (torchvision/ops/boxes.py:115-174 as used by
 * rpn.py:276-285 and roi_heads.py:693-714).
 *   seg_offsets_dev : DEVICE int32 [S + 1]; image_sizes_dev: DEVICE float [S, 2] (height, width)
 *   strict          : 0: score >= score_thresh (RPN), 1: score > score_thresh (RoIHeads)
 *   boxes_out       : device [N, 4] clipped boxes (may alias boxes)
 *   valid           : device uint8 [N], the `valid` input of b200det_nms_segmented */
int b200det_clip_filter_boxes(
    const void* boxes, const void* scores, int dtype, int64_t N,
    const int32_t* seg_offsets_dev, int64_t S, const float* image_sizes_dev,
    double min_size, double score_thresh, int strict, void* boxes_out,
    uint8_t* valid, void* stream);

/* The RPN side between the per-level top-k and NMS in ONE pass (torchvision/models/detection/rpn.py:258-285,
 * all images at once): candidate (b, j) takes anchor a = top_idx[b][j]; boxes_out = clip(proposals[b][a]),
 * scores_out = sigmoid(objectness[b][a]) (1 / (1 + exp(-x)) in fp32, as torch computes it on CUDA), levels_out =
 * the pyramid level of anchor a, valid = small-box (min_size) and score (>=) filters.  fp32 tensors.
 *   proposals [B, A, 4], objectness [B, A], top_idx int64 [B, P] (indices into A)
 *   image_sizes_dev: DEVICE float [B, 2] (height, width);  level_ends: HOST int[num_levels <= 8], cumulative
 *   anchors per level;  outputs: boxes_out [B*P, 4], scores_out [B*P], levels_out int64 [B*P], valid uint8 [B*P] */
int b200det_rpn_select(
    const void* proposals, const void* objectness, const int64_t* top_idx,
    int64_t B, int64_t A, int64_t P, const float* image_sizes_dev,
    const int32_t* level_ends, int num_levels, double min_size, double score_thresh,
    void* boxes_out, void* scores_out, int64_t* levels_out, uint8_t* valid, void* stream);

/* Top-k per pyramid level of the RPN objectness, all images and levels in ONE launch.  Replaces
 * RegionProposalNetwork._get_top_n_idx (torchvision/models/detection/rpn.py:231-240): the loop over levels of
 * `ob.topk(min(pre_nms_top_n, num_anchors), dim=1)`, `+ offset` and `cat`.
 *   scores      : device fp32 [B, A] (logits; NaN ranks first like torch.topk)
 *   level_ends  : HOST int[num_levels <= 8], cumulative anchors per level, the last one == A
 *   k           : 1 .. 2048 (B200DET_ENOSUP beyond: the caller keeps torch.topk)
 *   out_idx     : device int64 [B, P], P = sum over levels of min(k, anchors of the level); per level the indices
 *                 (into A) of its top-k scores in descending score order.  Equal scores come in ascending index
 *                 order -- torch.topk leaves the order of ties unspecified. */
int b200det_segmented_topk(
    const void* scores, int64_t B, int64_t A, const int32_t* level_ends,
    int num_levels, int64_t k, int64_t* out_idx, int64_t P, void* stream);

/* The RoIHeads side before NMS in ONE pass (torchvision/models/detection/roi_heads.py:682-716 with
 * BoxCoder.decode, models/detection/_utils.py:162-224): softmax over the NC class logits, the box of every class
 * c >= 1 decoded from box_regression in the reference's operation order, clipped to the proposal's image, and
 * the validity byte of the score (>) and small-box (min_size) filters.  Instance r * (NC - 1) + (c - 1).  fp32.
 *   class_logits [R, NC], box_regression [R, NC * 4], proposals [R, 4], image_of DEVICE int32 [R] (image of a
 *   proposal), image_sizes_dev DEVICE float [S, 2], weights HOST double[4] (wx, wy, ww, wh), bbox_xform_clip
 *   outputs: boxes_out [R * (NC - 1), 4], scores_out [R * (NC - 1)], valid uint8 [R * (NC - 1)] */
int b200det_roi_heads_decode(
    const void* class_logits, const void* box_regression, const void* proposals,
    const int32_t* image_of, const float* image_sizes_dev, int64_t R, int64_t NC,
    const double* weights, double bbox_xform_clip, double score_thresh, double min_size,
    void* boxes_out, void* scores_out, uint8_t* valid, void* stream);

/* Batched transposition dst[b][c][r] = src[b][r][c] of `batch` dense
 * [rows, cols] matrices: the NCHW <-> NHWC re-layout the dispatcher shim puts
 * around the channels_last kernels for callers that hold NCHW-contiguous
 * tensors, which is what the reference's kernels take
 * (input.contiguous(), cuda/roi_align_kernel.cu:384).
 *   NCHW -> NHWC: batch = N, rows = C, cols = H*W;  NHWC -> NCHW: rows = H*W, cols = C */
int b200det_relayout(
    const void* src, void* dst, int dtype, int64_t batch, int64_t rows,
    int64_t cols, void* stream);

/* torchvision::roi_pool (max pooling over integer bins).  Replaces
 * roi_pool_forward_kernel / roi_pool_backward_kernel
 * (torchvision/csrc/ops/cuda/roi_pool_kernel.cu:120-278; CPU twin
 * cpu/roi_pool_kernel.cpp:127-249).  Same geometry as the reference: round()
 * of the scaled corners, floor / ceil bin edges, empty bins give 0.
 *   argmax : device int32 [K, C, PH, PW] in the output's layout; h * W + w
 *            inside the channel plane, -1 for an empty bin (reference meaning)
 *   rois   : as for roi_align (arithmetic type of the features)
 * The backward zero-fills grad_input and honours arbitrary element strides of
 * `grad` (roi_pool_kernel.cu:255-258). */
int b200det_roi_pool_forward(
    const void* input, int in_layout, const void* rois, void* output,
    int32_t* argmax, int out_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, void* stream);
int b200det_roi_pool_backward(
    const void* grad, const int64_t* grad_strides, const int32_t* argmax,
    int argmax_layout, const void* rois, void* grad_input, int gi_layout,
    int dtype, int64_t N, int64_t C, int64_t H, int64_t W, int64_t K,
    int64_t PH, int64_t PW, void* stream);

/* torchvision::ps_roi_pool (position-sensitive average pooling, R-FCN).
 * Replaces ps_roi_pool_forward_kernel / ps_roi_pool_backward_kernel
 * (torchvision/csrc/ops/cuda/ps_roi_pool_kernel.cu:142-280; CPU twin
 * cpu/ps_roi_pool_kernel.cpp:148-260).  C must be a multiple of PH * PW; the
 * output has C / (PH * PW) channels; channel_mapping (int32, output's shape
 * and layout) records the input channel of every output element.  The
 * reference clips forward bins to [0, size - 1] and backward bins to
 * [0, size]; both are kept. */
int b200det_ps_roi_pool_forward(
    const void* input, int in_layout, const void* rois, void* output,
    int32_t* channel_mapping, int out_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, void* stream);
int b200det_ps_roi_pool_backward(
    const void* grad, const int64_t* grad_strides,
    const int32_t* channel_mapping, int mapping_layout, const void* rois,
    void* grad_input, int gi_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, void* stream);

/* torchvision::ps_roi_align (position-sensitive RoIAlign).  Replaces
 * ps_roi_align_forward_kernel / ps_roi_align_backward_kernel
 * (torchvision/csrc/ops/cuda/ps_roi_align_kernel.cu:293-444; CPU twin
 * cpu/ps_roi_align_kernel.cpp:302-431).  Geometry of roi_align with
 * aligned = 1 and no size clamp; count = grid_h * grid_w (no max(., 1)).
 * Arguments as b200det_ps_roi_pool_*, plus sampling_ratio (<= 0: adaptive). */
int b200det_ps_roi_align_forward(
    const void* input, int in_layout, const void* rois, void* output,
    int32_t* channel_mapping, int out_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, int sampling_ratio, void* stream);
int b200det_ps_roi_align_backward(
    const void* grad, const int64_t* grad_strides,
    const int32_t* channel_mapping, int mapping_layout, const void* rois,
    void* grad_input, int gi_layout, int dtype, int64_t N, int64_t C,
    int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, int sampling_ratio, void* stream);

/* ------------------------------------------------------------------------
 * Host-buffer entry points (what a CPU-tensor caller of the reference binds).
 * Inputs and outputs are HOST pointers; the call stages them through a
 * per-process device arena (grown on demand, pinned bounce buffers), runs the
 * kernels above on `device` and returns after the results are in host memory.
 * b200det_host_release() frees the arena.
 * --------------------------------------------------------------------- */

int b200det_roi_align_forward_host(
    const void* input, int in_layout, const void* rois, void* output,
    int out_layout, int dtype, int64_t N, int64_t C, int64_t H, int64_t W,
    int64_t K, int64_t PH, int64_t PW, double spatial_scale,
    int sampling_ratio, int aligned, int device);

int b200det_roi_align_backward_host(
    const void* grad, const int64_t* grad_strides, int64_t grad_span_elems,
    const void* rois, void* grad_input, int gi_layout, int dtype, int64_t N,
    int64_t C, int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
    double spatial_scale, int sampling_ratio, int aligned, int device);

/* Host-buffer MultiScaleRoIAlign (forward, backward): feats[l] / grad_feats[l],
 * rois, levels, output and grad are HOST pointers with the same meaning as in
 * the device entry points above.  `levels` may be NULL in the forward. */
int b200det_msroi_align_forward_host(
    const void* const* feats, const int64_t* heights, const int64_t* widths,
    const double* scales, int num_levels, int layout, const void* rois,
    int32_t* levels, void* output, int out_layout, int dtype, int64_t N,
    int64_t C, int64_t K, int64_t PH, int64_t PW, int sampling_ratio,
    int aligned, int k_min, int k_max, int canonical_scale,
    int canonical_level, double eps, int device);

int b200det_msroi_align_backward_host(
    const void* grad, const int64_t* grad_strides, int64_t grad_span_elems,
    const void* rois, const int32_t* levels, void* const* grad_feats,
    const int64_t* heights, const int64_t* widths, const double* scales,
    int num_levels, int layout, int dtype, int64_t N, int64_t C, int64_t K,
    int64_t PH, int64_t PW, int sampling_ratio, int aligned, int device);

int b200det_nms_host(
    const void* boxes, const void* scores, int dtype, int64_t N,
    double iou_threshold, int64_t* keep, int64_t* num_keep, int device);

int b200det_batched_nms_host(
    const void* boxes, const void* scores, const int64_t* idxs, int dtype,
    int64_t N, double iou_threshold, int mode, int64_t* keep,
    int64_t* num_keep, int device);

void b200det_host_release(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_DETECTION_OPS_H */
