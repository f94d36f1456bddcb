// c_api.cu -- extern "C" entry points declared in include/b200_detection_ops.h.
// Argument validation, layout bookkeeping and kernel selection live here; the
// kernels are in roi_align_sm100.cu / roi_align_generic.cu / nms_sm100.cu.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace b200det {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

size_t dtype_size(int dtype) {
  switch (dtype) {
    case B200DET_F32: return 4;
    case B200DET_F16: return 2;
    case B200DET_BF16: return 2;
    case B200DET_F64: return 8;
  }
  return 0;
}

static int check_common(const char* who, int dtype, int layout_a, int layout_b, int64_t N,
                        int64_t C, int64_t K, int64_t PH, int64_t PW) {
  if (dtype_size(dtype) == 0) return fail(B200DET_ENOSUP, "%s: unknown dtype %d", who, dtype);
  if ((layout_a != B200DET_NCHW && layout_a != B200DET_NHWC) ||
      (layout_b != B200DET_NCHW && layout_b != B200DET_NHWC))
    return fail(B200DET_EINVAL, "%s: unknown layout", who);
  if (N < 0 || C < 0 || K < 0 || PH < 0 || PW < 0)
    return fail(B200DET_EINVAL, "%s: negative size", who);
  if (PH > 1 << 20 || PW > 1 << 20 || C > 1 << 24)
    return fail(B200DET_EINVAL, "%s: size out of range", who);
  return 0;
}

static int fill_levels(const char* who, LevelArgs& lv, const void* const* feats,
                       const int64_t* heights, const int64_t* widths, const double* scales,
                       int num_levels, int layout, int64_t N, int64_t C) {
  if (num_levels < 1 || num_levels > B200DET_MAX_LEVELS)
    return fail(B200DET_EINVAL, "%s: num_levels=%d outside [1,%d]", who, num_levels, B200DET_MAX_LEVELS);
  lv.num_levels = num_levels;
  lv.layout = layout;
  lv.batch = N;
  lv.channels = C;
  for (int l = 0; l < num_levels; ++l) {
    if (heights[l] < 0 || widths[l] < 0 || heights[l] > INT32_MAX || widths[l] > INT32_MAX)
      return fail(B200DET_EINVAL, "%s: bad size of level %d", who, l);
    if (!feats[l] && N * C * heights[l] * widths[l] != 0)
      return fail(B200DET_EINVAL, "%s: null feature pointer for level %d", who, l);
    lv.ptr[l] = const_cast<void*>(feats[l]);
    lv.height[l] = heights[l];
    lv.width[l] = widths[l];
    lv.scale[l] = scales[l];
  }
  return 0;
}

static int forward_impl(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels,
                        void* output, int out_layout, int64_t K, int64_t PH, int64_t PW, int sr,
                        int aligned, cudaStream_t st, const PeerOut* peers = nullptr) {
  if (K * lv.channels * PH * PW == 0) return 0;
  if (!rois || !output) return fail(B200DET_EINVAL, "roi_align forward: null pointer");
  int taken = 0;
  int rc = roi_align_fwd_fast(lv, dtype, rois, levels, output, out_layout, K, (int)PH, (int)PW, sr,
                              aligned != 0, st, &taken, peers);
  if (rc != 0 || taken) return rc;
  if (peers && peers->n > 0)
    return fail(B200DET_ENOSUP, "msroi_align_forward_push: only the channels_last fast path pushes (C %% 128 == 0, "
                                "sampling_ratio 1 or 2, pooled_width * sampling_ratio <= 32, fp32 / fp16 / bf16)");
  rc = roi_align_fwd_anygrid(lv, dtype, rois, levels, output, out_layout, K, (int)PH, (int)PW, sr, aligned != 0, st,
                             &taken);
  if (rc != 0 || taken) return rc;
  return roi_align_fwd_generic(lv, dtype, rois, levels, output, out_layout, K, (int)PH, (int)PW, sr,
                               aligned != 0, st);
}

static int backward_impl(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                         const void* rois, const int32_t* levels, int64_t K, int64_t PH, int64_t PW,
                         int sr, int aligned, int flags, cudaStream_t st) {
  const bool zero_init = (flags & B200DET_BWD_ZERO_INIT) != 0;
  const bool deterministic = (flags & B200DET_BWD_DETERMINISTIC) != 0;
  const bool work = K * lv.channels * PH * PW != 0;
  if (work) {
    if (!rois || !grad || !gs) return fail(B200DET_EINVAL, "roi_align backward: null pointer");
    int taken = 0;
    int rc = 0;
    // Which kernel produces which level.  The write-once strip kernel (shared-memory accumulators, every
    // gradient byte stored once, deterministic) takes the levels whose maps are too big for L2 to act as the
    // accumulator: there a red costs a zero fill, a DRAM read-modify-write and a write-back per line.  The
    // vector-red scatter takes the rest (dense, L2-resident coarse levels).  Measured on B200
    // (profiles/r02_bwd_study.md) the strips do not beat the scatter on any BASELINE shape yet, so by default no
    // level is big enough (B200DET_STRIP_MIN_MB=<MB of one level, whole batch> lowers the bar); B200DET_BWD_STRIP=1
    // / 0 forces everything / nothing onto the strips; the deterministic flag means strips for every level.
    // (read per call, like B200DET_BWD_MODE: the parity tests switch them inside one process)
    const char* se = getenv("B200DET_BWD_STRIP");
    const char* sm = getenv("B200DET_STRIP_MIN_MB");
    const int strip_env = se ? atoi(se) : -1;
    const double strip_min_mb = sm ? atof(sm) : 1e30;
    const unsigned all_levels = (1u << lv.num_levels) - 1u;
    unsigned strip_levels = 0;
    if (deterministic || strip_env == 1) {
      strip_levels = all_levels;
    } else if (strip_env < 0) {
      for (int l = 0; l < lv.num_levels; ++l) {
        const double mb = (double)lv.batch * lv.channels * lv.height[l] * lv.width[l] * dtype_size(dtype) / 1048576.0;
        if (mb >= strip_min_mb) strip_levels |= 1u << l;
      }
    }
    unsigned done_levels = 0;
    if (strip_levels) {
      rc = roi_align_bwd_strip(lv, dtype, grad, gs, rois, levels, K, (int)PH, (int)PW, sr, aligned != 0, zero_init,
                               strip_levels, st, &taken);
      if (rc != 0) return rc;
      if (taken) done_levels = strip_levels;
      else if (deterministic)
        return fail(B200DET_ENOSUP,
                    "roi_align backward: the deterministic kernel needs channels_last tensors, C %% 128 == 0, "
                    "sampling_ratio 1 or 2 and a channel-contiguous gradient");
    }
    if (done_levels == all_levels) return 0;
    // fused zero fill + vector-red scatter for the remaining levels
    rc = roi_align_bwd_fast(lv, dtype, grad, gs, rois, levels, K, (int)PH, (int)PW, sr, aligned != 0, zero_init,
                            done_levels, st, &taken);
    if (rc != 0 || taken) return rc;
    if (done_levels)
      return fail(B200DET_ENOSUP, "roi_align backward: internal -- the scatter kernel refused a call the strip kernel took");
  }
  if (zero_init) {
    for (int l = 0; l < lv.num_levels; ++l) {
      const size_t bytes = (size_t)lv.batch * lv.channels * lv.height[l] * lv.width[l] * dtype_size(dtype);
      if (bytes) B200DET_CUDA_OK(cudaMemsetAsync(lv.ptr[l], 0, bytes, st));
    }
  }
  if (!work) return 0;
  if (!deterministic) {  // any sampling grid on channels_last maps (adds into the gradients zeroed above)
    int taken = 0;
    const int rc = roi_align_bwd_anygrid(lv, dtype, grad, gs, rois, levels, K, (int)PH, (int)PW, sr, aligned != 0, st,
                                         &taken);
    if (rc != 0 || taken) return rc;
  }
  return roi_align_bwd_generic(lv, dtype, grad, gs, rois, levels, K, (int)PH, (int)PW, sr,
                               aligned != 0, st);
}

}  // namespace b200det

using namespace b200det;

extern "C" {

int b200det_version(void) { return B200DET_VERSION; }
const char* b200det_last_error(void) { return g_err; }

int b200det_roi_align_forward(const void* input, int in_layout, const void* rois, void* output,
                              int out_layout, int dtype, int64_t N, int64_t C, int64_t H,
                              int64_t W, int64_t K, int64_t PH, int64_t PW, double spatial_scale,
                              int sampling_ratio, int aligned, void* stream) {
  int rc = check_common("roi_align_forward", dtype, in_layout, out_layout, N, C, K, PH, PW);
  if (rc) return rc;
  LevelArgs lv;
  const void* feats[1] = {input};
  rc = fill_levels("roi_align_forward", lv, feats, &H, &W, &spatial_scale, 1, in_layout, N, C);
  if (rc) return rc;
  return forward_impl(lv, dtype, rois, nullptr, output, out_layout, K, PH, PW, sampling_ratio,
                      aligned, (cudaStream_t)stream);
}

int b200det_roi_align_backward(const void* grad, const int64_t* grad_strides, const void* rois,
                               void* grad_input, int gi_layout, int dtype, int64_t N, int64_t C,
                               int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
                               double spatial_scale, int sampling_ratio, int aligned,
                               int flags, void* stream) {
  int rc = check_common("roi_align_backward", dtype, gi_layout, gi_layout, N, C, K, PH, PW);
  if (rc) return rc;
  LevelArgs lv;
  const void* feats[1] = {grad_input};
  rc = fill_levels("roi_align_backward", lv, feats, &H, &W, &spatial_scale, 1, gi_layout, N, C);
  if (rc) return rc;
  return backward_impl(lv, dtype, grad, grad_strides, rois, nullptr, K, PH, PW, sampling_ratio,
                       aligned, flags, (cudaStream_t)stream);
}

int b200det_fpn_level_map(const void* rois, int dtype, int64_t K, int k_min, int k_max,
                          int canonical_scale, int canonical_level, double eps, int32_t* levels,
                          void* stream) {
  if (K < 0) return fail(B200DET_EINVAL, "fpn_level_map: negative K");
  if (K == 0) return 0;
  if (!rois || !levels) return fail(B200DET_EINVAL, "fpn_level_map: null pointer");
  return fpn_level_map(rois, dtype, K, k_min, k_max, canonical_scale, canonical_level, eps, levels,
                       (cudaStream_t)stream);
}

int b200det_msroi_align_forward(const void* const* feats, const int64_t* heights,
                                const int64_t* widths, const double* scales, int num_levels,
                                int layout, const void* rois, int32_t* levels, int levels_given,
                                void* output, int out_layout, int dtype, int64_t N, int64_t C,
                                int64_t K, int64_t PH, int64_t PW, int sampling_ratio, int aligned,
                                int k_min, int k_max, int canonical_scale, int canonical_level,
                                double eps, void* stream) {
  int rc = check_common("msroi_align_forward", dtype, layout, out_layout, N, C, K, PH, PW);
  if (rc) return rc;
  if (!feats || !heights || !widths || !scales)
    return fail(B200DET_EINVAL, "msroi_align_forward: null level arrays");
  LevelArgs lv;
  rc = fill_levels("msroi_align_forward", lv, feats, heights, widths, scales, num_levels, layout, N, C);
  if (rc) return rc;
  const int32_t* lv_ids = nullptr;
  if (num_levels > 1) {
    if (K > 0 && !levels) return fail(B200DET_EINVAL, "msroi_align_forward: levels buffer required");
    if (!levels_given && K > 0) {
      rc = fpn_level_map(rois, dtype, K, k_min, k_max, canonical_scale, canonical_level, eps, levels,
                         (cudaStream_t)stream);
      if (rc) return rc;
    }
    lv_ids = levels;
  } else if (levels && !levels_given && K > 0) {
    // one level: every RoI maps to level 0 (callers keep the buffer for the backward)
    B200DET_CUDA_OK(cudaMemsetAsync(levels, 0, (size_t)K * sizeof(int32_t), (cudaStream_t)stream));
  }
  return forward_impl(lv, dtype, rois, lv_ids, output, out_layout, K, PH, PW, sampling_ratio,
                      aligned, (cudaStream_t)stream);
}

int b200det_msroi_align_forward_push(const void* const* feats, const int64_t* heights,
                                     const int64_t* widths, const double* scales, int num_levels,
                                     int layout, const void* rois, int32_t* levels, int levels_given,
                                     void* output, void* const* peer_outputs, int num_peers,
                                     int out_layout, int dtype, int64_t N, int64_t C, int64_t K,
                                     int64_t PH, int64_t PW, int sampling_ratio, int aligned, int k_min,
                                     int k_max, int canonical_scale, int canonical_level, double eps,
                                     void* stream) {
  if (num_peers < 0 || num_peers > kMaxPeerOut)
    return fail(B200DET_EINVAL, "msroi_align_forward_push: at most %d peer outputs", kMaxPeerOut);
  if (num_peers > 0 && !peer_outputs) return fail(B200DET_EINVAL, "msroi_align_forward_push: null peer_outputs");
  int rc = check_common("msroi_align_forward_push", dtype, layout, out_layout, N, C, K, PH, PW);
  if (rc) return rc;
  if (!feats || !heights || !widths || !scales)
    return fail(B200DET_EINVAL, "msroi_align_forward_push: null level arrays");
  PeerOut peers;
  peers.n = num_peers;
  for (int i = 0; i < num_peers; ++i) {
    if (!peer_outputs[i] || ((uintptr_t)peer_outputs[i]) % 16 != 0)
      return fail(B200DET_EINVAL, "msroi_align_forward_push: peer output %d is null or not 16-byte aligned", i);
    peers.delta[i] = (long long)((const char*)peer_outputs[i] - (const char*)output);
  }
  LevelArgs lv;
  rc = fill_levels("msroi_align_forward_push", lv, feats, heights, widths, scales, num_levels, layout, N, C);
  if (rc) return rc;
  const int32_t* lv_ids = nullptr;
  if (num_levels > 1) {
    if (K > 0 && !levels) return fail(B200DET_EINVAL, "msroi_align_forward_push: levels buffer required");
    if (!levels_given && K > 0) {
      rc = fpn_level_map(rois, dtype, K, k_min, k_max, canonical_scale, canonical_level, eps, levels,
                         (cudaStream_t)stream);
      if (rc) return rc;
    }
    lv_ids = levels;
  } else if (levels && !levels_given && K > 0) {
    B200DET_CUDA_OK(cudaMemsetAsync(levels, 0, (size_t)K * sizeof(int32_t), (cudaStream_t)stream));
  }
  return forward_impl(lv, dtype, rois, lv_ids, output, out_layout, K, PH, PW, sampling_ratio, aligned,
                      (cudaStream_t)stream, &peers);
}

// ---- peer-visible device memory (CUDA IPC) for the push gather --------------------------------------------
int b200det_peer_alloc(size_t bytes, void** ptr) {
  if (!ptr || bytes == 0) return fail(B200DET_EINVAL, "peer_alloc: null pointer or zero size");
  B200DET_CUDA_OK(cudaMalloc(ptr, bytes));
  return 0;
}

int b200det_peer_free(void* ptr) {
  if (ptr) B200DET_CUDA_OK(cudaFree(ptr));
  return 0;
}

int b200det_peer_export(const void* ptr, void* handle) {
  if (!ptr || !handle) return fail(B200DET_EINVAL, "peer_export: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == B200DET_PEER_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t h;
  B200DET_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
  memcpy(handle, &h, sizeof(h));
  return 0;
}

int b200det_peer_open(const void* handle, void** ptr) {
  if (!ptr || !handle) return fail(B200DET_EINVAL, "peer_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  B200DET_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int b200det_peer_close(void* ptr) {
  if (ptr) B200DET_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int b200det_msroi_align_backward(const void* grad, const int64_t* grad_strides, const void* rois,
                                 const int32_t* levels, void* const* grad_feats,
                                 const int64_t* heights, const int64_t* widths,
                                 const double* scales, int num_levels, int layout, int dtype,
                                 int64_t N, int64_t C, int64_t K, int64_t PH, int64_t PW,
                                 int sampling_ratio, int aligned, int flags, void* stream) {
  int rc = check_common("msroi_align_backward", dtype, layout, layout, N, C, K, PH, PW);
  if (rc) return rc;
  if (!grad_feats || !heights || !widths || !scales)
    return fail(B200DET_EINVAL, "msroi_align_backward: null level arrays");
  LevelArgs lv;
  rc = fill_levels("msroi_align_backward", lv, (const void* const*)grad_feats, heights, widths,
                   scales, num_levels, layout, N, C);
  if (rc) return rc;
  if (num_levels > 1 && K > 0 && !levels)
    return fail(B200DET_EINVAL, "msroi_align_backward: levels required");
  return backward_impl(lv, dtype, grad, grad_strides, rois, num_levels > 1 ? levels : nullptr, K, PH,
                       PW, sampling_ratio, aligned, flags, (cudaStream_t)stream);
}

size_t b200det_nms_workspace_bytes(int64_t N) { return nms_workspace_bytes(N); }

int b200det_nms(const void* boxes, const void* scores, int dtype, int64_t N, double iou_threshold,
                int64_t* keep, int64_t* num_keep, void* workspace, size_t workspace_bytes,
                void* stream) {
  if (N < 0) return fail(B200DET_EINVAL, "nms: negative N");
  if (!num_keep) return fail(B200DET_EINVAL, "nms: null num_keep");
  if (N > 0 && (!boxes || !scores || !keep || !workspace)) return fail(B200DET_EINVAL, "nms: null pointer");
  if (dtype_size(dtype) == 0) return fail(B200DET_ENOSUP, "nms: unknown dtype %d", dtype);
  return nms_device(boxes, scores, nullptr, dtype, N, iou_threshold, 0, keep, num_keep, workspace,
                    workspace_bytes, (cudaStream_t)stream);
}

int b200det_batched_nms(const void* boxes, const void* scores, const int64_t* idxs, int dtype,
                        int64_t N, double iou_threshold, int mode, int64_t* keep,
                        int64_t* num_keep, void* workspace, size_t workspace_bytes, void* stream) {
  if (N < 0) return fail(B200DET_EINVAL, "batched_nms: negative N");
  if (!num_keep) return fail(B200DET_EINVAL, "batched_nms: null num_keep");
  if (N > 0 && (!boxes || !scores || !idxs || !keep || !workspace))
    return fail(B200DET_EINVAL, "batched_nms: null pointer");
  if (dtype_size(dtype) == 0) return fail(B200DET_ENOSUP, "batched_nms: unknown dtype %d", dtype);
  int m;
  if (mode == B200DET_BNMS_AUTO) m = (4 * N > 20000) ? 2 : 1;  // boxes.py:72, CUDA rule
  else if (mode == B200DET_BNMS_COORD_TRICK) m = 1;
  else if (mode == B200DET_BNMS_VANILLA) m = 2;
  else return fail(B200DET_EINVAL, "batched_nms: unknown mode %d", mode);
  return nms_device(boxes, scores, idxs, dtype, N, iou_threshold, m, keep, num_keep, workspace,
                    workspace_bytes, (cudaStream_t)stream);
}

size_t b200det_nms_segmented_workspace_bytes(int64_t N, const int64_t* seg_offsets, int64_t S) {
  if (N < 0 || S < 0 || !seg_offsets) return 0;
  return nms_segmented_workspace_bytes(N, seg_offsets, S);
}

int b200det_nms_segmented(const void* boxes, const void* scores, const int64_t* idxs, const uint8_t* valid,
                          int dtype, int64_t N, const int64_t* seg_offsets, int64_t S, double iou_threshold,
                          int mode, int64_t top_n, int64_t* keep, int64_t* counts, void* workspace,
                          size_t workspace_bytes, void* stream) {
  if (N < 0 || S < 0) return fail(B200DET_EINVAL, "nms_segmented: negative N or S");
  if (S > 0 && !counts) return fail(B200DET_EINVAL, "nms_segmented: null counts");
  if (N > 0 && (!boxes || !scores || !keep || !workspace)) return fail(B200DET_EINVAL, "nms_segmented: null pointer");
  int m;
  if (!idxs) m = 0;
  else if (mode == B200DET_BNMS_AUTO) m = -1;
  else if (mode == B200DET_BNMS_COORD_TRICK) m = 1;
  else if (mode == B200DET_BNMS_VANILLA) m = 2;
  else return fail(B200DET_EINVAL, "nms_segmented: unknown mode %d", mode);
  const int tn = top_n > 0x7fffffffLL ? 0x7fffffff : (int)top_n;
  return nms_segmented_device(boxes, scores, idxs, valid, dtype, N, seg_offsets, S, iou_threshold, m, tn, keep, counts,
                              workspace, workspace_bytes, (cudaStream_t)stream);
}

int b200det_clip_filter_boxes(const void* boxes, const void* scores, int dtype, int64_t N,
                              const int32_t* seg_offsets_dev, int64_t S, const float* image_sizes_dev,
                              double min_size, double score_thresh, int strict, void* boxes_out, uint8_t* valid,
                              void* stream) {
  if (N < 0 || S < 0) return fail(B200DET_EINVAL, "clip_filter_boxes: negative N or S");
  if (N == 0) return 0;
  if (S == 0 || S > 0x7fffffffLL) return fail(B200DET_EINVAL, "clip_filter_boxes: N > 0 needs 1 <= S < 2^31 segments");
  if (!boxes || !scores || !seg_offsets_dev || !image_sizes_dev || !boxes_out || !valid)
    return fail(B200DET_EINVAL, "clip_filter_boxes: null pointer");
  return clip_filter_boxes(boxes, scores, dtype, N, seg_offsets_dev, (int)S, image_sizes_dev, min_size, score_thresh,
                           strict, boxes_out, valid, (cudaStream_t)stream);
}

int b200det_rpn_select(const void* proposals, const void* objectness, const int64_t* top_idx, int64_t B, int64_t A,
                       int64_t P, const float* image_sizes_dev, const int32_t* level_ends, int num_levels,
                       double min_size, double score_thresh, void* boxes_out, void* scores_out, int64_t* levels_out,
                       uint8_t* valid, void* stream) {
  if (B < 0 || A < 0 || P < 0) return fail(B200DET_EINVAL, "rpn_select: negative size");
  if (B * P == 0) return 0;
  if (B > 0x7fffffffLL || P > 0x7fffffffLL || num_levels < 1 || num_levels > 8)
    return fail(B200DET_EINVAL, "rpn_select: B, P < 2^31 and 1..8 levels expected");
  if (!proposals || !objectness || !top_idx || !image_sizes_dev || !level_ends || !boxes_out || !scores_out ||
      !levels_out || !valid)
    return fail(B200DET_EINVAL, "rpn_select: null pointer");
  return rpn_select((const float*)proposals, (const float*)objectness, top_idx, (int)B, A, (int)P, image_sizes_dev,
                    level_ends, num_levels, min_size, score_thresh, (float*)boxes_out, (float*)scores_out, levels_out,
                    valid, (cudaStream_t)stream);
}

int b200det_roi_heads_decode(const void* class_logits, const void* box_regression, const void* proposals,
                             const int32_t* image_of, const float* image_sizes_dev, int64_t R, int64_t NC,
                             const double* weights, double bbox_xform_clip, double score_thresh, double min_size,
                             void* boxes_out, void* scores_out, uint8_t* valid, void* stream) {
  if (R < 0 || NC < 0) return fail(B200DET_EINVAL, "roi_heads_decode: negative size");
  if (R == 0 || NC < 2) return 0;
  if (R > 0x7fffffffLL / 64 || NC > 65536) return fail(B200DET_EINVAL, "roi_heads_decode: R or NC too large");
  if (!class_logits || !box_regression || !proposals || !image_of || !image_sizes_dev || !weights || !boxes_out ||
      !scores_out || !valid)
    return fail(B200DET_EINVAL, "roi_heads_decode: null pointer");
  return roi_heads_decode((const float*)class_logits, (const float*)box_regression, (const float*)proposals, image_of,
                          image_sizes_dev, (int)R, (int)NC, weights, bbox_xform_clip, score_thresh, min_size,
                          (float*)boxes_out, (float*)scores_out, valid, (cudaStream_t)stream);
}

int b200det_segmented_topk(const void* scores, int64_t B, int64_t A, const int32_t* level_ends,
                           int num_levels, int64_t k, int64_t* out_idx, int64_t P, void* stream) {
  if (!level_ends) return fail(B200DET_EINVAL, "segmented_topk: null level_ends");
  return segmented_topk((const float*)scores, B, A, level_ends, num_levels, k, out_idx, P, (cudaStream_t)stream);
}

int b200det_relayout(const void* src, void* dst, int dtype, int64_t batch, int64_t rows, int64_t cols,
                     void* stream) {
  if (batch < 0 || rows < 0 || cols < 0) return fail(B200DET_EINVAL, "relayout: negative size");
  if (batch * rows * cols == 0) return 0;
  if (!src || !dst || src == dst) return fail(B200DET_EINVAL, "relayout: null or aliasing pointers");
  return relayout(src, dst, dtype, batch, rows, cols, (cudaStream_t)stream);
}

int b200det_roi_pool_forward(const void* input, int in_layout, const void* rois, void* output, int32_t* argmax,
                             int out_layout, int dtype, int64_t N, int64_t C, int64_t H, int64_t W, int64_t K,
                             int64_t PH, int64_t PW, double spatial_scale, void* stream) {
  int rc = check_common("roi_pool_forward", dtype, in_layout, out_layout, N, C, K, PH, PW);
  if (rc) return rc;
  if (H < 0 || W < 0 || H > INT32_MAX || W > INT32_MAX) return fail(B200DET_EINVAL, "roi_pool_forward: bad H or W");
  if (K * C * PH * PW == 0) return 0;
  if (!rois || !output || !argmax || (!input && N * C * H * W != 0))
    return fail(B200DET_EINVAL, "roi_pool_forward: null pointer");
  return roi_pool_forward(input, in_layout, rois, output, argmax, out_layout, dtype, N, C, H, W, K, (int)PH, (int)PW,
                          spatial_scale, (cudaStream_t)stream);
}

int b200det_roi_pool_backward(const void* grad, const int64_t* grad_strides, const int32_t* argmax, int argmax_layout,
                              const void* rois, void* grad_input, int gi_layout, int dtype, int64_t N, int64_t C,
                              int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW, void* stream) {
  int rc = check_common("roi_pool_backward", dtype, argmax_layout, gi_layout, N, C, K, PH, PW);
  if (rc) return rc;
  if (H < 0 || W < 0 || H > INT32_MAX || W > INT32_MAX) return fail(B200DET_EINVAL, "roi_pool_backward: bad H or W");
  if (N * C * H * W == 0) return 0;
  if (!grad_input) return fail(B200DET_EINVAL, "roi_pool_backward: null grad_input");
  if (K * C * PH * PW != 0 && (!grad || !grad_strides || !argmax || !rois))
    return fail(B200DET_EINVAL, "roi_pool_backward: null pointer");
  static const int64_t zero4[4] = {0, 0, 0, 0};
  return roi_pool_backward(grad, grad_strides ? grad_strides : zero4, argmax, argmax_layout, rois, grad_input, gi_layout,
                           dtype, N, C, H, W, K, (int)PH, (int)PW, (cudaStream_t)stream);
}

static int ps_check(const char* who, int dtype, int la, int lb, int64_t N, int64_t C, int64_t H, int64_t W, int64_t K,
                    int64_t PH, int64_t PW) {
  int rc = check_common(who, dtype, la, lb, N, C, K, PH, PW);
  if (rc) return rc;
  if (H < 0 || W < 0 || H > INT32_MAX || W > INT32_MAX) return fail(B200DET_EINVAL, "%s: bad H or W", who);
  if (PH < 1 || PW < 1 || C % (PH * PW) != 0)
    return fail(B200DET_EINVAL, "%s: input channels must be a multiple of pooling height * pooling width", who);
  return 0;
}

int b200det_ps_roi_pool_forward(const void* input, int in_layout, const void* rois, void* output,
                                int32_t* channel_mapping, int out_layout, int dtype, int64_t N, int64_t C, int64_t H,
                                int64_t W, int64_t K, int64_t PH, int64_t PW, double spatial_scale, void* stream) {
  int rc = ps_check("ps_roi_pool_forward", dtype, in_layout, out_layout, N, C, H, W, K, PH, PW);
  if (rc) return rc;
  if (K * C == 0) return 0;
  if (!rois || !output || !channel_mapping || (!input && N * C * H * W != 0))
    return fail(B200DET_EINVAL, "ps_roi_pool_forward: null pointer");
  return ps_roi_pool(true, input, nullptr, in_layout, rois, output, channel_mapping, out_layout, dtype, N, C, H, W, K,
                     (int)PH, (int)PW, spatial_scale, (cudaStream_t)stream);
}

int b200det_ps_roi_pool_backward(const void* grad, const int64_t* grad_strides, const int32_t* channel_mapping,
                                 int mapping_layout, const void* rois, void* grad_input, int gi_layout, int dtype,
                                 int64_t N, int64_t C, int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
                                 double spatial_scale, void* stream) {
  int rc = ps_check("ps_roi_pool_backward", dtype, gi_layout, mapping_layout, N, C, H, W, K, PH, PW);
  if (rc) return rc;
  if (N * C * H * W == 0) return 0;
  if (!grad_input) return fail(B200DET_EINVAL, "ps_roi_pool_backward: null grad_input");
  if (K * C != 0 && (!grad || !grad_strides || !channel_mapping || !rois))
    return fail(B200DET_EINVAL, "ps_roi_pool_backward: null pointer");
  static const int64_t zero4[4] = {0, 0, 0, 0};
  return ps_roi_pool(false, grad, grad_strides ? grad_strides : zero4, gi_layout, rois, grad_input,
                     const_cast<int32_t*>(channel_mapping), mapping_layout, dtype, N, C, H, W, K, (int)PH, (int)PW,
                     spatial_scale, (cudaStream_t)stream);
}

int b200det_ps_roi_align_forward(const void* input, int in_layout, const void* rois, void* output,
                                 int32_t* channel_mapping, int out_layout, int dtype, int64_t N, int64_t C, int64_t H,
                                 int64_t W, int64_t K, int64_t PH, int64_t PW, double spatial_scale,
                                 int sampling_ratio, void* stream) {
  int rc = ps_check("ps_roi_align_forward", dtype, in_layout, out_layout, N, C, H, W, K, PH, PW);
  if (rc) return rc;
  if (K * C == 0) return 0;
  if (!rois || !output || !channel_mapping || (!input && N * C * H * W != 0))
    return fail(B200DET_EINVAL, "ps_roi_align_forward: null pointer");
  return ps_roi_align(true, input, nullptr, in_layout, rois, output, channel_mapping, out_layout, dtype, N, C, H, W, K,
                      (int)PH, (int)PW, spatial_scale, sampling_ratio, (cudaStream_t)stream);
}

int b200det_ps_roi_align_backward(const void* grad, const int64_t* grad_strides, const int32_t* channel_mapping,
                                  int mapping_layout, const void* rois, void* grad_input, int gi_layout, int dtype,
                                  int64_t N, int64_t C, int64_t H, int64_t W, int64_t K, int64_t PH, int64_t PW,
                                  double spatial_scale, int sampling_ratio, void* stream) {
  int rc = ps_check("ps_roi_align_backward", dtype, gi_layout, mapping_layout, N, C, H, W, K, PH, PW);
  if (rc) return rc;
  if (N * C * H * W == 0) return 0;
  if (!grad_input) return fail(B200DET_EINVAL, "ps_roi_align_backward: null grad_input");
  if (K * C != 0 && (!grad || !grad_strides || !channel_mapping || !rois))
    return fail(B200DET_EINVAL, "ps_roi_align_backward: null pointer");
  static const int64_t zero4[4] = {0, 0, 0, 0};
  return ps_roi_align(false, grad, grad_strides ? grad_strides : zero4, gi_layout, rois, grad_input,
                      const_cast<int32_t*>(channel_mapping), mapping_layout, dtype, N, C, H, W, K, (int)PH, (int)PW,
                      spatial_scale, sampling_ratio, (cudaStream_t)stream);
}

}  // extern "C"
