// kernels.h -- internal launch interface between c_api.cu and the kernel files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_detection_ops.h"

namespace b200det {

// host-side description of the feature pyramid one call works on
struct LevelArgs {
  void* ptr[B200DET_MAX_LEVELS];
  int64_t height[B200DET_MAX_LEVELS];
  int64_t width[B200DET_MAX_LEVELS];
  double scale[B200DET_MAX_LEVELS];
  int num_levels;
  int layout;        // b200det_layout of every level
  int64_t batch;     // N
  int64_t channels;  // C
};

size_t dtype_size(int dtype);

// roi_align_generic.cu
int roi_align_fwd_generic(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels,
                          void* out, int out_layout, long long K, int PH, int PW, int sr,
                          bool aligned, cudaStream_t st);
int roi_align_bwd_generic(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                          const void* rois, const int32_t* levels, long long K, int PH, int PW,
                          int sr, bool aligned, cudaStream_t st);
int fpn_level_map(const void* rois, int dtype, long long K, int k_min, int k_max, int s0, int lvl0,
                  double eps, int32_t* levels, cudaStream_t st);

// Extra destinations of the forward's pooled rows (b200det_msroi_align_forward_push): byte offsets from `out` to
// the same row in up to 7 other buffers -- peer GPUs' gather buffers mapped over NVLink.
constexpr int kMaxPeerOut = 7;
struct PeerOut {
  long long delta[kMaxPeerOut];
  int n;
};

// roi_align_fwd_sm100.cu / roi_align_bwd_sm100.cu: returns 1 when it took the call, 0 when the generic path
// must run, <0 / >0 on error like the public API.  `peers` (may be null): every pooled bin is also stored there.
int roi_align_fwd_fast(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels,
                       void* out, int out_layout, long long K, int PH, int PW, int sr,
                       bool aligned, cudaStream_t st, int* taken, const PeerOut* peers = nullptr);
// The backward zero-fills the gradients itself when zero_init is set (fused with the scatter).  Levels named in
// skip_levels (bit l) are left alone: their RoIs are skipped and their maps not zero-filled (the strip kernel
// produces them).
int roi_align_bwd_fast(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                       const void* rois, const int32_t* levels, long long K, int PH, int PW,
                       int sr, bool aligned, bool zero_init, unsigned skip_levels, cudaStream_t st, int* taken);
void roi_align_bwd_release();
// roi_align_anygrid_sm100.cu: channels_last maps, any sampling grid (adaptive, > 2), C % 4 == 0: merged taps per
// axis, a channel vector per lane.  The backward only ADDS: the caller zero-fills the gradients first.
int roi_align_fwd_anygrid(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels,
                          void* out, int out_layout, long long K, int PH, int PW, int sr,
                          bool aligned, cudaStream_t st, int* taken);
int roi_align_bwd_anygrid(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                          const void* rois, const int32_t* levels, long long K, int PH, int PW,
                          int sr, bool aligned, cudaStream_t st, int* taken);
// roi_align_bwd_strip_sm100.cu: write-once backward (a warp owns a strip of one feature row; shared-memory
// accumulators, no atomics, no memset, deterministic); produces the levels named in level_mask (bit l) and
// leaves the others untouched
int roi_align_bwd_strip(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                        const void* rois, const int32_t* levels, long long K, int PH, int PW,
                        int sr, bool aligned, bool zero_init, unsigned level_mask, cudaStream_t st, int* taken);
void roi_align_bwd_strip_release();

// nms_sm100.cu
size_t nms_workspace_bytes(long long N);
int nms_device(const void* boxes, const void* scores, const int64_t* idxs, int dtype, long long N,
               double thr, int mode, int64_t* keep, int64_t* num_keep, void* ws, size_t ws_bytes,
               cudaStream_t st);

// segmented (per-image) NMS; seg_offsets is a HOST array of S + 1 box offsets; mode -1 auto / 0 plain / 1 trick / 2 vanilla
size_t nms_segmented_workspace_bytes(long long N, const int64_t* seg_offsets, long long S);
int nms_segmented_device(const void* boxes, const void* scores, const int64_t* idxs, const uint8_t* valid,
                         int dtype, long long N, const int64_t* seg_offsets, long long S, double thr, int mode,
                         int top_n, int64_t* keep, int64_t* counts, void* ws, size_t ws_bytes, cudaStream_t st);

// box_ops_sm100.cu: clip to the image + small-box and score filters as a validity byte (static shapes)
int clip_filter_boxes(const void* boxes, const void* scores, int dtype, long long n, const int32_t* off, int S,
                      const float* sizes, double min_size, double score_thresh, int strict, void* out,
                      uint8_t* valid, cudaStream_t st);

// topk_sm100.cu: top-k per pyramid level of the [B, A] objectness, indices in descending score order (ties by index)
int segmented_topk(const float* scores, long long B, long long A, const int32_t* level_ends, int num_levels,
                   long long k, int64_t* out_idx, long long P, cudaStream_t st);
// box_ops_sm100.cu: the RPN side between top-k and NMS in one pass (gather, sigmoid, level, clip, filters), fp32
int rpn_select(const float* proposals, const float* objectness, const int64_t* top_idx, int B, long long A, int P,
               const float* sizes, const int* level_ends, int num_levels, double min_size, double score_thresh,
               float* boxes, float* scores, int64_t* levels, uint8_t* valid, cudaStream_t st);
// box_ops_sm100.cu: the RoIHeads side before NMS in one pass (softmax, BoxCoder.decode, clip, filters), fp32
int roi_heads_decode(const float* logits, const float* reg, const float* props, const int32_t* image_of,
                     const float* sizes, int R, int NC, const double* weights, double xform_clip, double score_thresh,
                     double min_size, float* boxes, float* scores, uint8_t* valid, cudaStream_t st);

// relayout_sm100.cu: dst[b][c][r] = src[b][r][c] (NCHW -> NHWC: R = C, S = H*W; back: R = H*W, S = C)
int relayout(const void* src, void* dst, int dtype, long long B, long long R, long long S, cudaStream_t st);

// roi_pool_sm100.cu: torchvision::roi_pool; argmax = h * W + w inside the channel plane (-1: empty bin), stored in the
// output's layout; the backward zero-fills grad_in itself
int roi_pool_forward(const void* input, int in_layout, const void* rois, void* out, int32_t* argmax, int out_layout,
                     int dtype, long long N, long long C, long long H, long long W, long long K, int PH, int PW,
                     double scale, cudaStream_t st);
int roi_pool_backward(const void* grad, const int64_t* gs, const int32_t* argmax, int am_layout, const void* rois,
                      void* grad_in, int gi_layout, int dtype, long long N, long long C, long long H, long long W,
                      long long K, int PH, int PW, cudaStream_t st);

// ps_roi_pool_sm100.cu: torchvision::ps_roi_pool (position-sensitive average pooling); see the file for the argument roles
int ps_roi_pool(bool forward, const void* in_or_grad, const int64_t* gs, int in_layout, const void* rois,
                void* out_or_gi, int32_t* mapping, int out_layout, int dtype, long long N, long long C, long long H,
                long long W, long long K, int PH, int PW, double scale, cudaStream_t st);

// ps_roi_align_sm100.cu: torchvision::ps_roi_align; argument roles as ps_roi_pool
int ps_roi_align(bool forward, const void* in_or_grad, const int64_t* gs, int in_layout, const void* rois,
                 void* out_or_gi, int32_t* mapping, int out_layout, int dtype, long long N, long long C, long long H,
                 long long W, long long K, int PH, int PW, double scale, int sr, cudaStream_t st);

}  // namespace b200det
