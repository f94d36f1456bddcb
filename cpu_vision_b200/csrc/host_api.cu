// host_api.cu -- host-buffer entry points (b200det_*_host): what a caller that
// holds CPU tensors (the reference's CPU dispatch key,
// torchvision/csrc/ops/cpu/roi_align_kernel.cpp:282-386, cpu/nms_kernel.cpp:79-108)
// binds to.  Each call copies its operands into a per-process device arena,
// runs the device entry points of c_api.cu on a private stream and copies the
// results back before returning.  Pinned caller buffers get full PCIe rate;
// pageable ones work at whatever cudaMemcpyAsync's internal staging gives.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kMaxChunks = 16;

struct Arena {
  int device = -1;
  cudaStream_t stream = nullptr;      // kernels
  cudaStream_t s_in = nullptr;        // host -> device copies of the chunked entry points
  cudaStream_t s_out = nullptr;       // device -> host copies
  cudaEvent_t ev_in[kMaxChunks] = {}, ev_k[kMaxChunks] = {};
  void* stage = nullptr;              // pinned scratch (rebased RoIs)
  size_t stage_cap = 0;
  void* buf[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t cap[6] = {0, 0, 0, 0, 0, 0};
  void* pinned = nullptr;  // small bounce buffer for counts
};

Arena g_arena;
std::mutex g_mu;

void arena_destroy_aux() {
  if (g_arena.s_in) cudaStreamDestroy(g_arena.s_in);
  if (g_arena.s_out) cudaStreamDestroy(g_arena.s_out);
  for (int i = 0; i < kMaxChunks; ++i) {
    if (g_arena.ev_in[i]) cudaEventDestroy(g_arena.ev_in[i]);
    if (g_arena.ev_k[i]) cudaEventDestroy(g_arena.ev_k[i]);
  }
  if (g_arena.stage) cudaFreeHost(g_arena.stage);
}

int stage_get(size_t bytes, void** out) {
  if (g_arena.stage_cap < bytes) {
    if (g_arena.stage) B200DET_CUDA_OK(cudaFreeHost(g_arena.stage));
    g_arena.stage = nullptr;
    g_arena.stage_cap = 0;
    B200DET_CUDA_OK(cudaMallocHost(&g_arena.stage, bytes + bytes / 4 + 4096));
    g_arena.stage_cap = bytes + bytes / 4 + 4096;
  }
  *out = g_arena.stage;
  return 0;
}

// First RoI of every image when the rows are grouped by ascending, in-range image index
// (convert_boxes_to_roi_format order, ops/_utils.py:18-25); false otherwise.
bool image_starts(const void* rois, int dtype, int64_t K, int64_t N, std::vector<int64_t>& start) {
  start.assign((size_t)N + 1, K);
  int64_t prev = -1;
  for (int64_t k = 0; k < K; ++k) {
    const double bd = dtype == B200DET_F64 ? ((const double*)rois)[k * 5] : (double)((const float*)rois)[k * 5];
    const int64_t b = (int64_t)bd;
    if (b < 0 || b >= N || b < prev) return false;
    for (int64_t i = prev + 1; i <= b; ++i) start[(size_t)i] = k;
    prev = b;
  }
  return true;
}

// how many image chunks the host entry points pipeline over (H2D of chunk c+1, kernels of chunk c and
// D2H of chunk c-1 overlap; PCIe is full duplex).  B200DET_HOST_CHUNKS overrides (1 = no pipelining).
int host_chunks(int64_t N) {
  const char* e = getenv("B200DET_HOST_CHUNKS");  // read per call: the tests run 1, 3 and the default in one process
  int env = e ? atoi(e) : 16;
  if (env < 1) env = 1;
  if (env > kMaxChunks) env = kMaxChunks;
  return (int)(N < env ? (N > 0 ? N : 1) : env);
}

int arena_bind(int device) {
  if (g_arena.device != device) {
    if (g_arena.device >= 0) {
      cudaSetDevice(g_arena.device);
      for (int i = 0; i < 6; ++i) { if (g_arena.buf[i]) cudaFree(g_arena.buf[i]); g_arena.buf[i] = nullptr; g_arena.cap[i] = 0; }
      arena_destroy_aux();
      if (g_arena.stream) cudaStreamDestroy(g_arena.stream);
      if (g_arena.pinned) cudaFreeHost(g_arena.pinned);
      g_arena = Arena();
    }
    B200DET_CUDA_OK(cudaSetDevice(device));
    B200DET_CUDA_OK(cudaStreamCreateWithFlags(&g_arena.stream, cudaStreamNonBlocking));
    B200DET_CUDA_OK(cudaStreamCreateWithFlags(&g_arena.s_in, cudaStreamNonBlocking));
    B200DET_CUDA_OK(cudaStreamCreateWithFlags(&g_arena.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < kMaxChunks; ++i) {
      B200DET_CUDA_OK(cudaEventCreateWithFlags(&g_arena.ev_in[i], cudaEventDisableTiming));
      B200DET_CUDA_OK(cudaEventCreateWithFlags(&g_arena.ev_k[i], cudaEventDisableTiming));
    }
    B200DET_CUDA_OK(cudaMallocHost(&g_arena.pinned, 256));
    g_arena.device = device;
  } else {
    B200DET_CUDA_OK(cudaSetDevice(device));
  }
  return 0;
}

int arena_get(int slot, size_t bytes, void** out) {
  if (bytes == 0) bytes = 256;
  if (g_arena.cap[slot] < bytes) {
    if (g_arena.buf[slot]) B200DET_CUDA_OK(cudaFree(g_arena.buf[slot]));
    g_arena.buf[slot] = nullptr;
    g_arena.cap[slot] = 0;
    const size_t want = bytes + bytes / 8;
    B200DET_CUDA_OK(cudaMalloc(&g_arena.buf[slot], want));
    g_arena.cap[slot] = want;
  }
  *out = g_arena.buf[slot];
  return 0;
}

}  // namespace
}  // namespace b200det

using namespace b200det;

extern "C" {

int b200det_roi_align_forward_host(const void* input, int in_layout, const void* rois, void* output,
                                   int out_layout, int dtype, int64_t N, int64_t C, int64_t H,
                                   int64_t W, int64_t K, int64_t PH, int64_t PW,
                                   double spatial_scale, int sampling_ratio, int aligned,
                                   int device) {
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t es = dtype_size(dtype);
  if (!es) return fail(B200DET_ENOSUP, "roi_align_forward_host: unknown dtype %d", dtype);
  if (N < 0 || C < 0 || H < 0 || W < 0 || K < 0 || PH < 0 || PW < 0)
    return fail(B200DET_EINVAL, "roi_align_forward_host: negative size");
  int rc = arena_bind(device);
  if (rc) return rc;
  const size_t in_b = (size_t)N * C * H * W * es, roi_b = (size_t)K * 5 * (dtype == B200DET_F64 ? 8 : 4),
               out_b = (size_t)K * C * PH * PW * es;
  void *d_in, *d_roi, *d_out;
  if ((rc = arena_get(0, in_b, &d_in)) || (rc = arena_get(1, roi_b, &d_roi)) ||
      (rc = arena_get(2, out_b, &d_out)))
    return rc;
  cudaStream_t st = g_arena.stream;
  if (in_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_in, input, in_b, cudaMemcpyHostToDevice, st));
  if (roi_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_roi, rois, roi_b, cudaMemcpyHostToDevice, st));
  rc = b200det_roi_align_forward(d_in, in_layout, d_roi, d_out, out_layout, dtype, N, C, H, W, K, PH,
                                 PW, spatial_scale, sampling_ratio, aligned, st);
  if (rc) return rc;
  if (out_b) B200DET_CUDA_OK(cudaMemcpyAsync(output, d_out, out_b, cudaMemcpyDeviceToHost, st));
  B200DET_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

int b200det_roi_align_backward_host(const void* grad, const int64_t* grad_strides,
                                    int64_t grad_span_elems, const void* rois, void* grad_input,
                                    int gi_layout, int dtype, int64_t N, int64_t C, int64_t H,
                                    int64_t W, int64_t K, int64_t PH, int64_t PW,
                                    double spatial_scale, int sampling_ratio, int aligned,
                                    int device) {
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t es = dtype_size(dtype);
  if (!es) return fail(B200DET_ENOSUP, "roi_align_backward_host: unknown dtype %d", dtype);
  if (N < 0 || C < 0 || H < 0 || W < 0 || K < 0 || PH < 0 || PW < 0 || grad_span_elems < 0)
    return fail(B200DET_EINVAL, "roi_align_backward_host: negative size");
  int rc = arena_bind(device);
  if (rc) return rc;
  const size_t g_b = (size_t)grad_span_elems * es, roi_b = (size_t)K * 5 * (dtype == B200DET_F64 ? 8 : 4),
               gi_b = (size_t)N * C * H * W * es;
  void *d_g, *d_roi, *d_gi;
  if ((rc = arena_get(0, gi_b, &d_gi)) || (rc = arena_get(1, roi_b, &d_roi)) ||
      (rc = arena_get(2, g_b, &d_g)))
    return rc;
  cudaStream_t st = g_arena.stream;
  if (g_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_g, grad, g_b, cudaMemcpyHostToDevice, st));
  if (roi_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_roi, rois, roi_b, cudaMemcpyHostToDevice, st));
  rc = b200det_roi_align_backward(d_g, grad_strides, d_roi, d_gi, gi_layout, dtype, N, C, H, W, K,
                                  PH, PW, spatial_scale, sampling_ratio, aligned, 1, st);
  if (rc) return rc;
  if (gi_b) B200DET_CUDA_OK(cudaMemcpyAsync(grad_input, d_gi, gi_b, cudaMemcpyDeviceToHost, st));
  B200DET_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

// byte offsets of every level inside one arena slot (256-byte aligned)
static size_t level_offsets(const int64_t* heights, const int64_t* widths, int L, int64_t N,
                            int64_t C, size_t es, size_t* off) {
  size_t total = 0;
  for (int l = 0; l < L; ++l) {
    off[l] = total;
    total += ((size_t)N * C * heights[l] * widths[l] * es + 255) & ~(size_t)255;
  }
  return total;
}

int b200det_msroi_align_forward_host(const void* const* feats, const int64_t* heights,
                                     const int64_t* widths, const double* scales, int num_levels,
                                     int layout, const void* rois, int32_t* levels, void* output,
                                     int out_layout, int dtype, int64_t N, int64_t C, int64_t K,
                                     int64_t PH, int64_t PW, int sampling_ratio, int aligned,
                                     int k_min, int k_max, int canonical_scale,
                                     int canonical_level, double eps, int device) {
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t es = dtype_size(dtype);
  if (!es) return fail(B200DET_ENOSUP, "msroi_align_forward_host: unknown dtype %d", dtype);
  if (num_levels < 1 || num_levels > B200DET_MAX_LEVELS || !feats || !heights || !widths || !scales)
    return fail(B200DET_EINVAL, "msroi_align_forward_host: bad level description");
  if (N < 0 || C < 0 || K < 0 || PH < 0 || PW < 0)
    return fail(B200DET_EINVAL, "msroi_align_forward_host: negative size");
  int rc = arena_bind(device);
  if (rc) return rc;
  size_t off[B200DET_MAX_LEVELS];
  const size_t feat_b = level_offsets(heights, widths, num_levels, N, C, es, off);
  const size_t res = dtype == B200DET_F64 ? 8 : 4;
  const size_t roi_b = (size_t)K * 5 * res, row_b = (size_t)C * PH * PW * es, out_b = (size_t)K * row_b;
  void *d_feat, *d_roi, *d_out, *d_lv;
  if ((rc = arena_get(0, feat_b, &d_feat)) || (rc = arena_get(1, roi_b, &d_roi)) ||
      (rc = arena_get(2, out_b, &d_out)) || (rc = arena_get(3, (size_t)K * 4, &d_lv)))
    return rc;
  cudaStream_t st = g_arena.stream, s_in = g_arena.s_in, s_out = g_arena.s_out;
  const void* d_ptrs[B200DET_MAX_LEVELS];
  for (int l = 0; l < num_levels; ++l) d_ptrs[l] = (char*)d_feat + off[l];
  if (roi_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_roi, rois, roi_b, cudaMemcpyHostToDevice, st));

  // Pipeline over image chunks: the features of chunk c+1 go up while chunk c is pooled and the
  // pooled rows of chunk c-1 go down.  Needs RoIs grouped by image; otherwise one chunk.
  std::vector<int64_t> start;
  int chunks = host_chunks(N);
  if (chunks > 1 && (!rois || !image_starts(rois, dtype, K, N, start))) chunks = 1;
  if (chunks <= 1) start.assign((size_t)N + 1, K), start[0] = 0;
  for (int c = 0; c < chunks; ++c) {
    const int64_t i0 = N * c / chunks, i1 = N * (c + 1) / chunks;
    const int64_t k0 = chunks > 1 ? start[(size_t)i0] : 0, k1 = chunks > 1 ? start[(size_t)i1] : K;
    for (int l = 0; l < num_levels; ++l) {
      const size_t img_b = (size_t)C * heights[l] * widths[l] * es;
      if (img_b && i1 > i0)
        B200DET_CUDA_OK(cudaMemcpyAsync((char*)d_ptrs[l] + i0 * img_b, (const char*)feats[l] + i0 * img_b,
                                        (size_t)(i1 - i0) * img_b, cudaMemcpyHostToDevice, s_in));
    }
    B200DET_CUDA_OK(cudaEventRecord(g_arena.ev_in[c], s_in));
    B200DET_CUDA_OK(cudaStreamWaitEvent(st, g_arena.ev_in[c], 0));
    if (k1 > k0) {
      rc = b200det_msroi_align_forward(d_ptrs, heights, widths, scales, num_levels, layout,
                                       (const char*)d_roi + k0 * 5 * res, (int32_t*)d_lv + k0, 0,
                                       (char*)d_out + k0 * row_b, out_layout, dtype, N, C, k1 - k0, PH, PW,
                                       sampling_ratio, aligned, k_min, k_max, canonical_scale, canonical_level,
                                       eps, st);
      if (rc) return rc;
    }
    B200DET_CUDA_OK(cudaEventRecord(g_arena.ev_k[c], st));
    B200DET_CUDA_OK(cudaStreamWaitEvent(s_out, g_arena.ev_k[c], 0));
    if (k1 > k0 && row_b)
      B200DET_CUDA_OK(cudaMemcpyAsync((char*)output + k0 * row_b, (char*)d_out + k0 * row_b,
                                      (size_t)(k1 - k0) * row_b, cudaMemcpyDeviceToHost, s_out));
  }
  if (levels && K && num_levels > 1)
    B200DET_CUDA_OK(cudaMemcpyAsync(levels, d_lv, (size_t)K * 4, cudaMemcpyDeviceToHost, s_out));
  B200DET_CUDA_OK(cudaStreamSynchronize(s_out));
  B200DET_CUDA_OK(cudaStreamSynchronize(st));
  if (levels && num_levels == 1) memset(levels, 0, (size_t)K * 4);
  return 0;
}

int b200det_msroi_align_backward_host(const void* grad, const int64_t* grad_strides,
                                      int64_t grad_span_elems, const void* rois,
                                      const int32_t* levels, void* const* grad_feats,
                                      const int64_t* heights, const int64_t* widths,
                                      const double* scales, int num_levels, int layout, int dtype,
                                      int64_t N, int64_t C, int64_t K, int64_t PH, int64_t PW,
                                      int sampling_ratio, int aligned, int device) {
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t es = dtype_size(dtype);
  if (!es) return fail(B200DET_ENOSUP, "msroi_align_backward_host: unknown dtype %d", dtype);
  if (num_levels < 1 || num_levels > B200DET_MAX_LEVELS || !grad_feats || !heights || !widths || !scales)
    return fail(B200DET_EINVAL, "msroi_align_backward_host: bad level description");
  if (N < 0 || C < 0 || K < 0 || PH < 0 || PW < 0 || grad_span_elems < 0)
    return fail(B200DET_EINVAL, "msroi_align_backward_host: negative size");
  if (num_levels > 1 && K > 0 && !levels)
    return fail(B200DET_EINVAL, "msroi_align_backward_host: levels required");
  if (K > 0 && (!grad || !grad_strides || !rois))
    return fail(B200DET_EINVAL, "msroi_align_backward_host: null pointer");
  int rc = arena_bind(device);
  if (rc) return rc;
  size_t off[B200DET_MAX_LEVELS];
  const size_t feat_b = level_offsets(heights, widths, num_levels, N, C, es, off);
  const size_t res = dtype == B200DET_F64 ? 8 : 4;
  const size_t roi_b = (size_t)K * 5 * res, g_b = (size_t)grad_span_elems * es;
  void *d_feat, *d_roi, *d_g, *d_lv;
  if ((rc = arena_get(0, feat_b, &d_feat)) || (rc = arena_get(1, roi_b, &d_roi)) ||
      (rc = arena_get(2, g_b, &d_g)) || (rc = arena_get(3, (size_t)K * 4, &d_lv)))
    return rc;
  cudaStream_t st = g_arena.stream, s_in = g_arena.s_in, s_out = g_arena.s_out;
  char* d_ptrs[B200DET_MAX_LEVELS];
  for (int l = 0; l < num_levels; ++l) d_ptrs[l] = (char*)d_feat + off[l];

  // Pipeline over image chunks (see the forward).  Needs RoIs grouped by image and a gradient whose
  // rows are laid out one after the other (stride of k = size of a row block); otherwise one chunk.
  std::vector<int64_t> start;
  int chunks = host_chunks(N);
  const bool rows_dense = K > 0 && grad_strides[0] > 0 && grad_strides[0] * K == grad_span_elems;
  if (chunks > 1 && (!rows_dense || !image_starts(rois, dtype, K, N, start))) chunks = 1;
  if (levels && K) B200DET_CUDA_OK(cudaMemcpyAsync(d_lv, levels, (size_t)K * 4, cudaMemcpyHostToDevice, st));
  if (chunks <= 1) {
    if (roi_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_roi, rois, roi_b, cudaMemcpyHostToDevice, st));
    if (g_b) B200DET_CUDA_OK(cudaMemcpyAsync(d_g, grad, g_b, cudaMemcpyHostToDevice, st));
    rc = b200det_msroi_align_backward(d_g, grad_strides, d_roi, (const int32_t*)d_lv, (void* const*)d_ptrs,
                                      heights, widths, scales, num_levels, layout, dtype, N, C, K, PH, PW,
                                      sampling_ratio, aligned, B200DET_BWD_ZERO_INIT, st);
    if (rc) return rc;
    for (int l = 0; l < num_levels; ++l) {
      const size_t b = (size_t)N * C * heights[l] * widths[l] * es;
      if (b) B200DET_CUDA_OK(cudaMemcpyAsync(grad_feats[l], d_ptrs[l], b, cudaMemcpyDeviceToHost, st));
    }
    B200DET_CUDA_OK(cudaStreamSynchronize(st));
    return 0;
  }
  // each chunk is a call on its own images: batch indices rebased to the chunk (pinned staging copy)
  void* h_roi;
  if ((rc = stage_get(roi_b, &h_roi))) return rc;
  memcpy(h_roi, rois, roi_b);
  for (int c = 0; c < chunks; ++c) {
    const int64_t i0 = N * c / chunks;
    for (int64_t k = start[(size_t)i0]; k < start[(size_t)(N * (c + 1) / chunks)]; ++k) {
      if (dtype == B200DET_F64) ((double*)h_roi)[k * 5] -= (double)i0;
      else ((float*)h_roi)[k * 5] -= (float)i0;
    }
  }
  B200DET_CUDA_OK(cudaMemcpyAsync(d_roi, h_roi, roi_b, cudaMemcpyHostToDevice, st));
  const size_t gk_b = (size_t)grad_strides[0] * es;
  for (int c = 0; c < chunks; ++c) {
    const int64_t i0 = N * c / chunks, i1 = N * (c + 1) / chunks;
    const int64_t k0 = start[(size_t)i0], k1 = start[(size_t)i1];
    if (k1 > k0)
      B200DET_CUDA_OK(cudaMemcpyAsync((char*)d_g + k0 * gk_b, (const char*)grad + k0 * gk_b, (size_t)(k1 - k0) * gk_b,
                                      cudaMemcpyHostToDevice, s_in));
    B200DET_CUDA_OK(cudaEventRecord(g_arena.ev_in[c], s_in));
    B200DET_CUDA_OK(cudaStreamWaitEvent(st, g_arena.ev_in[c], 0));
    void* c_ptrs[B200DET_MAX_LEVELS];
    for (int l = 0; l < num_levels; ++l) c_ptrs[l] = d_ptrs[l] + (size_t)i0 * C * heights[l] * widths[l] * es;
    rc = b200det_msroi_align_backward((char*)d_g + k0 * gk_b, grad_strides, (char*)d_roi + k0 * 5 * res,
                                      (const int32_t*)d_lv + k0, c_ptrs, heights, widths, scales, num_levels,
                                      layout, dtype, i1 - i0, C, k1 - k0, PH, PW, sampling_ratio, aligned,
                                      B200DET_BWD_ZERO_INIT, st);
    if (rc) return rc;
    B200DET_CUDA_OK(cudaEventRecord(g_arena.ev_k[c], st));
    B200DET_CUDA_OK(cudaStreamWaitEvent(s_out, g_arena.ev_k[c], 0));
    for (int l = 0; l < num_levels; ++l) {
      const size_t img_b = (size_t)C * heights[l] * widths[l] * es;
      if (img_b && i1 > i0)
        B200DET_CUDA_OK(cudaMemcpyAsync((char*)grad_feats[l] + i0 * img_b, (char*)c_ptrs[l], (size_t)(i1 - i0) * img_b,
                                        cudaMemcpyDeviceToHost, s_out));
    }
  }
  B200DET_CUDA_OK(cudaStreamSynchronize(s_out));
  B200DET_CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

static int nms_host_common(const void* boxes, const void* scores, const int64_t* idxs, int dtype,
                           int64_t N, double thr, int mode, int64_t* keep, int64_t* num_keep,
                           int device) {
  std::lock_guard<std::mutex> lock(g_mu);
  const size_t es = dtype_size(dtype);
  if (!es) return fail(B200DET_ENOSUP, "nms_host: unknown dtype %d", dtype);
  if (N < 0) return fail(B200DET_EINVAL, "nms_host: negative N");
  if (!num_keep) return fail(B200DET_EINVAL, "nms_host: null num_keep");
  if (N == 0) { *num_keep = 0; return 0; }
  int rc = arena_bind(device);
  if (rc) return rc;
  const size_t ws_b = b200det_nms_workspace_bytes(N);
  void *d_boxes, *d_scores, *d_idx, *d_keep, *d_ws;
  if ((rc = arena_get(0, (size_t)N * 4 * es, &d_boxes)) || (rc = arena_get(1, (size_t)N * es, &d_scores)) ||
      (rc = arena_get(2, (size_t)N * 8, &d_idx)) || (rc = arena_get(3, (size_t)(N + 1) * 8, &d_keep)) ||
      (rc = arena_get(4, ws_b, &d_ws)))
    return rc;
  cudaStream_t st = g_arena.stream;
  B200DET_CUDA_OK(cudaMemcpyAsync(d_boxes, boxes, (size_t)N * 4 * es, cudaMemcpyHostToDevice, st));
  B200DET_CUDA_OK(cudaMemcpyAsync(d_scores, scores, (size_t)N * es, cudaMemcpyHostToDevice, st));
  int64_t* d_cnt = (int64_t*)d_keep + N;
  if (idxs) {
    B200DET_CUDA_OK(cudaMemcpyAsync(d_idx, idxs, (size_t)N * 8, cudaMemcpyHostToDevice, st));
    rc = b200det_batched_nms(d_boxes, d_scores, (const int64_t*)d_idx, dtype, N, thr, mode,
                             (int64_t*)d_keep, d_cnt, d_ws, ws_b, st);
  } else {
    rc = b200det_nms(d_boxes, d_scores, dtype, N, thr, (int64_t*)d_keep, d_cnt, d_ws, ws_b, st);
  }
  if (rc) return rc;
  // one copy brings back indices and the trailing count
  int64_t* bounce = nullptr;
  B200DET_CUDA_OK(cudaMemcpyAsync(g_arena.pinned, d_cnt, 8, cudaMemcpyDeviceToHost, st));
  B200DET_CUDA_OK(cudaStreamSynchronize(st));
  bounce = (int64_t*)g_arena.pinned;
  *num_keep = bounce[0];
  if (*num_keep > 0)
    B200DET_CUDA_OK(cudaMemcpy(keep, d_keep, (size_t)(*num_keep) * 8, cudaMemcpyDeviceToHost));
  return 0;
}

int b200det_nms_host(const void* boxes, const void* scores, int dtype, int64_t N,
                     double iou_threshold, int64_t* keep, int64_t* num_keep, int device) {
  return nms_host_common(boxes, scores, nullptr, dtype, N, iou_threshold, 0, keep, num_keep, device);
}

int b200det_batched_nms_host(const void* boxes, const void* scores, const int64_t* idxs, int dtype,
                             int64_t N, double iou_threshold, int mode, int64_t* keep,
                             int64_t* num_keep, int device) {
  if (!idxs && N > 0) return fail(B200DET_EINVAL, "batched_nms_host: null idxs");
  return nms_host_common(boxes, scores, idxs, dtype, N, iou_threshold, mode, keep, num_keep, device);
}

void b200det_host_release(void) {
  std::lock_guard<std::mutex> lock(g_mu);
  roi_align_bwd_release();
  roi_align_bwd_strip_release();
  if (g_arena.device < 0) return;
  cudaSetDevice(g_arena.device);
  for (int i = 0; i < 6; ++i) {
    if (g_arena.buf[i]) cudaFree(g_arena.buf[i]);
    g_arena.buf[i] = nullptr;
    g_arena.cap[i] = 0;
  }
  arena_destroy_aux();
  if (g_arena.stream) cudaStreamDestroy(g_arena.stream);
  if (g_arena.pinned) cudaFreeHost(g_arena.pinned);
  g_arena = Arena();
}

}  // extern "C"
