// box_ops_sm100.cu -- the elementwise box work that sits between the heads and NMS in the
// detection post-processing, fused into one pass over all images:
//   clip_boxes_to_image (torchvision/ops/boxes.py:143-174)  ->  remove_small_boxes (:115-140)
//   ->  score threshold (rpn.py:283-285 `>=`, roi_heads.py:709-710 `>`)
// The reference runs these per image and compacts after each filter (boxes[keep]); here shapes stay
// static: the clipped boxes are written once and a validity byte per box tells b200det_nms_segmented
// which boxes never enter NMS.  16 bytes in, 17 bytes out per box: HBM-bound, one thread per box.
#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

template <typename T>
__global__ void clip_filter_kernel(const T* __restrict__ boxes, const T* __restrict__ scores,
                                   const int32_t* __restrict__ off, int S, const float* __restrict__ sizes,
                                   float min_size, float score_thresh, int strict, long long n,
                                   T* __restrict__ out, uint8_t* __restrict__ valid) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int lo = 0, hi = S;  // off[lo] <= i < off[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((long long)__ldg(off + mid) <= i) lo = mid; else hi = mid;
  }
  const float h = __ldg(sizes + 2 * lo), w = __ldg(sizes + 2 * lo + 1);
  // clamp(min=0, max=size) in the boxes' dtype, like Tensor.clamp: a NaN coordinate stays NaN (fminf / fmaxf would
  // turn it into 0), so ws / hs are NaN, `>= min_size` is false and the box is removed like in the reference
  // (boxes.py:137-139,167-168)
  auto clampT = [](float v, float m) { return v != v ? v : fminf(fmaxf(v, 0.f), m); };
  const T x1 = Io<T>::to(clampT((float)Io<T>::ld(boxes + i * 4 + 0), w)), y1 = Io<T>::to(clampT((float)Io<T>::ld(boxes + i * 4 + 1), h));
  const T x2 = Io<T>::to(clampT((float)Io<T>::ld(boxes + i * 4 + 2), w)), y2 = Io<T>::to(clampT((float)Io<T>::ld(boxes + i * 4 + 3), h));
  out[i * 4 + 0] = x1; out[i * 4 + 1] = y1; out[i * 4 + 2] = x2; out[i * 4 + 3] = y2;
  // ws, hs = x2 - x1, y2 - y1 in the boxes' dtype; keep = (ws >= min_size) & (hs >= min_size)
  const float ws = (float)Io<T>::to(__fsub_rn((float)x2, (float)x1)), hs = (float)Io<T>::to(__fsub_rn((float)y2, (float)y1));
  const float sc = (float)Io<T>::ld(scores + i);
  const bool ok = ws >= min_size && hs >= min_size && (strict ? sc > score_thresh : sc >= score_thresh);
  valid[i] = ok ? 1 : 0;
}

}  // namespace

int clip_filter_boxes(const void* boxes, const void* scores, int dtype, long long n, const int32_t* off, int S,
                      const float* sizes, double min_size, double score_thresh, int strict, void* out,
                      uint8_t* valid, cudaStream_t st) {
  if (n == 0) return 0;
  const int th = 256;
  const unsigned bl = (unsigned)((n + th - 1) / th);
  switch (dtype) {
    case B200DET_F32:
      clip_filter_kernel<float><<<bl, th, 0, st>>>((const float*)boxes, (const float*)scores, off, S, sizes,
                                                   (float)min_size, (float)score_thresh, strict, n, (float*)out, valid);
      break;
    case B200DET_F16:
      clip_filter_kernel<__half><<<bl, th, 0, st>>>((const __half*)boxes, (const __half*)scores, off, S, sizes,
                                                    (float)min_size, (float)score_thresh, strict, n, (__half*)out, valid);
      break;
    case B200DET_BF16:
      clip_filter_kernel<__nv_bfloat16><<<bl, th, 0, st>>>((const __nv_bfloat16*)boxes, (const __nv_bfloat16*)scores, off,
                                                           S, sizes, (float)min_size, (float)score_thresh, strict, n,
                                                           (__nv_bfloat16*)out, valid);
      break;
    default:
      return fail(B200DET_ENOSUP, "clip_filter_boxes: dtype %d not supported", dtype);
  }
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b200det
