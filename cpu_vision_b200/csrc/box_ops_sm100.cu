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

// clamp(min=0, max=m) like Tensor.clamp: NaN stays NaN
__device__ __forceinline__ float clamp_keep_nan(float v, float m) { return v != v ? v : fminf(fmaxf(v, 0.f), m); }

// ---- RPN side: everything between the per-level top-k and NMS in ONE pass (rpn.py:258-285) ----------------
// candidate i = (image b, rank j): a = top_idx[b][j]; box = proposals[b][a]; score = sigmoid(objectness[b][a]);
// level = the level anchor a belongs to; clip to the image; small-box and score filters as a validity byte.
// Replaces three advanced-index gathers, torch.sigmoid, the level gather and the clip / filter kernel.
// sigmoid is torch's CUDA formula 1 / (1 + exp(-x)) in fp32 (aten/src/ATen/native/cuda/UnarySpecialOpsKernel.cu),
// so scores are bit-identical to the eager composition on the GPU.
struct LevelEnds { int n; int end[8]; };
__global__ void rpn_select_kernel(const float* __restrict__ proposals, const float* __restrict__ objectness,
                                  const int64_t* __restrict__ top_idx, int B, long long A, int P,
                                  const float* __restrict__ sizes, LevelEnds le, float min_size, float score_thresh,
                                  float* __restrict__ boxes, float* __restrict__ scores, int64_t* __restrict__ levels,
                                  uint8_t* __restrict__ valid) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)B * P) return;
  const int b = (int)(i / P);
  const long long a = __ldg(top_idx + i);
  const float4 bx = __ldg(reinterpret_cast<const float4*>(proposals) + (b * A + a));
  const float obj = __ldg(objectness + b * A + a);
  const float sc = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-obj)));
  int lvl = 0;
  while (lvl + 1 < le.n && a >= le.end[lvl]) ++lvl;
  const float h = __ldg(sizes + 2 * b), w = __ldg(sizes + 2 * b + 1);
  const float x1 = clamp_keep_nan(bx.x, w), y1 = clamp_keep_nan(bx.y, h), x2 = clamp_keep_nan(bx.z, w),
              y2 = clamp_keep_nan(bx.w, h);
  reinterpret_cast<float4*>(boxes)[i] = make_float4(x1, y1, x2, y2);
  scores[i] = sc;
  levels[i] = lvl;
  const float ws = __fsub_rn(x2, x1), hs = __fsub_rn(y2, y1);
  valid[i] = (ws >= min_size && hs >= min_size && sc >= score_thresh) ? 1 : 0;
}

// ---- RoIHeads side: softmax + BoxCoder.decode + clip + filters in ONE pass (roi_heads.py:682-716) ----------
// A warp per proposal r: softmax over the NC class logits (max and sum by warp reduction), then for every class
// c >= 1 the box decoded from box_regression[r][4c .. 4c+3] in the reference's operation order
// (models/detection/_utils.py:183-224; the divisions by the coder weights are multiplications by the fp32
// reciprocal, as torch's CUDA division by a Python scalar computes them), clipped to the image, with the
// validity byte of the score (>) and small-box (1e-2) filters.  Output instance index = r * (NC - 1) + (c - 1).
__global__ void roi_heads_decode_kernel(const float* __restrict__ logits, const float* __restrict__ reg,
                                        const float* __restrict__ props, const int32_t* __restrict__ image_of,
                                        const float* __restrict__ sizes, int R, int NC, float iwx, float iwy,
                                        float iww, float iwh, float xform_clip, float score_thresh, float min_size,
                                        float* __restrict__ boxes, float* __restrict__ scores,
                                        uint8_t* __restrict__ valid) {
  const int r = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= R) return;
  const float* lg = logits + (long long)r * NC;
  float m = -INFINITY;
  for (int c = lane; c < NC; c += 32) m = fmaxf(m, __ldg(lg + c));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float s = 0.f;
  for (int c = lane; c < NC; c += 32) s += expf(__fsub_rn(__ldg(lg + c), m));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float4 pb = __ldg(reinterpret_cast<const float4*>(props) + r);
  const float widths = __fsub_rn(pb.z, pb.x), heights = __fsub_rn(pb.w, pb.y);
  const float ctr_x = __fadd_rn(pb.x, __fmul_rn(0.5f, widths)), ctr_y = __fadd_rn(pb.y, __fmul_rn(0.5f, heights));
  const int img = __ldg(image_of + r);
  const float h = __ldg(sizes + 2 * img), w = __ldg(sizes + 2 * img + 1);
  for (int c = 1 + lane; c < NC; c += 32) {
    const float sc = __fdiv_rn(expf(__fsub_rn(__ldg(lg + c), m)), s);
    const float4 d = __ldg(reinterpret_cast<const float4*>(reg + (long long)r * NC * 4) + c);
    const float dx = __fmul_rn(d.x, iwx), dy = __fmul_rn(d.y, iwy);
    float dw = __fmul_rn(d.z, iww), dh = __fmul_rn(d.w, iwh);
    dw = dw != dw ? dw : fminf(dw, xform_clip);   // torch.clamp(max=...) keeps NaN
    dh = dh != dh ? dh : fminf(dh, xform_clip);
    const float pcx = __fadd_rn(__fmul_rn(dx, widths), ctr_x), pcy = __fadd_rn(__fmul_rn(dy, heights), ctr_y);
    const float pw = __fmul_rn(expf(dw), widths), ph = __fmul_rn(expf(dh), heights);
    const float cw = __fmul_rn(0.5f, pw), chh = __fmul_rn(0.5f, ph);
    const float x1 = clamp_keep_nan(__fsub_rn(pcx, cw), w), y1 = clamp_keep_nan(__fsub_rn(pcy, chh), h);
    const float x2 = clamp_keep_nan(__fadd_rn(pcx, cw), w), y2 = clamp_keep_nan(__fadd_rn(pcy, chh), h);
    const long long o = (long long)r * (NC - 1) + (c - 1);
    reinterpret_cast<float4*>(boxes)[o] = make_float4(x1, y1, x2, y2);
    scores[o] = sc;
    const float ws = __fsub_rn(x2, x1), hs = __fsub_rn(y2, y1);
    valid[o] = (ws >= min_size && hs >= min_size && sc > score_thresh) ? 1 : 0;
  }
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

int rpn_select(const float* proposals, const float* objectness, const int64_t* top_idx, int B, long long A, int P,
               const float* sizes, const int* level_ends, int num_levels, double min_size, double score_thresh,
               float* boxes, float* scores, int64_t* levels, uint8_t* valid, cudaStream_t st) {
  if ((long long)B * P == 0) return 0;
  LevelEnds le;
  le.n = num_levels;
  for (int l = 0; l < 8; ++l) le.end[l] = l < num_levels ? level_ends[l] : 0x7fffffff;
  const int th = 256;
  const unsigned bl = (unsigned)(((long long)B * P + th - 1) / th);
  rpn_select_kernel<<<bl, th, 0, st>>>(proposals, objectness, top_idx, B, A, P, sizes, le, (float)min_size,
                                       (float)score_thresh, boxes, scores, levels, valid);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

int roi_heads_decode(const float* logits, const float* reg, const float* props, const int32_t* image_of,
                     const float* sizes, int R, int NC, const double* weights, double xform_clip, double score_thresh,
                     double min_size, float* boxes, float* scores, uint8_t* valid, cudaStream_t st) {
  if (R == 0 || NC < 2) return 0;
  const int th = 256;  // 8 proposals per CTA
  const unsigned bl = (unsigned)(((long long)R * 32 + th - 1) / th);
  roi_heads_decode_kernel<<<bl, th, 0, st>>>(logits, reg, props, image_of, sizes, R, NC, 1.0f / (float)weights[0],
                                             1.0f / (float)weights[1], 1.0f / (float)weights[2],
                                             1.0f / (float)weights[3], (float)xform_clip, (float)score_thresh,
                                             (float)min_size, boxes, scores, valid);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b200det
