/*
 * detection_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C restatement of the reference's CPU algorithms for the
 * roi_align / nms hot path.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; the product (cpu_vision_b200/) never does.
 *
 * Parity is PINNED: tests/test_oracle.py checks every function here against
 *   (a) the reference's own known-answer vectors (SURVEY.md section 8a, K1-K7,
 *       taken from test/test_ops.py and test/test_onnx.py fixtures),
 *   (b) golden vectors generated from the reference's compiled CPU kernels
 *       and Python package (tests/golden/, made by tests/golden/make_golden.py),
 *   (c) oracle/_ref/libtvref.so -- the reference's unmodified
 *       cpu/roi_align_kernel.cpp and cpu/nms_kernel.cpp compiled where they lie
 *       (oracle/build.py) -- bit for bit on random inputs.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/build.py).
 * -ffp-contract=off matters: the reference's CPU build has no FMA contraction
 * (x86-64 baseline, setup.py passes no -march), and sample coordinates must
 * truncate identically.
 *
 * Citations are paths below /root/reference/torchvision/.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------
 * RoIAlign forward.  csrc/ops/cpu/roi_align_kernel.cpp:11-108 with the
 * pre-computed table of csrc/ops/cpu/roi_align_common.h:32-124 folded into the
 * loop (same arithmetic, same order: w1*v1 + w2*v2 + w3*v3 + w4*v4 summed over
 * iy, ix, then divided by count).
 * input NCHW contiguous, rois [K,5], output [K,C,PH,PW].
 * --------------------------------------------------------------------- */
#define DEFINE_ROI_ALIGN(T, SUFFIX, CEIL)                                                        \
  typedef struct { int pos1, pos2, pos3, pos4; T w1, w2, w3, w4; } precalc_##SUFFIX;              \
                                                                                                  \
  /* csrc/ops/cpu/roi_align_common.h:32-124 */                                                    \
  static void pre_calc_##SUFFIX(int height, int width, int ph_n, int pw_n, T start_h, T start_w,  \
                                T bin_h, T bin_w, int gh, int gw, precalc_##SUFFIX* pc) {         \
    int idx = 0;                                                                                  \
    for (int ph = 0; ph < ph_n; ph++)                                                             \
      for (int pw = 0; pw < pw_n; pw++)                                                           \
        for (int iy = 0; iy < gh; iy++) {                                                         \
          const T yy = start_h + ph * bin_h + (T)(iy + .5f) * bin_h / (T)gh;                      \
          for (int ix = 0; ix < gw; ix++) {                                                       \
            const T xx = start_w + pw * bin_w + (T)(ix + .5f) * bin_w / (T)gw;                    \
            T x = xx, y = yy;                                                                     \
            precalc_##SUFFIX e;                                                                   \
            if (y < -1.0 || y > height || x < -1.0 || x > width) {                                \
              memset(&e, 0, sizeof(e));                                                           \
              pc[idx++] = e;                                                                      \
              continue;                                                                           \
            }                                                                                     \
            if (y <= 0) y = 0;                                                                    \
            if (x <= 0) x = 0;                                                                    \
            int y_low = (int)y, x_low = (int)x, y_high, x_high;                                   \
            if (y_low >= height - 1) { y_high = y_low = height - 1; y = (T)y_low; }               \
            else y_high = y_low + 1;                                                              \
            if (x_low >= width - 1) { x_high = x_low = width - 1; x = (T)x_low; }                 \
            else x_high = x_low + 1;                                                              \
            const T ly = y - y_low, lx = x - x_low;                                               \
            const T hy = (T)(1. - ly), hx = (T)(1. - lx);                                         \
            e.pos1 = y_low * width + x_low;  e.pos2 = y_low * width + x_high;                     \
            e.pos3 = y_high * width + x_low; e.pos4 = y_high * width + x_high;                    \
            e.w1 = hy * hx; e.w2 = hy * lx; e.w3 = ly * hx; e.w4 = ly * lx;                       \
            pc[idx++] = e;                                                                        \
          }                                                                                       \
        }                                                                                         \
  }                                                                                               \
                                                                                                  \
  /* csrc/ops/cpu/roi_align_kernel.cpp:11-108 */                                                  \
  API void oracle_roi_align_forward_##SUFFIX(                                                     \
      const T* input, const T* rois, T* output, int n_rois, int channels, int height, int width,  \
      int pooled_height, int pooled_width, double spatial_scale_d, int sampling_ratio,            \
      int aligned) {                                                                              \
    const T spatial_scale = (T)spatial_scale_d;                                                   \
    for (int n = 0; n < n_rois; n++) {                                                            \
      const long index_n = (long)n * channels * pooled_width * pooled_height;                     \
      const T* r = rois + (long)n * 5;                                                            \
      const int batch = (int)r[0];                                                                \
      const T offset = aligned ? (T)0.5 : (T)0.0;                                                 \
      const T start_w = r[1] * spatial_scale - offset, start_h = r[2] * spatial_scale - offset;   \
      const T end_w = r[3] * spatial_scale - offset, end_h = r[4] * spatial_scale - offset;       \
      T roi_w = end_w - start_w, roi_h = end_h - start_h;                                         \
      if (!aligned) {                                                                             \
        roi_w = roi_w < (T)1. ? (T)1. : roi_w; /* std::max(roi_width, 1) */                       \
        roi_h = roi_h < (T)1. ? (T)1. : roi_h;                                                    \
      }                                                                                           \
      const T bin_h = roi_h / (T)pooled_height, bin_w = roi_w / (T)pooled_width;                  \
      const int gh = sampling_ratio > 0 ? sampling_ratio : (int)CEIL(roi_h / pooled_height);      \
      const int gw = sampling_ratio > 0 ? sampling_ratio : (int)CEIL(roi_w / pooled_width);       \
      const T count = (T)(gh * gw > 1 ? gh * gw : 1);                                             \
      long n_pc = (long)gh * gw * pooled_width * pooled_height;                                   \
      if (gh <= 0 || gw <= 0) n_pc = 0;                                                           \
      precalc_##SUFFIX* pc = (precalc_##SUFFIX*)malloc((n_pc > 0 ? n_pc : 1) * sizeof(*pc));      \
      if (n_pc > 0)                                                                               \
        pre_calc_##SUFFIX(height, width, pooled_height, pooled_width, start_h, start_w, bin_h,    \
                          bin_w, gh, gw, pc);                                                     \
      for (int c = 0; c < channels; c++) {                                                        \
        const long index_n_c = index_n + (long)c * pooled_width * pooled_height;                  \
        const T* in = input + ((long)batch * channels + c) * height * width;                      \
        long pi = 0;                                                                              \
        for (int ph = 0; ph < pooled_height; ph++)                                                \
          for (int pw = 0; pw < pooled_width; pw++) {                                             \
            T val = 0.;                                                                           \
            for (int iy = 0; iy < gh; iy++)                                                       \
              for (int ix = 0; ix < gw; ix++) {                                                   \
                const precalc_##SUFFIX e = pc[pi++];                                              \
                val += e.w1 * in[e.pos1] + e.w2 * in[e.pos2] + e.w3 * in[e.pos3] +                \
                       e.w4 * in[e.pos4];                                                         \
              }                                                                                   \
            val /= count;                                                                         \
            output[index_n_c + (long)ph * pooled_width + pw] = val;                               \
          }                                                                                       \
      }                                                                                           \
      free(pc);                                                                                   \
    }                                                                                             \
  }                                                                                               \
                                                                                                  \
  /* csrc/ops/cpu/roi_align_kernel.cpp:110-172 (bilinear_interpolate_gradient) and               \
   * :174-280 (backward loop).  grad has element strides (n,c,h,w); grad_input is NCHW           \
   * contiguous and must be zero on entry (the reference allocates it with at::zeros, :351). */  \
  API void oracle_roi_align_backward_##SUFFIX(                                                    \
      const T* grad, long gs_n, long gs_c, long gs_h, long gs_w, const T* rois, T* grad_input,    \
      int n_rois, int channels, int height, int width, int pooled_height, int pooled_width,       \
      double spatial_scale_d, int sampling_ratio, int aligned) {                                  \
    const T spatial_scale = (T)spatial_scale_d;                                                   \
    const long nthreads = (long)n_rois * channels * pooled_height * pooled_width;                 \
    for (long index = 0; index < nthreads; index++) {                                             \
      const int pw = (int)(index % pooled_width);                                                 \
      const int ph = (int)((index / pooled_width) % pooled_height);                               \
      const int c = (int)((index / pooled_width / pooled_height) % channels);                     \
      const long n = index / pooled_width / pooled_height / channels;                             \
      const T* r = rois + n * 5;                                                                  \
      const int batch = (int)r[0];                                                                \
      const T offset = aligned ? (T)0.5 : (T)0.0;                                                 \
      const T start_w = r[1] * spatial_scale - offset, start_h = r[2] * spatial_scale - offset;   \
      const T end_w = r[3] * spatial_scale - offset, end_h = r[4] * spatial_scale - offset;       \
      T roi_w = end_w - start_w, roi_h = end_h - start_h;                                         \
      if (!aligned) {                                                                             \
        roi_w = roi_w < (T)1. ? (T)1. : roi_w;                                                    \
        roi_h = roi_h < (T)1. ? (T)1. : roi_h;                                                    \
      }                                                                                           \
      const T bin_h = roi_h / (T)pooled_height, bin_w = roi_w / (T)pooled_width;                  \
      T* gi = grad_input + ((long)batch * channels + c) * height * width;                         \
      const T go = grad[n * gs_n + c * gs_c + ph * gs_h + pw * gs_w];                             \
      const int gh = sampling_ratio > 0 ? sampling_ratio : (int)CEIL(roi_h / pooled_height);      \
      const int gw = sampling_ratio > 0 ? sampling_ratio : (int)CEIL(roi_w / pooled_width);       \
      const T count = (T)(gh * gw);                                                               \
      for (int iy = 0; iy < gh; iy++) {                                                           \
        const T yy = start_h + ph * bin_h + (T)(iy + .5f) * bin_h / (T)gh;                        \
        for (int ix = 0; ix < gw; ix++) {                                                         \
          const T xx = start_w + pw * bin_w + (T)(ix + .5f) * bin_w / (T)gw;                      \
          T x = xx, y = yy;                                                                       \
          if (y < -1.0 || y > height || x < -1.0 || x > width) continue;                          \
          if (y <= 0) y = 0;                                                                      \
          if (x <= 0) x = 0;                                                                      \
          int y_low = (int)y, x_low = (int)x, y_high, x_high;                                     \
          if (y_low >= height - 1) { y_high = y_low = height - 1; y = (T)y_low; }                 \
          else y_high = y_low + 1;                                                                \
          if (x_low >= width - 1) { x_high = x_low = width - 1; x = (T)x_low; }                   \
          else x_high = x_low + 1;                                                                \
          const T ly = y - y_low, lx = x - x_low;                                                 \
          const T hy = (T)(1. - ly), hx = (T)(1. - lx);                                           \
          const T w1 = hy * hx, w2 = hy * lx, w3 = ly * hx, w4 = ly * lx;                         \
          const T g1 = go * w1 / count, g2 = go * w2 / count;                                     \
          const T g3 = go * w3 / count, g4 = go * w4 / count;                                     \
          gi[y_low * width + x_low] += g1;                                                        \
          gi[y_low * width + x_high] += g2;                                                       \
          gi[y_high * width + x_low] += g3;                                                       \
          gi[y_high * width + x_high] += g4;                                                      \
        }                                                                                         \
      }                                                                                           \
    }                                                                                             \
  }

DEFINE_ROI_ALIGN(float, f32, ceilf)
DEFINE_ROI_ALIGN(double, f64, ceil)

/* ------------------------------------------------------------------------
 * NMS.  csrc/ops/cpu/nms_kernel.cpp:9-77: areas = (x2-x1)*(y2-y1) (:28),
 * stable descending sort of the scores (:30-31), greedy double loop with
 * ovr = inter / (iarea + areas[j] - inter) and `ovr > iou_threshold` (:48-75).
 *
 * thr_is_float != 0 compares against (float)iou_threshold the way the CUDA
 * kernel being replaced does (csrc/ops/cuda/nms_kernel.cu:20,28); 0 keeps the
 * CPU kernel's double comparison.  They differ only when ovr == (float)thr
 * exactly and (float)thr > thr (SURVEY.md N-a).
 * Returns the number of kept boxes; keep[] receives original indices in
 * descending-score order.
 * --------------------------------------------------------------------- */
typedef struct { double key; int64_t idx; } sort_item;

static void stable_sort_desc(sort_item* a, sort_item* tmp, int64_t n) {
  /* bottom-up merge sort: stable, descending by key, NaN keys first like
   * torch.sort(descending=True) which treats NaN as the largest value */
  for (int64_t w = 1; w < n; w *= 2) {
    for (int64_t lo = 0; lo < n; lo += 2 * w) {
      int64_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int64_t i = lo, j = mid, k = lo;
      while (i < mid && j < hi) {
        const double x = a[i].key, y = a[j].key;
        /* take from the right run only if it is strictly greater */
        const int right_greater = (y != y) ? (x == x) : (x == x && y > x);
        if (right_greater) tmp[k++] = a[j++]; else tmp[k++] = a[i++];
      }
      while (i < mid) tmp[k++] = a[i++];
      while (j < hi) tmp[k++] = a[j++];
    }
    memcpy(a, tmp, (size_t)n * sizeof(sort_item));
  }
}

#define DEFINE_NMS(T, SUFFIX)                                                                     \
  API int64_t oracle_nms_##SUFFIX(const T* dets, const T* scores, int64_t n, double iou_threshold, \
                                  int thr_is_float, int64_t* keep) {                              \
    if (n == 0) return 0;                                                                         \
    T* areas = (T*)malloc((size_t)n * sizeof(T));                                                 \
    sort_item* order = (sort_item*)malloc((size_t)n * sizeof(sort_item));                         \
    sort_item* tmp = (sort_item*)malloc((size_t)n * sizeof(sort_item));                           \
    unsigned char* suppressed = (unsigned char*)calloc((size_t)n, 1);                             \
    for (int64_t i = 0; i < n; i++) {                                                             \
      areas[i] = (dets[4 * i + 2] - dets[4 * i + 0]) * (dets[4 * i + 3] - dets[4 * i + 1]);       \
      order[i].key = (double)scores[i];                                                           \
      order[i].idx = i;                                                                           \
    }                                                                                             \
    stable_sort_desc(order, tmp, n);                                                              \
    const float thr_f = (float)iou_threshold;                                                     \
    int64_t num = 0;                                                                              \
    for (int64_t _i = 0; _i < n; _i++) {                                                          \
      const int64_t i = order[_i].idx;                                                            \
      if (suppressed[i]) continue;                                                                \
      keep[num++] = i;                                                                            \
      const T ix1 = dets[4 * i], iy1 = dets[4 * i + 1], ix2 = dets[4 * i + 2],                    \
              iy2 = dets[4 * i + 3], iarea = areas[i];                                            \
      for (int64_t _j = _i + 1; _j < n; _j++) {                                                   \
        const int64_t j = order[_j].idx;                                                          \
        if (suppressed[j]) continue;                                                              \
        const T xx1 = ix1 > dets[4 * j] ? ix1 : dets[4 * j];                                      \
        const T yy1 = iy1 > dets[4 * j + 1] ? iy1 : dets[4 * j + 1];                              \
        const T xx2 = ix2 < dets[4 * j + 2] ? ix2 : dets[4 * j + 2];                              \
        const T yy2 = iy2 < dets[4 * j + 3] ? iy2 : dets[4 * j + 3];                              \
        const T w = (xx2 - xx1) > (T)0 ? (xx2 - xx1) : (T)0;                                      \
        const T h = (yy2 - yy1) > (T)0 ? (yy2 - yy1) : (T)0;                                      \
        const T inter = w * h;                                                                    \
        const T ovr = inter / (iarea + areas[j] - inter);                                         \
        const int over = thr_is_float ? (ovr > (T)thr_f) : ((double)ovr > iou_threshold);         \
        if (over) suppressed[j] = 1;                                                              \
      }                                                                                           \
    }                                                                                             \
    free(areas); free(order); free(tmp); free(suppressed);                                        \
    return num;                                                                                   \
  }

DEFINE_NMS(float, f32)
DEFINE_NMS(double, f64)

/* ------------------------------------------------------------------------
 * batched_nms, coordinate trick.  ops/boxes.py:78-95:
 *   max_coordinate = boxes.max(); offsets = idxs.to(boxes) * (max_coordinate + 1)
 *   boxes_for_nms = boxes + offsets[:, None]; nms(boxes_for_nms, scores, thr)
 * all in the boxes' dtype (fp32 here).
 * --------------------------------------------------------------------- */
API int64_t oracle_batched_nms_trick_f32(const float* boxes, const float* scores,
                                         const int64_t* idxs, int64_t n, double thr,
                                         int thr_is_float, int64_t* keep) {
  if (n == 0) return 0;
  float mx = boxes[0];
  /* Tensor.max() propagates NaN: one NaN coordinate makes every offset NaN */
  for (int64_t i = 1; i < 4 * n; i++) {
    if (boxes[i] != boxes[i] || mx != mx) mx = NAN;
    else mx = boxes[i] > mx ? boxes[i] : mx;
  }
  const float m1 = mx + 1.0f;
  float* shifted = (float*)malloc((size_t)n * 4 * sizeof(float));
  for (int64_t i = 0; i < n; i++) {
    const float off = (float)idxs[i] * m1;
    for (int k = 0; k < 4; k++) shifted[4 * i + k] = boxes[4 * i + k] + off;
  }
  const int64_t num = oracle_nms_f32(shifted, scores, n, thr, thr_is_float, keep);
  free(shifted);
  return num;
}

/* batched_nms, per-class loop.  ops/boxes.py:98-112: nms per class id on the
 * unshifted boxes, union of the kept indices, then ordered by descending
 * score (the reference's final sort is NOT stable; with distinct scores the
 * order is unique, with ties we break by ascending index). */
API int64_t oracle_batched_nms_vanilla_f32(const float* boxes, const float* scores,
                                           const int64_t* idxs, int64_t n, double thr,
                                           int thr_is_float, int64_t* keep) {
  if (n == 0) return 0;
  unsigned char* done = (unsigned char*)calloc((size_t)n, 1);
  unsigned char* kept = (unsigned char*)calloc((size_t)n, 1);
  float* cb = (float*)malloc((size_t)n * 4 * sizeof(float));
  float* cs = (float*)malloc((size_t)n * sizeof(float));
  int64_t* cidx = (int64_t*)malloc((size_t)n * sizeof(int64_t));
  int64_t* ck = (int64_t*)malloc((size_t)n * sizeof(int64_t));
  for (int64_t s = 0; s < n; s++) {
    if (done[s]) continue;
    const int64_t cls = idxs[s];
    int64_t m = 0;
    for (int64_t i = s; i < n; i++)
      if (idxs[i] == cls) {
        done[i] = 1;
        memcpy(cb + 4 * m, boxes + 4 * i, 4 * sizeof(float));
        cs[m] = scores[i];
        cidx[m++] = i;
      }
    const int64_t nk = oracle_nms_f32(cb, cs, m, thr, thr_is_float, ck);
    for (int64_t i = 0; i < nk; i++) kept[cidx[ck[i]]] = 1;
  }
  int64_t num = 0;
  for (int64_t i = 0; i < n; i++) num += kept[i];
  sort_item* it = (sort_item*)malloc((size_t)(num ? num : 1) * sizeof(sort_item));
  sort_item* tmp = (sort_item*)malloc((size_t)(num ? num : 1) * sizeof(sort_item));
  int64_t k = 0;
  for (int64_t i = 0; i < n; i++)
    if (kept[i]) { it[k].key = (double)scores[i]; it[k].idx = i; k++; }
  stable_sort_desc(it, tmp, num);
  for (int64_t i = 0; i < num; i++) keep[i] = it[i].idx;
  free(done); free(kept); free(cb); free(cs); free(cidx); free(ck); free(it); free(tmp);
  return num;
}

/* ------------------------------------------------------------------------
 * FPN level mapper.  ops/poolers.py:73-84 with box_area of ops/boxes.py:278-295:
 *   s = sqrt((x2-x1)*(y2-y1));
 *   lvl = floor(lvl0 + log2(s / s0) + eps); clamp to [k_min,k_max]; minus k_min
 * evaluated in fp32 (torch promotes the python ints to the tensor dtype).
 * rois: [K,5] rows (batch, x1, y1, x2, y2).
 * --------------------------------------------------------------------- */
API void oracle_fpn_level_map_f32(const float* rois, int64_t k, int k_min, int k_max, int s0,
                                  int lvl0, double eps, int64_t* levels) {
  const float epsf = (float)eps;
  for (int64_t i = 0; i < k; i++) {
    const float* r = rois + 5 * i;
    const float area = (r[3] - r[1]) * (r[4] - r[2]);
    const float s = sqrtf(area);
    float t = floorf(((float)lvl0 + log2f(s / (float)s0)) + epsf);
    if (t < (float)k_min) t = (float)k_min;
    if (t > (float)k_max) t = (float)k_max;
    levels[i] = (int64_t)t - k_min;
  }
}

/* ------------------------------------------------------------------------
 * MultiScaleRoIAlign.  ops/poolers.py:146-227: map every RoI to a level, run
 * roi_align (aligned=False unless told otherwise, :204-210) on that level's
 * feature map and place the result at the RoI's original position.
 * feats[l]: NCHW contiguous [N,C,H_l,W_l]; output [K,C,PH,PW].
 * --------------------------------------------------------------------- */
API void oracle_msroi_align_forward_f32(const float* const* feats, const int* heights,
                                        const int* widths, const double* scales, int num_levels,
                                        const float* rois, const int64_t* levels, float* output,
                                        int n_rois, int channels, int pooled_height,
                                        int pooled_width, int sampling_ratio, int aligned) {
  const long per = (long)channels * pooled_height * pooled_width;
  for (int n = 0; n < n_rois; n++) {
    const int l = num_levels > 1 ? (int)levels[n] : 0;
    oracle_roi_align_forward_f32(feats[l], rois + 5L * n, output + per * n, 1, channels, heights[l],
                                 widths[l], pooled_height, pooled_width, scales[l], sampling_ratio,
                                 aligned);
  }
}

API void oracle_msroi_align_backward_f32(const float* grad, long gs_n, long gs_c, long gs_h,
                                         long gs_w, const float* rois, const int64_t* levels,
                                         float* const* grad_feats, const int* heights,
                                         const int* widths, const double* scales, int num_levels,
                                         int n_rois, int channels, int pooled_height,
                                         int pooled_width, int sampling_ratio, int aligned) {
  for (int n = 0; n < n_rois; n++) {
    const int l = num_levels > 1 ? (int)levels[n] : 0;
    oracle_roi_align_backward_f32(grad + n * gs_n, gs_n, gs_c, gs_h, gs_w, rois + 5L * n,
                                  grad_feats[l], 1, channels, heights[l], widths[l], pooled_height,
                                  pooled_width, scales[l], sampling_ratio, aligned);
  }
}

/* ---- roi_pool ---------------------------------------------------
 * ops/../csrc/ops/cpu/roi_pool_kernel.cpp:16-83 (forward) and :85-125 (backward), NCHW,
 * same loops, same float -> int conversions. */
#include <float.h>

#define ORACLE_ROI_POOL(T, SFX)                                                                       \
  API void oracle_roi_pool_forward_##SFX(const T* input, const T* rois, T* output, int* argmax,      \
                                         int num_rois, int channels, int height, int width, int ph_n, \
                                         int pw_n, double scale_d) {                                  \
    const T spatial_scale = (T)scale_d;                                                               \
    for (int n = 0; n < num_rois; ++n) {                                                              \
      const T* r = rois + n * 5;                                                                      \
      int roi_batch_ind = (int)r[0];                                                                  \
      int roi_start_w = (int)round(r[1] * spatial_scale);                                             \
      int roi_start_h = (int)round(r[2] * spatial_scale);                                             \
      int roi_end_w = (int)round(r[3] * spatial_scale);                                               \
      int roi_end_h = (int)round(r[4] * spatial_scale);                                               \
      int roi_width = roi_end_w - roi_start_w + 1;                                                    \
      int roi_height = roi_end_h - roi_start_h + 1;                                                   \
      if (roi_width < 1) roi_width = 1;                                                               \
      if (roi_height < 1) roi_height = 1;                                                             \
      T bin_size_h = (T)roi_height / (T)ph_n;                                                         \
      T bin_size_w = (T)roi_width / (T)pw_n;                                                          \
      for (int ph = 0; ph < ph_n; ++ph)                                                               \
        for (int pw = 0; pw < pw_n; ++pw) {                                                           \
          int hstart = (int)floor((T)ph * bin_size_h);                                                \
          int wstart = (int)floor((T)pw * bin_size_w);                                                \
          int hend = (int)ceil((T)(ph + 1) * bin_size_h);                                             \
          int wend = (int)ceil((T)(pw + 1) * bin_size_w);                                             \
          hstart = hstart + roi_start_h; hstart = hstart < 0 ? 0 : (hstart > height ? height : hstart); \
          hend = hend + roi_start_h; hend = hend < 0 ? 0 : (hend > height ? height : hend);           \
          wstart = wstart + roi_start_w; wstart = wstart < 0 ? 0 : (wstart > width ? width : wstart); \
          wend = wend + roi_start_w; wend = wend < 0 ? 0 : (wend > width ? width : wend);             \
          int is_empty = (hend <= hstart) || (wend <= wstart);                                        \
          for (int c = 0; c < channels; ++c) {                                                        \
            T maxval = is_empty ? 0 : -FLT_MAX;                                                       \
            int maxidx = -1;                                                                          \
            const T* in = input + ((long)roi_batch_ind * channels + c) * height * width;              \
            for (int h = hstart; h < hend; ++h)                                                       \
              for (int w = wstart; w < wend; ++w) {                                                   \
                int idx = h * width + w;                                                              \
                if (in[idx] > maxval) { maxval = in[idx]; maxidx = idx; }                             \
              }                                                                                       \
            long o = (((long)n * channels + c) * ph_n + ph) * pw_n + pw;                              \
            output[o] = maxval;                                                                       \
            argmax[o] = maxidx;                                                                       \
          }                                                                                           \
        }                                                                                             \
    }                                                                                                 \
  }                                                                                                   \
  API void oracle_roi_pool_backward_##SFX(const T* grad, const int* argmax, const T* rois,            \
                                          T* grad_input, int num_rois, int channels, int height,     \
                                          int width, int ph_n, int pw_n, long n_stride,              \
                                          long c_stride, long h_stride, long w_stride) {             \
    /* grad_input must be zero on entry (at::zeros, roi_pool_kernel.cpp:207) */                      \
    for (int n = 0; n < num_rois; ++n) {                                                              \
      int roi_batch_ind = (int)rois[n * 5];                                                           \
      for (int c = 0; c < channels; ++c) {                                                            \
        T* gi = grad_input + ((long)roi_batch_ind * channels + c) * height * width;                   \
        const int* am = argmax + ((long)n * channels + c) * ph_n * pw_n;                              \
        for (int ph = 0; ph < ph_n; ++ph)                                                             \
          for (int pw = 0; pw < pw_n; ++pw) {                                                         \
            int a = am[ph * pw_n + pw];                                                               \
            if (a != -1) gi[a] += grad[n * n_stride + c * c_stride + ph * h_stride + pw * w_stride];  \
          }                                                                                           \
      }                                                                                               \
    }                                                                                                 \
  }

ORACLE_ROI_POOL(float, f32)
ORACLE_ROI_POOL(double, f64)

/* ---- ps_roi_pool -------------------------------------------------
 * csrc/ops/cpu/ps_roi_pool_kernel.cpp:14-83 (forward: bins clipped to [0, size - 1]) and :85-146
 * (backward: roundf, bins clipped to [0, size]), NCHW, same loops. */
#define ORACLE_PS_ROI_POOL(T, SFX)                                                                     \
  API void oracle_ps_roi_pool_forward_##SFX(const T* input, const T* rois, T* output, int* mapping,    \
                                            int num_rois, int channels, int height, int width,         \
                                            int ph_n, int pw_n, double scale_d) {                      \
    const T spatial_scale = (T)scale_d;                                                                \
    const int channels_out = channels / (ph_n * pw_n);                                                 \
    for (int n = 0; n < num_rois; ++n) {                                                               \
      const T* r = rois + n * 5;                                                                       \
      int b = (int)r[0];                                                                               \
      int sw = (int)round(r[1] * spatial_scale), sh = (int)round(r[2] * spatial_scale);                \
      int ew = (int)round(r[3] * spatial_scale), eh = (int)round(r[4] * spatial_scale);                \
      int rw = ew - sw; if (rw < 1) rw = 1;                                                            \
      int rh = eh - sh; if (rh < 1) rh = 1;                                                            \
      T bin_h = (T)rh / (T)ph_n, bin_w = (T)rw / (T)pw_n;                                              \
      int c_in = 0;                                                                                    \
      for (int c_out = 0; c_out < channels_out; ++c_out)                                               \
        for (int ph = 0; ph < ph_n; ++ph)                                                              \
          for (int pw = 0; pw < pw_n; ++pw) {                                                          \
            int hs = (int)floor((T)ph * bin_h) + sh, ws = (int)floor((T)pw * bin_w) + sw;              \
            int he = (int)ceil((T)(ph + 1) * bin_h) + sh, we = (int)ceil((T)(pw + 1) * bin_w) + sw;    \
            hs = hs < 0 ? 0 : (hs > height - 1 ? height - 1 : hs);                                     \
            he = he < 0 ? 0 : (he > height - 1 ? height - 1 : he);                                     \
            ws = ws < 0 ? 0 : (ws > width - 1 ? width - 1 : ws);                                       \
            we = we < 0 ? 0 : (we > width - 1 ? width - 1 : we);                                       \
            int is_empty = (he <= hs) || (we <= ws);                                                   \
            const T* in = input + ((long)b * channels + c_in) * height * width;                        \
            T sum = 0;                                                                                 \
            for (int h = hs; h < he; ++h)                                                              \
              for (int w = ws; w < we; ++w) sum += in[h * width + w];                                  \
            long o = (((long)n * channels_out + c_out) * ph_n + ph) * pw_n + pw;                       \
            T area = (T)((he - hs) * (we - ws));                                                       \
            output[o] = is_empty ? (T)0 : sum / area;                                                  \
            mapping[o] = c_in;                                                                         \
            c_in++;                                                                                    \
          }                                                                                            \
    }                                                                                                  \
  }                                                                                                    \
  API void oracle_ps_roi_pool_backward_##SFX(const T* grad, const int* mapping, const T* rois,         \
                                             T* grad_input, int num_rois, int channels, int height,    \
                                             int width, int ph_n, int pw_n, double scale_d) {          \
    const T spatial_scale = (T)scale_d;                                                                \
    const int channels_out = channels / (ph_n * pw_n);                                                 \
    for (int n = 0; n < num_rois; ++n) {                                                               \
      const T* r = rois + n * 5;                                                                       \
      int b = (int)r[0];                                                                               \
      int sw = (int)roundf(r[1] * spatial_scale), sh = (int)roundf(r[2] * spatial_scale);              \
      int ew = (int)roundf(r[3] * spatial_scale), eh = (int)roundf(r[4] * spatial_scale);              \
      int rw = ew - sw; if (rw < 1) rw = 1;                                                            \
      int rh = eh - sh; if (rh < 1) rh = 1;                                                            \
      T bin_h = (T)rh / (T)ph_n, bin_w = (T)rw / (T)pw_n;                                              \
      for (int ph = 0; ph < ph_n; ++ph)                                                                \
        for (int pw = 0; pw < pw_n; ++pw) {                                                            \
          int hs = (int)floor((T)ph * bin_h) + sh, ws = (int)floor((T)pw * bin_w) + sw;                \
          int he = (int)ceil((T)(ph + 1) * bin_h) + sh, we = (int)ceil((T)(pw + 1) * bin_w) + sw;      \
          hs = hs < 0 ? 0 : (hs > height ? height : hs);                                               \
          he = he < 0 ? 0 : (he > height ? height : he);                                               \
          ws = ws < 0 ? 0 : (ws > width ? width : ws);                                                 \
          we = we < 0 ? 0 : (we > width ? width : we);                                                 \
          int is_empty = (he <= hs) || (we <= ws);                                                     \
          for (int c_out = 0; c_out < channels_out; ++c_out) {                                         \
            long o = (((long)n * channels_out + c_out) * ph_n + ph) * pw_n + pw;                       \
            int c_in = mapping[o];                                                                     \
            T* gi = grad_input + ((long)b * channels + c_in) * height * width;                         \
            T area = (T)((he - hs) * (we - ws));                                                       \
            T diff = is_empty ? (T)0 : grad[o] / area;                                                 \
            for (int h = hs; h < he; ++h)                                                              \
              for (int w = ws; w < we; ++w) gi[h * width + w] += diff;                                 \
          }                                                                                            \
        }                                                                                              \
    }                                                                                                  \
  }

ORACLE_PS_ROI_POOL(float, f32)
ORACLE_PS_ROI_POOL(double, f64)

/* ---- ps_roi_align ------------------------------------------------
 * csrc/ops/cpu/ps_roi_align_kernel.cpp:9-60 (bilinear_interpolate), :62-140 (forward),
 * :142-300 (gradient weights + backward), NCHW, same loops and operation order. */
#define ORACLE_PS_ROI_ALIGN(T, SFX)                                                                    \
  static T ps_bilinear_##SFX(const T* in, int height, int width, T y, T x) {                           \
    if (y < -1.0 || y > height || x < -1.0 || x > width) return 0;                                     \
    if (y <= 0) y = 0;                                                                                 \
    if (x <= 0) x = 0;                                                                                 \
    int y_low = (int)y, x_low = (int)x, y_high, x_high;                                                \
    if (y_low >= height - 1) { y_high = y_low = height - 1; y = (T)y_low; } else y_high = y_low + 1;   \
    if (x_low >= width - 1) { x_high = x_low = width - 1; x = (T)x_low; } else x_high = x_low + 1;     \
    T ly = y - y_low, lx = x - x_low, hy = 1. - ly, hx = 1. - lx;                                      \
    T v1 = in[y_low * width + x_low], v2 = in[y_low * width + x_high];                                 \
    T v3 = in[y_high * width + x_low], v4 = in[y_high * width + x_high];                               \
    T w1 = hy * hx, w2 = hy * lx, w3 = ly * hx, w4 = ly * lx;                                          \
    return (w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4);                                                    \
  }                                                                                                    \
  API void oracle_ps_roi_align_forward_##SFX(const T* input, const T* rois, T* output, int* mapping,   \
                                             int num_rois, int channels, int height, int width,        \
                                             int ph_n, int pw_n, double scale_d, int sampling_ratio) { \
    const T spatial_scale = (T)scale_d;                                                                \
    const int channels_out = channels / (ph_n * pw_n);                                                 \
    for (int n = 0; n < num_rois; ++n) {                                                               \
      const T* r = rois + n * 5;                                                                       \
      int b = (int)r[0];                                                                               \
      T sw = r[1] * spatial_scale - (T)0.5, sh = r[2] * spatial_scale - (T)0.5;                        \
      T ew = r[3] * spatial_scale - (T)0.5, eh = r[4] * spatial_scale - (T)0.5;                        \
      T roi_w = ew - sw, roi_h = eh - sh;                                                              \
      T bin_h = roi_h / (T)ph_n, bin_w = roi_w / (T)pw_n;                                              \
      int c_in = 0;                                                                                    \
      for (int c_out = 0; c_out < channels_out; ++c_out)                                               \
        for (int ph = 0; ph < ph_n; ++ph)                                                              \
          for (int pw = 0; pw < pw_n; ++pw) {                                                          \
            long o = (((long)n * channels_out + c_out) * ph_n + ph) * pw_n + pw;                       \
            T hstart = (T)ph * bin_h + sh, wstart = (T)pw * bin_w + sw;                                \
            int gh = sampling_ratio > 0 ? sampling_ratio : (int)ceil(roi_h / ph_n);                    \
            int gw = sampling_ratio > 0 ? sampling_ratio : (int)ceil(roi_w / pw_n);                    \
            const T count = gh * gw;                                                                   \
            const T* in = input + ((long)b * channels + c_in) * height * width;                        \
            T sum = 0;                                                                                 \
            for (int iy = 0; iy < gh; iy++) {                                                          \
              const T y = hstart + (T)(iy + .5f) * bin_h / (T)gh;                                      \
              for (int ix = 0; ix < gw; ix++) {                                                        \
                const T x = wstart + (T)(ix + .5f) * bin_w / (T)gw;                                    \
                sum += ps_bilinear_##SFX(in, height, width, y, x);                                     \
              }                                                                                        \
            }                                                                                          \
            sum /= count;                                                                              \
            output[o] = sum;                                                                           \
            mapping[o] = c_in;                                                                         \
            c_in++;                                                                                    \
          }                                                                                            \
    }                                                                                                  \
  }                                                                                                    \
  API void oracle_ps_roi_align_backward_##SFX(const T* grad, const int* mapping, const T* rois,        \
                                              T* grad_input, int num_rois, int channels, int height,   \
                                              int width, int ph_n, int pw_n, double scale_d,           \
                                              int sampling_ratio) {                                    \
    const T spatial_scale = (T)scale_d;                                                                \
    const int channels_out = channels / (ph_n * pw_n);                                                 \
    const long nthreads = (long)num_rois * channels_out * ph_n * pw_n;                                 \
    for (long index = 0; index < nthreads; index++) {                                                  \
      int pw = index % pw_n, ph = (index / pw_n) % ph_n;                                               \
      long n = index / pw_n / ph_n / channels_out;                                                     \
      const T* r = rois + n * 5;                                                                       \
      int b = (int)r[0];                                                                               \
      T sw = r[1] * spatial_scale - (T)0.5, sh = r[2] * spatial_scale - (T)0.5;                        \
      T ew = r[3] * spatial_scale - (T)0.5, eh = r[4] * spatial_scale - (T)0.5;                        \
      T roi_w = ew - sw, roi_h = eh - sh;                                                              \
      T bin_h = roi_h / (T)ph_n, bin_w = roi_w / (T)pw_n;                                              \
      int c_in = mapping[index];                                                                       \
      T* gi = grad_input + ((long)b * channels + c_in) * height * width;                               \
      T hstart = (T)ph * bin_h + sh, wstart = (T)pw * bin_w + sw;                                      \
      const T go = grad[index];                                                                        \
      int gh = sampling_ratio > 0 ? sampling_ratio : (int)ceil(roi_h / ph_n);                          \
      int gw = sampling_ratio > 0 ? sampling_ratio : (int)ceil(roi_w / pw_n);                          \
      const T count = gh * gw;                                                                         \
      for (int iy = 0; iy < gh; iy++) {                                                                \
        T y = hstart + (T)(iy + .5f) * bin_h / (T)gh;                                                  \
        for (int ix = 0; ix < gw; ix++) {                                                              \
          T x = wstart + (T)(ix + .5f) * bin_w / (T)gw;                                                \
          T yy = y, xx = x;                                                                            \
          if (yy < -1.0 || yy > height || xx < -1.0 || xx > width) continue;                           \
          if (yy <= 0) yy = 0;                                                                         \
          if (xx <= 0) xx = 0;                                                                         \
          int y_low = (int)yy, x_low = (int)xx, y_high, x_high;                                        \
          if (y_low >= height - 1) { y_high = y_low = height - 1; yy = (T)y_low; } else y_high = y_low + 1; \
          if (x_low >= width - 1) { x_high = x_low = width - 1; xx = (T)x_low; } else x_high = x_low + 1;   \
          T ly = yy - y_low, lx = xx - x_low, hy = 1. - ly, hx = 1. - lx;                              \
          T w1 = hy * hx, w2 = hy * lx, w3 = ly * hx, w4 = ly * lx;                                    \
          gi[y_low * width + x_low] += go * w1 / count;                                                \
          gi[y_low * width + x_high] += go * w2 / count;                                               \
          gi[y_high * width + x_low] += go * w3 / count;                                               \
          gi[y_high * width + x_high] += go * w4 / count;                                              \
        }                                                                                              \
      }                                                                                                \
    }                                                                                                  \
  }

ORACLE_PS_ROI_ALIGN(float, f32)
ORACLE_PS_ROI_ALIGN(double, f64)
