// ref_wrap.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Compiles the reference's OWN CPU kernels, unmodified and from where they lie
// under /root/reference (never copied into this repo), into
// oracle/_ref/libtvref.so with a plain C ABI:
//   torchvision/csrc/ops/cpu/roi_align_kernel.cpp  (+ roi_align_common.h)
//   torchvision/csrc/ops/cpu/nms_kernel.cpp
// The two .cpp files are #included into this translation unit (their kernels
// sit in an anonymous namespace) with TORCH_LIBRARY_IMPL redirected to an
// unused static function, so loading the library registers nothing with the
// PyTorch dispatcher and it can live beside any torchvision build.
//
// Used by: tests (to validate oracle/detection_oracle.c bit for bit) and
// bench.py --impl reference / cpu_baseline (the reference's CPU path timed on
// the box's host cores).  The *_mt variants only shard independent work
// (RoIs for the forward, images for the backward) over std::threads; every
// shard runs the reference's single-threaded kernel as is.
//
// Build recipe: oracle/build.py (g++ -O3, no -march, like the reference's
// setup.py:118-131 which passes no architecture flags).
#include <ATen/ATen.h>
#include <ATen/Parallel.h>
#include <torch/library.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#undef TORCH_LIBRARY_IMPL
#define TVREF_CAT2(a, b) a##b
#define TVREF_CAT(a, b) TVREF_CAT2(a, b)
#define TORCH_LIBRARY_IMPL(ns, k, m) \
  [[maybe_unused]] static void TVREF_CAT(tvref_unused_registration_, __LINE__)(torch::Library & m)

#include "torchvision/csrc/ops/cpu/roi_align_kernel.cpp"
#include "torchvision/csrc/ops/cpu/nms_kernel.cpp"

namespace tvref = vision::ops;

#define API extern "C" __attribute__((visibility("default")))

template <typename T>
static void fwd(const T* input, const T* rois, T* output, int n_rois, int channels, int height,
                int width, int ph, int pw, double scale, int sr, int aligned) {
  if ((int64_t)n_rois * channels * ph * pw == 0) return;
  // cpu/roi_align_kernel.cpp:282-328 calls the impl exactly like this after
  // zero-filling the output
  std::memset(output, 0, sizeof(T) * (size_t)n_rois * channels * ph * pw);
  tvref::roi_align_forward_kernel_impl<T>(n_rois, input, scale, channels, height, width, ph, pw, sr,
                                          aligned != 0, rois, output);
}

template <typename T>
static void bwd(const T* grad, int64_t gs_n, int64_t gs_c, int64_t gs_h, int64_t gs_w, const T* rois,
                T* grad_input, int n_rois, int channels, int height, int width, int ph, int pw,
                double scale, int sr, int aligned) {
  // grad_input must already be zero (cpu/roi_align_kernel.cpp:351 at::zeros)
  tvref::roi_align_backward_kernel_impl<T>(n_rois * channels * ph * pw, grad, scale, channels, height,
                                           width, ph, pw, sr, aligned != 0, grad_input, rois,
                                           (int)gs_n, (int)gs_c, (int)gs_h, (int)gs_w);
}

API void tvref_roi_align_forward_f32(const float* input, const float* rois, float* output,
                                     int n_rois, int channels, int height, int width, int ph, int pw,
                                     double scale, int sr, int aligned) {
  fwd<float>(input, rois, output, n_rois, channels, height, width, ph, pw, scale, sr, aligned);
}
API void tvref_roi_align_forward_f64(const double* input, const double* rois, double* output,
                                     int n_rois, int channels, int height, int width, int ph, int pw,
                                     double scale, int sr, int aligned) {
  fwd<double>(input, rois, output, n_rois, channels, height, width, ph, pw, scale, sr, aligned);
}
API void tvref_roi_align_backward_f32(const float* grad, int64_t gs_n, int64_t gs_c, int64_t gs_h,
                                      int64_t gs_w, const float* rois, float* grad_input, int n_rois,
                                      int channels, int height, int width, int ph, int pw,
                                      double scale, int sr, int aligned) {
  bwd<float>(grad, gs_n, gs_c, gs_h, gs_w, rois, grad_input, n_rois, channels, height, width, ph, pw,
             scale, sr, aligned);
}
API void tvref_roi_align_backward_f64(const double* grad, int64_t gs_n, int64_t gs_c, int64_t gs_h,
                                      int64_t gs_w, const double* rois, double* grad_input,
                                      int n_rois, int channels, int height, int width, int ph, int pw,
                                      double scale, int sr, int aligned) {
  bwd<double>(grad, gs_n, gs_c, gs_h, gs_w, rois, grad_input, n_rois, channels, height, width, ph, pw,
              scale, sr, aligned);
}

// forward over `threads` host threads: RoIs are independent
API void tvref_roi_align_forward_f32_mt(const float* input, const float* rois, float* output,
                                        int n_rois, int channels, int height, int width, int ph,
                                        int pw, double scale, int sr, int aligned, int threads) {
  threads = std::max(1, std::min(threads, n_rois));
  std::vector<std::thread> pool;
  const int64_t per_roi = (int64_t)channels * ph * pw;
  for (int t = 0; t < threads; ++t) {
    const int lo = (int)((int64_t)n_rois * t / threads), hi = (int)((int64_t)n_rois * (t + 1) / threads);
    pool.emplace_back([=] {
      fwd<float>(input, rois + 5LL * lo, output + per_roi * lo, hi - lo, channels, height, width, ph,
                 pw, scale, sr, aligned);
    });
  }
  for (auto& th : pool) th.join();
}

// backward over host threads: images are independent (grad_input of image b is
// only touched by RoIs with batch index b); requires rois grouped by image.
// Returns the number of threads actually used, 0 if the rois were not grouped.
API int tvref_roi_align_backward_f32_mt(const float* grad, int64_t gs_n, int64_t gs_c, int64_t gs_h,
                                        int64_t gs_w, const float* rois, float* grad_input,
                                        int n_rois, int channels, int height, int width, int ph,
                                        int pw, double scale, int sr, int aligned, int threads) {
  std::vector<int> starts;  // first roi of every image run
  for (int i = 0; i < n_rois; ++i)
    if (i == 0 || (int)rois[5LL * i] != (int)rois[5LL * (i - 1)]) starts.push_back(i);
  starts.push_back(n_rois);
  std::vector<char> seen(1 << 20, 0);
  for (size_t r = 0; r + 1 < starts.size(); ++r) {
    const int b = (int)rois[5LL * starts[r]];
    if (b < 0 || b >= (int)seen.size() || seen[b]) return 0;
    seen[b] = 1;
  }
  const int runs = (int)starts.size() - 1;
  threads = std::max(1, std::min(threads, runs));
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    const int rlo = (int)((int64_t)runs * t / threads), rhi = (int)((int64_t)runs * (t + 1) / threads);
    const int lo = starts[rlo], hi = starts[rhi];
    if (hi <= lo) continue;
    pool.emplace_back([=] {
      bwd<float>(grad + gs_n * lo, gs_n, gs_c, gs_h, gs_w, rois + 5LL * lo, grad_input, hi - lo,
                 channels, height, width, ph, pw, scale, sr, aligned);
    });
  }
  for (auto& th : pool) th.join();
  return threads;
}

template <typename T>
static int64_t nms_t(const T* dets, const T* scores, int64_t n, double thr, int64_t* keep) {
  auto opt = at::TensorOptions().dtype(c10::CppTypeToScalarType<T>::value);
  at::Tensor d = at::from_blob(const_cast<T*>(dets), {n, 4}, opt);
  at::Tensor s = at::from_blob(const_cast<T*>(scores), {n}, opt);
  at::Tensor k = tvref::nms_kernel(d, s, thr);  // cpu/nms_kernel.cpp:79-108
  std::memcpy(keep, k.data_ptr<int64_t>(), sizeof(int64_t) * (size_t)k.numel());
  return k.numel();
}
API int64_t tvref_nms_f32(const float* dets, const float* scores, int64_t n, double thr, int64_t* keep) {
  return nms_t<float>(dets, scores, n, thr, keep);
}
API int64_t tvref_nms_f64(const double* dets, const double* scores, int64_t n, double thr, int64_t* keep) {
  return nms_t<double>(dets, scores, n, thr, keep);
}

API void tvref_set_num_threads(int n) { at::set_num_threads(n); }
