// ref_wrap_roi_pool.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The reference's OWN cpu/roi_pool_kernel.cpp, compiled unmodified from where it lies under
// /root/reference into oracle/_ref/libtvref.so (a translation unit of its own: its anonymous-namespace
// helpers have the same names as roi_align's).  See ref_wrap.cpp for the registration redirect.
#include <ATen/ATen.h>
#include <torch/library.h>

#include <cstdint>

#undef TORCH_LIBRARY_IMPL
#define TVREF_CAT2(a, b) a##b
#define TVREF_CAT(a, b) TVREF_CAT2(a, b)
#define TORCH_LIBRARY_IMPL(ns, k, m) \
  [[maybe_unused]] static void TVREF_CAT(tvref_unused_registration_, __LINE__)(torch::Library & m)

#include "torchvision/csrc/ops/cpu/roi_pool_kernel.cpp"

namespace tvref = vision::ops;
#define API extern "C" __attribute__((visibility("default")))

API void tvref_roi_pool_forward_f32(const float* input, const float* rois, float* output, int* argmax, int num_rois,
                                    int channels, int height, int width, int ph, int pw, double scale) {
  tvref::roi_pool_forward_kernel_impl<float>(input, (float)scale, channels, height, width, ph, pw, rois, num_rois,
                                             output, argmax);
}
API void tvref_roi_pool_forward_f64(const double* input, const double* rois, double* output, int* argmax,
                                    int num_rois, int channels, int height, int width, int ph, int pw,
                                    double scale) {
  tvref::roi_pool_forward_kernel_impl<double>(input, scale, channels, height, width, ph, pw, rois, num_rois, output,
                                              argmax);
}
API void tvref_roi_pool_backward_f32(const float* grad, const int* argmax, const float* rois, float* grad_input,
                                     int num_rois, int channels, int height, int width, int ph, int pw,
                                     int64_t ns, int64_t cs, int64_t hs, int64_t ws) {
  tvref::roi_pool_backward_kernel_impl<float>(grad, argmax, num_rois, channels, height, width, ph, pw, grad_input,
                                              rois, (int)ns, (int)cs, (int)hs, (int)ws);
}
API void tvref_roi_pool_backward_f64(const double* grad, const int* argmax, const double* rois, double* grad_input,
                                     int num_rois, int channels, int height, int width, int ph, int pw,
                                     int64_t ns, int64_t cs, int64_t hs, int64_t ws) {
  tvref::roi_pool_backward_kernel_impl<double>(grad, argmax, num_rois, channels, height, width, ph, pw, grad_input,
                                               rois, (int)ns, (int)cs, (int)hs, (int)ws);
}
