// ref_wrap_ps_roi_align.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// The reference's OWN cpu/ps_roi_align_kernel.cpp, compiled unmodified from where it lies under /root/reference
// into oracle/_ref/libtvref.so (own translation unit; see ref_wrap.cpp for the registration redirect).
#include <ATen/ATen.h>
#include <torch/library.h>

#undef TORCH_LIBRARY_IMPL
#define TVREF_CAT2(a, b) a##b
#define TVREF_CAT(a, b) TVREF_CAT2(a, b)
#define TORCH_LIBRARY_IMPL(ns, k, m) \
  [[maybe_unused]] static void TVREF_CAT(tvref_unused_registration_, __LINE__)(torch::Library & m)

#include "torchvision/csrc/ops/cpu/ps_roi_align_kernel.cpp"

namespace tvref = vision::ops;
#define API extern "C" __attribute__((visibility("default")))

#define WRAP(T, SFX)                                                                                              \
  API void tvref_ps_roi_align_forward_##SFX(const T* input, const T* rois, T* output, int* mapping, int num_rois, \
                                            int channels, int height, int width, int ph, int pw, double scale,    \
                                            int sr) {                                                             \
    tvref::ps_roi_align_forward_kernel_impl<T>(num_rois, input, (T)scale, channels, height, width, ph, pw, sr,    \
                                               rois, channels / (ph * pw), output, mapping);                      \
  }                                                                                                               \
  API void tvref_ps_roi_align_backward_##SFX(const T* grad, const int* mapping, const T* rois, T* grad_input,     \
                                             int num_rois, int channels, int height, int width, int ph, int pw,   \
                                             double scale, int sr) {                                              \
    const int co = channels / (ph * pw);                                                                          \
    tvref::ps_roi_align_backward_kernel_impl<T>(num_rois * co * ph * pw, grad, mapping, (T)scale, channels,       \
                                                height, width, ph, pw, sr, co, grad_input, rois);                 \
  }
WRAP(float, f32)
WRAP(double, f64)
