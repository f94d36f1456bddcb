// torch_shim.cpp -- thin PyTorch dispatcher shim over the C-ABI library.
//
// Registers the sm_100a kernels under the CUDA dispatch key of the reference's
// operators, mirroring
//   TORCH_LIBRARY_IMPL(torchvision, CUDA, m) in
//   torchvision/csrc/ops/cuda/roi_align_kernel.cu:470-477 and
//   torchvision/csrc/ops/cuda/nms_kernel.cu:169-171,
// with the argument checks of roi_align_kernel.cu:342-350,:408-415 and
// nms_kernel.cu:84-104.  The schemas (roi_align.cpp:124-129, nms.cpp:21-25)
// are NOT redefined when a torchvision build (the reference's CPU-only _C)
// already defined them; when none is loaded the shim defines them itself so
// the package also works stand-alone.  Extra fused entry points live in the
// `b200det` namespace.
//
// No kernel code here: every op forwards raw pointers to libb200det.so.
#include <ATen/ATen.h>
#include <ATen/autocast_mode.h>
#include <ATen/core/dispatch/Dispatcher.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <memory>
#include <vector>

#include "../../include/b200_detection_ops.h"

namespace {

int dtype_code(at::ScalarType t, const char* who) {
  switch (t) {
    case at::kFloat: return B200DET_F32;
    case at::kHalf: return B200DET_F16;
    case at::kBFloat16: return B200DET_BF16;
    case at::kDouble: return B200DET_F64;
    default: TORCH_CHECK(false, "\"", who, "\" not implemented for '", c10::toString(t), "'");
  }
  return -1;
}

void check_rc(int rc, const char* who) {
  TORCH_CHECK(rc == 0, who, " failed (", rc, "): ", b200det_last_error());
}

// NHWC fast path when the tensor is channels_last-dense and not also NCHW-dense
bool is_nhwc(const at::Tensor& t) {
  return t.dim() == 4 && t.is_contiguous(at::MemoryFormat::ChannelsLast) && !t.is_contiguous();
}

// rois travel in the arithmetic type of the features (include/b200_detection_ops.h)
at::Tensor arith_rois(const at::Tensor& rois, at::ScalarType feat) {
  return rois.to(feat == at::kDouble ? at::kDouble : at::kFloat).contiguous();
}

// The reference insists on rois.dtype == input.dtype (checkAllSameType,
// cuda/roi_align_kernel.cu:350).  We additionally accept fp32 rois with 16-bit
// features: that is what its autocast wrapper computes on.
void check_roi_dtype(const at::Tensor& feat, const at::Tensor& rois, const char* who) {
  const bool ok = rois.scalar_type() == feat.scalar_type() ||
                  (rois.scalar_type() == at::kFloat &&
                   (feat.scalar_type() == at::kHalf || feat.scalar_type() == at::kBFloat16));
  TORCH_CHECK(ok, "Expected tensor for argument #1 'input' to have the same type as tensor for argument #2 'rois'; but type ",
              feat.toString(), " does not equal ", rois.toString(), " (while checking arguments for ", who, ")");
}

void* cur_stream() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }

// ---- NCHW callers --------------------------------------------------------------------------
// The reference's kernels take NCHW (input.contiguous(), cuda/roi_align_kernel.cu:384) and that is
// what its detection models hand over.  The generic strided kernels serve it at the reference's
// parallelisation; the channels_last kernels are ~10x faster (cfg3: 9.4 -> 0.6 ms forward, 11.5 -> 1.0 ms
// backward), and a re-layout pass runs at HBM rate.  So when the shapes qualify for the fast kernels
// and the RoIs are not a handful, an NCHW caller is served by relayout -> fast kernel -> relayout,
// and still gets NCHW-contiguous results.  B200DET_NCHW_RELAYOUT=0 keeps the generic kernels.
bool fast_shape(int dt, int64_t C, int64_t PH, int64_t PW, int64_t sr) {
  if (dt == B200DET_F64 || C <= 0 || PH < 1 || PW < 1) return false;
  if (C % 128 == 0 && (sr == 1 || sr == 2) && PW * sr <= 32 && PH <= 32) return true;  // roi_align_{fwd,bwd}_sm100.cu
  // roi_align_anygrid_sm100.cu: adaptive / larger sampling grids, C a multiple of 4
  const char* e = getenv("B200DET_ANYGRID");
  return (!e || atoi(e) != 0) && C >= 32 && C % 4 == 0 && PH <= 64 && PW <= 64;
}
bool relayout_pays(int64_t K, int64_t PH, int64_t PW, int64_t feature_pixels) {
  const char* e = getenv("B200DET_NCHW_RELAYOUT");  // read per call (the tests compare both paths in one process)
  const bool on = !e || atoi(e) != 0;
  // a generic output element costs about as much as 16 feature elements of re-layout
  return on && K > 0 && K * PH * PW * 16 >= feature_pixels;
}
// dense NCHW -> the same values with channels_last strides
at::Tensor to_nhwc(const at::Tensor& t, int dt, const char* who) {
  const at::Tensor src = t.contiguous();
  at::Tensor dst = at::empty(src.sizes(), src.options().memory_format(at::MemoryFormat::ChannelsLast));
  check_rc(b200det_relayout(src.data_ptr(), dst.data_ptr(), dt, src.size(0), src.size(1), src.size(2) * src.size(3),
                            cur_stream()), who);
  return dst;
}
// channels_last-dense -> NCHW-contiguous
at::Tensor to_nchw(const at::Tensor& t, int dt, const char* who) {
  at::Tensor dst = at::empty(t.sizes(), t.options().memory_format(at::MemoryFormat::Contiguous));
  check_rc(b200det_relayout(t.data_ptr(), dst.data_ptr(), dt, t.size(0), t.size(2) * t.size(3), t.size(1),
                            cur_stream()), who);
  return dst;
}

// ---- torchvision::roi_align ------------------------------------------------
at::Tensor roi_align_forward_cuda(const at::Tensor& input, const at::Tensor& rois,
                                  double spatial_scale, int64_t pooled_height,
                                  int64_t pooled_width, int64_t sampling_ratio, bool aligned) {
  TORCH_CHECK(input.is_cuda(), "input must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(rois.size(1) == 5, "rois must have shape as Tensor[K, 5]");
  at::TensorArg input_t{input, "input", 1}, rois_t{rois, "rois", 2};
  at::CheckedFrom c = "roi_align_forward_kernel";
  at::checkAllSameGPU(c, {input_t, rois_t});
  check_roi_dtype(input, rois, c);
  TORCH_CHECK(input.dim() == 4, "input must be a 4-D tensor [N, C, H, W]");
  at::cuda::CUDAGuard device_guard(input.device());

  const int dt = dtype_code(input.scalar_type(), "roi_align_forward_kernel");
  const auto K = rois.size(0), C = input.size(1), H = input.size(2), W = input.size(3);
  const bool via_nhwc = !is_nhwc(input) && fast_shape(dt, C, pooled_height, pooled_width, sampling_ratio) &&
                        relayout_pays(K, pooled_height, pooled_width, input.size(0) * H * W);
  const bool nhwc = via_nhwc || is_nhwc(input);
  const at::Tensor input_ = via_nhwc ? to_nhwc(input, dt, "roi_align_forward_kernel")
                                     : (nhwc ? input : input.contiguous());
  const at::Tensor rois_ = arith_rois(rois, input.scalar_type());
  at::Tensor output = at::empty(
      {K, C, pooled_height, pooled_width},
      input.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  if (output.numel() == 0) return output;
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_roi_align_forward(input_.data_ptr(), layout, rois_.data_ptr(), output.data_ptr(),
                                     layout, dt, input.size(0), C, H, W, K, pooled_height,
                                     pooled_width, spatial_scale, (int)sampling_ratio, aligned,
                                     cur_stream()),
           "roi_align_forward_kernel");
  return via_nhwc ? to_nchw(output, dt, "roi_align_forward_kernel") : output;
}

at::Tensor roi_align_backward_cuda(const at::Tensor& grad, const at::Tensor& rois,
                                   double spatial_scale, int64_t pooled_height,
                                   int64_t pooled_width, int64_t batch_size, int64_t channels,
                                   int64_t height, int64_t width, int64_t sampling_ratio,
                                   bool aligned) {
  TORCH_CHECK(grad.is_cuda(), "grad must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  at::TensorArg grad_t{grad, "grad", 1}, rois_t{rois, "rois", 2};
  at::CheckedFrom c = "roi_align_backward_kernel";
  at::checkAllSameGPU(c, {grad_t, rois_t});
  check_roi_dtype(grad, rois, c);
  at::cuda::CUDAGuard device_guard(grad.device());

  const int dt = dtype_code(grad.scalar_type(), "roi_align_backward_kernel");
  // grad_input follows the memory format the gradient arrives in: a
  // channels_last forward produced a channels_last output, whose gradient
  // normally comes back channels_last as well.
  const bool via_nhwc = !is_nhwc(grad) && grad.dim() == 4 && grad.numel() > 0 &&
                        fast_shape(dt, channels, pooled_height, pooled_width, sampling_ratio) &&
                        relayout_pays(rois.size(0), pooled_height, pooled_width, batch_size * height * width);
  const bool nhwc = via_nhwc || is_nhwc(grad);
  at::Tensor grad_input = at::empty(
      {batch_size, channels, height, width},
      grad.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  if (grad.numel() == 0) return grad_input.zero_();
  const at::Tensor rois_ = arith_rois(rois, grad.scalar_type());
  const at::Tensor grad_ = via_nhwc ? to_nhwc(grad, dt, "roi_align_backward_kernel") : grad;
  const int64_t gs[4] = {grad_.stride(0), grad_.stride(1), grad_.stride(2), grad_.stride(3)};
  auto done = [&]() { return via_nhwc ? to_nchw(grad_input, dt, "roi_align_backward_kernel") : grad_input; };
  auto run = [&](int flags) {
    return b200det_roi_align_backward(grad_.data_ptr(), gs, rois_.data_ptr(), grad_input.data_ptr(),
                                      nhwc ? B200DET_NHWC : B200DET_NCHW, dt, batch_size, channels, height,
                                      width, rois.size(0), pooled_height, pooled_width, spatial_scale,
                                      (int)sampling_ratio, aligned, flags, cur_stream());
  };
  // torch.use_deterministic_algorithms(True): the tile-accumulating kernel is bitwise
  // reproducible; only where it does not apply does the reference's alert fire (:440)
  if (at::globalContext().deterministicAlgorithms()) {
    const int rc = run(B200DET_BWD_ZERO_INIT | B200DET_BWD_DETERMINISTIC);
    if (rc == 0) return done();
    TORCH_CHECK(rc == B200DET_ENOSUP, "roi_align_backward_kernel failed (", rc, "): ", b200det_last_error());
  }
  at::globalContext().alertNotDeterministic("roi_align_backward_kernel");
  check_rc(run(B200DET_BWD_ZERO_INIT), "roi_align_backward_kernel");
  return done();
}

// ---- torchvision::nms ---------------------------------------------------------
at::Tensor run_nms(const at::Tensor& dets, const at::Tensor& scores, const at::Tensor* idxs,
                   double iou_threshold, int mode) {
  const int dt = dtype_code(dets.scalar_type(), "nms_kernel");
  const at::Tensor dets_ = dets.contiguous();
  const at::Tensor scores_ = scores.to(dets.scalar_type()).contiguous();
  const int64_t N = dets.size(0);
  at::Tensor keep = at::empty({N + 1}, dets.options().dtype(at::kLong));
  const size_t ws_bytes = b200det_nms_workspace_bytes(N);
  at::Tensor ws = at::empty({(int64_t)ws_bytes}, dets.options().dtype(at::kByte));
  int64_t* keep_p = keep.data_ptr<int64_t>();
  int rc;
  if (idxs) {
    const at::Tensor idxs_ = idxs->to(at::kLong).contiguous();
    rc = b200det_batched_nms(dets_.data_ptr(), scores_.data_ptr(), idxs_.data_ptr<int64_t>(), dt, N,
                             iou_threshold, mode, keep_p, keep_p + N, ws.data_ptr(), ws_bytes,
                             cur_stream());
  } else {
    rc = b200det_nms(dets_.data_ptr(), scores_.data_ptr(), dt, N, iou_threshold, keep_p, keep_p + N,
                     ws.data_ptr(), ws_bytes, cur_stream());
  }
  check_rc(rc, "nms_kernel");
  // the op's result length is data dependent: one 8-byte read-back
  const int64_t n_keep = keep.narrow(0, N, 1).item<int64_t>();
  return keep.narrow(0, 0, n_keep);
}

at::Tensor nms_cuda(const at::Tensor& dets, const at::Tensor& scores, double iou_threshold) {
  TORCH_CHECK(dets.is_cuda(), "dets must be a CUDA tensor");
  TORCH_CHECK(scores.is_cuda(), "scores must be a CUDA tensor");
  TORCH_CHECK(dets.dim() == 2, "boxes should be a 2d tensor, got ", dets.dim(), "D");
  TORCH_CHECK(dets.size(1) == 4, "boxes should have 4 elements in dimension 1, got ", dets.size(1));
  TORCH_CHECK(scores.dim() == 1, "scores should be a 1d tensor, got ", scores.dim(), "D");
  TORCH_CHECK(dets.size(0) == scores.size(0),
              "boxes and scores should have same number of elements in ", "dimension 0, got ",
              dets.size(0), " and ", scores.size(0));
  at::cuda::CUDAGuard device_guard(dets.device());
  if (dets.numel() == 0) return at::empty({0}, dets.options().dtype(at::kLong));
  return run_nms(dets, scores, nullptr, iou_threshold, 0);
}

// ---- Autocast ------------------------------------------------------------------
// The reference's wrappers (csrc/ops/autocast/roi_align_kernel.cpp:12-31, autocast/nms_kernel.cpp:12-23) cast
// every argument to fp32, i.e. make an fp32 copy of the whole feature map, run the fp32 kernel and cast the
// result back.  The sm_100a kernels read fp16 / bf16 natively (fp32 arithmetic and accumulation, widened on load),
// which computes exactly ref_fp32(x.float(), rois.float()).to(x.dtype) without the copy.  So under
// torch.autocast("cuda") 16-bit features stay as they are, RoIs go to fp32 like in the reference, and fp32
// features take the fp32 path unchanged; the result has the input's dtype (test/test_ops.py:493-508).  These
// kernels are registered for the Autocast (CUDA) key and replace the reference's when its _C is loaded too.
at::Tensor roi_align_autocast_cuda(const at::Tensor& input, const at::Tensor& rois, double spatial_scale,
                                   int64_t pooled_height, int64_t pooled_width, int64_t sampling_ratio,
                                   bool aligned) {
  c10::impl::ExcludeDispatchKeyGuard no_autocast(c10::DispatchKey::Autocast);
  static auto op = c10::Dispatcher::singleton()
                       .findSchemaOrThrow("torchvision::roi_align", "")
                       .typed<at::Tensor(const at::Tensor&, const at::Tensor&, double, int64_t, int64_t, int64_t, bool)>();
  const auto it = input.scalar_type();
  const at::Tensor rois32 = at::autocast::cached_cast(at::kFloat, rois, c10::DeviceType::CUDA);
  if (input.is_cuda() && (it == at::kHalf || it == at::kBFloat16))
    return op.call(input, rois32, spatial_scale, pooled_height, pooled_width, sampling_ratio, aligned);
  return op
      .call(at::autocast::cached_cast(at::kFloat, input, c10::DeviceType::CUDA), rois32, spatial_scale,
            pooled_height, pooled_width, sampling_ratio, aligned)
      .to(it);
}

at::Tensor nms_autocast_cuda(const at::Tensor& dets, const at::Tensor& scores, double iou_threshold) {
  c10::impl::ExcludeDispatchKeyGuard no_autocast(c10::DispatchKey::Autocast);
  static auto op = c10::Dispatcher::singleton()
                       .findSchemaOrThrow("torchvision::nms", "")
                       .typed<at::Tensor(const at::Tensor&, const at::Tensor&, double)>();
  const auto dt = dets.scalar_type();
  // fp16 / bf16 boxes and scores of one dtype run natively: IoU in fp32 on the widened values, the same order of
  // scores -- what the fp32 cast would compute
  if (dets.is_cuda() && dt == scores.scalar_type() && (dt == at::kHalf || dt == at::kBFloat16))
    return op.call(dets, scores, iou_threshold);
  return op.call(at::autocast::cached_cast(at::kFloat, dets, c10::DeviceType::CUDA),
                 at::autocast::cached_cast(at::kFloat, scores, c10::DeviceType::CUDA), iou_threshold);
}

// ---- b200det::batched_nms -------------------------------------------------------
at::Tensor batched_nms_cuda(const at::Tensor& boxes, const at::Tensor& scores,
                            const at::Tensor& idxs, double iou_threshold, int64_t mode) {
  TORCH_CHECK(boxes.is_cuda() && scores.is_cuda() && idxs.is_cuda(), "batched_nms: CUDA tensors expected");
  TORCH_CHECK(boxes.dim() == 2 && boxes.size(1) == 4, "boxes should be a Tensor[N, 4]");
  TORCH_CHECK(scores.dim() == 1 && scores.size(0) == boxes.size(0), "scores should be a Tensor[N]");
  TORCH_CHECK(idxs.dim() == 1 && idxs.size(0) == boxes.size(0), "idxs should be a Tensor[N]");
  at::cuda::CUDAGuard device_guard(boxes.device());
  if (boxes.numel() == 0) return at::empty({0}, boxes.options().dtype(at::kLong));
  return run_nms(boxes, scores, &idxs, iou_threshold, (int)mode);
}

// ---- b200det::msroi_align ---------------------------------------------------------
std::tuple<at::Tensor, at::Tensor> msroi_align_forward_cuda(
    at::TensorList feats, const at::Tensor& rois, at::ArrayRef<double> scales,
    int64_t pooled_height, int64_t pooled_width, int64_t sampling_ratio, bool aligned,
    int64_t k_min, int64_t k_max, int64_t canonical_scale, int64_t canonical_level, double eps) {
  const int L = (int)feats.size();
  TORCH_CHECK(L >= 1 && L <= B200DET_MAX_LEVELS, "msroi_align: 1..", B200DET_MAX_LEVELS, " levels supported");
  TORCH_CHECK((int)scales.size() == L, "msroi_align: one scale per level expected");
  TORCH_CHECK(rois.is_cuda() && rois.dim() == 2 && rois.size(1) == 5, "rois must be a CUDA Tensor[K, 5]");
  const at::Tensor& f0 = feats[0];
  TORCH_CHECK(f0.is_cuda() && f0.dim() == 4, "features must be 4-D CUDA tensors");
  check_roi_dtype(f0, rois, "msroi_align");
  at::cuda::CUDAGuard device_guard(f0.device());
  const int dt = dtype_code(f0.scalar_type(), "msroi_align");
  int64_t pixels = 0;
  for (int l = 0; l < L; ++l) pixels += feats[l].dim() == 4 ? f0.size(0) * feats[l].size(2) * feats[l].size(3) : 0;
  const bool via_nhwc = !is_nhwc(f0) && fast_shape(dt, f0.size(1), pooled_height, pooled_width, sampling_ratio) &&
                        relayout_pays(rois.size(0), pooled_height, pooled_width, pixels);
  const bool nhwc = via_nhwc || is_nhwc(f0);
  std::vector<at::Tensor> hold(L);
  const void* ptrs[B200DET_MAX_LEVELS];
  int64_t hs[B200DET_MAX_LEVELS], ws[B200DET_MAX_LEVELS];
  double sc[B200DET_MAX_LEVELS];
  for (int l = 0; l < L; ++l) {
    const at::Tensor& f = feats[l];
    TORCH_CHECK(f.is_cuda() && f.dim() == 4 && f.scalar_type() == f0.scalar_type() &&
                    f.size(0) == f0.size(0) && f.size(1) == f0.size(1) && f.device() == f0.device(),
                "msroi_align: feature maps must agree in dtype, device, batch and channels");
    hold[l] = via_nhwc ? to_nhwc(f, dt, "msroi_align_forward")
                       : (nhwc ? f.contiguous(at::MemoryFormat::ChannelsLast) : f.contiguous());
    ptrs[l] = hold[l].data_ptr();
    hs[l] = f.size(2);
    ws[l] = f.size(3);
    sc[l] = scales[l];
  }
  const at::Tensor rois_ = arith_rois(rois, f0.scalar_type());
  const int64_t K = rois.size(0), C = f0.size(1);
  at::Tensor output = at::empty(
      {K, C, pooled_height, pooled_width},
      f0.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  at::Tensor levels = at::empty({K}, f0.options().dtype(at::kInt));
  if (output.numel() == 0) return {output, levels};
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_msroi_align_forward(ptrs, hs, ws, sc, L, layout, rois_.data_ptr(),
                                       levels.data_ptr<int32_t>(), 0, output.data_ptr(), layout, dt,
                                       f0.size(0), C, K, pooled_height, pooled_width,
                                       (int)sampling_ratio, aligned, (int)k_min, (int)k_max,
                                       (int)canonical_scale, (int)canonical_level, eps, cur_stream()),
           "msroi_align_forward");
  if (via_nhwc) output = to_nchw(output, dt, "msroi_align_forward");
  return {output, levels};
}

std::vector<at::Tensor> msroi_align_backward_cuda(
    const at::Tensor& grad, const at::Tensor& rois, const at::Tensor& levels,
    at::ArrayRef<double> scales, at::IntArrayRef heights, at::IntArrayRef widths,
    int64_t batch_size, int64_t channels, int64_t pooled_height, int64_t pooled_width,
    int64_t sampling_ratio, bool aligned, bool channels_last) {
  const int L = (int)scales.size();
  TORCH_CHECK(L >= 1 && L <= B200DET_MAX_LEVELS && (int)heights.size() == L && (int)widths.size() == L,
              "msroi_align_backward: bad level description");
  TORCH_CHECK(grad.is_cuda() && rois.is_cuda() && levels.is_cuda(), "msroi_align_backward: CUDA tensors expected");
  check_roi_dtype(grad, rois, "msroi_align_backward");
  at::cuda::CUDAGuard device_guard(grad.device());
  const int dt = dtype_code(grad.scalar_type(), "msroi_align_backward");
  std::vector<at::Tensor> out(L);
  void* ptrs[B200DET_MAX_LEVELS];
  int64_t hs[B200DET_MAX_LEVELS], ws[B200DET_MAX_LEVELS];
  double sc[B200DET_MAX_LEVELS];
  int64_t pixels = 0;
  for (int l = 0; l < L; ++l) pixels += batch_size * heights[l] * widths[l];
  const bool want_nhwc = channels_last;
  const bool via_nhwc = !want_nhwc && grad.dim() == 4 && grad.numel() > 0 && !is_nhwc(grad) &&
                        fast_shape(dt, channels, pooled_height, pooled_width, sampling_ratio) &&
                        relayout_pays(rois.size(0), pooled_height, pooled_width, pixels);
  channels_last = want_nhwc || via_nhwc;
  const at::Tensor grad_ = via_nhwc ? to_nhwc(grad, dt, "msroi_align_backward") : grad;
  for (int l = 0; l < L; ++l) {
    out[l] = at::empty({batch_size, channels, heights[l], widths[l]},
                       grad.options().memory_format(channels_last ? at::MemoryFormat::ChannelsLast
                                                                  : at::MemoryFormat::Contiguous));
    ptrs[l] = out[l].data_ptr();
    hs[l] = heights[l];
    ws[l] = widths[l];
    sc[l] = scales[l];
  }
  const at::Tensor rois_ = arith_rois(rois, grad.scalar_type());
  const at::Tensor levels_ = levels.to(at::kInt).contiguous();
  const int64_t gs[4] = {grad_.stride(0), grad_.stride(1), grad_.stride(2), grad_.stride(3)};
  auto done = [&]() {
    if (via_nhwc)
      for (int l = 0; l < L; ++l) out[l] = to_nchw(out[l], dt, "msroi_align_backward");
    return out;
  };
  auto run = [&](int flags) {
    return b200det_msroi_align_backward(grad_.data_ptr(), gs, rois_.data_ptr(), levels_.data_ptr<int32_t>(), ptrs,
                                        hs, ws, sc, L, channels_last ? B200DET_NHWC : B200DET_NCHW, dt,
                                        batch_size, channels, rois.size(0), pooled_height, pooled_width,
                                        (int)sampling_ratio, aligned, flags, cur_stream());
  };
  if (at::globalContext().deterministicAlgorithms()) {
    const int rc = run(B200DET_BWD_ZERO_INIT | B200DET_BWD_DETERMINISTIC);
    if (rc == 0) return done();
    TORCH_CHECK(rc == B200DET_ENOSUP, "msroi_align_backward failed (", rc, "): ", b200det_last_error());
  }
  at::globalContext().alertNotDeterministic("msroi_align_backward");
  check_rc(run(B200DET_BWD_ZERO_INIT), "msroi_align_backward");
  return done();
}

at::Tensor fpn_level_map_cuda(const at::Tensor& rois, int64_t k_min, int64_t k_max,
                              int64_t canonical_scale, int64_t canonical_level, double eps) {
  TORCH_CHECK(rois.is_cuda() && rois.dim() == 2 && rois.size(1) == 5, "rois must be a CUDA Tensor[K, 5]");
  at::cuda::CUDAGuard device_guard(rois.device());
  const at::Tensor rois_ = rois.to(at::kFloat).contiguous();
  at::Tensor levels = at::empty({rois.size(0)}, rois.options().dtype(at::kInt));
  check_rc(b200det_fpn_level_map(rois_.data_ptr(), B200DET_F32,
                                 rois.size(0), (int)k_min, (int)k_max, (int)canonical_scale,
                                 (int)canonical_level, eps, levels.data_ptr<int32_t>(), cur_stream()),
           "fpn_level_map");
  return levels;
}

// ---- torchvision::roi_pool ------------------------------------------------
// cuda/roi_pool_kernel.cu:120-278: same checks and messages; argmax is an int32 tensor with the
// reference's meaning.  Output and argmax follow the input's memory format.
std::tuple<at::Tensor, at::Tensor> roi_pool_forward_cuda(const at::Tensor& input, const at::Tensor& rois,
                                                         double spatial_scale, int64_t pooled_height,
                                                         int64_t pooled_width) {
  TORCH_CHECK(input.is_cuda(), "input must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(rois.size(1) == 5, "Tensor rois should have shape as Tensor[K, 5]");
  at::TensorArg input_t{input, "input", 1}, rois_t{rois, "rois", 2};
  at::CheckedFrom c = "roi_pool_forward_kernel";
  at::checkAllSameGPU(c, {input_t, rois_t});
  check_roi_dtype(input, rois, c);
  TORCH_CHECK(input.dim() == 4, "input must be a 4-D tensor [N, C, H, W]");
  at::cuda::CUDAGuard device_guard(input.device());
  const int dt = dtype_code(input.scalar_type(), "roi_pool_forward_kernel");
  const bool nhwc = is_nhwc(input);
  const at::Tensor input_ = nhwc ? input : input.contiguous();
  const at::Tensor rois_ = arith_rois(rois, input.scalar_type());
  const auto mf = nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous;
  at::Tensor output = at::empty({rois.size(0), input.size(1), pooled_height, pooled_width}, input.options().memory_format(mf));
  at::Tensor argmax = at::empty({rois.size(0), input.size(1), pooled_height, pooled_width},
                                input.options().dtype(at::kInt).memory_format(mf));
  if (output.numel() == 0) return {output, argmax};
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_roi_pool_forward(input_.data_ptr(), layout, rois_.data_ptr(), output.data_ptr(),
                                    argmax.data_ptr<int32_t>(), layout, dt, input.size(0), input.size(1), input.size(2),
                                    input.size(3), rois.size(0), pooled_height, pooled_width, spatial_scale,
                                    cur_stream()),
           "roi_pool_forward_kernel");
  return {output, argmax};
}

at::Tensor roi_pool_backward_cuda(const at::Tensor& grad, const at::Tensor& rois, const at::Tensor& argmax,
                                  double spatial_scale, int64_t pooled_height, int64_t pooled_width,
                                  int64_t batch_size, int64_t channels, int64_t height, int64_t width) {
  TORCH_CHECK(grad.is_cuda(), "grad must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(argmax.is_cuda(), "argmax must be a CUDA tensor");
  at::TensorArg grad_t{grad, "grad", 1}, rois_t{rois, "rois", 2}, argmax_t{argmax, "argmax", 3};
  at::CheckedFrom c = "roi_pool_backward_kernel";
  at::checkAllSameGPU(c, {grad_t, rois_t, argmax_t});
  check_roi_dtype(grad, rois, c);
  at::cuda::CUDAGuard device_guard(grad.device());
  const int dt = dtype_code(grad.scalar_type(), "roi_pool_backward_kernel");
  const bool am_nhwc = is_nhwc(argmax);
  const at::Tensor argmax_ = (am_nhwc ? argmax : argmax.contiguous()).to(at::kInt);
  const bool nhwc = am_nhwc;  // grad_input follows the forward's memory format, which argmax recorded
  at::Tensor grad_input = at::empty(
      {batch_size, channels, height, width},
      grad.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  if (grad.numel() == 0) return grad_input.zero_();
  const at::Tensor rois_ = arith_rois(rois, grad.scalar_type());
  const int64_t gs[4] = {grad.stride(0), grad.stride(1), grad.stride(2), grad.stride(3)};
  at::globalContext().alertNotDeterministic("roi_pool_backward_kernel");
  check_rc(b200det_roi_pool_backward(grad.data_ptr(), gs, argmax_.data_ptr<int32_t>(),
                                     am_nhwc ? B200DET_NHWC : B200DET_NCHW, rois_.data_ptr(), grad_input.data_ptr(),
                                     nhwc ? B200DET_NHWC : B200DET_NCHW, dt, batch_size, channels, height, width,
                                     rois.size(0), pooled_height, pooled_width, cur_stream()),
           "roi_pool_backward_kernel");
  return grad_input;
}

// ---- torchvision::ps_roi_pool ---------------------------------------------
// cuda/ps_roi_pool_kernel.cu:142-280: same checks and messages.
std::tuple<at::Tensor, at::Tensor> ps_roi_pool_forward_cuda(const at::Tensor& input, const at::Tensor& rois,
                                                            double spatial_scale, int64_t pooled_height,
                                                            int64_t pooled_width) {
  TORCH_CHECK(input.is_cuda(), "input must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(rois.size(1) == 5, "Tensor rois should have shape as Tensor[K, 5]");
  at::TensorArg input_t{input, "input", 1}, rois_t{rois, "rois", 2};
  at::CheckedFrom c = "ps_roi_pool_forward_kernel";
  at::checkAllSameGPU(c, {input_t, rois_t});
  check_roi_dtype(input, rois, c);
  TORCH_CHECK(input.dim() == 4, "input must be a 4-D tensor [N, C, H, W]");
  TORCH_CHECK(pooled_height > 0 && pooled_width > 0 && input.size(1) % (pooled_height * pooled_width) == 0,
              "input channels must be a multiple of pooling height * pooling width");
  at::cuda::CUDAGuard device_guard(input.device());
  const int dt = dtype_code(input.scalar_type(), "ps_roi_pool_forward_kernel");
  const bool nhwc = is_nhwc(input);
  const at::Tensor input_ = nhwc ? input : input.contiguous();
  const at::Tensor rois_ = arith_rois(rois, input.scalar_type());
  const auto mf = nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous;
  const int64_t c_out = input.size(1) / (pooled_height * pooled_width);
  at::Tensor output = at::empty({rois.size(0), c_out, pooled_height, pooled_width}, input.options().memory_format(mf));
  at::Tensor mapping = at::empty({rois.size(0), c_out, pooled_height, pooled_width},
                                 input.options().dtype(at::kInt).memory_format(mf));
  if (output.numel() == 0) return {output, mapping};
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_ps_roi_pool_forward(input_.data_ptr(), layout, rois_.data_ptr(), output.data_ptr(),
                                       mapping.data_ptr<int32_t>(), layout, dt, input.size(0), input.size(1),
                                       input.size(2), input.size(3), rois.size(0), pooled_height, pooled_width,
                                       spatial_scale, cur_stream()),
           "ps_roi_pool_forward_kernel");
  return {output, mapping};
}

at::Tensor ps_roi_pool_backward_cuda(const at::Tensor& grad, const at::Tensor& rois, const at::Tensor& channel_mapping,
                                     double spatial_scale, int64_t pooled_height, int64_t pooled_width,
                                     int64_t batch_size, int64_t channels, int64_t height, int64_t width) {
  TORCH_CHECK(grad.is_cuda(), "grad must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(channel_mapping.is_cuda(), "channel_mapping must be a CUDA tensor");
  at::TensorArg grad_t{grad, "grad", 1}, rois_t{rois, "rois", 2}, cm_t{channel_mapping, "channel_mapping", 3};
  at::CheckedFrom c = "ps_roi_pool_backward_kernel";
  at::checkAllSameGPU(c, {grad_t, rois_t, cm_t});
  check_roi_dtype(grad, rois, c);
  at::cuda::CUDAGuard device_guard(grad.device());
  const int dt = dtype_code(grad.scalar_type(), "ps_roi_pool_backward_kernel");
  const bool nhwc = is_nhwc(channel_mapping);
  const at::Tensor cm_ = (nhwc ? channel_mapping : channel_mapping.contiguous()).to(at::kInt);
  at::Tensor grad_input = at::empty(
      {batch_size, channels, height, width},
      grad.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  if (grad.numel() == 0) return grad_input.zero_();
  const at::Tensor rois_ = arith_rois(rois, grad.scalar_type());
  const int64_t gs[4] = {grad.stride(0), grad.stride(1), grad.stride(2), grad.stride(3)};
  at::globalContext().alertNotDeterministic("ps_roi_pool_backward_kernel");
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_ps_roi_pool_backward(grad.data_ptr(), gs, cm_.data_ptr<int32_t>(), layout, rois_.data_ptr(),
                                        grad_input.data_ptr(), layout, dt, batch_size, channels, height, width,
                                        rois.size(0), pooled_height, pooled_width, spatial_scale, cur_stream()),
           "ps_roi_pool_backward_kernel");
  return grad_input;
}

// ---- torchvision::ps_roi_align --------------------------------------------
// cuda/ps_roi_align_kernel.cu:293-444: same checks and messages.
std::tuple<at::Tensor, at::Tensor> ps_roi_align_forward_cuda(const at::Tensor& input, const at::Tensor& rois,
                                                             double spatial_scale, int64_t pooled_height,
                                                             int64_t pooled_width, int64_t sampling_ratio) {
  TORCH_CHECK(input.is_cuda(), "input must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(rois.size(1) == 5, "Tensor rois should have shape as Tensor[K, 5]");
  at::TensorArg input_t{input, "input", 1}, rois_t{rois, "rois", 2};
  at::CheckedFrom c = "ps_roi_align_forward_kernel";
  at::checkAllSameGPU(c, {input_t, rois_t});
  check_roi_dtype(input, rois, c);
  TORCH_CHECK(input.dim() == 4, "input must be a 4-D tensor [N, C, H, W]");
  TORCH_CHECK(pooled_height > 0 && pooled_width > 0 && input.size(1) % (pooled_height * pooled_width) == 0,
              "input channels must be a multiple of pooling height * pooling width");
  at::cuda::CUDAGuard device_guard(input.device());
  const int dt = dtype_code(input.scalar_type(), "ps_roi_align_forward_kernel");
  const bool nhwc = is_nhwc(input);
  const at::Tensor input_ = nhwc ? input : input.contiguous();
  const at::Tensor rois_ = arith_rois(rois, input.scalar_type());
  const auto mf = nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous;
  const int64_t c_out = input.size(1) / (pooled_height * pooled_width);
  at::Tensor output = at::empty({rois.size(0), c_out, pooled_height, pooled_width}, input.options().memory_format(mf));
  at::Tensor mapping = at::empty({rois.size(0), c_out, pooled_height, pooled_width},
                                 input.options().dtype(at::kInt).memory_format(mf));
  if (output.numel() == 0) return {output, mapping};
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_ps_roi_align_forward(input_.data_ptr(), layout, rois_.data_ptr(), output.data_ptr(),
                                        mapping.data_ptr<int32_t>(), layout, dt, input.size(0), input.size(1),
                                        input.size(2), input.size(3), rois.size(0), pooled_height, pooled_width,
                                        spatial_scale, (int)sampling_ratio, cur_stream()),
           "ps_roi_align_forward_kernel");
  return {output, mapping};
}

at::Tensor ps_roi_align_backward_cuda(const at::Tensor& grad, const at::Tensor& rois, const at::Tensor& channel_mapping,
                                      double spatial_scale, int64_t pooled_height, int64_t pooled_width,
                                      int64_t sampling_ratio, int64_t batch_size, int64_t channels, int64_t height,
                                      int64_t width) {
  TORCH_CHECK(grad.is_cuda(), "grad must be a CUDA tensor");
  TORCH_CHECK(rois.is_cuda(), "rois must be a CUDA tensor");
  TORCH_CHECK(channel_mapping.is_cuda(), "channel_mapping must be a CUDA tensor");
  at::TensorArg grad_t{grad, "grad", 1}, rois_t{rois, "rois", 2}, cm_t{channel_mapping, "channel_mapping", 3};
  at::CheckedFrom c = "ps_roi_align_backward_kernel";
  at::checkAllSameGPU(c, {grad_t, rois_t, cm_t});
  check_roi_dtype(grad, rois, c);
  at::cuda::CUDAGuard device_guard(grad.device());
  const int dt = dtype_code(grad.scalar_type(), "ps_roi_align_backward_kernel");
  const bool nhwc = is_nhwc(channel_mapping);
  const at::Tensor cm_ = (nhwc ? channel_mapping : channel_mapping.contiguous()).to(at::kInt);
  at::Tensor grad_input = at::empty(
      {batch_size, channels, height, width},
      grad.options().memory_format(nhwc ? at::MemoryFormat::ChannelsLast : at::MemoryFormat::Contiguous));
  if (grad.numel() == 0) return grad_input.zero_();
  const at::Tensor rois_ = arith_rois(rois, grad.scalar_type());
  const int64_t gs[4] = {grad.stride(0), grad.stride(1), grad.stride(2), grad.stride(3)};
  at::globalContext().alertNotDeterministic("ps_roi_align_backward_kernel");
  const int layout = nhwc ? B200DET_NHWC : B200DET_NCHW;
  check_rc(b200det_ps_roi_align_backward(grad.data_ptr(), gs, cm_.data_ptr<int32_t>(), layout, rois_.data_ptr(),
                                         grad_input.data_ptr(), layout, dt, batch_size, channels, height, width,
                                         rois.size(0), pooled_height, pooled_width, spatial_scale,
                                         (int)sampling_ratio, cur_stream()),
           "ps_roi_align_backward_kernel");
  return grad_input;
}

// ---- registration -------------------------------------------------------------------
int g_defined = 0;  // bit0: we defined the roi_align schemas, bit1: nms, bit2: roi_pool, bit3: ps_roi_pool, bit4: ps_roi_align
std::vector<std::unique_ptr<torch::Library>> g_libs;

bool has_schema(const char* name) {
  return c10::Dispatcher::singleton().findSchema({name, ""}).has_value();
}

struct Registrar {
  Registrar() {
    if (!has_schema("torchvision::roi_align")) {
      g_libs.emplace_back(new torch::Library(torch::Library::FRAGMENT, "torchvision", std::nullopt, __FILE__, __LINE__));
      g_libs.back()->def(
          "roi_align(Tensor input, Tensor rois, float spatial_scale, SymInt pooled_height, SymInt "
          "pooled_width, int sampling_ratio, bool aligned) -> Tensor");
      g_libs.back()->def(
          "_roi_align_backward(Tensor grad, Tensor rois, float spatial_scale, SymInt pooled_height, "
          "SymInt pooled_width, SymInt batch_size, SymInt channels, SymInt height, SymInt width, int "
          "sampling_ratio, bool aligned) -> Tensor");
      g_defined |= 1;
    }
    if (!has_schema("torchvision::nms")) {
      g_libs.emplace_back(new torch::Library(torch::Library::FRAGMENT, "torchvision", std::nullopt, __FILE__, __LINE__));
      g_libs.back()->def("nms(Tensor dets, Tensor scores, float iou_threshold) -> Tensor");
      g_defined |= 2;
    }
    if (!has_schema("torchvision::roi_pool")) {
      g_libs.emplace_back(new torch::Library(torch::Library::FRAGMENT, "torchvision", std::nullopt, __FILE__, __LINE__));
      g_libs.back()->def(
          "roi_pool(Tensor input, Tensor rois, float spatial_scale, SymInt pooled_height, SymInt pooled_width) -> "
          "(Tensor, Tensor)");
      g_libs.back()->def(
          "_roi_pool_backward(Tensor grad, Tensor rois, Tensor argmax, float spatial_scale, SymInt pooled_height, "
          "SymInt pooled_width, SymInt batch_size, SymInt channels, SymInt height, SymInt width) -> Tensor");
      g_defined |= 4;
    }
    if (!has_schema("torchvision::ps_roi_pool")) {
      g_libs.emplace_back(new torch::Library(torch::Library::FRAGMENT, "torchvision", std::nullopt, __FILE__, __LINE__));
      g_libs.back()->def(
          "ps_roi_pool(Tensor input, Tensor rois, float spatial_scale, SymInt pooled_height, SymInt pooled_width) -> "
          "(Tensor, Tensor)");
      g_libs.back()->def(
          "_ps_roi_pool_backward(Tensor grad, Tensor rois, Tensor channel_mapping, float spatial_scale, SymInt "
          "pooled_height, SymInt pooled_width, SymInt batch_size, SymInt channels, SymInt height, SymInt width) -> Tensor");
      g_defined |= 8;
    }
    if (!has_schema("torchvision::ps_roi_align")) {
      g_libs.emplace_back(new torch::Library(torch::Library::FRAGMENT, "torchvision", std::nullopt, __FILE__, __LINE__));
      g_libs.back()->def(
          "ps_roi_align(Tensor input, Tensor rois, float spatial_scale, SymInt pooled_height, SymInt pooled_width, "
          "int sampling_ratio) -> (Tensor, Tensor)");
      g_libs.back()->def(
          "_ps_roi_align_backward(Tensor grad, Tensor rois, Tensor channel_mapping, float spatial_scale, SymInt "
          "pooled_height, SymInt pooled_width, int sampling_ratio, SymInt batch_size, SymInt channels, SymInt height, "
          "SymInt width) -> Tensor");
      g_defined |= 16;
    }
    g_libs.emplace_back(new torch::Library(torch::Library::IMPL, "torchvision", c10::DispatchKey::CUDA, __FILE__, __LINE__));
    g_libs.back()->impl("roi_align", TORCH_FN(roi_align_forward_cuda));
    g_libs.back()->impl("_roi_align_backward", TORCH_FN(roi_align_backward_cuda));
    g_libs.back()->impl("nms", TORCH_FN(nms_cuda));
    g_libs.back()->impl("roi_pool", TORCH_FN(roi_pool_forward_cuda));
    g_libs.back()->impl("_roi_pool_backward", TORCH_FN(roi_pool_backward_cuda));
    g_libs.back()->impl("ps_roi_pool", TORCH_FN(ps_roi_pool_forward_cuda));
    g_libs.back()->impl("_ps_roi_pool_backward", TORCH_FN(ps_roi_pool_backward_cuda));
    g_libs.back()->impl("ps_roi_align", TORCH_FN(ps_roi_align_forward_cuda));
    g_libs.back()->impl("_ps_roi_align_backward", TORCH_FN(ps_roi_align_backward_cuda));
    g_libs.emplace_back(new torch::Library(torch::Library::IMPL, "torchvision", c10::DispatchKey::Autocast, __FILE__, __LINE__));
    g_libs.back()->impl("roi_align", TORCH_FN(roi_align_autocast_cuda));
    g_libs.back()->impl("nms", TORCH_FN(nms_autocast_cuda));
  }
};
Registrar g_registrar;

}  // namespace

TORCH_LIBRARY(b200det, m) {
  m.def(
      "msroi_align(Tensor[] feats, Tensor rois, float[] scales, int pooled_height, int pooled_width, "
      "int sampling_ratio, bool aligned, int k_min, int k_max, int canonical_scale, int "
      "canonical_level, float eps) -> (Tensor, Tensor)");
  m.def(
      "msroi_align_backward(Tensor grad, Tensor rois, Tensor levels, float[] scales, int[] heights, "
      "int[] widths, int batch_size, int channels, int pooled_height, int pooled_width, int "
      "sampling_ratio, bool aligned, bool channels_last) -> Tensor[]");
  m.def("batched_nms(Tensor boxes, Tensor scores, Tensor idxs, float iou_threshold, int mode) -> Tensor");
  m.def(
      "fpn_level_map(Tensor rois, int k_min, int k_max, int canonical_scale, int canonical_level, "
      "float eps) -> Tensor");
}

TORCH_LIBRARY_IMPL(b200det, CUDA, m) {
  m.impl("msroi_align", TORCH_FN(msroi_align_forward_cuda));
  m.impl("msroi_align_backward", TORCH_FN(msroi_align_backward_cuda));
  m.impl("batched_nms", TORCH_FN(batched_nms_cuda));
  m.impl("fpn_level_map", TORCH_FN(fpn_level_map_cuda));
}

extern "C" int b200det_shim_defined_schemas(void) { return g_defined; }
