// roi_pool_sm100.cu -- torchvision::roi_pool forward / backward (the sibling
// op with roi_align's skeleton).  Replaces roi_pool_forward_kernel_impl / roi_pool_backward_kernel_impl
// (torchvision/csrc/ops/cuda/roi_pool_kernel.cu:13-118; CPU twin cpu/roi_pool_kernel.cpp:16-125).
//
// One thread per pooled element, like the reference, but stride-addressed: with a channels_last feature map
// the flat index is decomposed channel-fastest, so a warp reads 32 consecutive channels of one pixel
// (one 128-byte line) where the NCHW decomposition reads 32 pixels of 32 planes.  Max pooling over
// integer bins -- HBM/L2-bound, no tensor cores.  argmax keeps the reference's meaning (h * W + w inside
// the channel plane, -1 for an empty bin) whatever the memory layout, so either backward can consume it.
#include <float.h>

#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kThreads = 256;

template <typename A>
struct PoolGeom {
  int batch, start_h, start_w;
  A bin_h, bin_w;
};

// roi_pool_kernel.cu:36-47: round() of the scaled corners, malformed RoIs forced to 1x1
template <typename A>
__device__ __forceinline__ PoolGeom<A> pool_geometry(const A* __restrict__ roi, A scale, int PH, int PW) {
  PoolGeom<A> g;
  g.batch = (int)__ldg(roi);
  const int sw = (int)round(Ar<A>::mul(__ldg(roi + 1), scale)), sh = (int)round(Ar<A>::mul(__ldg(roi + 2), scale));
  const int ew = (int)round(Ar<A>::mul(__ldg(roi + 3), scale)), eh = (int)round(Ar<A>::mul(__ldg(roi + 4), scale));
  const int rw = max(ew - sw + 1, 1), rh = max(eh - sh + 1, 1);
  g.start_h = sh;
  g.start_w = sw;
  g.bin_h = Ar<A>::div((A)rh, (A)PH);
  g.bin_w = Ar<A>::div((A)rw, (A)PW);
  return g;
}

template <typename T, bool CH_FASTEST>
__global__ void __launch_bounds__(kThreads) roi_pool_fwd_kernel(
    long long total, const T* __restrict__ input, long long sn, long long sc, long long sh, long long sw,
    const typename Io<T>::acc* __restrict__ rois, typename Io<T>::acc scale, int N, int C, int H, int W, int PH,
    int PW, T* __restrict__ out, int32_t* __restrict__ argmax, long long os_k, long long os_c, long long os_h,
    long long os_w) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    int pw, ph, c;
    long long k;
    if (CH_FASTEST) {
      c = (int)(index % C);
      pw = (int)((index / C) % PW);
      ph = (int)((index / C / PW) % PH);
      k = index / C / PW / PH;
    } else {
      pw = (int)(index % PW);
      ph = (int)((index / PW) % PH);
      c = (int)((index / PW / PH) % C);
      k = index / PW / PH / C;
    }
    const PoolGeom<A> g = pool_geometry<A>(rois + k * 5, scale, PH, PW);
    // roi_pool_kernel.cu:49-58: floor / ceil of the bin edges, offset, clip to the map
    int hstart = (int)floor(Ar<A>::mul((A)ph, g.bin_h)), wstart = (int)floor(Ar<A>::mul((A)pw, g.bin_w));
    int hend = (int)ceil(Ar<A>::mul((A)(ph + 1), g.bin_h)), wend = (int)ceil(Ar<A>::mul((A)(pw + 1), g.bin_w));
    hstart = min(max(hstart + g.start_h, 0), H);
    hend = min(max(hend + g.start_h, 0), H);
    wstart = min(max(wstart + g.start_w, 0), W);
    wend = min(max(wend + g.start_w, 0), W);
    const bool is_empty = (hend <= hstart) || (wend <= wstart) || (unsigned)g.batch >= (unsigned)N;
    A maxval = is_empty ? (A)0 : (A)-FLT_MAX;  // an empty pooling region is zero (:61)
    int maxidx = -1;                            // nothing pooled: nothing is back-propagated (:63)
    if (!is_empty) {
      const T* base = input + (long long)g.batch * sn + (long long)c * sc;
      for (int h = hstart; h < hend; ++h)
        for (int w = wstart; w < wend; ++w) {
          const A v = Io<T>::ld(base + h * sh + w * sw);
          if (v > maxval) {
            maxval = v;
            maxidx = h * W + w;
          }
        }
    }
    const long long o = k * os_k + c * os_c + ph * os_h + pw * os_w;
    out[o] = Io<T>::to(maxval);
    argmax[o] = maxidx;
  }
}

template <typename T, bool CH_FASTEST>
__global__ void __launch_bounds__(kThreads) roi_pool_bwd_kernel(
    long long total, const T* __restrict__ grad, long long gs_k, long long gs_c, long long gs_h, long long gs_w,
    const int32_t* __restrict__ argmax, long long as_k, long long as_c, long long as_h, long long as_w,
    const typename Io<T>::acc* __restrict__ rois, int N, int C, int W, int PH, int PW, T* __restrict__ grad_in,
    long long sn, long long sc, long long sh, long long sw) {
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    int pw, ph, c;
    long long k;
    if (CH_FASTEST) {
      c = (int)(index % C);
      pw = (int)((index / C) % PW);
      ph = (int)((index / C / PW) % PH);
      k = index / C / PW / PH;
    } else {
      pw = (int)(index % PW);
      ph = (int)((index / PW) % PH);
      c = (int)((index / PW / PH) % C);
      k = index / PW / PH / C;
    }
    const int b = (int)__ldg(rois + k * 5);
    const int am = argmax[k * as_k + c * as_c + ph * as_h + pw * as_w];
    if (am < 0 || (unsigned)b >= (unsigned)N) continue;  // roi_pool_kernel.cu:109
    const int h = am / W, w = am - h * W;
    atomicAdd(grad_in + (long long)b * sn + (long long)c * sc + h * sh + w * sw,
              grad[k * gs_k + c * gs_c + ph * gs_h + pw * gs_w]);
  }
}

inline void nchw_or_nhwc(int layout, long long C, long long H, long long W, long long* s) {
  if (layout == B200DET_NHWC) { s[0] = H * W * C; s[1] = 1; s[2] = W * C; s[3] = C; }
  else { s[0] = C * H * W; s[1] = H * W; s[2] = W; s[3] = 1; }
}

inline unsigned grid_for(long long total) {
  const long long b = (total + kThreads - 1) / kThreads;
  return (unsigned)(b > (1LL << 20) ? (1LL << 20) : (b < 1 ? 1 : b));
}

template <typename T>
int fwd_t(const void* input, int in_layout, const void* rois, void* out, int32_t* argmax, int out_layout, long long N,
          long long C, long long H, long long W, long long K, int PH, int PW, double scale, cudaStream_t st) {
  using A = typename Io<T>::acc;
  long long is[4], os[4];
  nchw_or_nhwc(in_layout, C, H, W, is);
  nchw_or_nhwc(out_layout, C, PH, PW, os);
  const long long total = K * C * PH * PW;
  if (in_layout == B200DET_NHWC)
    roi_pool_fwd_kernel<T, true><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)input, is[0], is[1], is[2], is[3], (const A*)rois, (A)scale, (int)N, (int)C, (int)H, (int)W,
        PH, PW, (T*)out, argmax, os[0], os[1], os[2], os[3]);
  else
    roi_pool_fwd_kernel<T, false><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)input, is[0], is[1], is[2], is[3], (const A*)rois, (A)scale, (int)N, (int)C, (int)H, (int)W,
        PH, PW, (T*)out, argmax, os[0], os[1], os[2], os[3]);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int bwd_t(const void* grad, const int64_t* gs, const int32_t* argmax, int am_layout, const void* rois, void* grad_in,
          int gi_layout, long long N, long long C, long long H, long long W, long long K, int PH, int PW,
          cudaStream_t st) {
  using A = typename Io<T>::acc;
  long long is[4], as[4];
  nchw_or_nhwc(gi_layout, C, H, W, is);
  nchw_or_nhwc(am_layout, C, PH, PW, as);
  B200DET_CUDA_OK(cudaMemsetAsync(grad_in, 0, (size_t)(N * C * H * W) * sizeof(T), st));
  const long long total = K * C * PH * PW;
  if (total == 0) return 0;
  if (gi_layout == B200DET_NHWC)
    roi_pool_bwd_kernel<T, true><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)grad, gs[0], gs[1], gs[2], gs[3], argmax, as[0], as[1], as[2], as[3], (const A*)rois, (int)N,
        (int)C, (int)W, PH, PW, (T*)grad_in, is[0], is[1], is[2], is[3]);
  else
    roi_pool_bwd_kernel<T, false><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)grad, gs[0], gs[1], gs[2], gs[3], argmax, as[0], as[1], as[2], as[3], (const A*)rois, (int)N,
        (int)C, (int)W, PH, PW, (T*)grad_in, is[0], is[1], is[2], is[3]);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int roi_pool_forward(const void* input, int in_layout, const void* rois, void* out, int32_t* argmax, int out_layout,
                     int dtype, long long N, long long C, long long H, long long W, long long K, int PH, int PW,
                     double scale, cudaStream_t st) {
  if (K * C * PH * PW == 0) return 0;
  if (H * W >= (1LL << 31)) return fail(B200DET_EINVAL, "roi_pool: H*W must fit in 31 bits (argmax is an int like the reference's)");
  switch (dtype) {
    case B200DET_F32: return fwd_t<float>(input, in_layout, rois, out, argmax, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_F16: return fwd_t<__half>(input, in_layout, rois, out, argmax, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_BF16: return fwd_t<__nv_bfloat16>(input, in_layout, rois, out, argmax, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_F64: return fwd_t<double>(input, in_layout, rois, out, argmax, out_layout, N, C, H, W, K, PH, PW, scale, st);
  }
  return fail(B200DET_ENOSUP, "roi_pool: unknown dtype %d", dtype);
}

int roi_pool_backward(const void* grad, const int64_t* gs, const int32_t* argmax, int am_layout, const void* rois,
                      void* grad_in, int gi_layout, int dtype, long long N, long long C, long long H, long long W,
                      long long K, int PH, int PW, cudaStream_t st) {
  if (N * C * H * W == 0) return 0;
  switch (dtype) {
    case B200DET_F32: return bwd_t<float>(grad, gs, argmax, am_layout, rois, grad_in, gi_layout, N, C, H, W, K, PH, PW, st);
    case B200DET_F16: return bwd_t<__half>(grad, gs, argmax, am_layout, rois, grad_in, gi_layout, N, C, H, W, K, PH, PW, st);
    case B200DET_BF16: return bwd_t<__nv_bfloat16>(grad, gs, argmax, am_layout, rois, grad_in, gi_layout, N, C, H, W, K, PH, PW, st);
    case B200DET_F64: return bwd_t<double>(grad, gs, argmax, am_layout, rois, grad_in, gi_layout, N, C, H, W, K, PH, PW, st);
  }
  return fail(B200DET_ENOSUP, "roi_pool: unknown dtype %d", dtype);
}

}  // namespace b200det
