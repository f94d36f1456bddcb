// ps_roi_pool_sm100.cu -- torchvision::ps_roi_pool forward / backward.
// Replaces ps_roi_pool_forward_kernel_impl / ps_roi_pool_backward_kernel_impl
// (torchvision/csrc/ops/cuda/ps_roi_pool_kernel.cu:13-140; CPU twin cpu/ps_roi_pool_kernel.cpp:14-146).
// Position-sensitive average pooling: output channel c_out of bin (ph, pw) averages input channel
// c_in = (c_out * PH + ph) * PW + pw over the integer bin.  One thread per pooled element, the bin summed in
// the reference's h-major order (same fp32 result), stride-addressed so NCHW and channels_last both work.
// The reference's asymmetry is kept: the forward clips bins to [0, size - 1], the backward to [0, size].
#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kThreads = 256;

template <typename A>
struct PsGeom {
  int batch, start_h, start_w;
  A bin_h, bin_w;
};

// ps_roi_pool_kernel.cu:33-44: round() of the scaled corners, too small RoIs forced to 1x1 (no +1 here)
template <typename A>
__device__ __forceinline__ PsGeom<A> ps_geometry(const A* __restrict__ roi, A scale, int PH, int PW) {
  PsGeom<A> g;
  g.batch = (int)__ldg(roi);
  const int sw = (int)round(Ar<A>::mul(__ldg(roi + 1), scale)), sh = (int)round(Ar<A>::mul(__ldg(roi + 2), scale));
  const int ew = (int)round(Ar<A>::mul(__ldg(roi + 3), scale)), eh = (int)round(Ar<A>::mul(__ldg(roi + 4), scale));
  g.start_h = sh;
  g.start_w = sw;
  g.bin_h = Ar<A>::div((A)max(eh - sh, 1), (A)PH);
  g.bin_w = Ar<A>::div((A)max(ew - sw, 1), (A)PW);
  return g;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) ps_roi_pool_fwd_kernel(
    long long total, const T* __restrict__ input, long long sn, long long sc, long long sh, long long sw,
    const typename Io<T>::acc* __restrict__ rois, typename Io<T>::acc scale, int N, int H, int W, int PH, int PW,
    int C_out, T* __restrict__ out, int32_t* __restrict__ mapping, long long os_k, long long os_c, long long os_h,
    long long os_w) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    const int pw = (int)(index % PW), ph = (int)((index / PW) % PH);
    const int c_out = (int)((index / PW / PH) % C_out);
    const long long k = index / PW / PH / C_out;
    const PsGeom<A> g = ps_geometry<A>(rois + k * 5, scale, PH, PW);
    int hstart = (int)floor(Ar<A>::mul((A)ph, g.bin_h)), wstart = (int)floor(Ar<A>::mul((A)pw, g.bin_w));
    int hend = (int)ceil(Ar<A>::mul((A)(ph + 1), g.bin_h)), wend = (int)ceil(Ar<A>::mul((A)(pw + 1), g.bin_w));
    hstart = min(max(hstart + g.start_h, 0), H - 1);
    hend = min(max(hend + g.start_h, 0), H - 1);
    wstart = min(max(wstart + g.start_w, 0), W - 1);
    wend = min(max(wend + g.start_w, 0), W - 1);
    const bool is_empty = (hend <= hstart) || (wend <= wstart) || (unsigned)g.batch >= (unsigned)N;
    const int c_in = (c_out * PH + ph) * PW + pw;
    A sum = 0;
    if (!is_empty) {
      const T* base = input + (long long)g.batch * sn + (long long)c_in * sc;
      for (int h = hstart; h < hend; ++h)
        for (int w = wstart; w < wend; ++w) sum = Ar<A>::add(sum, Io<T>::ld(base + h * sh + w * sw));
    }
    const A area = (A)((hend - hstart) * (wend - wstart));
    const long long o = k * os_k + c_out * os_c + ph * os_h + pw * os_w;
    out[o] = Io<T>::to(is_empty ? (A)0 : Ar<A>::div(sum, area));
    mapping[o] = c_in;
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) ps_roi_pool_bwd_kernel(
    long long total, const T* __restrict__ grad, long long gs_k, long long gs_c, long long gs_h, long long gs_w,
    const int32_t* __restrict__ mapping, long long ms_k, long long ms_c, long long ms_h, long long ms_w,
    const typename Io<T>::acc* __restrict__ rois, typename Io<T>::acc scale, int N, int H, int W, int PH, int PW,
    int C_out, T* __restrict__ grad_in, long long sn, long long sc, long long sh, long long sw) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    const int pw = (int)(index % PW), ph = (int)((index / PW) % PH);
    const int c_out = (int)((index / PW / PH) % C_out);
    const long long k = index / PW / PH / C_out;
    const PsGeom<A> g = ps_geometry<A>(rois + k * 5, scale, PH, PW);
    if ((unsigned)g.batch >= (unsigned)N) continue;
    int hstart = (int)floor(Ar<A>::mul((A)ph, g.bin_h)), wstart = (int)floor(Ar<A>::mul((A)pw, g.bin_w));
    int hend = (int)ceil(Ar<A>::mul((A)(ph + 1), g.bin_h)), wend = (int)ceil(Ar<A>::mul((A)(pw + 1), g.bin_w));
    hstart = min(max(hstart + g.start_h, 0), H);  // ps_roi_pool_kernel.cu:108-111: [0, size] here
    hend = min(max(hend + g.start_h, 0), H);
    wstart = min(max(wstart + g.start_w, 0), W);
    wend = min(max(wend + g.start_w, 0), W);
    if ((hend <= hstart) || (wend <= wstart)) continue;
    const int c_in = mapping[k * ms_k + c_out * ms_c + ph * ms_h + pw * ms_w];
    const A area = (A)((hend - hstart) * (wend - wstart));
    const A diff = Ar<A>::div((A)Io<T>::ld(grad + k * gs_k + c_out * gs_c + ph * gs_h + pw * gs_w), area);
    T* base = grad_in + (long long)g.batch * sn + (long long)c_in * sc;
    const T d = Io<T>::to(diff);
    for (int h = hstart; h < hend; ++h)
      for (int w = wstart; w < wend; ++w) atomicAdd(base + h * sh + w * sw, d);
  }
}

inline void strides_of(int layout, long long C, long long H, long long W, long long* s) {
  if (layout == B200DET_NHWC) { s[0] = H * W * C; s[1] = 1; s[2] = W * C; s[3] = C; }
  else { s[0] = C * H * W; s[1] = H * W; s[2] = W; s[3] = 1; }
}
inline unsigned grid_for(long long total) {
  const long long b = (total + kThreads - 1) / kThreads;
  return (unsigned)(b > (1LL << 20) ? (1LL << 20) : (b < 1 ? 1 : b));
}

template <typename T>
int run(bool forward, const void* in_or_grad, const int64_t* gs, int in_layout, const void* rois, void* out_or_gi,
        int32_t* mapping, int out_layout, long long N, long long C, long long H, long long W, long long K, int PH, int PW,
        double scale, cudaStream_t st) {
  using A = typename Io<T>::acc;
  const int C_out = (int)(C / (PH * PW));
  const long long total = K * C_out * PH * PW;
  long long is[4], os[4];
  strides_of(in_layout, C, H, W, is);
  strides_of(out_layout, C_out, PH, PW, os);
  if (forward) {
    ps_roi_pool_fwd_kernel<T><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)in_or_grad, is[0], is[1], is[2], is[3], (const A*)rois, (A)scale, (int)N, (int)H, (int)W, PH, PW,
        C_out, (T*)out_or_gi, mapping, os[0], os[1], os[2], os[3]);
  } else {
    B200DET_CUDA_OK(cudaMemsetAsync(out_or_gi, 0, (size_t)(N * C * H * W) * sizeof(T), st));
    if (total == 0) return 0;
    ps_roi_pool_bwd_kernel<T><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)in_or_grad, gs[0], gs[1], gs[2], gs[3], mapping, os[0], os[1], os[2], os[3], (const A*)rois,
        (A)scale, (int)N, (int)H, (int)W, PH, PW, C_out, (T*)out_or_gi, is[0], is[1], is[2], is[3]);
  }
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

// forward: in_layout / out_layout of input / (output, mapping).  backward: in_layout is grad_input's,
// out_layout the mapping's; `gs` the element strides of grad.
int ps_roi_pool(bool forward, const void* in_or_grad, const int64_t* gs, int in_layout, const void* rois,
                void* out_or_gi, int32_t* mapping, int out_layout, int dtype, long long N, long long C, long long H,
                long long W, long long K, int PH, int PW, double scale, cudaStream_t st) {
  switch (dtype) {
    case B200DET_F32: return run<float>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_F16: return run<__half>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_BF16: return run<__nv_bfloat16>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, st);
    case B200DET_F64: return run<double>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, st);
  }
  return fail(B200DET_ENOSUP, "ps_roi_pool: unknown dtype %d", dtype);
}

}  // namespace b200det
