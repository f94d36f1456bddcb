// ps_roi_align_sm100.cu -- torchvision::ps_roi_align forward / backward.
// Replaces ps_roi_align_forward_kernel_impl / ps_roi_align_backward_kernel_impl
// (torchvision/csrc/ops/cuda/ps_roi_align_kernel.cu:64-291; CPU twin cpu/ps_roi_align_kernel.cpp:62-300).
// Position-sensitive RoIAlign: output channel c_out of bin (ph, pw) averages bilinear samples of input channel
// c_in = (c_out * PH + ph) * PW + pw.  The geometry is roi_align's with aligned = True and no size clamp
// ("do not use rounding"), so common.cuh's roi_geometry / sample_coord / axis_tap serve it unchanged; the one
// difference is count = grid_h * grid_w without the max(., 1) (an empty adaptive grid gives 0 / 0 = NaN, like
// the reference).  One thread per pooled element, stride-addressed (NCHW and channels_last).
#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kThreads = 256;

template <typename T>
__global__ void __launch_bounds__(kThreads) ps_roi_align_fwd_kernel(
    long long total, const T* __restrict__ input, long long sn, long long sc, long long sh, long long sw,
    const typename Io<T>::acc* __restrict__ rois, typename Io<T>::acc scale, int N, int H, int W, int PH, int PW,
    int C_out, int sampling_ratio, T* __restrict__ out, int32_t* __restrict__ mapping, long long os_k, long long os_c,
    long long os_h, long long os_w) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    const int pw = (int)(index % PW), ph = (int)((index / PW) % PH);
    const int c_out = (int)((index / PW / PH) % C_out);
    const long long k = index / PW / PH / C_out;
    const RoiGeom<A> g = roi_geometry<A>(rois + k * 5, scale, PH, PW, sampling_ratio, true);
    const int c_in = (c_out * PH + ph) * PW + pw;
    const A count = (A)(g.grid_h * g.grid_w);
    A acc = 0;
    if ((unsigned)g.batch < (unsigned)N) {
      const T* base = input + (long long)g.batch * sn + (long long)c_in * sc;
      for (int iy = 0; iy < g.grid_h; ++iy) {
        const AxisTap<A> ty = axis_tap<A>(sample_coord<A>(g.start_h, g.bin_h, ph, iy, g.grid_h), H);
        for (int ix = 0; ix < g.grid_w; ++ix) {
          const AxisTap<A> tx = axis_tap<A>(sample_coord<A>(g.start_w, g.bin_w, pw, ix, g.grid_w), W);
          if (ty.valid && tx.valid) {
            const A v1 = Io<T>::ld(base + ty.lo * sh + tx.lo * sw), v2 = Io<T>::ld(base + ty.lo * sh + tx.hi * sw);
            const A v3 = Io<T>::ld(base + ty.hi * sh + tx.lo * sw), v4 = Io<T>::ld(base + ty.hi * sh + tx.hi * sw);
            const A w1 = ty.wl * tx.wl, w2 = ty.wl * tx.wh, w3 = ty.wh * tx.wl, w4 = ty.wh * tx.wh;
            acc += w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4;
          }
        }
      }
    }
    const long long o = k * os_k + c_out * os_c + ph * os_h + pw * os_w;
    out[o] = Io<T>::to(acc / count);
    mapping[o] = c_in;
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) ps_roi_align_bwd_kernel(
    long long total, const T* __restrict__ grad, long long gs_k, long long gs_c, long long gs_h, long long gs_w,
    const int32_t* __restrict__ mapping, long long ms_k, long long ms_c, long long ms_h, long long ms_w,
    const typename Io<T>::acc* __restrict__ rois, typename Io<T>::acc scale, int N, int H, int W, int PH, int PW,
    int C_out, int sampling_ratio, T* __restrict__ grad_in, long long sn, long long sc, long long sh, long long sw) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    const int pw = (int)(index % PW), ph = (int)((index / PW) % PH);
    const int c_out = (int)((index / PW / PH) % C_out);
    const long long k = index / PW / PH / C_out;
    const RoiGeom<A> g = roi_geometry<A>(rois + k * 5, scale, PH, PW, sampling_ratio, true);
    if ((unsigned)g.batch >= (unsigned)N) continue;
    const int c_in = mapping[k * ms_k + c_out * ms_c + ph * ms_h + pw * ms_w];
    const A go = Io<T>::ld(grad + k * gs_k + c_out * gs_c + ph * gs_h + pw * gs_w);
    const A count = (A)(g.grid_h * g.grid_w);
    T* base = grad_in + (long long)g.batch * sn + (long long)c_in * sc;
    for (int iy = 0; iy < g.grid_h; ++iy) {
      const AxisTap<A> ty = axis_tap<A>(sample_coord<A>(g.start_h, g.bin_h, ph, iy, g.grid_h), H);
      for (int ix = 0; ix < g.grid_w; ++ix) {
        const AxisTap<A> tx = axis_tap<A>(sample_coord<A>(g.start_w, g.bin_w, pw, ix, g.grid_w), W);
        if (!(ty.valid && tx.valid)) continue;
        // g_i = grad * w_i / count (cpu/ps_roi_align_kernel.cpp:283-286)
        atomicAdd(base + ty.lo * sh + tx.lo * sw, Io<T>::to(go * (ty.wl * tx.wl) / count));
        atomicAdd(base + ty.lo * sh + tx.hi * sw, Io<T>::to(go * (ty.wl * tx.wh) / count));
        atomicAdd(base + ty.hi * sh + tx.lo * sw, Io<T>::to(go * (ty.wh * tx.wl) / count));
        atomicAdd(base + ty.hi * sh + tx.hi * sw, Io<T>::to(go * (ty.wh * tx.wh) / count));
      }
    }
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
        double scale, int sr, cudaStream_t st) {
  using A = typename Io<T>::acc;
  const int C_out = (int)(C / (PH * PW));
  const long long total = K * C_out * PH * PW;
  long long is[4], os[4];
  strides_of(in_layout, C, H, W, is);
  strides_of(out_layout, C_out, PH, PW, os);
  if (forward) {
    ps_roi_align_fwd_kernel<T><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)in_or_grad, is[0], is[1], is[2], is[3], (const A*)rois, (A)scale, (int)N, (int)H, (int)W, PH, PW,
        C_out, sr, (T*)out_or_gi, mapping, os[0], os[1], os[2], os[3]);
  } else {
    B200DET_CUDA_OK(cudaMemsetAsync(out_or_gi, 0, (size_t)(N * C * H * W) * sizeof(T), st));
    if (total == 0) return 0;
    ps_roi_align_bwd_kernel<T><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)in_or_grad, gs[0], gs[1], gs[2], gs[3], mapping, os[0], os[1], os[2], os[3], (const A*)rois,
        (A)scale, (int)N, (int)H, (int)W, PH, PW, C_out, sr, (T*)out_or_gi, is[0], is[1], is[2], is[3]);
  }
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

// same argument roles as ps_roi_pool (ps_roi_pool_sm100.cu)
int ps_roi_align(bool forward, const void* in_or_grad, const int64_t* gs, int in_layout, const void* rois,
                 void* out_or_gi, int32_t* mapping, int out_layout, int dtype, long long N, long long C, long long H,
                 long long W, long long K, int PH, int PW, double scale, int sr, cudaStream_t st) {
  switch (dtype) {
    case B200DET_F32: return run<float>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, sr, st);
    case B200DET_F16: return run<__half>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, sr, st);
    case B200DET_BF16: return run<__nv_bfloat16>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, sr, st);
    case B200DET_F64: return run<double>(forward, in_or_grad, gs, in_layout, rois, out_or_gi, mapping, out_layout, N, C, H, W, K, PH, PW, scale, sr, st);
  }
  return fail(B200DET_ENOSUP, "ps_roi_align: unknown dtype %d", dtype);
}

}  // namespace b200det
