// roi_align_generic.cu -- stride-addressed RoIAlign forward/backward for every
// case the sm_100a kernels (roi_align_{fwd,bwd}_sm100.cu: sampling_ratio 1 / 2, C % 128 == 0;
// roi_align_anygrid_sm100.cu: any sampling grid, C % 4 == 0, both channels_last) do not take:
// fp64, NCHW feature maps, channel counts below 32 or not a multiple of 4, gradients
// that are not channel-contiguous.
//
// One thread per pooled output element, like the reference
// (torchvision/csrc/ops/cuda/roi_align_kernel.cu:68-143, :204-332), but
//  * the flat index is decomposed channel-fastest when the feature map is
//    NHWC, so a warp reads 32 consecutive channels of one tap (coalesced)
//    instead of 32 scattered pixels of one plane;
//  * every element of the output is written (zeros included), so no memset
//    pass precedes the kernel;
//  * one launch serves all FPN levels: the level of RoI k comes from levels[k]
//    and indexes a by-value table of feature pointers.
#include "common.cuh"
#include "kernels.h"

namespace b200det {

namespace {

constexpr int kThreads = 256;

template <typename T, bool CH_FASTEST>
__global__ void __launch_bounds__(kThreads) roi_align_fwd_generic_kernel(
    long long total, LevelTable<T> lv, const typename Io<T>::acc* __restrict__ rois,
    const int32_t* __restrict__ levels, T* __restrict__ out, long long os_k,
    long long os_c, long long os_h, long long os_w, int channels, int pooled_h,
    int pooled_w, int sampling_ratio, bool aligned) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    int pw, ph, c;
    long long k;
    if (CH_FASTEST) {
      c = (int)(index % channels);
      pw = (int)((index / channels) % pooled_w);
      ph = (int)((index / channels / pooled_w) % pooled_h);
      k = index / channels / pooled_w / pooled_h;
    } else {
      pw = (int)(index % pooled_w);
      ph = (int)((index / pooled_w) % pooled_h);
      c = (int)((index / pooled_w / pooled_h) % channels);
      k = index / pooled_w / pooled_h / channels;
    }
    const int l = levels ? levels[k] : 0;
    const int height = lv.height[l], width = lv.width[l];
    const RoiGeom<A> g = roi_geometry<A>(rois + k * 5, lv.scale[l], pooled_h, pooled_w,
                                            sampling_ratio, aligned);
    const T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + (long long)c * lv.sc[l];
    const long long sh = lv.sh[l], sw = lv.sw[l];

    // roi_align_kernel.cu:122: when the grid is empty the output is zero
    const int cnt = g.grid_h * g.grid_w;
    const A count = (A)(cnt > 1 ? cnt : 1);
    A acc = 0;
    for (int iy = 0; iy < g.grid_h; ++iy) {
      const AxisTap<A> ty = axis_tap<A>(sample_coord<A>(g.start_h, g.bin_h, ph, iy, g.grid_h), height);
      for (int ix = 0; ix < g.grid_w; ++ix) {
        const AxisTap<A> tx = axis_tap<A>(sample_coord<A>(g.start_w, g.bin_w, pw, ix, g.grid_w), width);
        if (ty.valid && tx.valid) {
          const A v1 = Io<T>::ld(base + ty.lo * sh + tx.lo * sw);
          const A v2 = Io<T>::ld(base + ty.lo * sh + tx.hi * sw);
          const A v3 = Io<T>::ld(base + ty.hi * sh + tx.lo * sw);
          const A v4 = Io<T>::ld(base + ty.hi * sh + tx.hi * sw);
          const A w1 = ty.wl * tx.wl, w2 = ty.wl * tx.wh, w3 = ty.wh * tx.wl, w4 = ty.wh * tx.wh;
          acc += w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4;
        }
      }
    }
    out[k * os_k + c * os_c + ph * os_h + pw * os_w] = Io<T>::to(acc / count);
  }
}

template <typename T> __device__ __forceinline__ void red_add(T* p, typename Io<T>::acc v);
template <> __device__ __forceinline__ void red_add<float>(float* p, float v) { atomicAdd(p, v); }
template <> __device__ __forceinline__ void red_add<double>(double* p, double v) { atomicAdd(p, v); }
template <> __device__ __forceinline__ void red_add<__half>(__half* p, float v) {
  atomicAdd(p, __float2half_rn(v));
}
template <> __device__ __forceinline__ void red_add<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  atomicAdd(p, __float2bfloat16_rn(v));
}

template <typename T, bool CH_FASTEST>
__global__ void __launch_bounds__(kThreads) roi_align_bwd_generic_kernel(
    long long total, const T* __restrict__ grad, long long gs_k, long long gs_c,
    long long gs_h, long long gs_w, LevelTable<T> lv, const typename Io<T>::acc* __restrict__ rois,
    const int32_t* __restrict__ levels, int channels, int pooled_h, int pooled_w,
    int sampling_ratio, bool aligned) {
  using A = typename Io<T>::acc;
  for (long long index = blockIdx.x * (long long)kThreads + threadIdx.x; index < total;
       index += (long long)gridDim.x * kThreads) {
    int pw, ph, c;
    long long k;
    if (CH_FASTEST) {
      c = (int)(index % channels);
      pw = (int)((index / channels) % pooled_w);
      ph = (int)((index / channels / pooled_w) % pooled_h);
      k = index / channels / pooled_w / pooled_h;
    } else {
      pw = (int)(index % pooled_w);
      ph = (int)((index / pooled_w) % pooled_h);
      c = (int)((index / pooled_w / pooled_h) % channels);
      k = index / pooled_w / pooled_h / channels;
    }
    const int l = levels ? levels[k] : 0;
    const int height = lv.height[l], width = lv.width[l];
    const RoiGeom<A> g = roi_geometry<A>(rois + k * 5, lv.scale[l], pooled_h, pooled_w,
                                            sampling_ratio, aligned);
    T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + (long long)c * lv.sc[l];
    const long long sh = lv.sh[l], sw = lv.sw[l];
    const A go = Io<T>::ld(grad + k * gs_k + c * gs_c + ph * gs_h + pw * gs_w);
    // roi_align_kernel.cu:266: no max(...,1) here; an empty grid adds nothing
    const A count = (A)(g.grid_h * g.grid_w);
    for (int iy = 0; iy < g.grid_h; ++iy) {
      const AxisTap<A> ty = axis_tap<A>(sample_coord<A>(g.start_h, g.bin_h, ph, iy, g.grid_h), height);
      for (int ix = 0; ix < g.grid_w; ++ix) {
        const AxisTap<A> tx = axis_tap<A>(sample_coord<A>(g.start_w, g.bin_w, pw, ix, g.grid_w), width);
        if (ty.valid && tx.valid) {
          // g_k = grad * w_k / count, roi_align_kernel.cu:298-301
          const A w1 = ty.wl * tx.wl, w2 = ty.wl * tx.wh, w3 = ty.wh * tx.wl, w4 = ty.wh * tx.wh;
          red_add<T>(base + ty.lo * sh + tx.lo * sw, Ar<A>::div(Ar<A>::mul(go, w1), count));
          red_add<T>(base + ty.lo * sh + tx.hi * sw, Ar<A>::div(Ar<A>::mul(go, w2), count));
          red_add<T>(base + ty.hi * sh + tx.lo * sw, Ar<A>::div(Ar<A>::mul(go, w3), count));
          red_add<T>(base + ty.hi * sh + tx.hi * sw, Ar<A>::div(Ar<A>::mul(go, w4), count));
        }
      }
    }
  }
}

// poolers.py:73-84.  Arithmetic in the boxes' dtype (fp32, or fp64 for fp64 boxes) in the order torch evaluates it
// on the reference's CPU path: area = (x2-x1)*(y2-y1) (box_area after _upcast, boxes.py:278-295), s = sqrt(area),
// floor(lvl0 + log2(s / s0) + eps) with a true division, clamp, minus k_min.
template <typename R> struct LvlMath;
template <> struct LvlMath<float> {
  static __device__ __forceinline__ float sqrt_(float v) { return sqrtf(v); }
  static __device__ __forceinline__ float log2_(float v) { return log2f(v); }
  static __device__ __forceinline__ float floor_(float v) { return floorf(v); }
};
template <> struct LvlMath<double> {
  static __device__ __forceinline__ double sqrt_(double v) { return sqrt(v); }
  static __device__ __forceinline__ double log2_(double v) { return log2(v); }
  static __device__ __forceinline__ double floor_(double v) { return floor(v); }
};
template <typename R>
__global__ void fpn_level_map_kernel(const R* __restrict__ rois, long long K, int k_min, int k_max, R s0, R lvl0,
                                     R eps, int32_t* __restrict__ levels) {
  const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (k >= K) return;
  const R* r = rois + k * 5;
  const R x1 = __ldg(r + 1), y1 = __ldg(r + 2), x2 = __ldg(r + 3), y2 = __ldg(r + 4);
  const R area = Ar<R>::mul(Ar<R>::sub(x2, x1), Ar<R>::sub(y2, y1));
  const R s = LvlMath<R>::sqrt_(area);
  R t = LvlMath<R>::floor_(Ar<R>::add(Ar<R>::add(lvl0, LvlMath<R>::log2_(Ar<R>::div(s, s0))), eps));
  // clamp; a NaN (negative area) lands on k_min where torch's int cast is undefined
  t = t >= (R)k_min ? t : (R)k_min;
  t = t <= (R)k_max ? t : (R)k_max;
  levels[k] = (int32_t)t - k_min;
}
template <typename T>
LevelTable<T> make_table(const LevelArgs& a) {
  LevelTable<T> t;
  t.num_levels = a.num_levels;
  for (int l = 0; l < a.num_levels; ++l) {
    t.ptr[l] = (T*)a.ptr[l];
    t.height[l] = (int)a.height[l];
    t.width[l] = (int)a.width[l];
    t.scale[l] = (typename Io<T>::acc)a.scale[l];
    const long long C = a.channels, H = a.height[l], W = a.width[l];
    if (a.layout == B200DET_NHWC) {
      t.sn[l] = H * W * C; t.sc[l] = 1; t.sh[l] = W * C; t.sw[l] = C;
    } else {
      t.sn[l] = C * H * W; t.sc[l] = H * W; t.sh[l] = W; t.sw[l] = 1;
    }
  }
  return t;
}

int grid_for(long long total) {
  // enough CTAs for 148 SMs x 8 resident blocks, grid-stride beyond that
  long long blocks = (total + kThreads - 1) / kThreads;
  const long long cap = 148LL * 32;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

template <typename T>
int fwd_generic_t(const LevelArgs& lv, const void* rois, const int32_t* levels, void* out,
                  int out_layout, long long K, int PH, int PW, int sr, bool aligned,
                  cudaStream_t st) {
  const long long C = lv.channels;
  const long long total = K * C * PH * PW;
  if (total == 0) return 0;
  long long os_k = C * PH * PW, os_c, os_h, os_w;
  if (out_layout == B200DET_NHWC) { os_c = 1; os_h = (long long)PW * C; os_w = C; }
  else { os_c = (long long)PH * PW; os_h = PW; os_w = 1; }
  auto table = make_table<T>(lv);
  const bool chf = lv.layout == B200DET_NHWC;
  if (chf)
    roi_align_fwd_generic_kernel<T, true><<<grid_for(total), kThreads, 0, st>>>(
        total, table, (const typename Io<T>::acc*)rois, levels, (T*)out, os_k, os_c, os_h, os_w, (int)C, PH, PW, sr, aligned);
  else
    roi_align_fwd_generic_kernel<T, false><<<grid_for(total), kThreads, 0, st>>>(
        total, table, (const typename Io<T>::acc*)rois, levels, (T*)out, os_k, os_c, os_h, os_w, (int)C, PH, PW, sr, aligned);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int bwd_generic_t(const LevelArgs& lv, const void* grad, const int64_t* gs, const void* rois,
                  const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned,
                  cudaStream_t st) {
  const long long C = lv.channels;
  const long long total = K * C * PH * PW;
  if (total == 0) return 0;
  auto table = make_table<T>(lv);
  // thread order follows the faster-moving side of the scatter target
  const bool chf = lv.layout == B200DET_NHWC;
  if (chf)
    roi_align_bwd_generic_kernel<T, true><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)grad, gs[0], gs[1], gs[2], gs[3], table, (const typename Io<T>::acc*)rois, levels, (int)C, PH, PW, sr, aligned);
  else
    roi_align_bwd_generic_kernel<T, false><<<grid_for(total), kThreads, 0, st>>>(
        total, (const T*)grad, gs[0], gs[1], gs[2], gs[3], table, (const typename Io<T>::acc*)rois, levels, (int)C, PH, PW, sr, aligned);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int roi_align_fwd_generic(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels,
                          void* out, int out_layout, long long K, int PH, int PW, int sr,
                          bool aligned, cudaStream_t st) {
  switch (dtype) {
    case B200DET_F32: return fwd_generic_t<float>(lv, rois, levels, out, out_layout, K, PH, PW, sr, aligned, st);
    case B200DET_F16: return fwd_generic_t<__half>(lv, rois, levels, out, out_layout, K, PH, PW, sr, aligned, st);
    case B200DET_BF16: return fwd_generic_t<__nv_bfloat16>(lv, rois, levels, out, out_layout, K, PH, PW, sr, aligned, st);
    case B200DET_F64: return fwd_generic_t<double>(lv, rois, levels, out, out_layout, K, PH, PW, sr, aligned, st);
  }
  return fail(B200DET_ENOSUP, "roi_align forward: unknown dtype %d", dtype);
}

int roi_align_bwd_generic(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs,
                          const void* rois, const int32_t* levels, long long K, int PH, int PW,
                          int sr, bool aligned, cudaStream_t st) {
  switch (dtype) {
    case B200DET_F32: return bwd_generic_t<float>(lv, grad, gs, rois, levels, K, PH, PW, sr, aligned, st);
    case B200DET_F16: return bwd_generic_t<__half>(lv, grad, gs, rois, levels, K, PH, PW, sr, aligned, st);
    case B200DET_BF16: return bwd_generic_t<__nv_bfloat16>(lv, grad, gs, rois, levels, K, PH, PW, sr, aligned, st);
    case B200DET_F64: return bwd_generic_t<double>(lv, grad, gs, rois, levels, K, PH, PW, sr, aligned, st);
  }
  return fail(B200DET_ENOSUP, "roi_align backward: unknown dtype %d", dtype);
}

int fpn_level_map(const void* rois, int dtype, long long K, int k_min, int k_max, int s0, int lvl0,
                  double eps, int32_t* levels, cudaStream_t st) {
  if (K == 0) return 0;
  const int th = 256;
  const int bl = (int)((K + th - 1) / th);
  if (dtype == B200DET_F64)
    fpn_level_map_kernel<double><<<bl, th, 0, st>>>((const double*)rois, K, k_min, k_max, (double)s0, (double)lvl0,
                                                    eps, levels);
  else
    fpn_level_map_kernel<float><<<bl, th, 0, st>>>((const float*)rois, K, k_min, k_max, (float)s0, (float)lvl0,
                                                   (float)eps, levels);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b200det
