// common.cuh -- shared device helpers for the sm_100a detection-head kernels.
//
// The RoI geometry below restates, operation for operation, what the
// reference computes per output element
// (torchvision/csrc/ops/cuda/roi_align_kernel.cu:89-133 and the CPU table
// builder torchvision/csrc/ops/cpu/roi_align_common.h:32-124).  All geometry
// arithmetic goes through Ar<A>::mul/add/sub/div, which map to the
// round-to-nearest intrinsics so that nvcc cannot contract a*b+c into an FMA:
// sample coordinates (and therefore the (int)y truncation) are bit-identical
// to the reference's CPU kernels, which are built without FMA contraction.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_detection_ops.h"

namespace b200det {

// ---- error plumbing (host) -------------------------------------------------
void set_last_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define B200DET_CUDA_OK(expr)                                              \
  do {                                                                     \
    cudaError_t _e = (expr);                                               \
    if (_e != cudaSuccess)                                                 \
      return ::b200det::fail((int)_e, "%s: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

// ---- I/O element traits ----------------------------------------------------
template <typename T> struct Io;
template <> struct Io<float> {
  using acc = float;
  static __device__ __forceinline__ float ld(const float* p) { return __ldg(p); }
  static __device__ __forceinline__ float to(float v) { return v; }
};
template <> struct Io<double> {
  using acc = double;
  static __device__ __forceinline__ double ld(const double* p) { return __ldg(p); }
  static __device__ __forceinline__ double to(double v) { return v; }
};
template <> struct Io<__half> {
  using acc = float;
  static __device__ __forceinline__ float ld(const __half* p) {
    return __half2float(__ldg(p));
  }
  static __device__ __forceinline__ __half to(float v) { return __float2half_rn(v); }
};
template <> struct Io<__nv_bfloat16> {
  using acc = float;
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) {
    return __bfloat162float(__ldg(p));
  }
  static __device__ __forceinline__ __nv_bfloat16 to(float v) {
    return __float2bfloat16_rn(v);
  }
};

// ---- contraction-free arithmetic ------------------------------------------
template <typename A> struct Ar;
template <> struct Ar<float> {
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float ceil_(float a) { return ceilf(a); }
};
template <> struct Ar<double> {
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
  static __device__ __forceinline__ double ceil_(double a) { return ceil(a); }
};

// ---- per-RoI geometry ------------------------------------------------------
template <typename A>
struct RoiGeom {
  int batch;        // (int)rois[k][0], truncation as roi_align_kernel.cu:90
  A start_h, start_w;
  A bin_h, bin_w;
  int grid_h, grid_w;  // samples per bin along h / w (may be <= 0: empty)
};

// roi_align_kernel.cu:89-118 / cpu/roi_align_kernel.cpp:31-58
template <typename A>
__device__ __forceinline__ RoiGeom<A> roi_geometry(
    const A* __restrict__ roi, A spatial_scale, int pooled_h, int pooled_w,
    int sampling_ratio, bool aligned) {
  RoiGeom<A> g;
  g.batch = (int)__ldg(roi + 0);
  const A offset = aligned ? (A)0.5 : (A)0.0;
  const A sw = Ar<A>::sub(Ar<A>::mul(__ldg(roi + 1), spatial_scale), offset);
  const A sh = Ar<A>::sub(Ar<A>::mul(__ldg(roi + 2), spatial_scale), offset);
  const A ew = Ar<A>::sub(Ar<A>::mul(__ldg(roi + 3), spatial_scale), offset);
  const A eh = Ar<A>::sub(Ar<A>::mul(__ldg(roi + 4), spatial_scale), offset);
  A roi_w = Ar<A>::sub(ew, sw);
  A roi_h = Ar<A>::sub(eh, sh);
  if (!aligned) {  // force malformed RoIs to be 1x1
    roi_w = roi_w > (A)1 ? roi_w : (A)1;   // max(roi_w, 1) incl. NaN -> 1 like fmax
    roi_h = roi_h > (A)1 ? roi_h : (A)1;
  }
  g.start_h = sh;
  g.start_w = sw;
  g.bin_h = Ar<A>::div(roi_h, (A)pooled_h);
  g.bin_w = Ar<A>::div(roi_w, (A)pooled_w);
  g.grid_h = sampling_ratio > 0 ? sampling_ratio
                                : (int)Ar<A>::ceil_(Ar<A>::div(roi_h, (A)pooled_h));
  g.grid_w = sampling_ratio > 0 ? sampling_ratio
                                : (int)Ar<A>::ceil_(Ar<A>::div(roi_w, (A)pooled_w));
  return g;
}

// One axis of a bilinear sample: indices of the two taps and their weights.
// valid == false reproduces the "empty" early-out of bilinear_interpolate
// (roi_align_kernel.cu:23-26) for this axis.
template <typename A>
struct AxisTap {
  int lo, hi;
  A wl;  // weight of tap `lo`  (hy / hx in the reference)
  A wh;  // weight of tap `hi`  (ly / lx)
  bool valid;
};

// coordinate of sample i of bin p: roi_align_kernel.cu:127-133
template <typename A>
__device__ __forceinline__ A sample_coord(A start, A bin, int p, int i, int grid) {
  const A a = Ar<A>::add(start, Ar<A>::mul((A)p, bin));
  const A b = Ar<A>::div(Ar<A>::mul((A)((float)i + .5f), bin), (A)grid);
  return Ar<A>::add(a, b);
}

// roi_align_kernel.cu:23-54 restricted to one axis (the function is separable)
template <typename A>
__device__ __forceinline__ AxisTap<A> axis_tap(A v, int size) {
  AxisTap<A> t;
  if (v < (A)-1.0 || v > (A)size) {
    t.lo = t.hi = 0;
    t.wl = t.wh = (A)0;
    t.valid = false;
    return t;
  }
  if (v <= (A)0) v = (A)0;
  int lo = (int)v;
  int hi;
  if (lo >= size - 1) {
    hi = lo = size - 1;
    v = (A)lo;
  } else {
    hi = lo + 1;
  }
  const A l = Ar<A>::sub(v, (A)lo);
  const A h = Ar<A>::sub((A)1., l);
  t.lo = lo;
  t.hi = hi;
  t.wl = h;
  t.wh = l;
  t.valid = true;
  return t;
}

// Up to B200DET_MAX_LEVELS feature maps addressed by element strides, so one
// kernel serves NCHW and NHWC, single level and FPN.
template <typename T>
struct LevelTable {
  T* ptr[B200DET_MAX_LEVELS];
  int height[B200DET_MAX_LEVELS];
  int width[B200DET_MAX_LEVELS];
  typename Io<T>::acc scale[B200DET_MAX_LEVELS];
  long long sn[B200DET_MAX_LEVELS];  // element strides of (n, c, h, w)
  long long sc[B200DET_MAX_LEVELS];
  long long sh[B200DET_MAX_LEVELS];
  long long sw[B200DET_MAX_LEVELS];
  int num_levels;
};

static inline int ceil_div_i64(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace b200det
