// roi_align_sm100.cu -- the sm_100a RoIAlign kernels for the detection-model
// case: channels_last feature maps, C a multiple of 128, sampling_ratio 1 or 2,
// pooled_width * sampling_ratio <= 32.  Everything else goes to
// roi_align_generic.cu.
//
// bilinear_interpolate (cuda/roi_align_kernel.cu:14-66) is separable: the
// weight of pixel (r, c) for a sample is wy(r) * wx(c).  Summed over the
// sampling grid of a bin that gives, for a pooled row ph,
//     out[ph][pw] = sum_c Wx[pw][c] * ( sum_r Wy[ph][r] * in[r][c] ).
//
// FORWARD.  CTA = one (RoI, 128-channel tile), 7 warps; a warp takes one pooled
// row at a time; a lane owns 4 consecutive channels, so every feature tap is one
// 16-byte (fp32) / 8-byte (bf16, fp16) load and a warp covers 512/256 contiguous
// bytes of one NHWC pixel.
//   * Warp 0 builds the x table once per CTA (it is the same for every pooled
//     row): per x sample its two tap weights and how many NEW feature columns
//     the sample needs (0, 1 or 2), plus the ordered list of those columns
//     ("column events").  It also prefetches into L2, with
//     cp.async.bulk.prefetch.L2, the footprint rows of a RoI a few hundred CTAs
//     ahead, so first-touch HBM latency is not paid by demand loads.
//   * Each warp derives the <= 2*SR feature rows of its pooled row with their
//     merged weights (1/count folded in; exact, count is a power of two).
//   * It then sweeps the x samples left to right keeping the vertical sums of
//     the two columns under the current sample in registers.  Each (row,
//     column) of the footprint is loaded once per pooled row; the loads of column
//     event e+1 are issued before the FMAs of event e (two register buffers).
//   * fp32 math is issued as packed FFMA2 (fma.rn.f32x2, new on sm_100): half
//     the FMA instructions of a scalar formulation, same rounding.
//
// BACKWARD is the transpose with a warp per PAIR of pooled rows (fewer atomics:
// the rows two adjacent pooled rows share are reduced once): per column it
// accumulates sum_pw Wx[pw][c] * grad[ph][pw] in registers and issues ONE vector
// red.global.add per (row, column) of the footprint (4 channels per lane, 128
// channels per warp instruction) instead of 4 scalar atomics per sample.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kFwdWarps = 7;
constexpr int kMaxSX = 32;  // x samples per pooled row handled by one warp
constexpr unsigned kFull = 0xffffffffu;

// ---- 4-channel vector I/O ---------------------------------------------------
template <typename T> struct V4;
template <> struct V4<float> {
  using raw = float4;
  static __device__ __forceinline__ raw ldraw(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ raw ldraw_stream(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ float4 cvt(raw r) { return r; }
  static __device__ __forceinline__ void st(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
  static __device__ __forceinline__ void red(float* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
};
template <> struct V4<__half> {
  using raw = uint2;
  static __device__ __forceinline__ raw ldraw(const __half* p) { return __ldg(reinterpret_cast<const uint2*>(p)); }
  static __device__ __forceinline__ raw ldraw_stream(const __half* p) { return __ldcs(reinterpret_cast<const uint2*>(p)); }
  static __device__ __forceinline__ float4 cvt(raw u) {
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
  }
  static __device__ __forceinline__ uint2 pack(float4 v) {
    const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<const unsigned*>(&a);
    u.y = *reinterpret_cast<const unsigned*>(&b);
    return u;
  }
  static __device__ __forceinline__ void st(__half* p, float4 v) { __stcs(reinterpret_cast<uint2*>(p), pack(v)); }
  static __device__ __forceinline__ void red(__half* p, float4 v) {
    const uint2 u = pack(v);
    asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(p), "r"(u.x), "r"(u.y) : "memory");
  }
};
template <> struct V4<__nv_bfloat16> {
  using raw = uint2;
  static __device__ __forceinline__ raw ldraw(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint2*>(p)); }
  static __device__ __forceinline__ raw ldraw_stream(const __nv_bfloat16* p) { return __ldcs(reinterpret_cast<const uint2*>(p)); }
  static __device__ __forceinline__ float4 cvt(raw u) {
    // bf16 -> fp32 is a 16-bit shift
    return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u),
                       __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
  }
  static __device__ __forceinline__ uint2 pack(float4 v) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<const unsigned*>(&a);
    u.y = *reinterpret_cast<const unsigned*>(&b);
    return u;
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float4 v) { __stcs(reinterpret_cast<uint2*>(p), pack(v)); }
  static __device__ __forceinline__ void red(__nv_bfloat16* p, float4 v) {
    const uint2 u = pack(v);
    asm volatile("red.global.add.noftz.v2.bf16x2 [%0], {%1, %2};" ::"l"(p), "r"(u.x), "r"(u.y) : "memory");
  }
};

__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// a += w * v on 4 channels as two packed f32x2 FMAs (FFMA2 on sm_100; each half is
// an IEEE fma.rn exactly like fmaf)
__device__ __forceinline__ void fma4(float4& a, float w, const float4& v) {
  unsigned long long ww, v0, v1, a0, a1;
  asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "f"(w));
  asm("mov.b64 %0, {%1, %2};" : "=l"(v0) : "f"(v.x), "f"(v.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(v1) : "f"(v.z), "f"(v.w));
  asm("mov.b64 %0, {%1, %2};" : "=l"(a0) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(a1) : "f"(a.z), "f"(a.w));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a0) : "l"(ww), "l"(v0));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a1) : "l"(ww), "l"(v1));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(a0));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a.z), "=f"(a.w) : "l"(a1));
}

// ---- x table (one warp, once per CTA) ------------------------------------------
// Lane j owns x sample j (pooled column j / SR, sample j % SR).  The sweep keeps
// a window of two feature columns (lo, hi = lo + 1, clamped like
// roi_align_kernel.cu:34-46).  A sample reaches its taps by leaving the window
// unchanged, sliding it right by one column (one new column) or reloading both.
struct XTable {
  float2 w[kMaxSX];        // (hx, lx): weights of the lo / hi column; 0 for an empty sample
  int colofs[2 * kMaxSX];  // element offsets (x * C) of the column events, in sweep order
  unsigned m_slide;        // bit j: sample j slides the window (1 new column)
  unsigned m_reload;       // bit j: sample j reloads the window (2 new columns)
  unsigned m_valid;        // bit j: sample j is not empty (roi_align_kernel.cu:23-26)
  int ncol;
};

template <int SR>
__device__ __forceinline__ void build_xtable(const RoiGeom<float>& g, int W, int C, int PW, int lane, XTable& xt) {
  const int SX = PW * SR;
  const int j = lane < SX ? lane : SX - 1;
  const float x = sample_coord<float>(g.start_w, g.bin_w, j / SR, j % SR, SR);
  const AxisTap<float> t = axis_tap<float>(x, W);
  // an empty sample parks the window at the edge it fell off (its weights are zero)
  const int xl = t.valid ? t.lo : (x < -1.0f ? 0 : W - 1);
  const int xh = min(xl + 1, W - 1);
  int pxl = __shfl_up_sync(kFull, xl, 1), pxh = __shfl_up_sync(kFull, xh, 1);
  if (lane == 0) { pxl = -2; pxh = -2; }
  int nnew = (xl == pxl) ? 0 : ((xl == pxh) ? 1 : 2);
  if (lane >= SX) nnew = 0;
  int incl = nnew;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += v;
  }
  const int base = incl - nnew;
  const unsigned m1 = __ballot_sync(kFull, nnew == 1), m2 = __ballot_sync(kFull, nnew == 2);
  const unsigned mv = __ballot_sync(kFull, t.valid && lane < SX);
  if (lane < SX) {
    xt.w[lane] = make_float2(t.wl, t.wh);
    if (nnew == 2) {
      xt.colofs[base] = xl * C;
      xt.colofs[base + 1] = xh * C;
    } else if (nnew == 1) {
      xt.colofs[base] = xh * C;
    }
  }
  if (lane == 31) {
    xt.ncol = incl;
    xt.m_slide = m1;
    xt.m_reload = m2;
    xt.m_valid = mv;
  }
}

// ---- L2 prefetch of a RoI footprint ----------------------------------------------------
// The rows of a RoI's footprint are contiguous NHWC segments (columns x channels).
// Lane r asks the L2 to fetch row r of the footprint; no registers or shared
// memory are tied up and nothing waits on it.
template <typename T>
__device__ __forceinline__ void prefetch_footprint(const LevelTable<T>& lv, const float* __restrict__ rois,
                                                   const int32_t* __restrict__ levels, long long k, int nimg, int PH,
                                                   int PW, int SR, bool aligned, int C, int lane) {
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  if (g.batch < 0 || g.batch >= nimg) return;
  const float y0 = g.start_h, y1 = g.start_h + g.bin_h * (float)PH;
  const float x0 = g.start_w, x1 = g.start_w + g.bin_w * (float)PW;
  if (!(y1 >= -1.f && y0 <= (float)H && x1 >= -1.f && x0 <= (float)W)) return;  // also drops NaNs
  const int ya = max((int)floorf(fmaxf(y0, 0.f)), 0), yb = min((int)floorf(fminf(y1, (float)H)) + 1, H - 1);
  const int xa = max((int)floorf(fmaxf(x0, 0.f)), 0), xb = min((int)floorf(fminf(x1, (float)W)) + 1, W - 1);
  if (xb < xa) return;
  const unsigned bytes = (unsigned)((xb - xa + 1) * C * (int)sizeof(T));  // multiple of 16: C % 128 == 0
  const T* img = lv.ptr[l] + (long long)g.batch * lv.sn[l];
  for (int y = ya + lane; y <= yb; y += 32) {
    const T* p = img + ((long long)y * W + xa) * C;
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
  }
}

// ---- forward ---------------------------------------------------------------------------
// Feature rows of ONE pooled row: 2*SR taps (lo/hi of each y sample), duplicates
// merged into the first occurrence.  Every lane computes the same values, so no
// shuffles or shared memory are involved.  off < 0 marks an unused slot.
template <int SR>
struct RowTaps {
  int off[2 * SR];    // row * W * C (element offset inside the image), or -1
  float w[2 * SR];    // merged weight, 1/count folded in
};

template <int SR>
__device__ __forceinline__ RowTaps<SR> row_taps(const RoiGeom<float>& g, int H, int row_stride, int ph) {
  RowTaps<SR> rt;
  int row[2 * SR];
#pragma unroll
  for (int iy = 0; iy < SR; ++iy) {
    const float y = sample_coord<float>(g.start_h, g.bin_h, ph, iy, SR);
    const AxisTap<float> t = axis_tap<float>(y, H);
    row[2 * iy] = t.valid ? t.lo : -1;
    row[2 * iy + 1] = t.valid ? t.hi : -1;
    rt.w[2 * iy] = t.wl * (1.0f / (float)(SR * SR));
    rt.w[2 * iy + 1] = t.wh * (1.0f / (float)(SR * SR));
  }
#pragma unroll
  for (int i = 1; i < 2 * SR; ++i) {
    bool merged = false;
#pragma unroll
    for (int j = 0; j < i; ++j) {
      if (!merged && row[i] >= 0 && row[i] == row[j]) {
        rt.w[j] += rt.w[i];
        merged = true;
      }
    }
    if (merged) row[i] = -1;
  }
#pragma unroll
  for (int i = 0; i < 2 * SR; ++i) rt.off[i] = row[i] >= 0 ? row[i] * row_stride : -1;
  return rt;
}

template <typename T, int SR>
__device__ __forceinline__ void issue_column(typename V4<T>::raw (&buf)[2 * SR], const RowTaps<SR>& rt,
                                             const T* __restrict__ base, int co) {
#pragma unroll
  for (int r = 0; r < 2 * SR; ++r)
    if (rt.off[r] >= 0) buf[r] = V4<T>::ldraw(base + (rt.off[r] + co));
}

template <typename T, int SR>
__device__ __forceinline__ float4 vertical(const typename V4<T>::raw (&buf)[2 * SR], const RowTaps<SR>& rt) {
  float4 v = f4zero();
#pragma unroll
  for (int r = 0; r < 2 * SR; ++r)
    if (rt.off[r] >= 0) fma4(v, rt.w[r], V4<T>::cvt(buf[r]));
  return v;
}

template <typename T, int SR>
__global__ void __launch_bounds__(kFwdWarps * 32, 4) roi_align_fwd_sm100_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels, T* __restrict__ out,
    int ntasks, int nimg, int C, int PH, int PW, bool aligned, int CT, int pf_dist) {
  __shared__ XTable xt;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int task = blockIdx.x;
  const int ct = task % CT;
  const long long k = task / CT;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  if (warp == 0) {
    build_xtable<SR>(g, W, C, PW, lane, xt);
  } else if (warp == 1 && ct == 0 && pf_dist > 0) {
    const long long kp = k + pf_dist;
    if (kp * CT < ntasks) prefetch_footprint<T>(lv, rois, levels, kp, nimg, PH, PW, SR, aligned, C, lane);
  }
  __syncthreads();
  const int ncol = xt.ncol;
  const unsigned m_slide = xt.m_slide, m_reload = xt.m_reload, m_valid = xt.m_valid;
  const int ch = ct * 128 + lane * 4;
  const T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
  T* obase = out + k * PH * PW * (long long)C + ch;

  for (int ph = warp; ph < PH; ph += kFwdWarps) {
    const RowTaps<SR> rt = row_taps<SR>(g, H, W * C, ph);
    T* o = obase + (long long)ph * PW * C;

    typename V4<T>::raw A[2 * SR], B[2 * SR];
    int e = 0;
    bool useA = true;
    if (ncol > 0) issue_column<T, SR>(A, rt, base, xt.colofs[0]);
    // vertical sum of column event e; the loads of event e+1 go out first
    auto consume = [&]() -> float4 {
      const int en = e + 1;
      const bool more = en < ncol;
      const int co_n = more ? xt.colofs[en] : 0;
      float4 v;
      if (useA) {
        if (more) issue_column<T, SR>(B, rt, base, co_n);
        v = vertical<T, SR>(A, rt);
      } else {
        if (more) issue_column<T, SR>(A, rt, base, co_n);
        v = vertical<T, SR>(B, rt);
      }
      useA = !useA;
      e = en;
      return v;
    };

    float4 lo = f4zero(), hi = f4zero();
    for (int pw = 0; pw < PW; ++pw) {
      float4 a = f4zero();
#pragma unroll
      for (int s = 0; s < SR; ++s) {
        const int j = pw * SR + s;
        const unsigned bit = 1u << j;
        if (m_reload & bit) {
          lo = consume();
          hi = consume();
        } else if (m_slide & bit) {
          lo = hi;
          hi = consume();
        }
        if (m_valid & bit) {
          const float2 w = xt.w[j];
          fma4(a, w.x, lo);
          fma4(a, w.y, hi);
        }
      }
      V4<T>::st(o + (long long)pw * C, a);
    }
  }
}

// ---- backward --------------------------------------------------------------------------
// Work item = one warp per (RoI, 128-channel tile, pair of pooled rows); the warps
// of a CTA are independent.
constexpr int kBwdWarps = 4;
constexpr int kMaxRows = 8;  // 2 pooled rows x 2 samples x 2 taps

struct __align__(16) BXEnt { int xl; int step; float hx; float lx; };
struct __align__(16) BYEnt { int rowoff; float w0; float w1; int pad; };

// Fills xt[0..SX) and yt[0..nrows) for (RoI geometry g, pooled rows 2*rp, 2*rp+1).
// step: 0 = same window as the previous sample, 1 = slid right by one column,
// 2 = new window.  Returns nrows.  All lanes of the warp must call it.
template <int SR>
__device__ __forceinline__ int build_bwd_tables(const RoiGeom<float>& g, int H, int W, int C, int PH, int PW,
                                                int rp, int lane, BXEnt* xt, BYEnt* yt) {
  const int SX = PW * SR;
  {
    const int j = lane < SX ? lane : SX - 1;
    const float x = sample_coord<float>(g.start_w, g.bin_w, j / SR, j % SR, SR);
    const AxisTap<float> t = axis_tap<float>(x, W);
    const int xl = t.valid ? t.lo : (x < -1.0f ? 0 : W - 1);
    const int prev = __shfl_up_sync(kFull, xl, 1);
    const int d = xl - prev;
    BXEnt e;
    e.xl = xl;
    e.step = (lane == 0) ? 2 : (d == 0 ? 0 : (d == 1 ? 1 : 2));
    e.hx = t.wl;  // 0 when !valid
    e.lx = t.wh;
    if (lane < SX) xt[lane] = e;
  }
  constexpr int nT = 4 * SR;
  int row = -1;
  float w = 0.f;
  if (lane < nT) {
    const int m = lane / (2 * SR);
    const int iy = (lane >> 1) % SR;
    const int ph = 2 * rp + m;
    if (ph < PH) {
      const float y = sample_coord<float>(g.start_h, g.bin_h, ph, iy, SR);
      const AxisTap<float> t = axis_tap<float>(y, H);
      if (t.valid) {
        row = (lane & 1) ? t.hi : t.lo;
        w = ((lane & 1) ? t.wh : t.wl) * (1.0f / (float)(SR * SR));
      }
    }
  }
  // merge taps that hit the same feature row; the first tap of a row leads
  bool leader = row >= 0;
  float w0 = 0.f, w1 = 0.f;
#pragma unroll
  for (int t = 0; t < nT; ++t) {
    const int rr = __shfl_sync(kFull, row, t);
    const float ww = __shfl_sync(kFull, w, t);
    if (rr == row && row >= 0) {
      if (t < lane) leader = false;
      if (t < 2 * SR) w0 += ww; else w1 += ww;
    }
  }
  const unsigned ball = __ballot_sync(kFull, leader);
  if (leader) {
    BYEnt e;
    e.rowoff = row * W * C;
    e.w0 = w0;
    e.w1 = w1;
    e.pad = 0;
    yt[__popc(ball & ((1u << lane) - 1u))] = e;
  }
  __syncwarp();
  return __popc(ball);
}

template <typename T>
__device__ __forceinline__ void flush_column(T* __restrict__ base, int coloff, const BYEnt* yt, int nrows,
                                             const float4& c0, const float4& c1) {
#pragma unroll 4
  for (int r = 0; r < nrows; ++r) {
    const BYEnt e = yt[r];
    float4 v;
    v.x = fmaf(e.w0, c0.x, e.w1 * c1.x);
    v.y = fmaf(e.w0, c0.y, e.w1 * c1.y);
    v.z = fmaf(e.w0, c0.z, e.w1 * c1.z);
    v.w = fmaf(e.w0, c0.w, e.w1 * c1.w);
    V4<T>::red(base + e.rowoff + coloff, v);
  }
}

template <typename T>
__device__ __forceinline__ float4 load_grad(const T* __restrict__ p, long long gs_c, bool vec) {
  if (vec) return V4<T>::cvt(V4<T>::ldraw_stream(p));
  return make_float4(Io<T>::ld(p), Io<T>::ld(p + gs_c), Io<T>::ld(p + 2 * gs_c), Io<T>::ld(p + 3 * gs_c));
}

template <typename T, int SR>
__global__ void __launch_bounds__(kBwdWarps * 32) roi_align_bwd_sm100_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels,
    const T* __restrict__ grad, long long gs_k, long long gs_c, long long gs_h, long long gs_w, bool vec,
    long long total_tasks, int C, int PH, int PW, bool aligned, int CT, int RP) {
  __shared__ BXEnt xtab[kBwdWarps][kMaxSX];
  __shared__ BYEnt ytab[kBwdWarps][kMaxRows];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long task = blockIdx.x * (long long)kBwdWarps + warp;
  if (task >= total_tasks) return;
  const int rp = (int)(task % RP);
  const int ct = (int)((task / RP) % CT);
  const long long k = task / RP / CT;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  BXEnt* xt = xtab[warp];
  BYEnt* yt = ytab[warp];
  const int nrows = build_bwd_tables<SR>(g, H, W, C, PH, PW, rp, lane, xt, yt);
  if (nrows == 0) return;

  const int ch = ct * 128 + lane * 4;
  T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
  const int ph0 = 2 * rp;
  const bool has1 = ph0 + 1 < PH;
  const T* g0p = grad + k * gs_k + (long long)ch * gs_c + (long long)ph0 * gs_h;
  const T* g1p = g0p + gs_h;

  float4 lo0 = f4zero(), lo1 = f4zero(), hi0 = f4zero(), hi1 = f4zero();
  int cur = 0;
  float4 gn0 = load_grad<T>(g0p, gs_c, vec);
  float4 gn1 = has1 ? load_grad<T>(g1p, gs_c, vec) : f4zero();
  for (int pw = 0; pw < PW; ++pw) {
    const float4 g0 = gn0, g1 = gn1;
    if (pw + 1 < PW) {  // the next pooled column's gradient is in flight during this one's FMAs
      gn0 = load_grad<T>(g0p + (long long)(pw + 1) * gs_w, gs_c, vec);
      if (has1) gn1 = load_grad<T>(g1p + (long long)(pw + 1) * gs_w, gs_c, vec);
    }
#pragma unroll
    for (int s = 0; s < SR; ++s) {
      const int j = pw * SR + s;
      const BXEnt e = xt[j];
      if (e.step != 0 && j > 0) {
        flush_column<T>(base, cur * C, yt, nrows, lo0, lo1);
        if (e.step == 2) {
          flush_column<T>(base, min(cur + 1, W - 1) * C, yt, nrows, hi0, hi1);
          lo0 = f4zero();
          lo1 = f4zero();
        } else {
          lo0 = hi0;
          lo1 = hi1;
        }
        hi0 = f4zero();
        hi1 = f4zero();
      }
      cur = e.xl;
      fma4(lo0, e.hx, g0);
      fma4(hi0, e.lx, g0);
      fma4(lo1, e.hx, g1);
      fma4(hi1, e.lx, g1);
    }
  }
  flush_column<T>(base, cur * C, yt, nrows, lo0, lo1);
  if (cur + 1 <= W - 1) flush_column<T>(base, (cur + 1) * C, yt, nrows, hi0, hi1);
}

template <typename T>
LevelTable<T> make_table(const LevelArgs& a) {
  LevelTable<T> t;
  t.num_levels = a.num_levels;
  for (int l = 0; l < B200DET_MAX_LEVELS; ++l) {
    const int s = l < a.num_levels ? l : 0;
    t.ptr[l] = (T*)a.ptr[s];
    t.height[l] = (int)a.height[s];
    t.width[l] = (int)a.width[s];
    t.scale[l] = (float)a.scale[s];
    const long long C = a.channels, H = a.height[s], W = a.width[s];
    t.sn[l] = H * W * C; t.sc[l] = 1; t.sh[l] = W * C; t.sw[l] = C;
  }
  return t;
}

bool eligible(const LevelArgs& lv, int dtype, int PH, int PW, int sr) {
  if (dtype == B200DET_F64) return false;
  if (lv.layout != B200DET_NHWC) return false;
  if (lv.channels % 128 != 0) return false;
  if (sr != 1 && sr != 2) return false;
  if (PW * sr > kMaxSX || PH < 1 || PW < 1) return false;
  const size_t es = dtype_size(dtype);
  for (int l = 0; l < lv.num_levels; ++l) {
    if (lv.height[l] < 1 || lv.width[l] < 1) return false;
    if (lv.height[l] * lv.width[l] * lv.channels >= (1LL << 31)) return false;  // 32-bit in-image offsets
    if (((uintptr_t)lv.ptr[l]) % (4 * es) != 0) return false;                   // vector alignment
  }
  return true;
}

// How many RoIs ahead the forward prefetches into L2.  Measured on B200 (kbench, r01): the
// prefetch LOSES (cfg3 fp32 fwd 0.70 ms off vs 0.78 ms at the old default, cfg1 0.09 vs 0.24 ms),
// so it is off unless B200DET_PF_DIST=<RoIs ahead> asks for it.
int prefetch_distance(int CT) {
  static int env = -2;
  if (env == -2) {
    const char* s = getenv("B200DET_PF_DIST");
    env = s ? atoi(s) : -1;
  }
  if (env >= 0) return env;
  (void)CT;
  return 0;
}

template <typename T>
int launch_fwd(const LevelArgs& lv, const void* rois, const int32_t* levels, void* out, long long tasks, int C,
               int PH, int PW, int sr, bool aligned, int CT, cudaStream_t st) {
  const LevelTable<T> tab = make_table<T>(lv);
  const int pf = prefetch_distance(CT);
  if (sr == 1)
    roi_align_fwd_sm100_kernel<T, 1><<<(unsigned)tasks, kFwdWarps * 32, 0, st>>>(
        tab, (const float*)rois, levels, (T*)out, (int)tasks, (int)lv.batch, C, PH, PW, aligned, CT, pf);
  else
    roi_align_fwd_sm100_kernel<T, 2><<<(unsigned)tasks, kFwdWarps * 32, 0, st>>>(
        tab, (const float*)rois, levels, (T*)out, (int)tasks, (int)lv.batch, C, PH, PW, aligned, CT, pf);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int launch_bwd(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
               const int32_t* levels, long long tasks, int C, int PH, int PW, int sr, bool aligned, int CT, int RP,
               cudaStream_t st) {
  const LevelTable<T> tab = make_table<T>(lv);
  const unsigned blocks = (unsigned)((tasks + kBwdWarps - 1) / kBwdWarps);
  if (sr == 1)
    roi_align_bwd_sm100_kernel<T, 1><<<blocks, kBwdWarps * 32, 0, st>>>(
        tab, (const float*)rois, levels, (const T*)grad, gs[0], gs[1], gs[2], gs[3], vec, tasks, C, PH, PW, aligned,
        CT, RP);
  else
    roi_align_bwd_sm100_kernel<T, 2><<<blocks, kBwdWarps * 32, 0, st>>>(
        tab, (const float*)rois, levels, (const T*)grad, gs[0], gs[1], gs[2], gs[3], vec, tasks, C, PH, PW, aligned,
        CT, RP);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int roi_align_fwd_fast(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels, void* out,
                       int out_layout, long long K, int PH, int PW, int sr, bool aligned, cudaStream_t st,
                       int* taken) {
  *taken = 0;
  if (!eligible(lv, dtype, PH, PW, sr) || out_layout != B200DET_NHWC) return 0;
  if (((uintptr_t)out) % (4 * dtype_size(dtype)) != 0) return 0;
  const int C = (int)lv.channels, CT = C / 128;
  const long long tasks = K * CT;
  if (tasks == 0) { *taken = 1; return 0; }
  if (tasks > 0x7fffffffLL) return 0;
  int rc = 0;
  switch (dtype) {
    case B200DET_F32: rc = launch_fwd<float>(lv, rois, levels, out, tasks, C, PH, PW, sr, aligned, CT, st); break;
    case B200DET_F16: rc = launch_fwd<__half>(lv, rois, levels, out, tasks, C, PH, PW, sr, aligned, CT, st); break;
    case B200DET_BF16: rc = launch_fwd<__nv_bfloat16>(lv, rois, levels, out, tasks, C, PH, PW, sr, aligned, CT, st); break;
    default: return 0;
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

int roi_align_bwd_fast(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                       const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned,
                       cudaStream_t st, int* taken) {
  *taken = 0;
  if (!eligible(lv, dtype, PH, PW, sr)) return 0;
  const size_t es = dtype_size(dtype);
  // vector loads of the gradient need unit channel stride and 4-element alignment of every row start
  const bool vec = gs[1] == 1 && gs[0] % 4 == 0 && gs[2] % 4 == 0 && gs[3] % 4 == 0 &&
                   ((uintptr_t)grad) % (4 * es) == 0;
  const int C = (int)lv.channels, CT = C / 128, RP = (PH + 1) / 2;
  const long long tasks = K * CT * RP;
  if (tasks == 0) { *taken = 1; return 0; }
  if ((tasks + kBwdWarps - 1) / kBwdWarps > 0x7fffffffLL) return 0;
  int rc = 0;
  switch (dtype) {
    case B200DET_F32: rc = launch_bwd<float>(lv, grad, gs, vec, rois, levels, tasks, C, PH, PW, sr, aligned, CT, RP, st); break;
    case B200DET_F16: rc = launch_bwd<__half>(lv, grad, gs, vec, rois, levels, tasks, C, PH, PW, sr, aligned, CT, RP, st); break;
    case B200DET_BF16: rc = launch_bwd<__nv_bfloat16>(lv, grad, gs, vec, rois, levels, tasks, C, PH, PW, sr, aligned, CT, RP, st); break;
    default: return 0;
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

}  // namespace b200det
