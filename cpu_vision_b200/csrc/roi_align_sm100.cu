// roi_align_sm100.cu -- the sm_100a RoIAlign BACKWARD kernel for the detection-model
// case (channels_last, C % 128 == 0, sampling_ratio 1 or 2, pooled_width *
// sampling_ratio <= 32); the forward lives in roi_align_fwd_sm100.cu, everything
// else in roi_align_generic.cu.
//
// The backward is the transpose of the separable forward with a warp per PAIR of
// pooled rows (fewer atomics: the rows two adjacent pooled rows share are reduced
// once): per column it accumulates sum_pw Wx[pw][c] * grad[ph][pw] in registers and
// issues ONE vector red.global.add per (row, column) of the footprint (4 channels
// per lane, 128 channels per warp instruction) instead of 4 scalar atomics per
// sample (cuda/roi_align_kernel.cu:303-327).
#include "sm100_common.cuh"

namespace b200det {
namespace {

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
int launch_bwd(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
               const int32_t* levels, long long tasks, int C, int PH, int PW, int sr, bool aligned, int CT, int RP,
               cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
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

int roi_align_bwd_fast(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                       const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned,
                       cudaStream_t st, int* taken) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr)) return 0;
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
