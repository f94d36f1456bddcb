// sm100_common.cuh -- pieces shared by the sm_100a RoIAlign forward and backward
// kernels (roi_align_fwd_sm100.cu, roi_align_sm100.cu): packed f32x2 arithmetic
// (FFMA2 / FMUL2, new on sm_100), the per-lane channel vector, level tables.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace b200det {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kMaxSX = 32;  // x samples per pooled row (pooled_width * sampling_ratio)
constexpr int kMaxPH = 32;  // pooled rows a CTA tabulates

// ---- packed pair of fp32 in one 64-bit register -------------------------------
typedef unsigned long long f32x2;

__device__ __forceinline__ f32x2 pack2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
// each half is an IEEE fma.rn / mul.rn exactly like fmaf / __fmul_rn
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// one vector reduction per 16 bytes: red.global.add.v4.f32 (sm_90+)
__device__ __forceinline__ void red_v4_f32(float* p, f32x2 a, f32x2 b) {
  float x, y, z, w;
  unpack2(a, x, y);
  unpack2(b, z, w);
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

// ---- the channels one lane owns at one pixel ------------------------------------
// CH channels per lane (4 or 8), held as CH/2 packed pairs once converted.
// Raw<T, CH> is what one pixel's load leaves in registers.
template <typename T, int CH> struct LaneVec;

template <> struct LaneVec<float, 4> {
  struct raw { ulonglong2 a; };
  static __device__ __forceinline__ raw ld(const float* p) { return {__ldg(reinterpret_cast<const ulonglong2*>(p))}; }
  static __device__ __forceinline__ raw ld_stream(const float* p) { return {__ldcs(reinterpret_cast<const ulonglong2*>(p))}; }
  static __device__ __forceinline__ void pairs(const raw& r, f32x2 (&v)[2]) { v[0] = r.a.x; v[1] = r.a.y; }
  static __device__ __forceinline__ void st(float* p, const f32x2 (&v)[2]) {
    __stcs(reinterpret_cast<ulonglong2*>(p), make_ulonglong2(v[0], v[1]));
  }
  static __device__ __forceinline__ void red(float* p, const f32x2 (&v)[2]) { red_v4_f32(p, v[0], v[1]); }
};
template <> struct LaneVec<float, 8> {
  // channels [4*lane, +4) and [128 + 4*lane, +4) of a 256-channel tile: two
  // fully coalesced 512-byte warp rows
  struct raw { ulonglong2 a, b; };
  static __device__ __forceinline__ raw ld(const float* p) {
    return {__ldg(reinterpret_cast<const ulonglong2*>(p)), __ldg(reinterpret_cast<const ulonglong2*>(p + 128))};
  }
  static __device__ __forceinline__ raw ld_stream(const float* p) {
    return {__ldcs(reinterpret_cast<const ulonglong2*>(p)), __ldcs(reinterpret_cast<const ulonglong2*>(p + 128))};
  }
  static __device__ __forceinline__ void pairs(const raw& r, f32x2 (&v)[4]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.b.x; v[3] = r.b.y;
  }
  static __device__ __forceinline__ void st(float* p, const f32x2 (&v)[4]) {
    __stcs(reinterpret_cast<ulonglong2*>(p), make_ulonglong2(v[0], v[1]));
    __stcs(reinterpret_cast<ulonglong2*>(p + 128), make_ulonglong2(v[2], v[3]));
  }
  static __device__ __forceinline__ void red(float* p, const f32x2 (&v)[4]) {
    red_v4_f32(p, v[0], v[1]);
    red_v4_f32(p + 128, v[2], v[3]);
  }
};

// 16-bit element <-> fp32 pair conversions
template <typename T> struct Cvt16;
template <> struct Cvt16<__nv_bfloat16> {
  // bf16 -> fp32 is a 16-bit shift
  static __device__ __forceinline__ f32x2 widen(unsigned u) {
    return pack2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  }
  static __device__ __forceinline__ unsigned narrow(f32x2 v) {
    float a, b;
    unpack2(v, a, b);
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const unsigned*>(&h);
  }
  static __device__ __forceinline__ void red2(__nv_bfloat16* p, unsigned a, unsigned b) {
    asm volatile("red.global.add.noftz.v2.bf16x2 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
  }
  static __device__ __forceinline__ void red4(__nv_bfloat16* p, unsigned a, unsigned b, unsigned c, unsigned d) {
    asm volatile("red.global.add.noftz.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
  }
};
template <> struct Cvt16<__half> {
  static __device__ __forceinline__ f32x2 widen(unsigned u) {
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u));
    return pack2(f.x, f.y);
  }
  static __device__ __forceinline__ unsigned narrow(f32x2 v) {
    float a, b;
    unpack2(v, a, b);
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const unsigned*>(&h);
  }
  static __device__ __forceinline__ void red2(__half* p, unsigned a, unsigned b) {
    asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
  }
  static __device__ __forceinline__ void red4(__half* p, unsigned a, unsigned b, unsigned c, unsigned d) {
    asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
  }
};

template <typename T> struct LaneVec16x4 {
  struct raw { uint2 a; };
  static __device__ __forceinline__ raw ld(const T* p) { return {__ldg(reinterpret_cast<const uint2*>(p))}; }
  static __device__ __forceinline__ raw ld_stream(const T* p) { return {__ldcs(reinterpret_cast<const uint2*>(p))}; }
  static __device__ __forceinline__ void pairs(const raw& r, f32x2 (&v)[2]) {
    v[0] = Cvt16<T>::widen(r.a.x); v[1] = Cvt16<T>::widen(r.a.y);
  }
  static __device__ __forceinline__ void st(T* p, const f32x2 (&v)[2]) {
    __stcs(reinterpret_cast<uint2*>(p), make_uint2(Cvt16<T>::narrow(v[0]), Cvt16<T>::narrow(v[1])));
  }
  static __device__ __forceinline__ void red(T* p, const f32x2 (&v)[2]) {
    Cvt16<T>::red2(p, Cvt16<T>::narrow(v[0]), Cvt16<T>::narrow(v[1]));
  }
};
template <typename T> struct LaneVec16x8 {
  // channels [8*lane, +8): one 16-byte load, 512 contiguous bytes per warp
  struct raw { uint4 a; };
  static __device__ __forceinline__ raw ld(const T* p) { return {__ldg(reinterpret_cast<const uint4*>(p))}; }
  static __device__ __forceinline__ raw ld_stream(const T* p) { return {__ldcs(reinterpret_cast<const uint4*>(p))}; }
  static __device__ __forceinline__ void pairs(const raw& r, f32x2 (&v)[4]) {
    v[0] = Cvt16<T>::widen(r.a.x); v[1] = Cvt16<T>::widen(r.a.y);
    v[2] = Cvt16<T>::widen(r.a.z); v[3] = Cvt16<T>::widen(r.a.w);
  }
  static __device__ __forceinline__ void st(T* p, const f32x2 (&v)[4]) {
    __stcs(reinterpret_cast<uint4*>(p), make_uint4(Cvt16<T>::narrow(v[0]), Cvt16<T>::narrow(v[1]),
                                                    Cvt16<T>::narrow(v[2]), Cvt16<T>::narrow(v[3])));
  }
  static __device__ __forceinline__ void red(T* p, const f32x2 (&v)[4]) {
    Cvt16<T>::red4(p, Cvt16<T>::narrow(v[0]), Cvt16<T>::narrow(v[1]), Cvt16<T>::narrow(v[2]), Cvt16<T>::narrow(v[3]));
  }
};
template <> struct LaneVec<__half, 4> : LaneVec16x4<__half> {};
template <> struct LaneVec<__half, 8> : LaneVec16x8<__half> {};
template <> struct LaneVec<__nv_bfloat16, 4> : LaneVec16x4<__nv_bfloat16> {};
template <> struct LaneVec<__nv_bfloat16, 8> : LaneVec16x8<__nv_bfloat16> {};

// ---- the lane's CH channels of one pixel, float2 flavour --------------------------------------
// Same channel assignment as LaneVec, but loaded as float4 / 32-bit words and handed out as float2 for the sm_100
// packed-fp32 intrinsics (__ffma2_rn, __fmul2_rn, __fadd2_rn).  With these ptxas updates accumulators in place
// and uses loaded registers directly; with LaneVec's opaque 64-bit asm operands it copied every packed sum back
// (12 MOVs per sample in the strip sweep, half of the forward's SASS): measured in round 2.
template <typename T> struct Cvt2;
template <> struct Cvt2<__nv_bfloat16> {
  static __device__ __forceinline__ float2 widen(unsigned u) {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  }
  static __device__ __forceinline__ unsigned narrow(float2 v) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
    return *reinterpret_cast<const unsigned*>(&h);
  }
};
template <> struct Cvt2<__half> {
  static __device__ __forceinline__ float2 widen(unsigned u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }
  static __device__ __forceinline__ unsigned narrow(float2 v) {
    const __half2 h = __floats2half2_rn(v.x, v.y);
    return *reinterpret_cast<const unsigned*>(&h);
  }
};

template <typename T, int CH> struct PixVec;
template <int CH> struct PixVec<float, CH> {
  struct raw { float4 v[CH / 4]; };
  static __device__ __forceinline__ void ld(raw& r, const float* p) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h) r.v[h] = __ldg(reinterpret_cast<const float4*>(p + 128 * h));
  }
  static __device__ __forceinline__ void ld_stream(raw& r, const float* p) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h) r.v[h] = __ldcs(reinterpret_cast<const float4*>(p + 128 * h));
  }
  static __device__ __forceinline__ void f2(const raw& r, float2 (&o)[CH / 2]) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h) {
      o[2 * h] = make_float2(r.v[h].x, r.v[h].y);
      o[2 * h + 1] = make_float2(r.v[h].z, r.v[h].w);
    }
  }
  static __device__ __forceinline__ void st(float* p, const float2 (&v)[CH / 2]) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h)
      __stcs(reinterpret_cast<float4*>(p + 128 * h), make_float4(v[2 * h].x, v[2 * h].y, v[2 * h + 1].x, v[2 * h + 1].y));
  }
  // one vector reduction per 16 bytes: red.global.add.v4.f32 (sm_90+)
  static __device__ __forceinline__ void red(float* p, const float2 (&v)[CH / 2]) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + 128 * h), "f"(v[2 * h].x), "f"(v[2 * h].y),
                   "f"(v[2 * h + 1].x), "f"(v[2 * h + 1].y) : "memory");
  }
};
template <typename T, int CH> struct PixVec {   // 16-bit elements: CH * 2 contiguous bytes per lane
  struct raw { unsigned u[CH / 2]; };
  static __device__ __forceinline__ void ld(raw& r, const T* p) {
    if (CH == 8) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
      r.u[0] = v.x; r.u[1] = v.y; r.u[CH / 2 - 2] = v.z; r.u[CH / 2 - 1] = v.w;
    } else {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
      r.u[0] = v.x; r.u[1] = v.y;
    }
  }
  static __device__ __forceinline__ void ld_stream(raw& r, const T* p) {
    if (CH == 8) {
      const uint4 v = __ldcs(reinterpret_cast<const uint4*>(p));
      r.u[0] = v.x; r.u[1] = v.y; r.u[CH / 2 - 2] = v.z; r.u[CH / 2 - 1] = v.w;
    } else {
      const uint2 v = __ldcs(reinterpret_cast<const uint2*>(p));
      r.u[0] = v.x; r.u[1] = v.y;
    }
  }
  static __device__ __forceinline__ void f2(const raw& r, float2 (&o)[CH / 2]) {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i) o[i] = Cvt2<T>::widen(r.u[i]);
  }
  static __device__ __forceinline__ void st(T* p, const float2 (&v)[CH / 2]) {
    if (CH == 8)
      __stcs(reinterpret_cast<uint4*>(p), make_uint4(Cvt2<T>::narrow(v[0]), Cvt2<T>::narrow(v[1]),
                                                      Cvt2<T>::narrow(v[CH / 2 - 2]), Cvt2<T>::narrow(v[CH / 2 - 1])));
    else
      __stcs(reinterpret_cast<uint2*>(p), make_uint2(Cvt2<T>::narrow(v[0]), Cvt2<T>::narrow(v[1])));
  }
  // packed 16-bit reductions: red.global.add.noftz.v4 / v2 .bf16x2 / .f16x2
  static __device__ __forceinline__ void red(T* p, const float2 (&v)[CH / 2]) {
    if (CH == 8)
      Cvt16<T>::red4(p, Cvt2<T>::narrow(v[0]), Cvt2<T>::narrow(v[1]), Cvt2<T>::narrow(v[CH / 2 - 2]),
                     Cvt2<T>::narrow(v[CH / 2 - 1]));
    else
      Cvt16<T>::red2(p, Cvt2<T>::narrow(v[0]), Cvt2<T>::narrow(v[1]));
  }
};

// first channel a lane owns inside its tile (the second half of LaneVec<float, 8>
// sits 128 channels further)
template <typename T, int CH>
__device__ __forceinline__ int lane_channel(int lane) {
  return (sizeof(T) == 4) ? lane * 4 : lane * CH;
}

// ---- x axis tables -----------------------------------------------------------------------
constexpr int kMaxEvents = 2 * kMaxSX;

struct __align__(16) XTab {
  float4 samp[kMaxSX];                  // (hx, lx, valid as 0/1 bits, -): weights of the lo / hi column
  int colofs[kMaxEvents + 2];           // element offset (x * C) of column event e; 2 entries of padding
  unsigned char ev_end[kMaxEvents + 4]; // [e + 1]: samples complete once event e arrived; [0]: before any
  int ncol;
};

// ---- x axis (one warp; lane j owns x sample j) ------------------------------------------
// The sweep keeps a window of two feature columns (lo, hi = min(lo + 1, W - 1),
// clamped like roi_align_kernel.cu:34-46).  A sample finds its taps by leaving
// the window as it is, shifting it by one column (one event) or refilling it
// (two events).  Empty samples (roi_align_kernel.cu:23-26) fetch nothing.
template <int SR>
__device__ __forceinline__ void build_x(const RoiGeom<float>& g, int W, int C, int PW, int lane, XTab& tb) {
  const int SX = PW * SR;
  const bool active = lane < SX;
  const int j = active ? lane : SX - 1;
  const float x = sample_coord<float>(g.start_w, g.bin_w, j / SR, j % SR, SR);
  const AxisTap<float> t = axis_tap<float>(x, W);
  const bool valid = t.valid && active;
  const int xl = t.lo, xh = min(t.lo + 1, W - 1);
  // window left by the last valid sample before this one
  const unsigned vmask = __ballot_sync(kFull, valid);
  const unsigned before = vmask & ((1u << lane) - 1u);
  const int p = before ? 31 - __clz(before) : 0;
  int pxl = __shfl_sync(kFull, xl, p), pxh = __shfl_sync(kFull, xh, p);
  if (!before) { pxl = -2; pxh = -2; }
  const int nnew = !valid ? 0 : (xl == pxl ? 0 : (xl == pxh ? 1 : 2));
  int incl = nnew;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += v;
  }
  const int base = incl - nnew;
  const int ncol = __shfl_sync(kFull, incl, 31);
  if (active) {
    tb.samp[lane] = make_float4(t.wl, t.wh, __int_as_float(valid ? 1 : 0), 0.f);
    if (nnew == 2) {
      tb.colofs[base] = xl * C;
      tb.colofs[base + 1] = xh * C;
    } else if (nnew == 1) {
      tb.colofs[base] = xh * C;
    }
  }
  // sample j is complete after event last = incl - 1 (-1: needs no event at all)
  const int last = incl - 1;
  int nxt = __shfl_down_sync(kFull, last, 1);
  if (lane == SX - 1) nxt = ncol;
  if (active) {
    for (int e = last; e < nxt; ++e) tb.ev_end[e + 1] = (unsigned char)(lane + 1);
    if (lane == 0)
      for (int e = -1; e < last; ++e) tb.ev_end[e + 1] = 0;
  }
  if (lane == 0) tb.ncol = ncol;
  __syncwarp();
  if (lane < 2 && ncol > 0) tb.colofs[ncol + lane] = tb.colofs[ncol - 1];
}

// ---- host helpers --------------------------------------------------------------------
template <typename T>
inline LevelTable<T> make_nhwc_table(const LevelArgs& a) {
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

// what the sm_100a kernels handle; everything else runs in roi_align_generic.cu
inline bool sm100_eligible(const LevelArgs& lv, int dtype, int PH, int PW, int sr) {
  if (dtype == B200DET_F64) return false;
  if (lv.layout != B200DET_NHWC) return false;
  if (lv.channels % 128 != 0) return false;
  if (sr != 1 && sr != 2) return false;
  if (PW * sr > kMaxSX || PH < 1 || PW < 1 || PH > kMaxPH) return false;
  const size_t es = dtype_size(dtype);
  for (int l = 0; l < lv.num_levels; ++l) {
    if (lv.height[l] < 1 || lv.width[l] < 1) return false;
    if (lv.height[l] * lv.width[l] * lv.channels >= (1LL << 31)) return false;  // 32-bit in-image offsets
    if (((uintptr_t)lv.ptr[l]) % 16 != 0) return false;                         // 16-byte vector alignment
  }
  (void)es;
  return true;
}

}  // namespace b200det
