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
};
template <> struct LaneVec<__half, 4> : LaneVec16x4<__half> {};
template <> struct LaneVec<__half, 8> : LaneVec16x8<__half> {};
template <> struct LaneVec<__nv_bfloat16, 4> : LaneVec16x4<__nv_bfloat16> {};
template <> struct LaneVec<__nv_bfloat16, 8> : LaneVec16x8<__nv_bfloat16> {};

// first channel a lane owns inside its tile (the second half of LaneVec<float, 8>
// sits 128 channels further)
template <typename T, int CH>
__device__ __forceinline__ int lane_channel(int lane) {
  return (sizeof(T) == 4) ? lane * 4 : lane * CH;
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
