// roi_align_fast.cu -- the sm_100a RoIAlign kernels for the detection-model
// case: channels_last feature maps, C a multiple of 128, sampling_ratio 1 or 2,
// pooled_width * sampling_ratio <= 32.  Everything else goes to
// roi_align_generic.cu.
//
// Work item = one WARP per (RoI, 128-channel tile, pair of pooled rows).  A lane
// owns 4 consecutive channels, so every feature tap is one 16-byte (fp32) or
// 8-byte (bf16/fp16) load and a warp covers 512/256 contiguous bytes.
//
// bilinear_interpolate (cuda/roi_align_kernel.cu:14-66) is separable: the
// weight of pixel (r, c) for a sample is wy(r) * wx(c).  Summed over the
// sampling grid of a bin that gives, for a pooled row ph,
//     out[ph][pw] = sum_c Wx[pw][c] * ( sum_r Wy[ph][r] * in[r][c] ).
// The warp therefore
//   * builds, once, the <= 4*sr distinct feature rows its two pooled rows
//     touch with their merged weights Wy (1/count folded in; exact, the count
//     is a power of two), and the per-sample x taps -- both tables staged in
//     shared memory and read back as one broadcast LDS.128 per entry;
//   * sweeps the x samples left to right keeping the vertical sums of the two
//     columns under the current sample in registers ("lo"/"hi").  A sample
//     that stays on the same pixel reuses both, one that moves right by one
//     pixel reuses one, so every (row, column) of the RoI's footprint is loaded
//     ONCE per row pair instead of once per sample tap: for a 14x14 mask-head
//     RoI that is ~4x fewer L1/L2 requests than one gather per tap, which is
//     what keeps the kernel on the HBM roofline instead of the L1 one.
// The backward is the exact transpose: per column it accumulates
// sum_pw Wx[pw][c] * grad[ph][pw] in registers and issues ONE vector
// red.global.add per (row, column) of the footprint (warp-aggregated: 4
// channels per instruction, 128 channels per warp instruction) instead of 4
// scalar atomics per sample.
#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kWarps = 4;
constexpr int kMaxSX = 32;    // x samples per pooled row handled by one warp
constexpr int kMaxRows = 8;   // 2 pooled rows x 2 samples x 2 taps

struct __align__(16) XEnt { int xl; int step; float hx; float lx; };
struct __align__(16) YEnt { int rowoff; float w0; float w1; int pad; };

// ---- 4-channel vector I/O ---------------------------------------------------
template <typename T> struct V4;
template <> struct V4<float> {
  static __device__ __forceinline__ float4 ld(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ void st(float* p, float4 v) { __stcs(reinterpret_cast<float4*>(p), v); }
  static __device__ __forceinline__ void red(float* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
};
template <> struct V4<__half> {
  static __device__ __forceinline__ float4 ld(const __half* p) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
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
  static __device__ __forceinline__ float4 ld(const __nv_bfloat16* p) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
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
__device__ __forceinline__ void fma4(float4& a, float w, const float4& v) {
  a.x = fmaf(w, v.x, a.x); a.y = fmaf(w, v.y, a.y); a.z = fmaf(w, v.z, a.z); a.w = fmaf(w, v.w, a.w);
}

// ---- per-warp tables ----------------------------------------------------------
// Fills xt[0..SX) and yt[0..nrows) for (RoI geometry g, pooled rows 2*rp, 2*rp+1).
// Returns nrows.  All lanes of the warp must call it.
__device__ __forceinline__ int build_tables(const RoiGeom<float>& g, int H, int W, int C, int PH, int PW,
                                            int sr, int rp, int lane, XEnt* xt, YEnt* yt) {
  const int SX = PW * sr;
  // x samples: lane j owns sample j
  {
    const int j = lane < SX ? lane : SX - 1;
    const float x = sample_coord<float>(g.start_w, g.bin_w, j / sr, j % sr, sr);
    const AxisTap<float> t = axis_tap<float>(x, W);
    // an empty sample keeps the window where the sweep already is
    const int xl = t.valid ? t.lo : (x < -1.0f ? 0 : W - 1);
    const int prev = __shfl_up_sync(0xffffffffu, xl, 1);
    const int d = xl - prev;
    XEnt e;
    e.xl = xl;
    e.step = (lane == 0) ? 2 : (d == 0 ? 0 : (d == 1 ? 1 : 2));
    e.hx = t.wl;   // 0 when !valid
    e.lx = t.wh;
    if (lane < SX) xt[lane] = e;
  }
  // y taps: lane t < 4*sr owns tap (pooled row m, sample iy, lo/hi)
  const int nT = 4 * sr;
  int row = -1;
  float w = 0.f;
  int m = 0;
  if (lane < nT) {
    m = lane / (2 * sr);
    const int iy = (lane >> 1) % sr;
    const int ph = 2 * rp + m;
    if (ph < PH) {
      const float y = sample_coord<float>(g.start_h, g.bin_h, ph, iy, sr);
      const AxisTap<float> t = axis_tap<float>(y, H);
      if (t.valid) {
        row = (lane & 1) ? t.hi : t.lo;
        w = ((lane & 1) ? t.wh : t.wl) * (1.0f / (float)(sr * sr));
      }
    }
  }
  // merge taps that hit the same feature row; the first tap of a row leads
  bool leader = row >= 0;
  float w0 = 0.f, w1 = 0.f;
  for (int t = 0; t < nT; ++t) {
    const int rr = __shfl_sync(0xffffffffu, row, t);
    const float ww = __shfl_sync(0xffffffffu, w, t);
    const int mm = __shfl_sync(0xffffffffu, m, t);
    if (rr == row && row >= 0) {
      if (t < lane) leader = false;
      if (mm == 0) w0 += ww; else w1 += ww;
    }
  }
  const unsigned ball = __ballot_sync(0xffffffffu, leader);
  if (leader) {
    YEnt e;
    e.rowoff = row * W * C;
    e.w0 = w0;
    e.w1 = w1;
    e.pad = 0;
    yt[__popc(ball & ((1u << lane) - 1u))] = e;
  }
  __syncwarp();
  return __popc(ball);
}

// vertical sums of column `col` for the two pooled rows
template <typename T>
__device__ __forceinline__ void load_column(const T* __restrict__ base, int coloff, const YEnt* yt, int nrows,
                                            float4& v0, float4& v1) {
  v0 = f4zero();
  v1 = f4zero();
#pragma unroll 4
  for (int r = 0; r < nrows; ++r) {
    const YEnt e = yt[r];
    const float4 v = V4<T>::ld(base + e.rowoff + coloff);
    fma4(v0, e.w0, v);
    fma4(v1, e.w1, v);
  }
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32) roi_align_fwd_fast_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels,
    T* __restrict__ out, long long total_tasks, int C, int PH, int PW, int sr, bool aligned, int CT, int RP) {
  __shared__ XEnt xtab[kWarps][kMaxSX];
  __shared__ YEnt ytab[kWarps][kMaxRows];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long task = blockIdx.x * (long long)kWarps + warp;
  if (task >= total_tasks) return;
  const int rp = (int)(task % RP);
  const int ct = (int)((task / RP) % CT);
  const long long k = task / RP / CT;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, sr, aligned);
  XEnt* xt = xtab[warp];
  YEnt* yt = ytab[warp];
  const int nrows = build_tables(g, H, W, C, PH, PW, sr, rp, lane, xt, yt);

  const int ch = ct * 128 + lane * 4;
  const T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
  const int ph0 = 2 * rp;
  const bool has1 = ph0 + 1 < PH;
  T* o0 = out + ((k * PH + ph0) * PW) * (long long)C + ch;
  T* o1 = o0 + (long long)PW * C;

  float4 lo0 = f4zero(), lo1 = f4zero(), hi0 = f4zero(), hi1 = f4zero(), a0 = f4zero(), a1 = f4zero();
  const int SX = PW * sr;
  for (int j = 0; j < SX; ++j) {
    const XEnt e = xt[j];
    if (e.step != 0) {
      const int xh = min(e.xl + 1, W - 1);
      if (e.step == 2) {
        load_column<T>(base, e.xl * C, yt, nrows, lo0, lo1);
      } else {
        lo0 = hi0;
        lo1 = hi1;
      }
      load_column<T>(base, xh * C, yt, nrows, hi0, hi1);
    }
    fma4(a0, e.hx, lo0);
    fma4(a0, e.lx, hi0);
    fma4(a1, e.hx, lo1);
    fma4(a1, e.lx, hi1);
    if ((j + 1) % sr == 0) {
      const int pw = j / sr;
      V4<T>::st(o0 + (long long)pw * C, a0);
      if (has1) V4<T>::st(o1 + (long long)pw * C, a1);
      a0 = f4zero();
      a1 = f4zero();
    }
  }
}

// ---- backward --------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void flush_column(T* __restrict__ base, int coloff, const YEnt* yt, int nrows,
                                             const float4& c0, const float4& c1) {
#pragma unroll 4
  for (int r = 0; r < nrows; ++r) {
    const YEnt e = yt[r];
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
  if (vec) return V4<T>::ld(p);
  return make_float4(Io<T>::ld(p), Io<T>::ld(p + gs_c), Io<T>::ld(p + 2 * gs_c), Io<T>::ld(p + 3 * gs_c));
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32) roi_align_bwd_fast_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels,
    const T* __restrict__ grad, long long gs_k, long long gs_c, long long gs_h, long long gs_w, bool vec,
    long long total_tasks, int C, int PH, int PW, int sr, bool aligned, int CT, int RP) {
  __shared__ XEnt xtab[kWarps][kMaxSX];
  __shared__ YEnt ytab[kWarps][kMaxRows];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long task = blockIdx.x * (long long)kWarps + warp;
  if (task >= total_tasks) return;
  const int rp = (int)(task % RP);
  const int ct = (int)((task / RP) % CT);
  const long long k = task / RP / CT;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, sr, aligned);
  XEnt* xt = xtab[warp];
  YEnt* yt = ytab[warp];
  const int nrows = build_tables(g, H, W, C, PH, PW, sr, rp, lane, xt, yt);
  if (nrows == 0) return;

  const int ch = ct * 128 + lane * 4;
  T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
  const int ph0 = 2 * rp;
  const bool has1 = ph0 + 1 < PH;
  const T* g0p = grad + k * gs_k + (long long)ch * gs_c + (long long)ph0 * gs_h;
  const T* g1p = g0p + gs_h;

  float4 lo0 = f4zero(), lo1 = f4zero(), hi0 = f4zero(), hi1 = f4zero(), g0 = f4zero(), g1 = f4zero();
  int cur = 0;
  const int SX = PW * sr;
  for (int j = 0; j < SX; ++j) {
    const XEnt e = xt[j];
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
    if (j % sr == 0) {
      const int pw = j / sr;
      g0 = load_grad<T>(g0p + (long long)pw * gs_w, gs_c, vec);
      g1 = has1 ? load_grad<T>(g1p + (long long)pw * gs_w, gs_c, vec) : f4zero();
    }
    fma4(lo0, e.hx, g0);
    fma4(hi0, e.lx, g0);
    fma4(lo1, e.hx, g1);
    fma4(hi1, e.lx, g1);
  }
  flush_column<T>(base, cur * C, yt, nrows, lo0, lo1);
  if (cur + 1 <= W - 1) flush_column<T>(base, (cur + 1) * C, yt, nrows, hi0, hi1);
}

template <typename T>
LevelTable<T> make_table(const LevelArgs& a) {
  LevelTable<T> t;
  t.num_levels = a.num_levels;
  for (int l = 0; l < a.num_levels; ++l) {
    t.ptr[l] = (T*)a.ptr[l];
    t.height[l] = (int)a.height[l];
    t.width[l] = (int)a.width[l];
    t.scale[l] = (float)a.scale[l];
    const long long C = a.channels, H = a.height[l], W = a.width[l];
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

}  // namespace

int roi_align_fwd_fast(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels, void* out,
                       int out_layout, long long K, int PH, int PW, int sr, bool aligned, cudaStream_t st,
                       int* taken) {
  *taken = 0;
  if (!eligible(lv, dtype, PH, PW, sr) || out_layout != B200DET_NHWC) return 0;
  if (((uintptr_t)out) % (4 * dtype_size(dtype)) != 0) return 0;
  const int C = (int)lv.channels, CT = C / 128, RP = (PH + 1) / 2;
  const long long tasks = K * CT * RP;
  if (tasks == 0) { *taken = 1; return 0; }
  const long long blocks = (tasks + kWarps - 1) / kWarps;
  if (blocks > 0x7fffffffLL) return 0;
  const float* r = (const float*)rois;
  switch (dtype) {
    case B200DET_F32:
      roi_align_fwd_fast_kernel<float><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<float>(lv), r, levels, (float*)out, tasks, C, PH, PW, sr, aligned, CT, RP);
      break;
    case B200DET_F16:
      roi_align_fwd_fast_kernel<__half><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<__half>(lv), r, levels, (__half*)out, tasks, C, PH, PW, sr, aligned, CT, RP);
      break;
    case B200DET_BF16:
      roi_align_fwd_fast_kernel<__nv_bfloat16><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<__nv_bfloat16>(lv), r, levels, (__nv_bfloat16*)out, tasks, C, PH, PW, sr, aligned, CT, RP);
      break;
    default: return 0;
  }
  B200DET_CUDA_OK(cudaGetLastError());
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
  const long long blocks = (tasks + kWarps - 1) / kWarps;
  if (blocks > 0x7fffffffLL) return 0;
  const float* r = (const float*)rois;
  switch (dtype) {
    case B200DET_F32:
      roi_align_bwd_fast_kernel<float><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<float>(lv), r, levels, (const float*)grad, gs[0], gs[1], gs[2], gs[3], vec, tasks, C, PH, PW,
          sr, aligned, CT, RP);
      break;
    case B200DET_F16:
      roi_align_bwd_fast_kernel<__half><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<__half>(lv), r, levels, (const __half*)grad, gs[0], gs[1], gs[2], gs[3], vec, tasks, C, PH, PW,
          sr, aligned, CT, RP);
      break;
    case B200DET_BF16:
      roi_align_bwd_fast_kernel<__nv_bfloat16><<<(unsigned)blocks, kWarps * 32, 0, st>>>(
          make_table<__nv_bfloat16>(lv), r, levels, (const __nv_bfloat16*)grad, gs[0], gs[1], gs[2], gs[3], vec,
          tasks, C, PH, PW, sr, aligned, CT, RP);
      break;
    default: return 0;
  }
  B200DET_CUDA_OK(cudaGetLastError());
  *taken = 1;
  return 0;
}

}  // namespace b200det
