// roi_align_anygrid_sm100.cu -- sm_100a RoIAlign forward / backward for channels_last maps with ANY sampling grid:
// adaptive grids (sampling_ratio <= 0: ceil(roi_h / pooled_h) x ceil(roi_w / pooled_w) samples per bin, the
// default of torchvision.ops.roi_align and what every forward of the reference's test/test_ops.py uses,
// cuda/roi_align_kernel.cu:112-115), sampling_ratio > 2, wide outputs (pooled_width * sampling_ratio > 32) and
// channel counts that are a multiple of 4 but not of 128.  roi_align_{fwd,bwd}_sm100.cu keep sampling_ratio 1 / 2
// with C % 128 == 0; what neither takes (fp64, NCHW without a re-layout, C % 4 != 0) stays in roi_align_generic.cu.
//
// The reference walks the samples: 4 taps per sample, grid_h * grid_w samples per bin, one thread per output
// element.  bilinear_interpolate (cuda/roi_align_kernel.cu:14-66) is separable, and the samples of a bin sit at
// most one pixel apart when the grid is adaptive, so neighbouring samples hit the same pixels again and again:
// a bin of g x g samples touches (g + 1) x (g + 1) pixels, not 4 g^2.  Here the taps of a bin are MERGED per axis
// first -- per pooled row the distinct feature rows with their summed weights, per pooled column the distinct
// feature columns -- and a bin is then
//     out[ph][pw] = ( sum_r sum_c  Wy[ph][r] * Wx[pw][c] * in[r][c] ) / count
// over its merged rows and columns: one 16-byte load per lane and pixel, (g + 1)^2 instead of 4 g^2 of them, no
// per-sample tap arithmetic in the loop.  The backward is the transpose: one vector red per lane and pixel.
//
// CTA = one (RoI, channel tile), 4 warps.  A lane owns 4 or 8 channels of every pixel, a warp covers up to 1 KB of
// contiguous NHWC bytes; lanes beyond C idle (C only has to be a multiple of 4).  Warp 0 merges the x axis, warp 1
// the y axis (lane = bin: count, warp scan, fill), then every warp takes bins.  The tables hold kEnt entries per
// axis; a RoI whose merged taps do not fit (a 700-pixel RoI on a stride-1 map) is walked sample by sample by the
// same CTA, so nothing depends on a bound the host would have to guess from data it cannot see.
#include <stdlib.h>

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kAgWarps = 4;
constexpr int kEnt = 768;      // merged taps per axis a CTA tabulates (adaptive grids need <= size + 2 * bins)
constexpr int kMaxBins = 64;   // pooled_height, pooled_width

struct __align__(16) AxisTab {
  int off[kEnt];               // element offset of the row (row * W * C) / column (col * C)
  float w[kEnt];               // summed weight of that row / column over the samples of the bin
  short start[kMaxBins + 1];   // bin p owns entries [start[p], start[p + 1])
};

struct __align__(16) AgTables {
  AxisTab x, y;
  int direct;                  // the merged taps of this RoI do not fit: walk the samples
};

// The merged taps of bin p along one axis.  Samples ascend, so a tap row is the last entry, the one before it, or
// new.  FILL = false only counts.  roi_align_kernel.cu:23-54 per axis is axis_tap(); empty samples contribute nothing.
template <bool FILL>
__device__ __forceinline__ int merge_bin(float start, float bin, int p, int grid, int size, int stride, int* off,
                                         float* w) {
  int n = 0, r_last = -1, r_prev = -1;
  auto add = [&](int row, float wt) {
    if (n > 0 && row == r_last) {
      if (FILL) w[n - 1] += wt;
    } else if (n > 1 && row == r_prev) {
      if (FILL) w[n - 2] += wt;
    } else {
      if (FILL) { off[n] = row * stride; w[n] = wt; }
      r_prev = r_last;
      r_last = row;
      ++n;
    }
  };
  for (int i = 0; i < grid; ++i) {
    const AxisTap<float> t = axis_tap<float>(sample_coord<float>(start, bin, p, i, grid), size);
    if (!t.valid) continue;
    add(t.lo, t.wl);
    add(t.hi, t.wh);
  }
  return n;
}

// one warp tabulates one axis; returns false (to every lane) when the entries do not fit
__device__ __forceinline__ bool build_axis(float start, float bin, int bins, int grid, int size, int stride, int lane,
                                           AxisTab& tb) {
  int total = 0;
  for (int p0 = 0; p0 < bins; p0 += 32) {
    const int p = p0 + lane;
    const int n = p < bins ? merge_bin<false>(start, bin, p, grid, size, stride, nullptr, nullptr) : 0;
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, o);
      if (lane >= o) incl += v;
    }
    const int first = total + incl - n;
    total += __shfl_sync(kFull, incl, 31);
    if (total > kEnt) return false;   // uniform: `total` is the same in every lane
    if (p < bins) {
      tb.start[p] = (short)first;
      merge_bin<true>(start, bin, p, grid, size, stride, tb.off + first, tb.w + first);
    }
  }
  if (lane == 0) tb.start[bins] = (short)total;
  return true;
}

struct AgGeom {
  RoiGeom<float> g;
  int l, H, W;
};

template <typename T>
__device__ __forceinline__ AgGeom ag_setup(const LevelTable<T>& lv, const float* __restrict__ rois,
                                           const int32_t* __restrict__ levels, long long k, int PH, int PW, int sr,
                                           bool aligned, int C, AgTables& tb) {
  AgGeom a;
  a.l = levels ? __ldg(levels + k) : 0;
  a.H = lv.height[a.l];
  a.W = lv.width[a.l];
  a.g = roi_geometry<float>(rois + k * 5, lv.scale[a.l], PH, PW, sr, aligned);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) tb.direct = 0;
  __syncthreads();
  if (warp == 0) {
    if (!build_axis(a.g.start_w, a.g.bin_w, PW, a.g.grid_w, a.W, C, lane, tb.x) && lane == 0) tb.direct = 1;
  } else if (warp == 1) {
    if (!build_axis(a.g.start_h, a.g.bin_h, PH, a.g.grid_h, a.H, a.W * C, lane, tb.y) && lane == 0) tb.direct = 1;
  }
  __syncthreads();
  return a;
}

// ---- forward ---------------------------------------------------------------------------------
template <typename T, int CH>
__global__ void __launch_bounds__(kAgWarps * 32) roi_align_fwd_anygrid_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels, T* __restrict__ out, int C,
    int PH, int PW, int sr, bool aligned, int CT) {
  using PV = PixVec<T, CH>;
  constexpr int H2 = CH / 2;
  __shared__ AgTables tb;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ct = blockIdx.x % CT;
  const long long k = blockIdx.x / CT;
  const AgGeom a = ag_setup<T>(lv, rois, levels, k, PH, PW, sr, aligned, C, tb);
  const int ch = ct * (32 * CH) + lane_channel<T, CH>(lane);
  if (ch >= C) return;  // C is a multiple of 4 (8 for 8-channel lanes of 16-bit data): a lane is in or out as a whole
  const T* base = lv.ptr[a.l] + (long long)a.g.batch * lv.sn[a.l] + ch;
  T* obase = out + k * PH * PW * (long long)C + ch;
  const int cnt = a.g.grid_h * a.g.grid_w;
  const float count = (float)(cnt > 1 ? cnt : 1);  // roi_align_kernel.cu:122
  const bool direct = tb.direct != 0;
  for (int b = warp; b < PH * PW; b += kAgWarps) {
    const int ph = b / PW, pw = b - ph * PW;
    float2 acc[H2];
#pragma unroll
    for (int i = 0; i < H2; ++i) acc[i] = make_float2(0.f, 0.f);
    if (!direct) {
      const int y0 = tb.y.start[ph], y1 = tb.y.start[ph + 1];
      const int x0 = tb.x.start[pw], x1 = tb.x.start[pw + 1];
      for (int iy = y0; iy < y1; ++iy) {
        const T* rp = base + tb.y.off[iy];
        const float wy = tb.y.w[iy];
#pragma unroll 4
        for (int ix = x0; ix < x1; ++ix) {
          typename PV::raw v;
          PV::ld(v, rp + tb.x.off[ix]);
          const float w = wy * tb.x.w[ix];
          float2 t[H2];
          PV::f2(v, t);
          const float2 w2 = make_float2(w, w);
#pragma unroll
          for (int i = 0; i < H2; ++i) acc[i] = __ffma2_rn(w2, t[i], acc[i]);
        }
      }
    } else {  // sample by sample, cuda/roi_align_kernel.cu:124-139 with a channel vector per lane
      const int rs = a.W * C;
      for (int iy = 0; iy < a.g.grid_h; ++iy) {
        const AxisTap<float> ty = axis_tap<float>(sample_coord<float>(a.g.start_h, a.g.bin_h, ph, iy, a.g.grid_h), a.H);
        if (!ty.valid) continue;
        for (int ix = 0; ix < a.g.grid_w; ++ix) {
          const AxisTap<float> tx = axis_tap<float>(sample_coord<float>(a.g.start_w, a.g.bin_w, pw, ix, a.g.grid_w), a.W);
          if (!tx.valid) continue;
          typename PV::raw v[4];
          PV::ld(v[0], base + ty.lo * rs + tx.lo * C);
          PV::ld(v[1], base + ty.lo * rs + tx.hi * C);
          PV::ld(v[2], base + ty.hi * rs + tx.lo * C);
          PV::ld(v[3], base + ty.hi * rs + tx.hi * C);
          const float w[4] = {ty.wl * tx.wl, ty.wl * tx.wh, ty.wh * tx.wl, ty.wh * tx.wh};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float2 t[H2];
            PV::f2(v[q], t);
            const float2 w2 = make_float2(w[q], w[q]);
#pragma unroll
            for (int i = 0; i < H2; ++i) acc[i] = __ffma2_rn(w2, t[i], acc[i]);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < H2; ++i) acc[i] = make_float2(acc[i].x / count, acc[i].y / count);
    PV::st(obase + (long long)b * C, acc);
  }
}

// ---- backward --------------------------------------------------------------------------------
// grad[k][ph][pw][:] / count, times Wy * Wx, reduced into every pixel of the bin's merged footprint
// (roi_align_kernel.cu:298-327 adds grad * w_i / count per tap; the merged weights are the sums of the w_i).
// The gradient buffers are zeroed by the caller (cudaMemsetAsync) before this kernel.
template <typename T, int CH>
__global__ void __launch_bounds__(kAgWarps * 32) roi_align_bwd_anygrid_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels, const T* __restrict__ grad,
    long long gs_k, long long gs_h, long long gs_w, int N, int C, int PH, int PW, int sr, bool aligned, int CT) {
  using PV = PixVec<T, CH>;
  constexpr int H2 = CH / 2;
  __shared__ AgTables tb;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ct = blockIdx.x % CT;
  const long long k = blockIdx.x / CT;
  const AgGeom a = ag_setup<T>(lv, rois, levels, k, PH, PW, sr, aligned, C, tb);
  const int ch = ct * (32 * CH) + lane_channel<T, CH>(lane);
  if (ch >= C) return;
  if (a.g.batch < 0 || a.g.batch >= N) return;  // a batch index outside the gradient buffers writes nothing
  T* base = lv.ptr[a.l] + (long long)a.g.batch * lv.sn[a.l] + ch;
  const T* gbase = grad + k * gs_k + ch;
  const float count = (float)(a.g.grid_h * a.g.grid_w);  // roi_align_kernel.cu:266: no max(., 1) here
  const bool direct = tb.direct != 0;
  for (int b = warp; b < PH * PW; b += kAgWarps) {
    const int ph = b / PW, pw = b - ph * PW;
    typename PV::raw gv;
    PV::ld_stream(gv, gbase + ph * gs_h + pw * gs_w);
    float2 g[H2];
    PV::f2(gv, g);
#pragma unroll
    for (int i = 0; i < H2; ++i) g[i] = make_float2(g[i].x / count, g[i].y / count);
    if (!direct) {
      const int y0 = tb.y.start[ph], y1 = tb.y.start[ph + 1];
      const int x0 = tb.x.start[pw], x1 = tb.x.start[pw + 1];
      for (int iy = y0; iy < y1; ++iy) {
        T* rp = base + tb.y.off[iy];
        const float wy = tb.y.w[iy];
#pragma unroll 4
        for (int ix = x0; ix < x1; ++ix) {
          const float w = wy * tb.x.w[ix];
          const float2 w2 = make_float2(w, w);
          float2 v[H2];
#pragma unroll
          for (int i = 0; i < H2; ++i) v[i] = __fmul2_rn(w2, g[i]);
          PV::red(rp + tb.x.off[ix], v);
        }
      }
    } else {
      const int rs = a.W * C;
      for (int iy = 0; iy < a.g.grid_h; ++iy) {
        const AxisTap<float> ty = axis_tap<float>(sample_coord<float>(a.g.start_h, a.g.bin_h, ph, iy, a.g.grid_h), a.H);
        if (!ty.valid) continue;
        for (int ix = 0; ix < a.g.grid_w; ++ix) {
          const AxisTap<float> tx = axis_tap<float>(sample_coord<float>(a.g.start_w, a.g.bin_w, pw, ix, a.g.grid_w), a.W);
          if (!tx.valid) continue;
          const float w[4] = {ty.wl * tx.wl, ty.wl * tx.wh, ty.wh * tx.wl, ty.wh * tx.wh};
          const int o[4] = {ty.lo * rs + tx.lo * C, ty.lo * rs + tx.hi * C, ty.hi * rs + tx.lo * C, ty.hi * rs + tx.hi * C};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 w2 = make_float2(w[q], w[q]);
            float2 v[H2];
#pragma unroll
            for (int i = 0; i < H2; ++i) v[i] = __fmul2_rn(w2, g[i]);
            PV::red(base + o[q], v);
          }
        }
      }
    }
  }
}

// ---- host side -------------------------------------------------------------------------------
bool anygrid_enabled() {
  const char* e = getenv("B200DET_ANYGRID");  // read per call: the parity tests compare this path with the generic one
  return !e || atoi(e) != 0;
}

// channels per lane: 8 when the lane's two 16-byte halves (fp32: 128 channels apart) / its 16 bytes (16-bit) exist
int lane_channels(int dtype, int C) {
  if (dtype == B200DET_F32) return C % 256 == 0 ? 8 : 4;
  return C % 8 == 0 ? 8 : 4;
}

bool anygrid_eligible(const LevelArgs& lv, int dtype, int PH, int PW) {
  if (dtype != B200DET_F32 && dtype != B200DET_F16 && dtype != B200DET_BF16) return false;
  if (lv.layout != B200DET_NHWC) return false;
  if (lv.channels < 32 || lv.channels % 4 != 0) return false;   // below 32 channels most lanes would idle
  if (PH < 1 || PW < 1 || PH > kMaxBins || PW > kMaxBins) return false;
  for (int l = 0; l < lv.num_levels; ++l) {
    if (lv.height[l] < 1 || lv.width[l] < 1) return false;
    if (lv.height[l] * lv.width[l] * lv.channels >= (1LL << 31)) return false;  // 32-bit in-image offsets
    if (((uintptr_t)lv.ptr[l]) % 16 != 0) return false;
  }
  return true;
}

template <typename T, int CH>
int launch_fwd_ag(const LevelArgs& lv, const void* rois, const int32_t* levels, void* out, long long K, int C, int PH,
                  int PW, int sr, bool aligned, cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CT = (C + 32 * CH - 1) / (32 * CH);
  roi_align_fwd_anygrid_kernel<T, CH><<<(unsigned)(K * CT), kAgWarps * 32, 0, st>>>(
      tab, (const float*)rois, levels, (T*)out, C, PH, PW, sr, aligned, CT);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T, int CH>
int launch_bwd_ag(const LevelArgs& lv, const void* rois, const int32_t* levels, const void* grad, const int64_t* gs,
                  long long K, int C, int PH, int PW, int sr, bool aligned, cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CT = (C + 32 * CH - 1) / (32 * CH);
  roi_align_bwd_anygrid_kernel<T, CH><<<(unsigned)(K * CT), kAgWarps * 32, 0, st>>>(
      tab, (const float*)rois, levels, (const T*)grad, gs[0], gs[2], gs[3], (int)lv.batch, C, PH, PW, sr, aligned, CT);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int roi_align_fwd_anygrid(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels, void* out,
                          int out_layout, long long K, int PH, int PW, int sr, bool aligned, cudaStream_t st,
                          int* taken) {
  *taken = 0;
  if (!anygrid_enabled() || !anygrid_eligible(lv, dtype, PH, PW) || out_layout != B200DET_NHWC) return 0;
  if (((uintptr_t)out) % 16 != 0) return 0;
  const int C = (int)lv.channels;
  if (K == 0) { *taken = 1; return 0; }
  const int ch = lane_channels(dtype, C);
  if (K * ((C + 32 * ch - 1) / (32 * ch)) > 0x7fffffffLL) return 0;
  int rc = 0;
#define B200DET_AG_FWD(T)                                                                               \
  rc = ch == 8 ? launch_fwd_ag<T, 8>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st)              \
               : launch_fwd_ag<T, 4>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st)
  switch (dtype) {
    case B200DET_F32: B200DET_AG_FWD(float); break;
    case B200DET_F16: B200DET_AG_FWD(__half); break;
    case B200DET_BF16: B200DET_AG_FWD(__nv_bfloat16); break;
    default: return 0;
  }
#undef B200DET_AG_FWD
  if (rc) return rc;
  *taken = 1;
  return 0;
}

// The gradient buffers must already be zero (the caller's memset): this kernel only adds.
int roi_align_bwd_anygrid(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                          const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned, cudaStream_t st,
                          int* taken) {
  *taken = 0;
  if (!anygrid_enabled() || !anygrid_eligible(lv, dtype, PH, PW)) return 0;
  if (K == 0 || lv.batch == 0) return 0;
  const int C = (int)lv.channels;
  const int ch = lane_channels(dtype, C);
  // one vector load of the lane's channels per bin: unit channel stride, every bin start aligned to the vector
  const long long vec = dtype == B200DET_F32 ? 4 : ch;
  if (gs[1] != 1 || gs[0] % vec != 0 || gs[2] % vec != 0 || gs[3] % vec != 0 || ((uintptr_t)grad) % 16 != 0) return 0;
  if (K * ((C + 32 * ch - 1) / (32 * ch)) > 0x7fffffffLL) return 0;
  int rc = 0;
#define B200DET_AG_BWD(T)                                                                               \
  rc = ch == 8 ? launch_bwd_ag<T, 8>(lv, rois, levels, grad, gs, K, C, PH, PW, sr, aligned, st)         \
               : launch_bwd_ag<T, 4>(lv, rois, levels, grad, gs, K, C, PH, PW, sr, aligned, st)
  switch (dtype) {
    case B200DET_F32: B200DET_AG_BWD(float); break;
    case B200DET_F16: B200DET_AG_BWD(__half); break;
    case B200DET_BF16: B200DET_AG_BWD(__nv_bfloat16); break;
    default: return 0;
  }
#undef B200DET_AG_BWD
  if (rc) return rc;
  *taken = 1;
  return 0;
}

}  // namespace b200det
