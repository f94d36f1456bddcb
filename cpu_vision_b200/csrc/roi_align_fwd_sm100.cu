// roi_align_fwd_sm100.cu -- the sm_100a RoIAlign FORWARD kernel for the
// detection-model case: channels_last feature maps, C a multiple of 128,
// sampling_ratio 1 or 2, pooled_width * sampling_ratio <= 32, pooled_height <= 32.
// Everything else goes to roi_align_generic.cu.
//
// bilinear_interpolate (cuda/roi_align_kernel.cu:14-66) is separable: the weight
// of pixel (r, c) for a sample is wy(r) * wx(c).  Summed over the sampling grid
// of a bin that gives, for a pooled row ph,
//     out[ph][pw] = sum_c Wx[pw][c] * ( sum_r Wy[ph][r] * in[r][c] ).
//
// CTA = one (RoI, channel tile), 4 warps (measured best of 2/4/7/8 on B200).  A lane owns 8 channels of every pixel
// (4 when C is not a multiple of 256), so a feature tap is one or two 16-byte
// loads and a warp covers 1 KB / 512 B of contiguous NHWC bytes.
//   * warp 0 tabulates the x axis once per CTA: per x sample its two tap weights,
//     and the ordered list of feature columns the sweep has to fetch ("column
//     events") with, per event, how many samples are complete once it arrived.
//   * warp 1 tabulates the y axis: per pooled row the <= 2*SR distinct feature
//     rows with their merged weights (1/count folded in; exact, count is 1 or 4).
//   * every warp then takes pooled rows (ph = warp, warp + 4, ...).  It sweeps
//     the column events left to right: the vertical sum of column e goes into
//     one of two register vectors X / Y (alternating, so the pair always holds
//     the two columns under the current sample and nothing is ever moved); the
//     registers of a consumed row are refilled at once with the column two events
//     ahead, so 2 x rows loads per warp are always in flight.  The code is
//     specialised on the number of distinct rows, so there is no predication in
//     the loop.  Each (row, column) of the footprint is loaded once per pooled
//     row; overlapping pooled rows of the same CTA meet in L1.
//   * fp32 math is packed FFMA2/FMUL2 (fma.rn.f32x2, new on sm_100): half the
//     FP instructions of a scalar formulation, same rounding.
#include <stdlib.h>

#include "sm100_common.cuh"

namespace b200det {
namespace {

// warps per CTA (>= 2: warps 0 and 1 build the tables); 2..4 were measured (profiles/r01_kbench_fwd_v3_warps_sweep.log), 4 wins.
// Register budget: 128 per thread for fp32 (16 warps/SM); 96 for 16-bit I/O (20 warps/SM) -- measured
// on B200: fp32 loses 15% at 96 registers (spills), bf16 gains 6%.
constexpr int kFwdWarps = 4;
int fwd_warps() { return kFwdWarps; }
struct __align__(16) FwdTables {
  XTab x;
  int y_nr[kMaxPH];                     // distinct feature rows of pooled row ph
  int y_off[kMaxPH][4];                 // row * W * C
  float y_w[kMaxPH][4];                 // merged weight / count
};

// ---- y axis (warp 1; lane = pooled row) --------------------------------------------------
// 2*SR taps (lo / hi of each y sample) in ascending order; equal neighbours are
// merged and empty samples dropped, so the sweep sees nr <= 2*SR distinct rows.
template <int SR>
__device__ __forceinline__ void build_y(const RoiGeom<float>& g, int H, int row_stride, int PH, int lane,
                                        FwdTables& tb) {
  if (lane >= PH) return;
  int off[4] = {0, 0, 0, 0};
  float w[4] = {0.f, 0.f, 0.f, 0.f};
  int n = 0, last_row = -1;
  auto put = [&](int row, float wt) {
    if (n > 0 && row == last_row) {
      // merge into slot n - 1
      if (n == 1) w[0] += wt; else if (n == 2) w[1] += wt; else if (n == 3) w[2] += wt; else w[3] += wt;
      return;
    }
    const int o = row * row_stride;
    if (n == 0) { off[0] = o; w[0] = wt; } else if (n == 1) { off[1] = o; w[1] = wt; }
    else if (n == 2) { off[2] = o; w[2] = wt; } else { off[3] = o; w[3] = wt; }
    last_row = row;
    ++n;
  };
#pragma unroll
  for (int iy = 0; iy < SR; ++iy) {
    const float y = sample_coord<float>(g.start_h, g.bin_h, lane, iy, SR);
    const AxisTap<float> t = axis_tap<float>(y, H);
    if (t.valid) {
      put(t.lo, t.wl * (1.0f / (float)(SR * SR)));
      put(t.hi, t.wh * (1.0f / (float)(SR * SR)));
    }
  }
  tb.y_nr[lane] = n;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    tb.y_off[lane][i] = off[i];
    tb.y_w[lane][i] = w[i];
  }
}

// ---- one pooled row -----------------------------------------------------------------------
// a pooled bin leaves the SM: one streaming store to the output and, in the push variant, one more to the same row
// of every peer buffer (peer GPUs' memory mapped over NVLink: the all-gather happens while the rows are produced)
template <typename T, int CH, bool MULTI>
__device__ __forceinline__ void store_bin(T* p, const float2 (&v)[CH / 2], const PeerOut& peers) {
  PixVec<T, CH>::st(p, v);
  if (MULTI) {
    for (int d = 0; d < peers.n; ++d)
      PixVec<T, CH>::st(reinterpret_cast<T*>(reinterpret_cast<char*>(p) + peers.delta[d]), v);
  }
}

template <typename T, int CH, int SR, int NR, bool MULTI>
__device__ __forceinline__ void sweep_row(const FwdTables& tb, const T* __restrict__ base, T* __restrict__ o,
                                          int ph, int C, int SX, const PeerOut& peers) {
  using PV = PixVec<T, CH>;
  constexpr int H2 = CH / 2;
  const T* rp[NR];
  float wv[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    rp[r] = base + tb.y_off[ph][r];
    wv[r] = tb.y_w[ph][r];
  }
  typename PV::raw A[NR], B[NR];
  float2 X[H2], Y[H2], acc[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X[i] = make_float2(0.f, 0.f); Y[i] = X[i]; acc[i] = X[i]; }

  // v = sum_r w_r * buf[r]; as soon as a row of `buf` is consumed its registers
  // are refilled with the same row of the column two events ahead (offset co2)
  auto vertical_refill = [&](typename PV::raw (&buf)[NR], float2 (&v)[H2], int co2) {
    float2 t[H2];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      PV::f2(buf[r], t);
      const float2 w2 = make_float2(wv[r], wv[r]);
#pragma unroll
      for (int i = 0; i < H2; ++i) v[i] = (r == 0) ? __fmul2_rn(w2, t[i]) : __ffma2_rn(w2, t[i], v[i]);
      PV::ld(buf[r], rp[r] + co2);
    }
  };
  int j = 0;
  // samples j .. end-1 read the window (lo, hi); a bin is stored when its last sample is done
  auto samples = [&](int end, const float2 (&lo)[H2], const float2 (&hi)[H2]) {
    for (; j < end; ++j) {
      const float4 s = tb.x.samp[j];
      if (__float_as_int(s.z)) {
        const float2 hx = make_float2(s.x, s.x), lx = make_float2(s.y, s.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) acc[i] = __ffma2_rn(lx, hi[i], __ffma2_rn(hx, lo[i], acc[i]));
      }
      if (SR == 1 || (j & 1)) {
        store_bin<T, CH, MULTI>(o, acc, peers);
        o += C;
#pragma unroll
        for (int i = 0; i < H2; ++i) acc[i] = make_float2(0.f, 0.f);
      }
    }
  };

  const int ncol = tb.x.ncol;
  samples(tb.x.ev_end[0], X, Y);
  if (ncol > 0) {
    // colofs[] is padded with two copies of its last entry, so the refills past the
    // end re-read a column that is already in L1 instead of branching
#pragma unroll
    for (int r = 0; r < NR; ++r) PV::ld(A[r], rp[r] + tb.x.colofs[0]);
#pragma unroll
    for (int r = 0; r < NR; ++r) PV::ld(B[r], rp[r] + tb.x.colofs[1]);
    int e = 0;
    while (true) {
      // event e sits in A and becomes X: window = (lo Y, hi X)
      vertical_refill(A, X, tb.x.colofs[e + 2]);
      samples(tb.x.ev_end[e + 1], Y, X);
      if (++e >= ncol) break;
      // event e sits in B and becomes Y: window = (lo X, hi Y)
      vertical_refill(B, Y, tb.x.colofs[e + 2]);
      samples(tb.x.ev_end[e + 1], X, Y);
      if (++e >= ncol) break;
    }
  }
  samples(SX, X, Y);  // nothing left unless the tables disagree; keeps every bin written
}

template <typename T, int CH, int SR, bool MULTI>
__global__ void __launch_bounds__(128, sizeof(T) == 2 ? 5 : 4) roi_align_fwd_sm100_kernel(
    LevelTable<T> lv, const float* __restrict__ rois, const int32_t* __restrict__ levels, T* __restrict__ out,
    int C, int PH, int PW, bool aligned, int CT, const __grid_constant__ PeerOut peers) {
  __shared__ FwdTables tb;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ct = blockIdx.x % CT;
  const long long k = blockIdx.x / CT;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  if (warp == 0) build_x<SR>(g, W, C, PW, lane, tb.x);
  else if (warp == 1) build_y<SR>(g, H, W * C, PH, lane, tb);
  __syncthreads();
  const int ch = ct * (32 * CH) + lane_channel<T, CH>(lane);
  const T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
  T* obase = out + k * PH * PW * (long long)C + ch;
  const int SX = PW * SR;
  const int nwarps = blockDim.x >> 5;
  for (int ph = warp; ph < PH; ph += nwarps) {
    T* o = obase + (long long)ph * PW * C;
    switch (tb.y_nr[ph]) {
      case 1: sweep_row<T, CH, SR, 1, MULTI>(tb, base, o, ph, C, SX, peers); break;
      case 2: sweep_row<T, CH, SR, 2, MULTI>(tb, base, o, ph, C, SX, peers); break;
      case 3: if (SR == 2) sweep_row<T, CH, SR, 3, MULTI>(tb, base, o, ph, C, SX, peers); break;
      case 4: if (SR == 2) sweep_row<T, CH, SR, 4, MULTI>(tb, base, o, ph, C, SX, peers); break;
      default: {  // every y sample of the row is empty: the bins are zero
        float2 z[CH / 2];
#pragma unroll
        for (int i = 0; i < CH / 2; ++i) z[i] = make_float2(0.f, 0.f);
        for (int pw = 0; pw < PW; ++pw) store_bin<T, CH, MULTI>(o + (long long)pw * C, z, peers);
      }
    }
  }
}

template <typename T, int CH, bool MULTI>
int launch_fwd(const LevelArgs& lv, const void* rois, const int32_t* levels, void* out, long long K, int C, int PH,
               int PW, int sr, bool aligned, cudaStream_t st, const PeerOut& peers) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CT = C / (32 * CH);
  const unsigned grid = (unsigned)(K * CT);
  const int nw = fwd_warps();
  if (sr == 1)
    roi_align_fwd_sm100_kernel<T, CH, 1, MULTI><<<grid, nw * 32, 0, st>>>(tab, (const float*)rois, levels, (T*)out,
                                                                                C, PH, PW, aligned, CT, peers);
  else
    roi_align_fwd_sm100_kernel<T, CH, 2, MULTI><<<grid, nw * 32, 0, st>>>(tab, (const float*)rois, levels, (T*)out,
                                                                                C, PH, PW, aligned, CT, peers);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int launch_fwd_t(const LevelArgs& lv, const void* rois, const int32_t* levels, void* out, long long K, int C, int PH,
                 int PW, int sr, bool aligned, cudaStream_t st, const PeerOut* peers) {
  PeerOut none;
  none.n = 0;
  if (peers && peers->n > 0) {
    if (C % 256 == 0) return launch_fwd<T, 8, true>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, *peers);
    return launch_fwd<T, 4, true>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, *peers);
  }
  if (C % 256 == 0) return launch_fwd<T, 8, false>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, none);
  return launch_fwd<T, 4, false>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, none);
}

}  // namespace

int roi_align_fwd_fast(const LevelArgs& lv, int dtype, const void* rois, const int32_t* levels, void* out,
                       int out_layout, long long K, int PH, int PW, int sr, bool aligned, cudaStream_t st,
                       int* taken, const PeerOut* peers) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr) || out_layout != B200DET_NHWC) return 0;
  if (((uintptr_t)out) % 16 != 0) return 0;
  const int C = (int)lv.channels;
  if (K == 0) { *taken = 1; return 0; }
  if (K * (C / 128) > 0x7fffffffLL) return 0;
  int rc = 0;
  switch (dtype) {
    case B200DET_F32: rc = launch_fwd_t<float>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, peers); break;
    case B200DET_F16: rc = launch_fwd_t<__half>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, peers); break;
    case B200DET_BF16: rc = launch_fwd_t<__nv_bfloat16>(lv, rois, levels, out, K, C, PH, PW, sr, aligned, st, peers); break;
    default: return 0;
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

}  // namespace b200det
