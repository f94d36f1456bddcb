// roi_align_bwd_sm100.cu -- the sm_100a RoIAlign BACKWARD kernel for the
// detection-model case (channels_last, C % 128 == 0, sampling_ratio 1 or 2,
// pooled_width * sampling_ratio <= 32); the forward lives in
// roi_align_fwd_sm100.cu, everything else in roi_align_generic.cu.
//
// Math.  The backward is the transpose of the separable forward
//     grad_in[r][c] += sum_ph Wy[ph][r] * ( sum_pw Wx[pw][c] * grad[ph][pw] ).
// A warp owns one task = (RoI, channel tile, PAIR of pooled rows); a lane owns 8
// channels (4 when C % 256 != 0).  It sweeps the x samples left to right keeping,
// per pooled row of the pair, the partial sums of the two feature columns under
// the current sample in registers (X / Y alternate as in the forward, nothing is
// moved).  When a column leaves the window it is flushed: ONE vector
// red.global.add per distinct feature row of the pair (w0 * acc0 + w1 * acc1;
// 16 bytes per lane, 512 B per warp instruction) instead of the reference's 4
// scalar atomics per sample and channel (cuda/roi_align_kernel.cu:303-327).
//
// Memory system.  The zero fill is FUSED into the same persistent launch as a phase in front:
// chunks of ~125 KB are claimed from a second counter and a warp scatters only after every chunk is
// reported done (no co-residency assumption: a warp only ever waits for chunks that running warps have
// already claimed).  A segmented schedule (RoIs counting-sorted by image / level / band of rows, fill just
// in time) was measured in round 1 and lost (profiles/r01_bwd_rows_segments.md); it is gone.  Levels whose
// map is too big and too sparsely covered for L2 to act as the accumulator are left to the write-once
// strip kernel (roi_align_bwd_strip_sm100.cu): `skip_levels` names them, their RoIs are skipped here and
// their maps are not zero-filled here.
#include <stdlib.h>

#include <mutex>

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kBwdWarps = 4;
constexpr int kMaxRows = 8;        // 2 pooled rows x 2 samples x 2 taps
constexpr int kPlanSlots = 8;      // calls in flight on different streams before a slot is reused

// per-call scheduling state; the three hot words sit in different 128-byte lines: 2368 warps hammer them
struct BwdPlan {
  int task_ctr;                    // next scatter task
  int pad0[31];
  int zero_ctr;                    // next zero chunk
  int pad1[31];
  int zero_done;                   // chunks finished
  int pad2[31];
};

struct __align__(16) BYEnt { int rowoff; float w0; float w1; int pad; };

struct BwdWarpTables {
  XTab x;
  BYEnt y[kMaxRows];
};

// ---- y axis of a pair of pooled rows ----------------------------------------------------
// Lane t < 4*SR computes one tap (pooled row 2*rp + t / (2*SR), sample, lo/hi); taps
// on the same feature row are merged (first occurrence leads) into (rowoff, w0, w1)
// with w0 / w1 the weights towards the first / second pooled row.  Returns the
// number of distinct rows.  All lanes of the warp must call it.
template <int SR>
__device__ __forceinline__ int build_y_pair(const RoiGeom<float>& g, int H, int row_stride, int PH, int rp, int lane,
                                            BYEnt* yt) {
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
    e.rowoff = row * row_stride;
    e.w0 = w0;
    e.w1 = w1;
    e.pad = 0;
    yt[__popc(ball & ((1u << lane) - 1u))] = e;
  }
  return __popc(ball);
}

// ---- upstream gradient of one pooled bin (the lane's channels) ---------------------------
// first channel (relative to the lane's first) of element i of the lane vector
template <typename T, int CH>
__device__ __forceinline__ int lane_elem_channel(int i) {
  return (sizeof(T) == 4 && CH == 8 && i >= 4) ? 124 + i : i;
}

template <typename T, int CH, bool VEC> struct GBuf;
template <typename T, int CH> struct GBuf<T, CH, true> {   // unit channel stride: vector loads, converted on use
  typename PixVec<T, CH>::raw r;
  __device__ __forceinline__ void load(const T* p, long long) { PixVec<T, CH>::ld_stream(r, p); }
  __device__ __forceinline__ void get(float2 (&v)[CH / 2]) const { PixVec<T, CH>::f2(r, v); }
};
template <typename T, int CH> struct GBuf<T, CH, false> {  // arbitrary strides (roi_align_kernel.cu:253-256)
  float2 v_[CH / 2];
  __device__ __forceinline__ void load(const T* p, long long gs_c) {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i)
      v_[i] = make_float2((float)Io<T>::ld(p + lane_elem_channel<T, CH>(2 * i) * gs_c),
                          (float)Io<T>::ld(p + lane_elem_channel<T, CH>(2 * i + 1) * gs_c));
  }
  __device__ __forceinline__ void get(float2 (&v)[CH / 2]) const {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i) v[i] = v_[i];
  }
};

// ---- one task: (RoI, channel tile, pair of pooled rows) ------------------------------------
template <typename T, int CH, int SR, bool VEC>
__device__ __forceinline__ void bwd_sweep(const BwdWarpTables& tb, int nrows, T* __restrict__ base,
                                          const T* __restrict__ g0p, const T* __restrict__ g1p, bool has1,
                                          long long gs_c, long long gs_w, int PW) {
  using PV = PixVec<T, CH>;
  constexpr int H2 = CH / 2;
  const float2 zero2 = make_float2(0.f, 0.f);
  float2 X0[H2], X1[H2], Y0[H2], Y1[H2], G0[H2], G1[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X0[i] = X1[i] = Y0[i] = Y1[i] = G0[i] = G1[i] = zero2; }
  GBuf<T, CH, VEC> n0, n1;
  n0.load(g0p, gs_c);
  if (has1) n1.load(g1p, gs_c);

  int j = 0;
  // samples j .. end-1 add into the window (lo, hi); the gradient of the next pooled
  // column is requested as soon as the current one has been taken over
  auto samples = [&](int end, float2 (&lo0)[H2], float2 (&hi0)[H2], float2 (&lo1)[H2], float2 (&hi1)[H2]) {
#pragma unroll 1
    for (; j < end; ++j) {
      if (SR == 1 || !(j & 1)) {
        n0.get(G0);
        if (has1) n1.get(G1);
        const int pwn = j / SR + 1;
        if (pwn < PW) {
          n0.load(g0p + pwn * gs_w, gs_c);
          if (has1) n1.load(g1p + pwn * gs_w, gs_c);
        }
      }
      const float4 s = tb.x.samp[j];
      if (__float_as_int(s.z)) {
        const float2 hx = make_float2(s.x, s.x), lx = make_float2(s.y, s.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          lo0[i] = __ffma2_rn(hx, G0[i], lo0[i]);
          hi0[i] = __ffma2_rn(lx, G0[i], hi0[i]);
          lo1[i] = __ffma2_rn(hx, G1[i], lo1[i]);
          hi1[i] = __ffma2_rn(lx, G1[i], hi1[i]);
        }
      }
    }
  };
  // a column leaves the window: one vector red per distinct feature row
  auto flush = [&](float2 (&a0)[H2], float2 (&a1)[H2], int co) {
#pragma unroll 1
    for (int r = 0; r < nrows; ++r) {
      const BYEnt e = tb.y[r];
      const float2 w0 = make_float2(e.w0, e.w0), w1 = make_float2(e.w1, e.w1);
      float2 v[H2];
#pragma unroll
      for (int i = 0; i < H2; ++i) v[i] = __ffma2_rn(w0, a0[i], __fmul2_rn(w1, a1[i]));
      PV::red(base + (e.rowoff + co), v);
    }
#pragma unroll
    for (int i = 0; i < H2; ++i) { a0[i] = zero2; a1[i] = zero2; }
  };

  const int ncol = tb.x.ncol;
  samples(tb.x.ev_end[0], Y0, X0, Y1, X1);  // leading empty samples only advance the gradient buffers
  int e = 0;
  while (e < ncol) {
    // even event: its column accumulates in X, which still holds column e-2
    if (e >= 2) flush(X0, X1, tb.x.colofs[e - 2]);
    samples(tb.x.ev_end[e + 1], Y0, X0, Y1, X1);
    if (++e >= ncol) break;
    // odd event: same with Y
    if (e >= 2) flush(Y0, Y1, tb.x.colofs[e - 2]);
    samples(tb.x.ev_end[e + 1], X0, Y0, X1, Y1);
    ++e;
  }
  // the last two columns are still in registers
  if (ncol >= 2) {
    if ((ncol - 2) & 1) flush(Y0, Y1, tb.x.colofs[ncol - 2]); else flush(X0, X1, tb.x.colofs[ncol - 2]);
  }
  if ((ncol - 1) & 1) flush(Y0, Y1, tb.x.colofs[ncol - 1]); else flush(X0, X1, tb.x.colofs[ncol - 1]);
}

template <typename T, int CH, int SR>
__device__ __forceinline__ void bwd_task(const LevelTable<T>& lv, const float* __restrict__ rois,
                                         const int32_t* __restrict__ levels, const T* __restrict__ grad,
                                         long long gs_k, long long gs_c, long long gs_h, long long gs_w, bool vec,
                                         long long k, int ct, int rp, int N, int C, int PH, int PW, bool aligned,
                                         BwdWarpTables& tb, int lane) {
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  if ((unsigned)g.batch >= (unsigned)N) return;  // the reference has no check (roi_align_kernel.cu:90); we skip
  build_x<SR>(g, W, C, PW, lane, tb.x);
  const int nrows = build_y_pair<SR>(g, H, W * C, PH, rp, lane, tb.y);
  __syncwarp();
  if (nrows > 0 && tb.x.ncol > 0) {
    const int ch = ct * (32 * CH) + lane_channel<T, CH>(lane);
    T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
    const int ph0 = 2 * rp;
    const bool has1 = ph0 + 1 < PH;
    const T* g0p = grad + k * gs_k + (long long)ch * gs_c + (long long)ph0 * gs_h;
    const T* g1p = g0p + gs_h;
    if (vec) bwd_sweep<T, CH, SR, true>(tb, nrows, base, g0p, g1p, has1, gs_c, gs_w, PW);
    else bwd_sweep<T, CH, SR, false>(tb, nrows, base, g0p, g1p, has1, gs_c, gs_w, PW);
  }
  __syncwarp();  // the tables are rewritten by the next task
}

// ---- row-gather formulation (MODE 1) ---------------------------------------------------------
// The pair sweep above sends a feature row to L2 once per pooled-row pair that taps it: for the
// mask-head shape (14x14 bins, 28 samples over 4..28 pixels) that is 4.1 GB of reds for a 2.9 GB
// footprint.  Here a warp owns FEATURE rows of the RoI instead: for row r it gathers
//     G[pw] = sum_{ph taps r} Wy[ph][r] * grad[ph][pw]
// (<= 4 pooled rows per pass, next pooled column's loads in flight), spreads G over the two feature
// columns under each x sample exactly like the pair sweep, and flushes every column of the row with
// ONE red: each footprint pixel of a RoI leaves the SM exactly once.  The upstream gradient is re-read
// by the 2-3 feature rows a pooled row taps; those re-reads hit L1/L2, not HBM.
// Task = (RoI, channel tile, one of kRowChunks contiguous chunks of the RoI's feature rows).
constexpr int kRowChunks = 4;

template <typename T, int CH, int SR, int NR>
__device__ __forceinline__ void rows_sweep(const XTab& xt, T* __restrict__ rowbase, const T* const (&gp)[NR],
                                           const float (&wy)[NR], long long gs_w, int PW) {
  using PV = PixVec<T, CH>;
  constexpr int H2 = CH / 2;
  const float2 zero2 = make_float2(0.f, 0.f);
  typename PV::raw n[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) PV::ld(n[r], gp[r]);
  float2 X[H2], Y[H2], G[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X[i] = Y[i] = G[i] = zero2; }
  int j = 0;
  auto samples = [&](int end, float2 (&lo)[H2], float2 (&hi)[H2]) {
#pragma unroll 1
    for (; j < end; ++j) {
      if (SR == 1 || !(j & 1)) {
        const int pwn = j / SR + 1;
        const long long nxt = (long long)min(pwn, PW - 1) * gs_w;  // past the end: re-read the last column (L1)
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          float2 t[H2];
          PV::f2(n[r], t);
          const float2 w2 = make_float2(wy[r], wy[r]);
#pragma unroll
          for (int i = 0; i < H2; ++i) G[i] = (r == 0) ? __fmul2_rn(w2, t[i]) : __ffma2_rn(w2, t[i], G[i]);
          PV::ld(n[r], gp[r] + nxt);
        }
      }
      const float4 s = xt.samp[j];
      if (__float_as_int(s.z)) {
        const float2 hx = make_float2(s.x, s.x), lx = make_float2(s.y, s.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          lo[i] = __ffma2_rn(hx, G[i], lo[i]);
          hi[i] = __ffma2_rn(lx, G[i], hi[i]);
        }
      }
    }
  };
  auto flush = [&](float2 (&a)[H2], int co) {
    PV::red(rowbase + co, a);
#pragma unroll
    for (int i = 0; i < H2; ++i) a[i] = zero2;
  };
  const int ncol = xt.ncol;
  samples(xt.ev_end[0], Y, X);
  int e = 0;
  while (e < ncol) {
    if (e >= 2) flush(X, xt.colofs[e - 2]);
    samples(xt.ev_end[e + 1], Y, X);
    if (++e >= ncol) break;
    if (e >= 2) flush(Y, xt.colofs[e - 2]);
    samples(xt.ev_end[e + 1], X, Y);
    ++e;
  }
  if (ncol >= 2) {
    if ((ncol - 2) & 1) flush(Y, xt.colofs[ncol - 2]); else flush(X, xt.colofs[ncol - 2]);
  }
  if ((ncol - 1) & 1) flush(Y, xt.colofs[ncol - 1]); else flush(X, xt.colofs[ncol - 1]);
}

// takes the NR lowest pooled rows out of pmask (lane p holds the weight wp of pooled row p) and sweeps them
template <typename T, int CH, int SR, int NR>
__device__ __forceinline__ void rows_pass(const XTab& xt, T* __restrict__ rowbase, const T* __restrict__ gbase,
                                          long long gs_h, long long gs_w, int PW, float wp, unsigned& pmask) {
  const T* gp[NR];
  float wy[NR];
#pragma unroll
  for (int i = 0; i < NR; ++i) {
    const int ph = __ffs(pmask) - 1;
    pmask &= pmask - 1;
    gp[i] = gbase + ph * gs_h;
    wy[i] = __shfl_sync(kFull, wp, ph);
  }
  rows_sweep<T, CH, SR, NR>(xt, rowbase, gp, wy, gs_w, PW);
}

template <typename T, int CH, int SR>
__device__ __forceinline__ void bwd_rows_task(const LevelTable<T>& lv, const float* __restrict__ rois,
                                              const int32_t* __restrict__ levels, const T* __restrict__ grad,
                                              long long gs_k, long long gs_h, long long gs_w, long long k, int ct,
                                              int chunk, int N, int C, int PH, int PW, bool aligned,
                                              BwdWarpTables& tb, int lane) {
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + k * 5, lv.scale[l], PH, PW, SR, aligned);
  if ((unsigned)g.batch >= (unsigned)N) return;
  build_x<SR>(g, W, C, PW, lane, tb.x);
  // lane j owns y sample j (PH * SR <= 32): its two taps, 1/count folded in
  int lo = -1, hi = -1;
  float wl = 0.f, wh = 0.f;
  if (lane < PH * SR) {
    const AxisTap<float> t = axis_tap<float>(sample_coord<float>(g.start_h, g.bin_h, lane / SR, lane % SR, SR), H);
    if (t.valid) {
      lo = t.lo; hi = t.hi;
      wl = t.wl * (1.0f / (float)(SR * SR));
      wh = t.wh * (1.0f / (float)(SR * SR));
    }
  }
  const int r_first = __reduce_min_sync(kFull, lo >= 0 ? lo : 0x7fffffff);
  const int r_last = __reduce_max_sync(kFull, hi);
  __syncwarp();
  if (r_last >= 0 && tb.x.ncol > 0) {
    const int nr = r_last - r_first + 1;
    const int cs = (nr + kRowChunks - 1) / kRowChunks;
    const int ra = r_first + chunk * cs, rb = min(ra + cs, r_last + 1);
    const int ch = ct * (32 * CH) + lane_channel<T, CH>(lane);
    T* base = lv.ptr[l] + (long long)g.batch * lv.sn[l] + ch;
    const T* gbase = grad + k * gs_k + ch;
    const long long row_stride = (long long)W * C;
    for (int r = ra; r < rb; ++r) {
      const bool m = (lo == r) || (hi == r);
      const float w = (lo == r ? wl : 0.f) + (hi == r ? wh : 0.f);
      // lane p < PH: merged weight of pooled row p towards feature row r
      const unsigned sm = __ballot_sync(kFull, m);
      float wp = 0.f;
#pragma unroll
      for (int i = 0; i < SR; ++i) wp += __shfl_sync(kFull, w, (lane * SR + i) & 31);
      const bool mp = lane < PH && ((sm >> (lane * SR)) & ((1u << SR) - 1u)) != 0u;
      unsigned pmask = __ballot_sync(kFull, mp);
      T* rowbase = base + r * row_stride;
      while (pmask) {
        switch (min(__popc(pmask), 4)) {
          case 1: rows_pass<T, CH, SR, 1>(tb.x, rowbase, gbase, gs_h, gs_w, PW, wp, pmask); break;
          case 2: rows_pass<T, CH, SR, 2>(tb.x, rowbase, gbase, gs_h, gs_w, PW, wp, pmask); break;
          case 3: rows_pass<T, CH, SR, 3>(tb.x, rowbase, gbase, gs_h, gs_w, PW, wp, pmask); break;
          default: rows_pass<T, CH, SR, 4>(tb.x, rowbase, gbase, gs_h, gs_w, PW, wp, pmask); break;
        }
      }
    }
  }
  __syncwarp();  // the tables are rewritten by the next task
}

// ---- persistent kernel -----------------------------------------------------------------------
struct BwdArgs {
  const float* rois;
  const int32_t* levels;
  const void* grad;
  long long gs_k, gs_c, gs_h, gs_w;
  BwdPlan* plan;        // nullptr: static task assignment, the caller has zeroed
  int ztotal;           // zero chunks of the whole call; 0: no zero fill
  long long K;
  int N, C, PH, PW, CT, RP;
  int zero_chunk;       // bytes per zero chunk (multiple of 512); 0: no zero fill
  unsigned skip_levels; // bit l: the RoIs of level l are not this launch's (another launch or the strip kernel has them)
  unsigned nozero_levels;  // bit l: the map of level l is not zero-filled by this launch
  bool vec, aligned;
  int mode;             // 0: pooled-row pair sweep, 1: feature-row gather
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// writes zero chunk z (levels in order, skipped levels have none) and reports it done; all lanes participate
template <typename T>
__device__ __forceinline__ void zero_chunk(const LevelTable<T>& lv, const BwdArgs& a, int z, int lane) {
  char* p = nullptr;
  long long bytes = 0;
  int c = z;
  for (int l = 0; l < lv.num_levels; ++l) {
    if ((a.nozero_levels >> l) & 1) continue;
    bytes = (long long)a.N * lv.sn[l] * (long long)sizeof(T);
    const int n = (int)((bytes + a.zero_chunk - 1) / a.zero_chunk);
    if (c < n) { p = reinterpret_cast<char*>(lv.ptr[l]); break; }
    c -= n;
  }
  if (p) {
    const long long off = (long long)c * a.zero_chunk;
    const int len = (int)min((long long)a.zero_chunk, bytes - off);  // multiple of 256: C % 128 == 0
    p += off;
    const uint4 zz = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll 4
    for (int o = lane * 16; o < len; o += 512) *reinterpret_cast<uint4*>(p + o) = zz;
  }
  __threadfence();
  __syncwarp();
  if (lane == 0) atomicAdd(&a.plan->zero_done, 1);
}

template <typename T, int CH, int SR, int MODE>
__global__ void __launch_bounds__(kBwdWarps * 32, 4) roi_align_bwd_sm100_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                              const __grid_constant__ BwdArgs a) {
  __shared__ BwdWarpTables tabs[kBwdWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  BwdWarpTables& tb = tabs[warp];
  const T* grad = reinterpret_cast<const T*>(a.grad);
  const int tasks_per_roi = a.RP * a.CT;  // MODE 1: RP = kRowChunks
  const long long total = a.K * tasks_per_roi;

  auto run_task = [&](long long t) {
    const int rp = (int)(t % a.RP);
    const int ct = (int)((t / a.RP) % a.CT);
    const long long k = t / tasks_per_roi;
    if (a.skip_levels && ((a.skip_levels >> (a.levels ? __ldg(a.levels + k) : 0)) & 1u)) return;
    if (MODE == 1)
      bwd_rows_task<T, CH, SR>(lv, a.rois, a.levels, grad, a.gs_k, a.gs_h, a.gs_w, k, ct, rp, a.N, a.C, a.PH, a.PW,
                               a.aligned, tb, lane);
    else
      bwd_task<T, CH, SR>(lv, a.rois, a.levels, grad, a.gs_k, a.gs_c, a.gs_h, a.gs_w, a.vec, k, ct, rp, a.N, a.C,
                          a.PH, a.PW, a.aligned, tb, lane);
  };

  if (a.plan == nullptr) {  // static assignment, no zero fill (the caller did it)
    for (long long t = blockIdx.x * (long long)kBwdWarps + warp; t < total; t += (long long)gridDim.x * kBwdWarps)
      run_task(t);
    return;
  }

  // Zero chunks and tasks are handed out by one global counter each.  A warp first helps until every chunk
  // is claimed, waits once until all of them are reported done -- claimed chunks belong to running warps, so
  // this needs no co-residency -- and then scatters.
  BwdPlan* plan = a.plan;
  if (a.ztotal > 0) {
    while (true) {
      int z = 0;
      if (lane == 0) z = atomicAdd(&plan->zero_ctr, 1);
      z = __shfl_sync(kFull, z, 0);
      if (z >= a.ztotal) break;
      zero_chunk<T>(lv, a, z, lane);
    }
    if (lane == 0)
      while (ld_acquire(&plan->zero_done) < a.ztotal) __nanosleep(64);
    __syncwarp();
    __threadfence();
  }
  while (true) {
    int t = 0;
    if (lane == 0) t = atomicAdd(&plan->task_ctr, 1);
    t = __shfl_sync(kFull, t, 0);
    if (t >= total) break;
    run_task(t);
  }
}

// ---- host side ---------------------------------------------------------------------------------
// The plan ring is the one allocation the scatter path makes on its own (kPlanSlots x 384 bytes per
// device, first backward call).  Slots are handed out under a mutex; every slot carries the event of its
// last user and the next user's stream waits for it, so calls on different streams or from different host
// threads never share a live plan.  If the ring cannot be had (e.g. the first call happens inside a
// stream capture) the call falls back to memset + static task assignment.
struct PlanSlot {
  cudaEvent_t done = nullptr;
  bool used = false;
};
struct DeviceScratch {
  char* buf = nullptr;
  PlanSlot slot[kPlanSlots];
  unsigned next = 0;
  int sms = 0;
  bool tried = false;
};
DeviceScratch g_scratch[64];
std::mutex g_scratch_mutex;

constexpr size_t kPlanBytes = (sizeof(BwdPlan) + 255) & ~(size_t)255;

// plan slot for this call (its stream already waits for the slot's previous user); nullptr when it cannot be had
BwdPlan* plan_slot(cudaStream_t st, int* sms, PlanSlot** out) {
  int dev = 0;
  *sms = 148;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  DeviceScratch& s = g_scratch[dev];
  if (!s.tried) {
    s.tried = true;
    cudaDeviceGetAttribute(&s.sms, cudaDevAttrMultiProcessorCount, dev);
    void* p = nullptr;
    if (cudaMalloc(&p, kPlanBytes * kPlanSlots) != cudaSuccess) (void)cudaGetLastError();
    else s.buf = (char*)p;
  }
  if (s.sms > 0) *sms = s.sms;
  if (!s.buf) return nullptr;
  const unsigned i = s.next++ % kPlanSlots;
  PlanSlot& sl = s.slot[i];
  if (!sl.done && cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming) != cudaSuccess) {
    (void)cudaGetLastError();
    sl.done = nullptr;
    return nullptr;
  }
  if (sl.used && cudaStreamWaitEvent(st, sl.done, 0) != cudaSuccess) {
    (void)cudaGetLastError();
    return nullptr;
  }
  *out = &sl;
  return (BwdPlan*)(s.buf + kPlanBytes * i);
}

int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

template <typename T, int CH, int SR, int MODE>
int launch_bwd_m(const LevelTable<T>& tab, const BwdArgs& a, int sms, cudaStream_t st) {
  int per_sm = 0;
  B200DET_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, roi_align_bwd_sm100_kernel<T, CH, SR, MODE>,
                                                                kBwdWarps * 32, 0));
  if (per_sm < 1) per_sm = 1;
  const long long total = a.K * a.RP * a.CT;
  long long blocks = (long long)sms * per_sm;
  const long long need = (total + kBwdWarps - 1) / kBwdWarps;
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  roi_align_bwd_sm100_kernel<T, CH, SR, MODE><<<(unsigned)blocks, kBwdWarps * 32, 0, st>>>(tab, a);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T, int CH, int SR>
int launch_bwd_k(const LevelTable<T>& tab, const BwdArgs& a, int sms, cudaStream_t st) {
  return a.mode == 1 ? launch_bwd_m<T, CH, SR, 1>(tab, a, sms, st) : launch_bwd_m<T, CH, SR, 0>(tab, a, sms, st);
}

template <typename T>
int launch_bwd_one(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
                   const int32_t* levels, long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init,
                   unsigned skip_levels, unsigned nozero_levels, int mode, cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CH = C % 256 == 0 ? 8 : 4;
  BwdArgs a;
  a.rois = (const float*)rois;
  a.levels = levels;
  a.grad = grad;
  a.gs_k = gs[0]; a.gs_c = gs[1]; a.gs_h = gs[2]; a.gs_w = gs[3];
  a.K = K;
  a.N = (int)lv.batch; a.C = C; a.PH = PH; a.PW = PW;
  a.CT = C / (32 * CH);
  a.skip_levels = skip_levels;
  a.nozero_levels = nozero_levels;
  a.mode = mode;
  a.RP = a.mode == 1 ? kRowChunks : (PH + 1) / 2;
  a.vec = vec; a.aligned = aligned;
  a.plan = nullptr; a.zero_chunk = 0; a.ztotal = 0;

  int sms = 148;
  PlanSlot* slot = nullptr;
  const long long total = K * a.RP * a.CT;
  BwdPlan* plan = total < 0x7fffffffLL ? plan_slot(st, &sms, &slot) : nullptr;
  size_t fill_bytes = 0;
  for (int l = 0; l < lv.num_levels; ++l)
    if (!((nozero_levels >> l) & 1)) fill_bytes += (size_t)lv.batch * lv.height[l] * lv.width[l] * C * sizeof(T);
  if (plan) {
    B200DET_CUDA_OK(cudaMemsetAsync(plan, 0, sizeof(BwdPlan), st));
    a.plan = plan;
    if (zero_init && fill_bytes) {
      // the fill is a phase of its own in front, and big chunks are faster (measured: 16 KB 1.07 ms, 27 KB
      // 1.00, 51 KB 0.96 for cfg3 fp32), so about four chunks per resident warp
      long long zero_chunk = ((long long)((double)fill_bytes / ((double)sms * 64.0)) + 4095) / 4096 * 4096;
      if (zero_chunk < 16384) zero_chunk = 16384;
      if (zero_chunk > (1 << 20)) zero_chunk = 1 << 20;
      a.zero_chunk = (int)zero_chunk;
      long long zt = 0;
      for (int l = 0; l < lv.num_levels; ++l) {
        if ((nozero_levels >> l) & 1) continue;
        const long long bytes = (long long)a.N * lv.height[l] * lv.width[l] * C * (long long)sizeof(T);
        zt += (bytes + zero_chunk - 1) / zero_chunk;
      }
      if (zt >= 0x7fffffffLL) return 1;  // cannot happen: >= 16 KB per chunk
      a.ztotal = (int)zt;
    }
  } else if (zero_init) {
    for (int l = 0; l < lv.num_levels; ++l) {
      const size_t bytes = (size_t)lv.batch * lv.channels * lv.height[l] * lv.width[l] * sizeof(T);
      if (bytes && !((nozero_levels >> l) & 1)) B200DET_CUDA_OK(cudaMemsetAsync(lv.ptr[l], 0, bytes, st));
    }
  }
  int rc;
  if (CH == 8)
    rc = sr == 1 ? launch_bwd_k<T, 8, 1>(tab, a, sms, st) : launch_bwd_k<T, 8, 2>(tab, a, sms, st);
  else
    rc = sr == 1 ? launch_bwd_k<T, 4, 1>(tab, a, sms, st) : launch_bwd_k<T, 4, 2>(tab, a, sms, st);
  if (slot) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (cudaEventRecord(slot->done, st) == cudaSuccess) slot->used = true;
    else (void)cudaGetLastError();
  }
  return rc;
}

// Two task shapes (file header): the pooled-row PAIR sweep (MODE 0) and the feature-ROW gather (MODE 1, one red
// per footprint pixel but the upstream gradient re-read by every feature row a pooled row taps).  Measured on
// B200 (profiles/r01_bwd_rows_segments.md): the row gather wins for box-head shapes (7x7 bins, samples >= 1 pixel
// apart: a pooled-row pair spreads over up to 8 feature rows) and loses for the mask-head shape (14x14: the
// re-reads evict the gradient lines the reds work on), so it is taken when a RoI has at most 16 y samples.
// Choosing the shape per pyramid level (row gather for the dense coarse levels only, two launches) was measured
// in round 2 and lost as well (0.97 -> 1.12-1.19 ms on cfg3, profiles/r02_bwd_study.md).  B200DET_BWD_MODE=0 / 1
// forces one shape (tests/test_roi_align_gpu.py runs both on every fast-path case).
template <typename T>
int launch_bwd(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
               const int32_t* levels, long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init,
               unsigned skip_levels, cudaStream_t st) {
  const int mode_env = env_int("B200DET_BWD_MODE", -1);
  const bool rows_ok = vec && PH * sr <= 32;
  const int mode = rows_ok && (mode_env == 1 || (mode_env < 0 && PH * sr <= 16)) ? 1 : 0;
  return launch_bwd_one<T>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, skip_levels,
                           skip_levels, mode, st);
}

}  // namespace

int roi_align_bwd_fast(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                       const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned, bool zero_init,
                       unsigned skip_levels, cudaStream_t st, int* taken) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr)) return 0;
  if (K == 0 || lv.batch == 0) return 0;  // nothing to scatter: the caller's memset path handles it
  const size_t es = dtype_size(dtype);
  // vector loads of the gradient need unit channel stride and 16-byte alignment of every row start
  const long long per16 = 16 / (long long)es;
  const bool vec = gs[1] == 1 && gs[0] % per16 == 0 && gs[2] % per16 == 0 && gs[3] % per16 == 0 &&
                   ((uintptr_t)grad) % 16 == 0;
  const int C = (int)lv.channels;
  int rc = 0;
  switch (dtype) {
    case B200DET_F32: rc = launch_bwd<float>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, skip_levels, st); break;
    case B200DET_F16: rc = launch_bwd<__half>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, skip_levels, st); break;
    case B200DET_BF16: rc = launch_bwd<__nv_bfloat16>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, skip_levels, st); break;
    default: return 0;
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

void roi_align_bwd_release() {
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < 64; ++d) {
    DeviceScratch& s = g_scratch[d];
    if (!s.buf && !s.tried) continue;
    cudaSetDevice(d);
    if (s.buf) cudaFree(s.buf);
    for (int i = 0; i < kPlanSlots; ++i)
      if (s.slot[i].done) cudaEventDestroy(s.slot[i].done);
    s = DeviceScratch();
  }
  cudaSetDevice(cur);
}

}  // namespace b200det
