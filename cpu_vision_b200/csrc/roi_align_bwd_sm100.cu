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
// Memory system.  With a separate memset the reds of a 1.2 GB gradient miss L2:
// every line is written as zeros, read back from HBM by the first red and
// written again (ncu, r01: 61% of red sectors missed, +0.9 GB of reads).  So the
// zero fill is FUSED and PHASED: RoIs arrive grouped by image (the reference's
// convert_boxes_to_roi_format, ops/_utils.py:18-25, builds them that way); images
// are cut into groups whose gradient fits in about half the L2; a persistent
// grid works group by group: the zero chunks of group g+1 are interleaved, one
// per finished task, with the scatter tasks of group g, and nobody starts
// group g+1 before all of its chunks are written.  Reds then hit zero lines that
// are still dirty in L2 and every gradient line goes to HBM once.  Progress needs
// no co-residency: a warp only ever waits for chunks that running warps have
// already claimed.  RoIs that are not grouped by image (a tiny plan kernel
// checks) are handled as one group: still one launch, same results.
#include <stdlib.h>

#include <vector>

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kBwdWarps = 4;
constexpr int kMaxRows = 8;        // 2 pooled rows x 2 samples x 2 taps
constexpr int kMaxSegs = 32;       // segments (level, band of rows) per image
constexpr int kMaxGroups = 8192;   // images x segments a plan can index; more -> single group
constexpr int kPlanSlots = 8;      // calls in flight on different streams before a slot is reused

// A GROUP is one segment of one image: a band of rows of one level, i.e. one contiguous byte range of
// the gradient, small enough that two of them sit in L2 together (B200DET_BWD_GROUP_MB).  A RoI
// belongs to the group that holds the LAST feature row it taps; everything else it taps lies in
// groups that come earlier in processing order, so it is already zero-filled.
struct SegTable {
  int nseg;                             // per image; 0: no segmentation (one group for everything)
  int level[kMaxSegs], row0[kMaxSegs], rows[kMaxSegs];
  int first_seg[B200DET_MAX_LEVELS], band_rows[B200DET_MAX_LEVELS], nbands[B200DET_MAX_LEVELS];
  int zfirst[kMaxSegs + 1];             // zero chunks of an image before segment s; [nseg]: per image
  // levels too big to keep in L2 between fill and scatter are zero-filled up front, all images at once
  int up_mask;                          // bit l: level l is filled up front
  int zup;                              // chunks of the up-front fill (numbered before all others)
  int up_first[B200DET_MAX_LEVELS + 1]; // first up-front chunk of level l
};

// per-call scheduling state, followed in memory by gid[Kcap], order[Kcap], gsorted[Kcap]
struct BwdPlan {
  // the three hot words sit in different 128-byte lines: 2368 warps hammer them
  int task_ctr;                    // next scatter task (global, in group order)
  int pad0[31];
  int zero_ctr;                    // next zero chunk (global, in group order)
  int pad1[31];
  int up_done;                     // up-front chunks finished
  int pad2[31];
  int zero_done[kMaxGroups];       // per group: chunks finished
  int count[kMaxGroups];           // RoIs of the group
  int start[kMaxGroups + 8];       // exclusive scan of count
  int cursor[kMaxGroups];          // fill positions of the counting sort
};

struct __align__(16) BYEnt { int rowoff; float w0; float w1; int pad; };

struct BwdWarpTables {
  XTab x;
  BYEnt y[kMaxRows];
};

// ---- plan: counting sort of the RoIs by group ----------------------------------------------
// (1) group id of every RoI + histogram, (2) exclusive scan (one CTA), (3) fill order[].
template <int SR>
__global__ void bwd_plan_gid_kernel(LevelTable<float> lv, SegTable sg, const float* __restrict__ rois,
                                    const int32_t* __restrict__ levels, int K, int N, int PH, int PW, bool aligned,
                                    int* __restrict__ gid, BwdPlan* __restrict__ plan) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int l = levels ? __ldg(levels + k) : 0;
  const RoiGeom<float> g = roi_geometry<float>(rois + (long long)k * 5, lv.scale[l], PH, PW, SR, aligned);
  int r_last = -1;
  for (int s = 0; s < PH * SR; ++s) {
    const AxisTap<float> t = axis_tap<float>(sample_coord<float>(g.start_h, g.bin_h, s / SR, s % SR, SR), lv.height[l]);
    if (t.valid) r_last = max(r_last, t.hi);
  }
  int id = 0;  // RoIs that scatter nothing (bad image index, no valid sample) can sit anywhere
  if ((unsigned)g.batch < (unsigned)N && r_last >= 0)
    id = g.batch * sg.nseg + sg.first_seg[l] + min(r_last / sg.band_rows[l], sg.nbands[l] - 1);
  gid[k] = id;
  atomicAdd(&plan->count[id], 1);
}

__global__ void __launch_bounds__(1024) bwd_plan_scan_kernel(BwdPlan* __restrict__ plan, int G) {
  __shared__ int wsum[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int PER = kMaxGroups / 1024;
  int v[PER], tot = 0;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int g = tid * PER + i;
    v[i] = g < G ? plan->count[g] : 0;
    tot += v[i];
  }
  int incl = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = wsum[lane], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(kFull, wi, o);
      if (lane >= o) wi += t;
    }
    wsum[lane] = wi - w;
  }
  __syncthreads();
  int run = wsum[warp] + incl - tot;
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int g = tid * PER + i;
    if (g < G) { plan->start[g] = run; plan->cursor[g] = run; }
    run += v[i];
    if (g == G - 1) plan->start[G] = run;
  }
}

__global__ void bwd_plan_fill_kernel(const int* __restrict__ gid, int K, BwdPlan* __restrict__ plan,
                                     int* __restrict__ order, int* __restrict__ gsorted) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int g = gid[k];
  const int pos = atomicAdd(&plan->cursor[g], 1);
  order[pos] = k;
  gsorted[pos] = g;
}

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
  typename LaneVec<T, CH>::raw r;
  __device__ __forceinline__ void load(const T* p, long long) { r = LaneVec<T, CH>::ld_stream(p); }
  __device__ __forceinline__ void get(f32x2 (&v)[CH / 2]) const { LaneVec<T, CH>::pairs(r, v); }
};
template <typename T, int CH> struct GBuf<T, CH, false> {  // arbitrary strides (roi_align_kernel.cu:253-256)
  f32x2 v_[CH / 2];
  __device__ __forceinline__ void load(const T* p, long long gs_c) {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i)
      v_[i] = pack2((float)Io<T>::ld(p + lane_elem_channel<T, CH>(2 * i) * gs_c),
                    (float)Io<T>::ld(p + lane_elem_channel<T, CH>(2 * i + 1) * gs_c));
  }
  __device__ __forceinline__ void get(f32x2 (&v)[CH / 2]) const {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i) v[i] = v_[i];
  }
};

// ---- one task: (RoI, channel tile, pair of pooled rows) ------------------------------------
template <typename T, int CH, int SR, bool VEC>
__device__ __forceinline__ void bwd_sweep(const BwdWarpTables& tb, int nrows, T* __restrict__ base,
                                          const T* __restrict__ g0p, const T* __restrict__ g1p, bool has1,
                                          long long gs_c, long long gs_w, int PW) {
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  f32x2 X0[H2], X1[H2], Y0[H2], Y1[H2], G0[H2], G1[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X0[i] = X1[i] = Y0[i] = Y1[i] = G0[i] = G1[i] = 0ull; }
  GBuf<T, CH, VEC> n0, n1;
  n0.load(g0p, gs_c);
  if (has1) n1.load(g1p, gs_c);

  int j = 0;
  // samples j .. end-1 add into the window (lo, hi); the gradient of the next pooled
  // column is requested as soon as the current one has been taken over
  auto samples = [&](int end, f32x2 (&lo0)[H2], f32x2 (&hi0)[H2], f32x2 (&lo1)[H2], f32x2 (&hi1)[H2]) {
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
        const f32x2 hx = pack2(s.x, s.x), lx = pack2(s.y, s.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          lo0[i] = fma2(hx, G0[i], lo0[i]);
          hi0[i] = fma2(lx, G0[i], hi0[i]);
          lo1[i] = fma2(hx, G1[i], lo1[i]);
          hi1[i] = fma2(lx, G1[i], hi1[i]);
        }
      }
    }
  };
  // a column leaves the window: one vector red per distinct feature row
  auto flush = [&](f32x2 (&a0)[H2], f32x2 (&a1)[H2], int co) {
#pragma unroll 1
    for (int r = 0; r < nrows; ++r) {
      const BYEnt e = tb.y[r];
      const f32x2 w0 = pack2(e.w0, e.w0), w1 = pack2(e.w1, e.w1);
      f32x2 v[H2];
#pragma unroll
      for (int i = 0; i < H2; ++i) v[i] = fma2(w0, a0[i], mul2(w1, a1[i]));
      LV::red(base + (e.rowoff + co), v);
    }
#pragma unroll
    for (int i = 0; i < H2; ++i) { a0[i] = 0ull; a1[i] = 0ull; }
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
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  f32x2 wv[NR];
  typename LV::raw n[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    wv[r] = pack2(wy[r], wy[r]);
    n[r] = LV::ld(gp[r]);
  }
  f32x2 X[H2], Y[H2], G[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X[i] = Y[i] = G[i] = 0ull; }
  int j = 0;
  auto samples = [&](int end, f32x2 (&lo)[H2], f32x2 (&hi)[H2]) {
#pragma unroll 1
    for (; j < end; ++j) {
      if (SR == 1 || !(j & 1)) {
        const int pwn = j / SR + 1;
        const long long nxt = (long long)min(pwn, PW - 1) * gs_w;  // past the end: re-read the last column (L1)
#pragma unroll
        for (int r = 0; r < NR; ++r) {
          f32x2 t[H2];
          LV::pairs(n[r], t);
#pragma unroll
          for (int i = 0; i < H2; ++i) G[i] = (r == 0) ? mul2(wv[0], t[i]) : fma2(wv[r], t[i], G[i]);
          n[r] = LV::ld(gp[r] + nxt);
        }
      }
      const float4 s = xt.samp[j];
      if (__float_as_int(s.z)) {
        const f32x2 hx = pack2(s.x, s.x), lx = pack2(s.y, s.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          lo[i] = fma2(hx, G[i], lo[i]);
          hi[i] = fma2(lx, G[i], hi[i]);
        }
      }
    }
  };
  auto flush = [&](f32x2 (&a)[H2], int co) {
    LV::red(rowbase + co, a);
#pragma unroll
    for (int i = 0; i < H2; ++i) a[i] = 0ull;
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
  BwdPlan* plan;        // nullptr: one group, static task assignment, caller has zeroed
  const int* order;     // RoIs sorted by group (segmented plans only)
  const int* gsorted;   // group of order[p]
  int ztotal;           // zero chunks of the whole call; 0: no zero fill
  int zlook, zper;      // the fill runs up to zlook groups ahead of the scatter, at most zper chunks per task
  long long K;
  int N, C, PH, PW, CT, RP;
  int ngroups;          // 1: everything is one group
  int zero_chunk;       // bytes per zero chunk (multiple of 512); 0: no zero fill
  bool vec, aligned;
  int mode;             // 0: pooled-row pair sweep, 1: feature-row gather
  SegTable sg;
};

// pacing reads need no ordering (and ld.acquire costs an L1 invalidation each time)
__device__ __forceinline__ int ld_relaxed(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <typename T>
__device__ __forceinline__ long long seg_bytes(const LevelTable<T>& lv, const BwdArgs& a, int seg) {
  const int l = a.sg.level[seg];
  return (long long)a.sg.rows[seg] * lv.sh[l] * (long long)sizeof(T);
}

// writes zero chunk z (global numbering, group order) and reports it done; all lanes participate
template <typename T>
__device__ __forceinline__ void zero_chunk(const LevelTable<T>& lv, const BwdArgs& a, int z, int lane) {
  char* p = nullptr;
  long long bytes = 0;
  int c = z, g = 0;
  if (a.ngroups == 1) {
    for (int l = 0; l < lv.num_levels; ++l) {
      bytes = (long long)a.N * lv.sn[l] * (long long)sizeof(T);
      const int n = (int)((bytes + a.zero_chunk - 1) / a.zero_chunk);
      if (c < n) { p = reinterpret_cast<char*>(lv.ptr[l]); break; }
      c -= n;
    }
  } else if (z < a.sg.zup) {
    g = -1;
    for (int l = 0; l < lv.num_levels; ++l)
      if (((a.sg.up_mask >> l) & 1) && z < a.sg.up_first[l + 1]) {
        c = z - a.sg.up_first[l];
        bytes = (long long)a.N * lv.sn[l] * (long long)sizeof(T);
        p = reinterpret_cast<char*>(lv.ptr[l]);
        break;
      }
  } else {
    const int cpi = a.sg.zfirst[a.sg.nseg];
    const int zz = z - a.sg.zup;
    const int img = zz / cpi;
    c = zz - img * cpi;
    int seg = 0;
    while (seg + 1 < a.sg.nseg && c >= a.sg.zfirst[seg + 1]) ++seg;
    c -= a.sg.zfirst[seg];
    g = img * a.sg.nseg + seg;
    const int l = a.sg.level[seg];
    bytes = seg_bytes<T>(lv, a, seg);
    p = reinterpret_cast<char*>(lv.ptr[l] + (long long)img * lv.sn[l] + (long long)a.sg.row0[seg] * lv.sh[l]);
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
  if (lane == 0) atomicAdd(g < 0 ? &a.plan->up_done : &a.plan->zero_done[g], 1);
}

// zero chunks before group g (g may be ngroups: all of them)
__device__ __forceinline__ int zero_chunks_before(const BwdArgs& a, int g) {
  if (a.ngroups == 1) return g > 0 ? a.ztotal : 0;
  if (g >= a.ngroups) return a.ztotal;
  const int img = g / a.sg.nseg;
  return a.sg.zup + img * a.sg.zfirst[a.sg.nseg] + a.sg.zfirst[g - img * a.sg.nseg];
}

template <typename T, int CH, int SR, int MODE>
__global__ void __launch_bounds__(kBwdWarps * 32, 4) roi_align_bwd_sm100_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                              const __grid_constant__ BwdArgs a) {
  __shared__ BwdWarpTables tabs[kBwdWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  BwdWarpTables& tb = tabs[warp];
  const T* grad = reinterpret_cast<const T*>(a.grad);
  const int tasks_per_roi = a.RP * a.CT;  // MODE 1: RP = kRowChunks
  const bool sorted = a.ngroups > 1;
  const long long total = a.K * tasks_per_roi;

  auto run_task = [&](long long t) {
    const int rp = (int)(t % a.RP);
    const int ct = (int)((t / a.RP) % a.CT);
    const long long p = t / tasks_per_roi;
    const long long k = sorted ? (long long)__ldg(a.order + p) : p;
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

  // Tasks and zero chunks are both handed out in group order by one global counter each.  Before a
  // warp scatters a task of group g it (1) helps until every chunk of groups <= g is claimed, (2) waits
  // until those groups are reported done -- claimed chunks belong to running warps, so this needs no
  // co-residency -- and after the task it (3) writes one chunk if the fill is less than a group ahead.
  BwdPlan* plan = a.plan;
  const bool zeroing = a.ztotal > 0;
  int done_upto = 0;  // groups [0, done_upto) are known to be zero-filled
  bool all_claimed = false;
  auto claim_zero = [&]() {
    int z = 0;
    if (lane == 0) z = atomicAdd(&plan->zero_ctr, 1);
    z = __shfl_sync(kFull, z, 0);
    if (z < a.ztotal) zero_chunk<T>(lv, a, z, lane);
    return z;
  };
  while (true) {
    int t = 0;
    if (lane == 0) t = atomicAdd(&plan->task_ctr, 1);
    t = __shfl_sync(kFull, t, 0);
    if (t >= total) break;
    const int g = sorted ? __ldg(a.gsorted + t / tasks_per_roi) : 0;
    if (zeroing) {
      if (done_upto <= g) {
        const int need = zero_chunks_before(a, g + 1);
        while (ld_relaxed(&plan->zero_ctr) < need) claim_zero();
        all_claimed = need >= a.ztotal;
        if (lane == 0) {
          if (done_upto == 0 && sorted)
            while (ld_acquire(&plan->up_done) < a.sg.zup) __nanosleep(64);
          for (int q = done_upto; q <= g; ++q) {
            const int n = zero_chunks_before(a, q + 1) - zero_chunks_before(a, q);
            while (ld_acquire(&plan->zero_done[q]) < n) __nanosleep(64);
          }
        }
        __syncwarp();
        __threadfence();
        done_upto = g + 1;
      }
    }
    run_task(t);
    if (zeroing && !all_claimed) {
      const int ahead = zero_chunks_before(a, min(g + 1 + a.zlook, a.ngroups));
      for (int i = 0; i < a.zper; ++i) {
        const int zc = ld_relaxed(&plan->zero_ctr);
        all_claimed = zc >= a.ztotal;
        if (zc >= ahead) break;
        claim_zero();
      }
    }
  }
  // whatever the RoIs did not reach (groups without RoIs at the end)
  if (zeroing && !all_claimed)
    while (ld_relaxed(&plan->zero_ctr) < a.ztotal) claim_zero();
}

// ---- host side ---------------------------------------------------------------------------------
// The plan ring is the one allocation this library makes on its own (kPlanSlots x (130 KB + 12 bytes
// per RoI) per device, first backward call, regrown when a call brings more RoIs).  If it cannot be
// had (e.g. the first call happens inside a stream capture) the call falls back to memset + static
// task assignment.
struct DeviceScratch {
  char* buf = nullptr;
  size_t slot_bytes = 0;
  long long kcap = 0;
  unsigned next = 0;
  int l2_bytes = 0;
  int sms = 0;
  bool tried = false;
};
DeviceScratch g_scratch[64];
std::vector<void*>* g_outgrown = nullptr;  // freed at release: kernels in flight may still use them

constexpr size_t kPlanHead = (sizeof(BwdPlan) + 255) & ~(size_t)255;

DeviceScratch* scratch_for_current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  DeviceScratch& s = g_scratch[dev];
  if (!s.tried) {
    s.tried = true;
    cudaDeviceGetAttribute(&s.l2_bytes, cudaDevAttrL2CacheSize, dev);
    cudaDeviceGetAttribute(&s.sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return &s;
}

// plan slot for a call with K RoIs: head, then gid[K], order[K]; nullptr when memory cannot be had
BwdPlan* plan_slot(DeviceScratch* s, long long K) {
  if (!s) return nullptr;
  if (!s->buf || K > s->kcap) {
    const long long kcap = K + K / 2 + 4096;
    const size_t slot = kPlanHead + (((size_t)kcap * 12 + 255) & ~(size_t)255);
    void* p = nullptr;
    if (cudaMalloc(&p, slot * kPlanSlots) != cudaSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    if (s->buf) {
      if (!g_outgrown) g_outgrown = new std::vector<void*>();
      g_outgrown->push_back(s->buf);
    }
    s->buf = (char*)p;
    s->slot_bytes = slot;
    s->kcap = kcap;
  }
  return (BwdPlan*)(s->buf + s->slot_bytes * (s->next++ % kPlanSlots));
}

int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

// Cuts every level whose map (one image) is at most `up_bytes` into bands of at most `budget` bytes; bigger
// levels are filled up front and their RoIs ride with the image's last segment.  False when nothing is
// left to segment or it needs too many segments.
bool make_segments(const LevelArgs& lv, size_t es, double budget, double up_bytes, SegTable* sg) {
  sg->nseg = 0;
  sg->up_mask = 0;
  for (int li = 0; li < lv.num_levels; ++li) {
    const int l = lv.num_levels - 1 - li;  // coarse levels first
    const long long H = lv.height[l];
    const double row_bytes = (double)lv.width[l] * (double)lv.channels * (double)es;
    if (row_bytes * (double)H > up_bytes) {
      sg->up_mask |= 1 << l;
      continue;
    }
    long long band = (long long)(budget / row_bytes);
    if (band < 1) band = 1;
    long long nb = (H + band - 1) / band;
    band = (H + nb - 1) / nb;  // equal bands
    nb = (H + band - 1) / band;
    if (sg->nseg + nb > kMaxSegs) return false;
    sg->first_seg[l] = sg->nseg;
    sg->band_rows[l] = (int)band;
    sg->nbands[l] = (int)nb;
    for (long long b = 0; b < nb; ++b) {
      const int i = sg->nseg++;
      sg->level[i] = l;
      sg->row0[i] = (int)(b * band);
      sg->rows[i] = (int)((b + 1) * band <= H ? band : H - b * band);
    }
  }
  if (sg->nseg == 0) return false;
  for (int l = 0; l < B200DET_MAX_LEVELS; ++l)
    if (l >= lv.num_levels || ((sg->up_mask >> l) & 1)) {
      sg->first_seg[l] = sg->nseg - 1; sg->band_rows[l] = 0x7fffffff; sg->nbands[l] = 1;
    }
  return true;
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
int launch_bwd(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
               const int32_t* levels, long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init,
               cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CH = (C % 256 == 0 && env_int("B200DET_BWD_CH", 8) == 8) ? 8 : 4;
  BwdArgs a;
  a.rois = (const float*)rois;
  a.levels = levels;
  a.grad = grad;
  a.gs_k = gs[0]; a.gs_c = gs[1]; a.gs_h = gs[2]; a.gs_w = gs[3];
  a.K = K;
  a.N = (int)lv.batch; a.C = C; a.PH = PH; a.PW = PW;
  a.CT = C / (32 * CH);
  // Row gather (one red per footprint pixel) needs vector gradient loads and a lane per y sample.
  // Measured on B200 (profiles/r01_bwd_rows_segments.md): it wins for box-head shapes (7x7 bins, samples
  // >= 1 pixel apart: a pooled-row pair spreads over up to 8 feature rows) and loses for the mask-head
  // shape (14x14: the upstream gradient is re-read 2.5x and half of those re-reads miss L2), so by
  // default it is taken when a RoI has at most 16 y samples.  B200DET_BWD_MODE=0 / 1 forces one.
  const int mode_env = env_int("B200DET_BWD_MODE", -1);
  const bool rows_ok = vec && PH * sr <= 32;
  a.mode = rows_ok && (mode_env == 1 || (mode_env < 0 && PH * sr <= 16)) ? 1 : 0;
  a.RP = a.mode == 1 ? kRowChunks : (PH + 1) / 2;
  a.vec = vec; a.aligned = aligned;
  a.plan = nullptr; a.order = nullptr; a.gsorted = nullptr; a.ngroups = 1; a.zero_chunk = 0; a.ztotal = 0; a.zlook = 1; a.zper = 1;
  a.sg.nseg = 0; a.sg.zup = 0; a.sg.up_mask = 0;

  DeviceScratch* s = scratch_for_current_device();
  const int sms = s && s->sms > 0 ? s->sms : 148;
  const long long total = K * a.RP * a.CT;
  size_t image_bytes = 0;
  for (int l = 0; l < lv.num_levels; ++l) image_bytes += (size_t)lv.height[l] * lv.width[l] * C * sizeof(T);
  const bool dynamic = total < 0x7fffffffLL && K < 0x7fffffffLL && lv.batch <= 0x7fffffffLL &&
                       env_int("B200DET_BWD_FUSED", 1) != 0;
  BwdPlan* plan = dynamic ? plan_slot(s, K) : nullptr;
  if (plan) {
    // Segmented mode (default; B200DET_BWD_PHASED=0 turns it off): RoIs are counting-sorted by
    // (image, level, band of rows) and the persistent grid works group by group, zero-filling group
    // g+1 while it scatters group g, so the reds land on lines that are still dirty in L2 and every
    // gradient line goes to HBM once.  A band is B200DET_BWD_GROUP_MB megabytes (default 20: the 126 MB
    // L2 behaves like ~60 MB for data every SM touches, profiles/r01_bwd_phased.md, and two groups plus the
    // streaming upstream gradient share it).
    const double budget = (double)env_int("B200DET_BWD_GROUP_MB", 20) * 1048576.0;
    bool seg = env_int("B200DET_BWD_PHASED", 0) != 0 && (double)image_bytes * (double)a.N > budget &&
               make_segments(lv, sizeof(T), budget, (double)env_int("B200DET_BWD_UP_MB", 32) * 1048576.0, &a.sg);
    if (seg && (long long)a.sg.nseg * a.N > kMaxGroups) seg = false;
    if (!seg) a.sg.nseg = 0;
    a.ngroups = seg ? a.sg.nseg * a.N : 1;
    B200DET_CUDA_OK(cudaMemsetAsync(plan, 0, sizeof(BwdPlan), st));
    if (seg) {
      int* gid = reinterpret_cast<int*>(reinterpret_cast<char*>(plan) + kPlanHead);
      int* order = gid + s->kcap;
      int* gsorted = order + s->kcap;
      const LevelTable<float> ftab = make_nhwc_table<float>(lv);
      const int th = 128, bl = (int)((K + th - 1) / th);
      if (sr == 1)
        bwd_plan_gid_kernel<1><<<bl, th, 0, st>>>(ftab, a.sg, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned,
                                                  gid, plan);
      else
        bwd_plan_gid_kernel<2><<<bl, th, 0, st>>>(ftab, a.sg, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned,
                                                  gid, plan);
      bwd_plan_scan_kernel<<<1, 1024, 0, st>>>(plan, a.ngroups);
      bwd_plan_fill_kernel<<<bl, th, 0, st>>>(gid, (int)K, plan, order, gsorted);
      B200DET_CUDA_OK(cudaGetLastError());
      a.order = order;
      a.gsorted = gsorted;
    }
    a.plan = plan;
    if (zero_init) {
      // one zero chunk per ~0.75 tasks so the fill hides behind the scatter of the previous group
      a.zlook = env_int("B200DET_BWD_ZLOOK", 2);
      a.zper = env_int("B200DET_BWD_ZPER", 3);
      // segmented: one zero chunk per ~0.75 tasks so the fill keeps pace with the scatter; one group: the
      // fill is a phase of its own in front, and big chunks are faster (measured: 16 KB 1.07 ms, 27 KB 1.00,
      // 51 KB 0.96 for cfg3 fp32), so about four chunks per resident warp
      double zc = seg ? (double)image_bytes * (double)a.N /
                            (0.01 * env_int("B200DET_BWD_ZTASKS_PCT", 75) * (double)(total > 1 ? total : 1))
                      : (double)image_bytes * (double)a.N / ((double)sms * 16.0 * env_int("B200DET_BWD_ZPER_WARP", 4));
      long long zero_chunk = ((long long)zc + 4095) / 4096 * 4096;
      if (zero_chunk < 16384) zero_chunk = 16384;
      if (zero_chunk > (1 << 20)) zero_chunk = 1 << 20;
      a.zero_chunk = (int)zero_chunk;
      long long zt = 0;
      if (seg) {
        a.sg.up_first[0] = 0;
        for (int l = 0; l < B200DET_MAX_LEVELS; ++l) {
          long long n = 0;
          if (l < lv.num_levels && ((a.sg.up_mask >> l) & 1))
            n = ((long long)a.N * lv.height[l] * lv.width[l] * C * (long long)sizeof(T) + zero_chunk - 1) / zero_chunk;
          if (a.sg.up_first[l] + n >= 0x7fffffffLL) return 1;
          a.sg.up_first[l + 1] = a.sg.up_first[l] + (int)n;
        }
        a.sg.zup = a.sg.up_first[B200DET_MAX_LEVELS];
        for (int i = 0; i < a.sg.nseg; ++i) {
          const long long bytes = (long long)a.sg.rows[i] * lv.width[a.sg.level[i]] * C * (long long)sizeof(T);
          a.sg.zfirst[i] = (int)zt;
          zt += (bytes + zero_chunk - 1) / zero_chunk;
        }
        a.sg.zfirst[a.sg.nseg] = (int)zt;
        zt = zt * a.N + a.sg.zup;
      } else {
        for (int l = 0; l < lv.num_levels; ++l) {
          const long long bytes = (long long)a.N * lv.height[l] * lv.width[l] * C * (long long)sizeof(T);
          zt += (bytes + zero_chunk - 1) / zero_chunk;
        }
      }
      if (zt >= 0x7fffffffLL) return 1;  // cannot happen: >= 16 KB per chunk
      a.ztotal = (int)zt;
    }
  } else if (zero_init) {
    for (int l = 0; l < lv.num_levels; ++l) {
      const size_t bytes = (size_t)lv.batch * lv.channels * lv.height[l] * lv.width[l] * sizeof(T);
      if (bytes) B200DET_CUDA_OK(cudaMemsetAsync(lv.ptr[l], 0, bytes, st));
    }
  }
  if (CH == 8)
    return sr == 1 ? launch_bwd_k<T, 8, 1>(tab, a, sms, st) : launch_bwd_k<T, 8, 2>(tab, a, sms, st);
  return sr == 1 ? launch_bwd_k<T, 4, 1>(tab, a, sms, st) : launch_bwd_k<T, 4, 2>(tab, a, sms, st);
}

}  // namespace

int roi_align_bwd_fast(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                       const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned, bool zero_init,
                       cudaStream_t st, int* taken) {
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
    case B200DET_F32: rc = launch_bwd<float>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st); break;
    case B200DET_F16: rc = launch_bwd<__half>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st); break;
    case B200DET_BF16: rc = launch_bwd<__nv_bfloat16>(lv, grad, gs, vec, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st); break;
    default: return 0;
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

void roi_align_bwd_release() {
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < 64; ++d)
    if (g_scratch[d].buf) {
      cudaSetDevice(d);
      cudaFree(g_scratch[d].buf);
      g_scratch[d] = DeviceScratch();
    }
  cudaSetDevice(cur);
  if (g_outgrown) {
    for (void* p : *g_outgrown) cudaFree(p);
    g_outgrown->clear();
  }
}

}  // namespace b200det
