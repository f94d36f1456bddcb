// roi_align_bwd_strip_sm100.cu -- write-once RoIAlign BACKWARD for sm_100a
// (channels_last, C % 128 == 0, sampling_ratio 1 or 2, channel-contiguous gradient).
//
// Why.  The scatter formulation (roi_align_bwd_sm100.cu, and the reference's scalar
// atomics, cuda/roi_align_kernel.cu:303-327) pushes every RoI's footprint through the
// SM->L2 red path: 4.2 GB of reds for the 1.2 GB gradient of cfg3, 61 % of them missing L2
// (a DRAM read-modify-write each), on top of a zero fill.  Here every gradient byte is
// produced ON CHIP and stored to HBM exactly once: no memset, no atomics, no
// read-modify-write in DRAM.
//
// Ownership.  The unit of work is a STRIP: one feature row r of one image and level, a
// range of kTW columns, one channel tile.  A strip belongs to ONE WARP for its whole life:
// the warp keeps the strip's fp32 accumulators in its private slice of shared memory
// (kTW x 32*CH floats), adds the contribution of every RoI that taps the strip, stores the
// finished strip with streaming 16-byte stores and takes the next strip from a global
// counter.  Warps never share an accumulator, so there are no shared-memory atomics (fp32
// ATOMS is a CAS loop on sm_100), no barriers anywhere in the kernel, and the dynamic queue
// balances at warp granularity (the earlier tile kernel tied 8 warps to a tile and idled at
// 12 % occupancy behind __syncthreads).  Strips are queued coarse level first (busy strips
// early), inside a level image by image, column range by column range, row by row: the
// ~2400 strips in flight cover neighbouring rows and columns, so the upstream gradient of a
// RoI is pulled from HBM once and re-read (2-3 feature rows tap a pooled row) from L2.
//
// One (RoI, strip) task, transposing the separable forward:
//     G[pw]        = sum_{ph taps r} Wy[ph][r] * grad[ph][pw]        (<= kNRP pooled rows per pass,
//                                                                    loads two columns ahead)
//     acc[r][c]   += sum_{pw, x sample taps c} Wx * G[pw]            (two-column register window,
//                                                                    a column goes to shared memory
//                                                                    once, when the window leaves it)
// Summation order is fixed (RoIs in ascending index, pooled rows ascending), so the result is
// deterministic: this kernel also serves B200DET_BWD_DETERMINISTIC / torch.use_deterministic_algorithms
// (the reference only has a pure-PyTorch path for that, ops/roi_align.py:35-200,251-255).
#include <stdlib.h>

#include <mutex>
#include <vector>

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kSW = 16;              // warps per CTA = strips in flight per SM
constexpr int kNRP = 4;              // pooled rows gathered per pass
constexpr int kMaxImagesS = 4096;
constexpr int kStripSlots = 4;       // plan buffers in flight (different streams) before one is reused
constexpr int kTileRows = 16;        // a TILE (kTileRows rows x TW columns) shares one list of the RoIs that overlap it

template <int CH> struct StripShape { static constexpr int TW = CH == 8 ? 12 : 24; };  // 12 KB of accumulators per warp

struct StripPlan {
  int job_ctr;
  int unsorted;
  int pad[2];
  int img_start[kMaxImagesS + 1];
};

// what the plan kernel leaves per RoI: the geometry evaluated once and the footprint box
struct __align__(16) RoiRec {
  float start_h, bin_h, start_w, bin_w;
  unsigned rows, cols;   // feature rows [ya, yb] / columns [xa, xb] that receive taps: lo | hi << 16
  int level, batch;
};

// tiles of all levels and images, numbered level by level (level 0 first), image, tile row, tile column
struct TileGeom {
  int base[B200DET_MAX_LEVELS + 1];
  int tiles_y[B200DET_MAX_LEVELS], tiles_x[B200DET_MAX_LEVELS];
  int tw;
  unsigned level_mask;   // bit l: this launch produces level l (the others belong to the scatter kernel)
};

struct StripArgs {
  const void* grad;
  long long gs_k;
  int gs_h, gs_w;        // element strides of the upstream gradient inside a RoI (channel stride is 1)
  StripPlan* plan;
  const RoiRec* recs;
  long long K;
  int N, C, PH, PW, CT;
  int num_jobs;
  int job_base[B200DET_MAX_LEVELS + 1];   // first job of each level in processing order (coarse levels first)
  int strips_x[B200DET_MAX_LEVELS];
  TileGeom tg;
  const int* tile_ofs;   // [tiles + 1] exclusive scan of the per-tile RoI counts
  const int2* entries;   // (RoI, its rows ya | yb << 16), tile by tile, ascending RoI index inside a tile
  int entry_cap;         // entries that fit; a tile whose list does not fit scans the image's records instead
  bool accumulate;
};

// sample i of bin p (roi_align_kernel.cu:127-133); the division by the grid size is a multiplication by
// 1, 1/2: bit-identical for these powers of two
template <int SR>
__device__ __forceinline__ float sample_coord_sr(float start, float bin, int p, int i) {
  const float a = __fadd_rn(start, __fmul_rn((float)p, bin));
  const float b = __fmul_rn(__fmul_rn((float)i + .5f, bin), 1.0f / (float)SR);
  return __fadd_rn(a, b);
}

// ---- plan: per-RoI record + first RoI of every image ----------------------------------------
template <int SR>
__global__ void strip_plan_kernel(LevelTable<float> lv, const float* __restrict__ rois,
                                  const int32_t* __restrict__ levels, int K, int N, int PH, int PW, bool aligned,
                                  StripPlan* __restrict__ plan, RoiRec* __restrict__ recs, TileGeom tg,
                                  int* __restrict__ tile_cnt) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + (long long)k * 5, lv.scale[l], PH, PW, SR, aligned);
  int ya = 0x7fff, yb = -1, xa = 0x7fff, xb = -1;
  for (int s = 0; s < PH * SR; ++s) {
    const AxisTap<float> t = axis_tap<float>(sample_coord_sr<SR>(g.start_h, g.bin_h, s / SR, s % SR), H);
    if (t.valid) { ya = min(ya, t.lo); yb = max(yb, t.hi); }
  }
  for (int s = 0; s < PW * SR; ++s) {
    const AxisTap<float> t = axis_tap<float>(sample_coord_sr<SR>(g.start_w, g.bin_w, s / SR, s % SR), W);
    if (t.valid) { xa = min(xa, t.lo); xb = max(xb, t.hi); }
  }
  const int b = g.batch;
  const bool in_range = b >= 0 && b < N;
  if (!in_range || yb < 0 || xb < 0) { ya = 1; yb = 0; xa = 1; xb = 0; }  // never overlaps a strip
  RoiRec rc;
  rc.start_h = g.start_h; rc.bin_h = g.bin_h; rc.start_w = g.start_w; rc.bin_w = g.bin_w;
  rc.rows = (unsigned)ya | ((unsigned)yb << 16);
  rc.cols = (unsigned)xa | ((unsigned)xb << 16);
  rc.level = l;
  rc.batch = b;
  recs[k] = rc;
  if (in_range && yb >= ya && xb >= xa && ((tg.level_mask >> l) & 1u)) {  // one more RoI in the list of every tile its footprint overlaps
    const int t0 = tg.base[l] + b * tg.tiles_y[l] * tg.tiles_x[l];
    for (int ty = ya / kTileRows; ty <= yb / kTileRows; ++ty)
      for (int tx = xa / tg.tw; tx <= xb / tg.tw; ++tx) atomicAdd(tile_cnt + t0 + ty * tg.tiles_x[l] + tx, 1);
  }
  const int prev = k > 0 ? (int)__ldg(rois + (long long)(k - 1) * 5) : -1;
  if (!in_range || b < prev) {  // not grouped by image: every strip scans all K records
    plan->unsorted = 1;
    return;
  }
  for (int i = prev + 1; i <= b; ++i) plan->img_start[i] = k;
  if (k == K - 1)
    for (int i = b + 1; i <= N; ++i) plan->img_start[i] = K;
}

// exclusive scan of the tile counts, in place; [T] receives the total (one CTA: T is a few thousand)
__global__ void __launch_bounds__(1024) strip_tile_scan_kernel(int* __restrict__ cnt, int T) {
  __shared__ int wsum[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int per = (T + 1023) / 1024;
  const int i0 = tid * per, i1 = min(i0 + per, T);
  int tot = 0;
  for (int i = i0; i < i1; ++i) tot += cnt[i];
  int incl = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const int w = wsum[lane];
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(kFull, wi, o);
      if (lane >= o) wi += t;
    }
    wsum[lane] = wi - w;
  }
  __syncthreads();
  int run = wsum[warp] + incl - tot;
  for (int i = i0; i < i1; ++i) {
    const int c = cnt[i];
    cnt[i] = run;
    run += c;
  }
  if (tid == 1023) cnt[T] = run;
}

// one warp per tile writes the tile's list in ascending RoI index (ordered compaction of the image's records)
__global__ void strip_tile_fill_kernel(const LevelTable<float> lv, TileGeom tg, const StripPlan* __restrict__ plan,
                                       const RoiRec* __restrict__ recs, int K, int T,
                                       const int* __restrict__ tile_ofs, int2* __restrict__ entries, int cap) {
  const int t = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (t >= T) return;
  int pos = tile_ofs[t];
  const int end = tile_ofs[t + 1];
  if (end == pos || end > cap) return;  // empty, or the lists overflow: this tile's strips scan instead
  int l = 0;
  while (t >= tg.base[l + 1]) ++l;
  int rem = t - tg.base[l];
  const int tx = rem % tg.tiles_x[l];
  rem /= tg.tiles_x[l];
  const int ty = rem % tg.tiles_y[l];
  const int b = rem / tg.tiles_y[l];
  const int y0 = ty * kTileRows, y1 = y0 + kTileRows - 1, x0 = tx * tg.tw, x1 = x0 + tg.tw - 1;
  int k0 = 0, k1 = K;
  if (!plan->unsorted) { k0 = plan->img_start[b]; k1 = plan->img_start[b + 1]; }
  for (int ks = k0; ks < k1 && pos < end; ks += 32) {
    const int k = ks + lane;
    bool hit = false;
    uint4 bx = make_uint4(0u, 0u, 0u, 0u);
    if (k < k1) {
      bx = __ldg(reinterpret_cast<const uint4*>(recs + k) + 1);
      const int ya = bx.x & 0xffff, yb = bx.x >> 16, xa = bx.y & 0xffff, xb = bx.y >> 16;
      hit = (int)bx.z == l && (int)bx.w == b && ya <= y1 && yb >= y0 && xa <= x1 && xb >= x0 && yb >= ya;
    }
    const unsigned m = __ballot_sync(kFull, hit);
    if (hit) entries[pos + __popc(m & ((1u << lane) - 1u))] = make_int2(k, (int)bx.x);
    pos += __popc(m);
  }
}

__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// ---- shared-memory accumulators -----------------------------------------------------------
// One pixel holds 32*CH floats in a lane-centric order: lane i keeps its first four channels at
// [4i, 4i+4) and (CH == 8) its other four at [128 + 4i, ...), whatever channels those are for T;
// every access is a conflict-free 16-byte one (acc_add below).

// ---- x axis of a RoI restricted to the strip's columns [x0, x1] (lane j owns x sample j) ----
// Same window / event scheme as build_x (sm100_common.cuh): the sweep keeps two feature columns
// (lo, hi = lo + 1) in registers; a sample finds its taps by leaving the window as it is, shifting
// it by one column (one EVENT: a new column enters) or restarting it (two events).  A sample
// whose two taps both fall outside the strip, or that is empty (roi_align_kernel.cu:23-26), has zero
// weights and needs no event: it adds nothing and moves nothing.  Columns are float offsets into
// the warp's accumulators; the column left of the strip (-PX: only the right tap of a sample is
// inside) and the one right of it (tw * PX: only the left tap, or the clamped last column of the
// map, whose right weight is 0) exist as events like any other and are dropped by the flush.
struct __align__(16) SXTab {
  float2 w[kMaxSX];                      // weights of the lo / hi column of sample j
  int colofs[kMaxEvents + 2];            // accumulator offset of the column that enters at event e
  unsigned char ev_end[kMaxEvents + 4];  // [e + 1]: samples below this index are complete once event e arrived; [0]: before any
  int ncol;
};

template <int SR>
__device__ __forceinline__ unsigned build_x_strip(float start_w, float bin_w, int W, int x0, int x1, int PW, int lane,
                                                  int px_floats, SXTab& tb) {
  const int SX = PW * SR;
  const bool active = lane < SX;
  const int j = active ? lane : SX - 1;
  const AxisTap<float> t = axis_tap<float>(sample_coord_sr<SR>(start_w, bin_w, j / SR, j % SR), W);
  const int xl = t.lo;
  const bool valid = active && t.valid && xl >= x0 - 1 && xl <= x1;
  const unsigned vmask = __ballot_sync(kFull, valid);
  const unsigned before = vmask & ((1u << lane) - 1u);
  const int pxl = __shfl_sync(kFull, xl, before ? 31 - __clz(before) : lane);
  // x samples run left to right for proper RoIs; anything else (x2 < x1 with aligned=True) restarts the window
  const int d = xl - pxl;
  const int nnew = !valid ? 0 : (!before ? 2 : (d == 0 ? 0 : (d == 1 ? 1 : 2)));
  int incl = nnew;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(kFull, incl, o);
    if (lane >= o) incl += v;
  }
  const int base = incl - nnew;
  const int ncol = __shfl_sync(kFull, incl, 31);
  const int last = incl - 1;  // sample j is complete after event last (-1: needs no event at all)
  int nxt = __shfl_down_sync(kFull, last, 1);
  if (lane == SX - 1) nxt = ncol;
  if (active) {
    tb.w[lane] = valid ? make_float2(t.wl, t.wh) : make_float2(0.f, 0.f);
    const int co = (xl - x0) * px_floats;
    if (nnew == 2) {
      tb.colofs[base] = co;
      tb.colofs[base + 1] = co + px_floats;
    } else if (nnew == 1) {
      tb.colofs[base] = co + px_floats;
    }
    for (int e = last; e < nxt; ++e) tb.ev_end[e + 1] = (unsigned char)(lane + 1);
    if (lane == 0)
      for (int e = -1; e < last; ++e) tb.ev_end[e + 1] = 0;
  }
  if (lane == 0) tb.ncol = ncol;
  return vmask;
}

// ---- one pass of a (RoI, strip) task: pooled rows ph0 .. ph0 + nr - 1, x samples [j0, j1) ----
// Event-driven like the scatter kernel's row sweep (a new column entering the two-column window is an event).
// Packed fp32 pairs are float2 + the sm_100 intrinsics (__ffma2_rn ...): with them ptxas updates the
// accumulators in place, with opaque 64-bit asm operands it copied every sum back (12 MOVs per sample).
template <int CH>
__device__ __forceinline__ void acc_add(float* px, const float2 (&v)[CH / 2]) {
#pragma unroll
  for (int h = 0; h < CH / 4; ++h) {
    float4* p = reinterpret_cast<float4*>(px + h * 128);
    const float4 a = *p;
    const float2 s0 = __fadd2_rn(make_float2(a.x, a.y), v[2 * h]);
    const float2 s1 = __fadd2_rn(make_float2(a.z, a.w), v[2 * h + 1]);
    *p = make_float4(s0.x, s0.y, s1.x, s1.y);
  }
}

// the lane's CH channels of one upstream-gradient bin, as loaded and as packed fp32 pairs (float4 /
// 32-bit words rather than LaneVec's 64-bit raw words: ptxas then uses the loaded registers in place)
template <typename T, int CH> struct GVec;
template <int CH> struct GVec<float, CH> {
  struct raw { float4 v[CH / 4]; };
  static __device__ __forceinline__ void ld(raw& r, const float* p) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h) r.v[h] = __ldg(reinterpret_cast<const float4*>(p + 128 * h));
  }
  static __device__ __forceinline__ void f2(const raw& r, float2 (&o)[CH / 2]) {
#pragma unroll
    for (int h = 0; h < CH / 4; ++h) {
      o[2 * h] = make_float2(r.v[h].x, r.v[h].y);
      o[2 * h + 1] = make_float2(r.v[h].z, r.v[h].w);
    }
  }
};
template <typename T> struct Widen16;
template <> struct Widen16<__nv_bfloat16> {
  static __device__ __forceinline__ float2 f(unsigned u) {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  }
};
template <> struct Widen16<__half> {
  static __device__ __forceinline__ float2 f(unsigned u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }
};
template <typename T, int CH> struct GVec {   // 16-bit elements: CH * 2 contiguous bytes per lane
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
  static __device__ __forceinline__ void f2(const raw& r, float2 (&o)[CH / 2]) {
#pragma unroll
    for (int i = 0; i < CH / 2; ++i) o[i] = Widen16<T>::f(r.u[i]);
  }
};

// keeps a value in a register: without it the compiler rebuilds every load address from the kernel
// parameters (five instructions per load instead of one IMAD.WIDE)
template <typename P> __device__ __forceinline__ P* keep(P* p) {
  asm volatile("" : "+l"(p));
  return p;
}
__device__ __forceinline__ int keep(int v) {
  asm volatile("" : "+r"(v));
  return v;
}

template <typename T, int CH, int SR, int NRP>
__device__ __forceinline__ void strip_sweep(const SXTab& xt, float* __restrict__ acc, int tw,
                                            const T* __restrict__ gp_, int gs_h_, int gs_w, int nr,
                                            const float (&wy)[NRP], int j0, int j1) {
  using GV = GVec<T, CH>;
  constexpr int H2 = CH / 2;
  constexpr int PX = 32 * CH;
  const T* gp = keep(gp_);
  const int gs_h = keep(gs_h_);
  const int pw_last = (j1 - 1) / SR;
  typename GV::raw n[NRP] = {};
  {
    const int o = (j0 / SR) * gs_w;
#pragma unroll
    for (int r = 0; r < NRP; ++r)
      if (r < nr) GV::ld(n[r], gp + (r * gs_h + o));
  }
  float2 X[H2], Y[H2], G[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X[i] = make_float2(0.f, 0.f); Y[i] = X[i]; G[i] = X[i]; }
  const int lim = tw * PX;
  int j = j0;
  // samples j .. end - 1 add into the window (lo, hi); the gradient of the next pooled column is
  // requested as soon as the current one has been taken over
  auto samples = [&](int end, float2 (&lo)[H2], float2 (&hi)[H2]) {
    end = min(end, j1);
#pragma unroll 1
    for (; j < end; ++j) {
      if (SR == 1 || !(j & 1)) {
        const int o = min(j / SR + 1, pw_last) * gs_w;
#pragma unroll
        for (int r = 0; r < NRP; ++r) {
          float2 t[H2];
          GV::f2(n[r], t);
          const float2 w2 = make_float2(wy[r], wy[r]);
#pragma unroll
          for (int i = 0; i < H2; ++i) G[i] = (r == 0) ? __fmul2_rn(w2, t[i]) : __ffma2_rn(w2, t[i], G[i]);
          if (r < nr) GV::ld(n[r], gp + (r * gs_h + o));
        }
      }
      const float2 w = xt.w[j];
      const float2 wl = make_float2(w.x, w.x), wh = make_float2(w.y, w.y);
#pragma unroll
      for (int i = 0; i < H2; ++i) {
        lo[i] = __ffma2_rn(wl, G[i], lo[i]);
        hi[i] = __ffma2_rn(wh, G[i], hi[i]);
      }
    }
  };
  auto flush = [&](const float2 (&a)[H2], int ofs) {
    if ((unsigned)ofs < (unsigned)lim) acc_add<CH>(acc + ofs, a);
  };
  // One loop, fixed roles: X is the window's left column, Y its right one.  With e events consumed the
  // window holds the columns of events e - 2 and e - 1; event e pushes the left column out (it is complete:
  // one add to shared memory) and opens a new right column.
  const int ncol = xt.ncol;
  int e = 0;
#pragma unroll 1
  while (true) {
    samples(xt.ev_end[e], X, Y);
    if (e >= ncol) break;
    if (e >= 2) flush(X, xt.colofs[e - 2]);
#pragma unroll
    for (int i = 0; i < H2; ++i) { X[i] = Y[i]; Y[i] = make_float2(0.f, 0.f); }
    ++e;
  }
  if (ncol >= 2) flush(X, xt.colofs[ncol - 2]);
  if (ncol >= 1) flush(Y, xt.colofs[ncol - 1]);
}

// ---- persistent strip kernel ----------------------------------------------------------------
template <typename T, int CH, int SR>
__global__ void __launch_bounds__(kSW * 32, 1) roi_align_bwd_strip_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                          const __grid_constant__ StripArgs a) {
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  constexpr int PX = 32 * CH;  // floats per pixel
  constexpr int TW = StripShape<CH>::TW;
  extern __shared__ __align__(16) float smem[];  // [kSW][TW][PX] accumulators, then [kSW] x tables
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* acc = smem + warp * (TW * PX) + lane * 4;  // this lane's slot of column 0
  SXTab& xt = reinterpret_cast<SXTab*>(smem + kSW * TW * PX)[warp];
  {
    const ulonglong2 z = make_ulonglong2(0ull, 0ull);
    for (int o = 0; o < TW * (CH / 4); ++o) *reinterpret_cast<ulonglong2*>(acc + o * 128) = z;
  }
  const bool single = a.plan->unsorted != 0;
  const T* grad = reinterpret_cast<const T*>(a.grad);
  const float inv_count = 1.0f / (float)(SR * SR);

  // (Handing a CTA kSW neighbouring rows at a time raises the L1 hit rate from 5 % to 30 % but runs 10 % slower:
  // warps that work on the same RoIs stall on the same loads at the same time.  Measured, not kept.)
  while (true) {
    int job = 0;
    if (lane == 0) job = atomicAdd(&a.plan->job_ctr, 1);
    job = __shfl_sync(kFull, job, 0);
    if (job >= a.num_jobs) break;
    // decode: levels coarse to fine; inside a level image, column range, row, channel tile (fastest)
    int li = 0;
    while (job >= a.job_base[li + 1]) ++li;
    const int l = lv.num_levels - 1 - li;
    const int H = lv.height[l], W = lv.width[l];
    int rem = job - a.job_base[li];
    const int ct = rem % a.CT;
    rem /= a.CT;
    const int r = rem % H;
    rem /= H;
    const int tx = rem % a.strips_x[l];
    const int b = rem / a.strips_x[l];
    const int x0 = tx * TW;
    const int tw = min(TW, W - x0);
    const int x1 = x0 + tw - 1;
    const int ch = ct * PX + lane_channel<T, CH>(lane);

    // the RoIs that can tap this strip: the list of its tile (only the row range is left to test), or, when the
    // lists did not fit, the records of its image
    const int tile = a.tg.base[l] + (b * a.tg.tiles_y[l] + r / kTileRows) * a.tg.tiles_x[l] + tx;
    int i0 = __ldg(a.tile_ofs + tile), i1 = __ldg(a.tile_ofs + tile + 1);
    const bool listed = i1 <= a.entry_cap;
    if (!listed) {
      i0 = 0; i1 = (int)a.K;
      if (!single) { i0 = a.plan->img_start[b]; i1 = a.plan->img_start[b + 1]; }
    }
    bool dirty = false;
    for (int is = i0; is < i1; is += 32) {
      const int i = is + lane;
      bool hit = false;
      int kmine = i;
      if (i < i1) {
        if (listed) {
          const int2 en = __ldg(a.entries + i);
          kmine = en.x;
          hit = (en.y & 0xffff) <= r && (int)((unsigned)en.y >> 16) >= r;
        } else {
          const uint4 bx = __ldg(reinterpret_cast<const uint4*>(a.recs + i) + 1);
          const int ya = bx.x & 0xffff, yb = bx.x >> 16, xa = bx.y & 0xffff, xb = bx.y >> 16;
          hit = (int)bx.z == l && (int)bx.w == b && ya <= r && yb >= r && xa <= x1 && xb >= x0;
        }
      }
      unsigned m = __ballot_sync(kFull, hit);
      while (m) {
        const int kk = __shfl_sync(kFull, kmine, __ffs(m) - 1);
        m &= m - 1;
        const float4 geo = __ldg(reinterpret_cast<const float4*>(a.recs + kk));  // start_h, bin_h, start_w, bin_w
        // pooled rows that tap feature row r, and with which weight (lane = pooled row)
        float wph = 0.f;
        bool touched = false;
        if (lane < a.PH) {
#pragma unroll
          for (int iy = 0; iy < SR; ++iy) {
            const AxisTap<float> t = axis_tap<float>(sample_coord_sr<SR>(geo.x, geo.y, lane, iy), H);
            if (t.valid) {
              if (t.lo == r) { wph += t.wl; touched = true; }
              if (t.hi == r) { wph += t.wh; touched = true; }
            }
          }
        }
        wph *= inv_count;
        const unsigned pm = __ballot_sync(kFull, touched);
        if (!pm) continue;
        __syncwarp();  // the previous task's sweep is done with the table
        const unsigned vmask = build_x_strip<SR>(geo.z, geo.w, W, x0, x1, a.PW, lane, PX, xt);
        __syncwarp();
        if (!vmask) continue;
        dirty = true;
        // x samples of the pooled columns that hold a valid sample
        const int j0 = ((__ffs(vmask) - 1) / SR) * SR, j1 = ((31 - __clz(vmask)) / SR + 1) * SR;
        // the pooled rows that tap one feature row are consecutive (samples are monotone in y)
        const int ph_lo = __ffs(pm) - 1, ph_hi = 31 - __clz(pm);
        const T* gbase = grad + (long long)kk * a.gs_k + ch;
        const int nph = ph_hi - ph_lo + 1;
        if (nph <= 2) {  // the common case for RoIs of 14 pixels and more: half the registers, half the gather
          float wy[2];
          wy[0] = __shfl_sync(kFull, wph, ph_lo);
          wy[1] = nph == 2 ? __shfl_sync(kFull, wph, ph_hi) : 0.f;
          strip_sweep<T, CH, SR, 2>(xt, acc, tw, gbase + ph_lo * a.gs_h, a.gs_h, a.gs_w, nph, wy, j0, j1);
        } else {
          for (int ph0 = ph_lo; ph0 <= ph_hi; ph0 += kNRP) {
            float wy[kNRP];
#pragma unroll
            for (int i = 0; i < kNRP; ++i) {
              const float w = __shfl_sync(kFull, wph, (ph0 + i) & 31);
              wy[i] = ph0 + i <= ph_hi ? w : 0.f;
            }
            strip_sweep<T, CH, SR, kNRP>(xt, acc, tw, gbase + ph0 * a.gs_h, a.gs_h, a.gs_w,
                                         min(kNRP, ph_hi - ph0 + 1), wy, j0, j1);
          }
        }
      }
    }
    // store the strip: every gradient byte goes to HBM once; the accumulators are left zeroed
    T* out = lv.ptr[l] + (long long)b * lv.sn[l] + ((long long)r * W + x0) * a.C + ch;
    if (dirty) {
      for (int c = 0; c < tw; ++c) {
        f32x2 v[H2];
#pragma unroll
        for (int h = 0; h < CH / 4; ++h) {
          ulonglong2* p = reinterpret_cast<ulonglong2*>(acc + c * PX + h * 128);
          const ulonglong2 f = *p;
          *p = make_ulonglong2(0ull, 0ull);
          v[2 * h] = f.x;
          v[2 * h + 1] = f.y;
        }
        T* p = out + (long long)c * a.C;
        if (a.accumulate) {
          f32x2 old[H2];
          LV::pairs(LV::ld_stream(p), old);
#pragma unroll
          for (int i = 0; i < H2; ++i) v[i] = add2(v[i], old[i]);
        }
        LV::st(p, v);
      }
    } else if (!a.accumulate) {
      f32x2 v[H2];
#pragma unroll
      for (int i = 0; i < H2; ++i) v[i] = 0ull;
      for (int c = 0; c < tw; ++c) LV::st(out + (long long)c * a.C, v);
    }
  }
}

// ---- host side ---------------------------------------------------------------------------------
// Plan memory (job counter, first RoI of every image, 32 bytes per RoI) is the one allocation this
// path makes on its own: a ring of kStripSlots buffers per device, grown on demand.  A slot is
// handed out under a mutex and carries the event of its last user; the next user's stream waits for it,
// so calls on different streams or from different host threads never share a live plan.
struct StripSlot {
  char* buf = nullptr;
  size_t cap = 0;
  cudaEvent_t done = nullptr;
  bool used = false;
};
struct StripScratch {
  StripSlot slot[kStripSlots];
  unsigned next = 0;
  int sms = 0;
};
StripScratch g_sscratch[64];
std::mutex g_smutex;
std::vector<void*>* g_sgarbage = nullptr;  // outgrown buffers, freed at release

int senv_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

StripSlot* strip_slot(size_t bytes, cudaStream_t st, int* sms) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_smutex);
  StripScratch& s = g_sscratch[dev];
  if (s.sms == 0) cudaDeviceGetAttribute(&s.sms, cudaDevAttrMultiProcessorCount, dev);
  *sms = s.sms > 0 ? s.sms : 148;
  StripSlot& sl = s.slot[s.next++ % kStripSlots];
  if (!sl.done && cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming) != cudaSuccess) {
    (void)cudaGetLastError();
    sl.done = nullptr;
    return nullptr;
  }
  if (sl.used && cudaStreamWaitEvent(st, sl.done, 0) != cudaSuccess) {
    (void)cudaGetLastError();
    return nullptr;
  }
  if (sl.cap < bytes) {
    void* p = nullptr;
    const size_t want = bytes + bytes / 2 + (64 << 10);
    if (cudaMalloc(&p, want) != cudaSuccess) {  // e.g. inside a stream capture: the caller falls back
      (void)cudaGetLastError();
      return nullptr;
    }
    if (sl.buf) {
      if (!g_sgarbage) g_sgarbage = new std::vector<void*>();
      g_sgarbage->push_back(sl.buf);  // a kernel in flight may still read it
    }
    sl.buf = (char*)p;
    sl.cap = want;
  }
  return &sl;
}

template <typename T, int CH, int SR>
int launch_strip_k(const LevelTable<T>& tab, const StripArgs& a, int sms, cudaStream_t st) {
  constexpr int TW = StripShape<CH>::TW;
  const int smem = kSW * (TW * 32 * CH * (int)sizeof(float) + (int)sizeof(SXTab));
  B200DET_CUDA_OK(cudaFuncSetAttribute(roi_align_bwd_strip_kernel<T, CH, SR>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  long long blocks = ((long long)a.num_jobs + kSW - 1) / kSW;
  if (blocks > sms) blocks = sms;
  roi_align_bwd_strip_kernel<T, CH, SR><<<(unsigned)blocks, kSW * 32, smem, st>>>(tab, a);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int launch_strip(const LevelArgs& lv, const void* grad, const int64_t* gs, const void* rois, const int32_t* levels,
                 long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init, unsigned level_mask,
                 cudaStream_t st, int* taken) {
  const int CH = (C % 256 == 0 && senv_int("B200DET_STRIP_CH", 8) == 8) ? 8 : 4;
  const int TW = CH == 8 ? StripShape<8>::TW : StripShape<4>::TW;
  StripArgs a;
  a.grad = grad;
  a.gs_k = gs[0]; a.gs_h = (int)gs[2]; a.gs_w = (int)gs[3];
  a.K = K;
  a.N = (int)lv.batch; a.C = C; a.PH = PH; a.PW = PW;
  a.CT = C / (32 * CH);
  a.accumulate = !zero_init;
  long long jobs = 0;
  for (int li = 0; li < lv.num_levels; ++li) {
    const int l = lv.num_levels - 1 - li;
    a.strips_x[l] = (int)((lv.width[l] + TW - 1) / TW);
    a.job_base[li] = (int)jobs;
    if ((level_mask >> l) & 1u) jobs += (long long)a.strips_x[l] * lv.height[l] * lv.batch * a.CT;
    if (jobs > 0x7fffffffLL) return 0;
  }
  for (int li = lv.num_levels; li <= B200DET_MAX_LEVELS; ++li) a.job_base[li] = (int)jobs;
  a.num_jobs = (int)jobs;
  // tiles (kTileRows x TW) and their RoI lists
  long long tiles = 0;
  a.tg.tw = TW;
  a.tg.level_mask = level_mask;
  for (int l = 0; l < B200DET_MAX_LEVELS; ++l) {
    a.tg.base[l] = (int)tiles;
    a.tg.tiles_y[l] = a.tg.tiles_x[l] = 1;
    if (l < lv.num_levels) {
      a.tg.tiles_y[l] = (int)((lv.height[l] + kTileRows - 1) / kTileRows);
      a.tg.tiles_x[l] = a.strips_x[l];
      tiles += (long long)a.tg.tiles_y[l] * a.tg.tiles_x[l] * lv.batch;
    }
  }
  a.tg.base[B200DET_MAX_LEVELS] = (int)tiles;
  if (tiles > 0x3fffffffLL) return 0;
  const int NT = (int)tiles;
  a.entry_cap = (int)(K * 16 + 4096);
  int sms = 148;
  const size_t plan_bytes = (sizeof(StripPlan) + 255) & ~(size_t)255;
  const size_t rec_bytes = ((size_t)K * sizeof(RoiRec) + 255) & ~(size_t)255;
  const size_t ofs_bytes = ((size_t)(NT + 1) * sizeof(int) + 255) & ~(size_t)255;
  StripSlot* sl = strip_slot(plan_bytes + ofs_bytes + rec_bytes + (size_t)a.entry_cap * sizeof(int2), st, &sms);
  if (!sl) return 0;
  a.plan = (StripPlan*)sl->buf;
  int* tile_ofs = (int*)(sl->buf + plan_bytes);
  a.tile_ofs = tile_ofs;
  a.recs = (const RoiRec*)(sl->buf + plan_bytes + ofs_bytes);
  a.entries = (const int2*)(sl->buf + plan_bytes + ofs_bytes + rec_bytes);
  B200DET_CUDA_OK(cudaMemsetAsync(a.plan, 0, plan_bytes + ofs_bytes, st));  // plan header and tile counts
  const LevelTable<float> ftab = make_nhwc_table<float>(lv);
  const int thr = 128, blk = (int)((K + thr - 1) / thr);
  if (sr == 1)
    strip_plan_kernel<1><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                              (RoiRec*)a.recs, a.tg, tile_ofs);
  else
    strip_plan_kernel<2><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                              (RoiRec*)a.recs, a.tg, tile_ofs);
  strip_tile_scan_kernel<<<1, 1024, 0, st>>>(tile_ofs, NT);
  strip_tile_fill_kernel<<<(unsigned)((NT + 3) / 4), 128, 0, st>>>(ftab, a.tg, a.plan, a.recs, (int)K, NT, tile_ofs,
                                                                   (int2*)a.entries, a.entry_cap);
  B200DET_CUDA_OK(cudaGetLastError());
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  int rc;
  if (CH == 8) rc = sr == 1 ? launch_strip_k<T, 8, 1>(tab, a, sms, st) : launch_strip_k<T, 8, 2>(tab, a, sms, st);
  else rc = sr == 1 ? launch_strip_k<T, 4, 1>(tab, a, sms, st) : launch_strip_k<T, 4, 2>(tab, a, sms, st);
  {
    std::lock_guard<std::mutex> lock(g_smutex);
    if (cudaEventRecord(sl->done, st) == cudaSuccess) sl->used = true;
    else (void)cudaGetLastError();
  }
  if (rc) return rc;
  *taken = 1;
  return 0;
}

}  // namespace

int roi_align_bwd_strip(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                        const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned, bool zero_init,
                        unsigned level_mask, cudaStream_t st, int* taken) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr)) return 0;
  level_mask &= (1u << lv.num_levels) - 1u;
  if (!level_mask) return 0;
  if (K == 0 || lv.batch == 0 || K > 0x7fffffffLL / 64 || lv.batch > kMaxImagesS) return 0;
  const size_t es = dtype_size(dtype);
  const long long per16 = 16 / (long long)es;
  const bool vec = gs[1] == 1 && gs[0] % per16 == 0 && gs[2] % per16 == 0 && gs[3] % per16 == 0 &&
                   ((uintptr_t)grad) % 16 == 0;
  if (!vec) return 0;  // strided gradients take the scatter kernel
  // 32-bit offsets inside one RoI's gradient block
  if (gs[2] < 0 || gs[3] < 0 || (gs[2] * PH + gs[3] * PW) >= (1LL << 30)) return 0;
  for (int l = 0; l < lv.num_levels; ++l)
    if (lv.height[l] > 0x7ffe || lv.width[l] > 0x7ffe) return 0;
  const int C = (int)lv.channels;
  switch (dtype) {
    case B200DET_F32: return launch_strip<float>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, level_mask, st, taken);
    case B200DET_F16: return launch_strip<__half>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, level_mask, st, taken);
    case B200DET_BF16: return launch_strip<__nv_bfloat16>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, level_mask, st, taken);
  }
  return 0;
}

void roi_align_bwd_strip_release() {
  std::lock_guard<std::mutex> lock(g_smutex);
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < 64; ++d)
    for (int i = 0; i < kStripSlots; ++i) {
      StripSlot& sl = g_sscratch[d].slot[i];
      if (sl.buf || sl.done) {
        cudaSetDevice(d);
        if (sl.buf) cudaFree(sl.buf);
        if (sl.done) cudaEventDestroy(sl.done);
        sl = StripSlot();
      }
    }
  cudaSetDevice(cur);
  if (g_sgarbage) {
    for (void* p : *g_sgarbage) cudaFree(p);
    g_sgarbage->clear();
  }
}

}  // namespace b200det
