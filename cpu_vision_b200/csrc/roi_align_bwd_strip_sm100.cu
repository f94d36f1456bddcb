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
                                  StripPlan* __restrict__ plan, RoiRec* __restrict__ recs) {
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
  const int prev = k > 0 ? (int)__ldg(rois + (long long)(k - 1) * 5) : -1;
  if (!in_range || b < prev) {  // not grouped by image: every strip scans all K records
    plan->unsorted = 1;
    return;
  }
  for (int i = prev + 1; i <= b; ++i) plan->img_start[i] = k;
  if (k == K - 1)
    for (int i = b + 1; i <= N; ++i) plan->img_start[i] = K;
}

__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// ---- shared-memory accumulators -----------------------------------------------------------
// One pixel holds 32*CH floats in a lane-centric order: lane i keeps its first four channels at
// [4i, 4i+4) and (CH == 8) its other four at [128 + 4i, ...), whatever channels those are for T;
// every access is a conflict-free 16-byte one.
template <int CH>
__device__ __forceinline__ void acc_add(float* px, const f32x2 (&v)[CH / 2]) {
#pragma unroll
  for (int h = 0; h < CH / 4; ++h) {
    ulonglong2* p = reinterpret_cast<ulonglong2*>(px + h * 128);
    ulonglong2 a = *p;
    a.x = add2(a.x, v[2 * h]);
    a.y = add2(a.y, v[2 * h + 1]);
    *p = a;
  }
}

// ---- x axis of a RoI restricted to the strip's columns [x0, x1] (lane j owns x sample j) ----
// The sweep keeps a window of two columns (c, c + 1).  Entry j = (weight of column c, weight of
// column c + 1, float offset of column c relative to x0 in the accumulators, shift): shift 0 keeps the window of the previous valid sample,
// 1 moves it one column to the right, 2 restarts it at c.  A sample whose two taps both fall
// outside the strip, or that is empty (roi_align_kernel.cu:23-26), is all zeros: it adds nothing and
// moves nothing.  c may be -1 (only the right tap is inside) and c + 1 may be tw (only the left
// one, or the clamped last column where the right weight is 0): the flush skips those.
template <int SR>
__device__ __forceinline__ unsigned build_x_strip(float start_w, float bin_w, int W, int x0, int x1, int PW, int lane,
                                                  int px_floats, int4* __restrict__ samp) {
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
  const int shift = !valid ? 0 : (!before ? 2 : (d == 0 ? 0 : (d == 1 ? 1 : 2)));
  // integers, so that c = -1 (all ones, a NaN pattern) cannot be touched by a floating-point select
  if (active)
    samp[lane] = valid ? make_int4(__float_as_int(t.wl), __float_as_int(t.wh), (xl - x0) * px_floats, shift)
                       : make_int4(0, 0, 0, 0);
  return vmask;
}

// ---- one pass of a (RoI, strip) task: pooled rows ph0 .. ph0 + nr - 1, pooled columns [pw_a, pw_b] ----
template <typename T, int CH, int SR>
__device__ __forceinline__ void strip_sweep(const int4* __restrict__ samp, float* __restrict__ acc, int tw,
                                            const T* __restrict__ gp, int gs_h, int gs_w, int nr,
                                            const float (&wy)[kNRP], int pw_a, int pw_b) {
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  constexpr int PX = 32 * CH;
  typename LV::raw A[kNRP] = {}, B[kNRP] = {};
  const int pw_b2 = min(pw_a + 1, pw_b);
#pragma unroll
  for (int r = 0; r < kNRP; ++r)
    if (r < nr) A[r] = LV::ld(gp + (r * gs_h + pw_a * gs_w));
#pragma unroll
  for (int r = 0; r < kNRP; ++r)
    if (r < nr) B[r] = LV::ld(gp + (r * gs_h + pw_b2 * gs_w));
  f32x2 wv[kNRP];
#pragma unroll
  for (int r = 0; r < kNRP; ++r) wv[r] = pack2(wy[r], wy[r]);
  f32x2 X[H2], Y[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { X[i] = 0ull; Y[i] = 0ull; }
  // window = columns (c, c + 1), tracked as the float offset co = c * PX of column c; both accumulators are zero
  int co = -4 * PX;
  const int lim = tw * PX;
  auto flush = [&](const f32x2 (&a)[H2], int ofs) {
    if ((unsigned)ofs < (unsigned)lim) acc_add<CH>(acc + ofs, a);
  };
  // pooled column pw: G = sum_r wy_r * grad[ph_r][pw]; its registers are refilled with column pw + 2
  auto bin = [&](typename LV::raw (&buf)[kNRP], int pw) {
    f32x2 G[H2], t[H2];
    const int nxt = min(pw + 2, pw_b) * gs_w;
#pragma unroll
    for (int r = 0; r < kNRP; ++r) {
      LV::pairs(buf[r], t);
#pragma unroll
      for (int i = 0; i < H2; ++i) G[i] = (r == 0) ? mul2(wv[0], t[i]) : fma2(wv[r], t[i], G[i]);
      if (r < nr) buf[r] = LV::ld(gp + (r * gs_h + nxt));
    }
#pragma unroll
    for (int s = 0; s < SR; ++s) {
      const int4 e = samp[pw * SR + s];
      const int shift = e.w;
      if (shift) {  // the window moves: the outgoing column is complete
        // (the new offset is formed before the flushes on purpose: ptxas 12.9 otherwise assigns the old c + 1 and
        // the new c to one register and flushes Y one window too far -- seen in the SASS, the PTX was right)
        const int co_old = co;
        co = shift == 2 ? e.z : co + PX;
        flush(X, co_old);
        if (shift == 2) {
          flush(Y, co_old + PX);
#pragma unroll
          for (int i = 0; i < H2; ++i) { X[i] = 0ull; Y[i] = 0ull; }
        } else {
#pragma unroll
          for (int i = 0; i < H2; ++i) { X[i] = Y[i]; Y[i] = 0ull; }
        }
      }
      const float fl = __int_as_float(e.x), fh = __int_as_float(e.y);
      const f32x2 wl = pack2(fl, fl), wh = pack2(fh, fh);
#pragma unroll
      for (int i = 0; i < H2; ++i) {
        X[i] = fma2(wl, G[i], X[i]);
        Y[i] = fma2(wh, G[i], Y[i]);
      }
    }
  };
#pragma unroll 1
  for (int pw = pw_a; pw <= pw_b; pw += 2) {
    bin(A, pw);
    if (pw + 1 <= pw_b) bin(B, pw + 1);
  }
  flush(X, co);
  flush(Y, co + PX);
}

// ---- persistent strip kernel ----------------------------------------------------------------
template <typename T, int CH, int SR>
__global__ void __launch_bounds__(kSW * 32, 1) roi_align_bwd_strip_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                          const __grid_constant__ StripArgs a) {
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  constexpr int PX = 32 * CH;  // floats per pixel
  constexpr int TW = StripShape<CH>::TW;
  extern __shared__ __align__(16) float smem[];  // [kSW][TW][PX] accumulators, then [kSW][32] x-sample entries
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* acc = smem + warp * (TW * PX) + lane * 4;  // this lane's slot of column 0
  int4* samp = reinterpret_cast<int4*>(smem + kSW * TW * PX) + warp * 32;
  {
    const ulonglong2 z = make_ulonglong2(0ull, 0ull);
    for (int o = 0; o < TW * (CH / 4); ++o) *reinterpret_cast<ulonglong2*>(acc + o * 128) = z;
  }
  const bool single = a.plan->unsorted != 0;
  const T* grad = reinterpret_cast<const T*>(a.grad);
  const float inv_count = 1.0f / (float)(SR * SR);

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

    int k0 = 0, k1 = (int)a.K;
    if (!single) { k0 = a.plan->img_start[b]; k1 = a.plan->img_start[b + 1]; }
    bool dirty = false;
    for (int ks = k0; ks < k1; ks += 32) {
      const int k = ks + lane;
      bool hit = false;
      if (k < k1) {
        const uint4 bx = __ldg(reinterpret_cast<const uint4*>(a.recs + k) + 1);
        const int ya = bx.x & 0xffff, yb = bx.x >> 16, xa = bx.y & 0xffff, xb = bx.y >> 16;
        hit = (int)bx.z == l && (int)bx.w == b && ya <= r && yb >= r && xa <= x1 && xb >= x0;
      }
      unsigned m = __ballot_sync(kFull, hit);
      while (m) {
        const int kk = ks + __ffs(m) - 1;
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
        __syncwarp();  // the previous task's sweep is done with samp[]
        const unsigned vmask = build_x_strip<SR>(geo.z, geo.w, W, x0, x1, a.PW, lane, PX, samp);
        __syncwarp();
        if (!vmask) continue;
        dirty = true;
        const int pw_a = (__ffs(vmask) - 1) / SR, pw_b = (31 - __clz(vmask)) / SR;
        // the pooled rows that tap one feature row are consecutive (samples are monotone in y)
        const int ph_lo = __ffs(pm) - 1, ph_hi = 31 - __clz(pm);
        const T* gbase = grad + (long long)kk * a.gs_k + ch;
        for (int ph0 = ph_lo; ph0 <= ph_hi; ph0 += kNRP) {
          float wy[kNRP];
#pragma unroll
          for (int i = 0; i < kNRP; ++i) {
            const float w = __shfl_sync(kFull, wph, (ph0 + i) & 31);
            wy[i] = ph0 + i <= ph_hi ? w : 0.f;
          }
          strip_sweep<T, CH, SR>(samp, acc, tw, gbase + ph0 * a.gs_h, a.gs_h, a.gs_w, min(kNRP, ph_hi - ph0 + 1), wy,
                                 pw_a, pw_b);
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
  const int smem = kSW * (TW * 32 * CH * (int)sizeof(float) + 32 * (int)sizeof(float4));
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
                 long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init, cudaStream_t st,
                 int* taken) {
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
    jobs += (long long)a.strips_x[l] * lv.height[l] * lv.batch * a.CT;
    if (jobs > 0x7fffffffLL) return 0;
  }
  for (int li = lv.num_levels; li <= B200DET_MAX_LEVELS; ++li) a.job_base[li] = (int)jobs;
  a.num_jobs = (int)jobs;
  int sms = 148;
  const size_t plan_bytes = (sizeof(StripPlan) + 255) & ~(size_t)255;
  StripSlot* sl = strip_slot(plan_bytes + (size_t)K * sizeof(RoiRec), st, &sms);
  if (!sl) return 0;
  a.plan = (StripPlan*)sl->buf;
  a.recs = (const RoiRec*)(sl->buf + plan_bytes);
  B200DET_CUDA_OK(cudaMemsetAsync(a.plan, 0, sizeof(StripPlan), st));
  const LevelTable<float> ftab = make_nhwc_table<float>(lv);
  const int thr = 128, blk = (int)((K + thr - 1) / thr);
  if (sr == 1)
    strip_plan_kernel<1><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                              (RoiRec*)a.recs);
  else
    strip_plan_kernel<2><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                              (RoiRec*)a.recs);
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
                        cudaStream_t st, int* taken) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr)) return 0;
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
    case B200DET_F32: return launch_strip<float>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
    case B200DET_F16: return launch_strip<__half>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
    case B200DET_BF16: return launch_strip<__nv_bfloat16>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
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
