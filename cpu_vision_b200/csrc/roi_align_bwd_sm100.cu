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

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kBwdWarps = 4;
constexpr int kMaxRows = 8;        // 2 pooled rows x 2 samples x 2 taps
constexpr int kMaxImages = 4096;   // images a plan can index; more -> single group
constexpr int kPlanSlots = 8;      // calls in flight on different streams before a slot is reused

struct BwdPlan {
  int unsorted;                    // set when rois are not grouped by ascending, in-range image index
  int pad[3];
  int img_start[kMaxImages + 1];   // first RoI of image i (valid when !unsorted)
  int task_ctr[kMaxImages];        // per group: next scatter task
  int zero_ctr[kMaxImages];        // per group: next zero chunk
  int zero_done[kMaxImages];       // per group: chunks finished
};

struct __align__(16) BYEnt { int rowoff; float w0; float w1; int pad; };

struct BwdWarpTables {
  XTab x;
  BYEnt y[kMaxRows];
};

// ---- plan ---------------------------------------------------------------------------------
// One thread per RoI; the plan was zeroed by a memset before.
__global__ void bwd_plan_kernel(const float* __restrict__ rois, int K, int N, BwdPlan* __restrict__ plan) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int b = (int)__ldg(rois + (long long)k * 5);
  const int prev = k > 0 ? (int)__ldg(rois + (long long)(k - 1) * 5) : -1;
  if (b < 0 || b >= N || b < prev) {
    plan->unsorted = 1;
    return;
  }
  for (int i = prev + 1; i <= b; ++i) plan->img_start[i] = k;
  if (k == K - 1)
    for (int i = b + 1; i <= N; ++i) plan->img_start[i] = K;
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

// ---- persistent kernel -----------------------------------------------------------------------
struct BwdArgs {
  const float* rois;
  const int32_t* levels;
  const void* grad;
  long long gs_k, gs_c, gs_h, gs_w;
  BwdPlan* plan;        // nullptr: one group, static task assignment, caller has zeroed
  long long K;
  int N, C, PH, PW, CT, RP;
  int ipg;              // images per group
  int zero_chunk;       // bytes per zero chunk (multiple of 512); 0: no zero fill
  bool vec, aligned;
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <typename T>
struct GroupView {
  int i0, i1;           // images [i0, i1)
  long long k0, k1;     // RoIs [k0, k1)
  int nchunks;
};

template <typename T>
__device__ __forceinline__ GroupView<T> group_view(const LevelTable<T>& lv, const BwdArgs& a, bool single, int g) {
  GroupView<T> v;
  if (single) {
    v.i0 = 0; v.i1 = a.N; v.k0 = 0; v.k1 = a.K;
  } else {
    v.i0 = g * a.ipg;
    v.i1 = min(a.N, v.i0 + a.ipg);
    v.k0 = a.plan->img_start[v.i0];
    v.k1 = a.plan->img_start[v.i1];
  }
  v.nchunks = 0;
  if (a.zero_chunk > 0)
    for (int l = 0; l < lv.num_levels; ++l) {
      const long long bytes = (long long)(v.i1 - v.i0) * lv.sn[l] * (long long)sizeof(T);
      v.nchunks += (int)((bytes + a.zero_chunk - 1) / a.zero_chunk);
    }
  return v;
}

// writes zero chunk c of the group; all lanes participate
template <typename T>
__device__ __forceinline__ void zero_chunk(const LevelTable<T>& lv, const BwdArgs& a, const GroupView<T>& v, int c,
                                           int lane) {
  int l = 0;
  long long bytes = 0;
  for (; l < lv.num_levels; ++l) {
    bytes = (long long)(v.i1 - v.i0) * lv.sn[l] * (long long)sizeof(T);
    const int n = (int)((bytes + a.zero_chunk - 1) / a.zero_chunk);
    if (c < n) break;
    c -= n;
  }
  if (l >= lv.num_levels) return;
  const long long off = (long long)c * a.zero_chunk;
  const int len = (int)min((long long)a.zero_chunk, bytes - off);  // multiple of 256: C % 128 == 0
  char* p = reinterpret_cast<char*>(lv.ptr[l] + (long long)v.i0 * lv.sn[l]) + off;
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll 4
  for (int o = lane * 16; o < len; o += 512) *reinterpret_cast<uint4*>(p + o) = z;
}

template <typename T, int CH, int SR>
__global__ void __launch_bounds__(kBwdWarps * 32, 4) roi_align_bwd_sm100_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                              const __grid_constant__ BwdArgs a) {
  __shared__ BwdWarpTables tabs[kBwdWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  BwdWarpTables& tb = tabs[warp];
  const T* grad = reinterpret_cast<const T*>(a.grad);
  const int tasks_per_roi = a.RP * a.CT;

  auto run_task = [&](long long k0, long long t) {
    const int rp = (int)(t % a.RP);
    const int ct = (int)((t / a.RP) % a.CT);
    const long long k = k0 + t / tasks_per_roi;
    bwd_task<T, CH, SR>(lv, a.rois, a.levels, grad, a.gs_k, a.gs_c, a.gs_h, a.gs_w, a.vec, k, ct, rp, a.N, a.C,
                        a.PH, a.PW, a.aligned, tb, lane);
  };

  if (a.plan == nullptr) {  // static assignment, no zero fill (the caller did it)
    const long long total = a.K * tasks_per_roi;
    for (long long t = blockIdx.x * (long long)kBwdWarps + warp; t < total; t += (long long)gridDim.x * kBwdWarps)
      run_task(0, t);
    return;
  }

  BwdPlan* plan = a.plan;
  const bool single = plan->unsorted != 0 || a.ipg >= a.N;
  const int ngroups = single ? 1 : (a.N + a.ipg - 1) / a.ipg;
  // claims zero chunks of group g while there are any; `once`: at most one
  auto zero_some = [&](const GroupView<T>& v, int g, bool once) {
    while (true) {
      int c = 0;
      if (lane == 0) c = atomicAdd(&plan->zero_ctr[g], 1);
      c = __shfl_sync(kFull, c, 0);
      if (c >= v.nchunks) return;
      zero_chunk<T>(lv, a, v, c, lane);
      __threadfence();
      __syncwarp();
      if (lane == 0) atomicAdd(&plan->zero_done[g], 1);
      if (once) return;
    }
  };

  GroupView<T> cur = group_view<T>(lv, a, single, 0);
  if (cur.nchunks > 0) zero_some(cur, 0, false);
  for (int g = 0; g < ngroups; ++g) {
    GroupView<T> nxt = cur;
    const bool more = g + 1 < ngroups;
    if (more) nxt = group_view<T>(lv, a, single, g + 1);
    if (cur.nchunks > 0) {  // every chunk of this group has been claimed by a running warp: wait for them
      if (lane == 0)
        while (ld_acquire(&plan->zero_done[g]) < cur.nchunks) __nanosleep(64);
      __syncwarp();
      __threadfence();
    }
    const long long ntasks = (cur.k1 - cur.k0) * tasks_per_roi;
    while (true) {
      int t = 0;
      if (lane == 0) t = atomicAdd(&plan->task_ctr[g], 1);
      t = __shfl_sync(kFull, t, 0);
      if (t >= ntasks) break;
      run_task(cur.k0, t);
      if (more && nxt.nchunks > 0) zero_some(nxt, g + 1, true);
    }
    if (more && nxt.nchunks > 0) zero_some(nxt, g + 1, false);
    cur = nxt;
  }
}

// ---- host side ---------------------------------------------------------------------------------
struct DeviceScratch {
  BwdPlan* plans = nullptr;  // kPlanSlots plans
  unsigned next = 0;
  int l2_bytes = 0;
  int sms = 0;
  bool tried = false;
};
DeviceScratch g_scratch[64];

// The plan ring is the one allocation this library makes on its own (2 MB per device,
// first backward call).  If it cannot be had (e.g. the first call happens inside a stream
// capture) the call falls back to memset + static task assignment.
DeviceScratch* scratch_for_current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  DeviceScratch& s = g_scratch[dev];
  if (!s.tried) {
    s.tried = true;
    cudaDeviceGetAttribute(&s.l2_bytes, cudaDevAttrL2CacheSize, dev);
    cudaDeviceGetAttribute(&s.sms, cudaDevAttrMultiProcessorCount, dev);
    void* p = nullptr;
    if (cudaMalloc(&p, sizeof(BwdPlan) * kPlanSlots) == cudaSuccess) s.plans = (BwdPlan*)p;
    else (void)cudaGetLastError();
  }
  return &s;
}

int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

template <typename T, int CH, int SR>
int launch_bwd_k(const LevelTable<T>& tab, const BwdArgs& a, int sms, cudaStream_t st) {
  int per_sm = 0;
  B200DET_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, roi_align_bwd_sm100_kernel<T, CH, SR>,
                                                                kBwdWarps * 32, 0));
  if (per_sm < 1) per_sm = 1;
  const long long total = a.K * a.RP * a.CT;
  long long blocks = (long long)sms * per_sm;
  const long long need = (total + kBwdWarps - 1) / kBwdWarps;
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  roi_align_bwd_sm100_kernel<T, CH, SR><<<(unsigned)blocks, kBwdWarps * 32, 0, st>>>(tab, a);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int launch_bwd(const LevelArgs& lv, const void* grad, const int64_t* gs, bool vec, const void* rois,
               const int32_t* levels, long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init,
               cudaStream_t st) {
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  const int CH = (C % 256 == 0) ? 8 : 4;
  BwdArgs a;
  a.rois = (const float*)rois;
  a.levels = levels;
  a.grad = grad;
  a.gs_k = gs[0]; a.gs_c = gs[1]; a.gs_h = gs[2]; a.gs_w = gs[3];
  a.K = K;
  a.N = (int)lv.batch; a.C = C; a.PH = PH; a.PW = PW;
  a.CT = C / (32 * CH);
  a.RP = (PH + 1) / 2;
  a.vec = vec; a.aligned = aligned;
  a.plan = nullptr; a.ipg = a.N; a.zero_chunk = 0;

  DeviceScratch* s = scratch_for_current_device();
  const int sms = s && s->sms > 0 ? s->sms : 148;
  const long long total = K * a.RP * a.CT;
  size_t image_bytes = 0;
  for (int l = 0; l < lv.num_levels; ++l) image_bytes += (size_t)lv.height[l] * lv.width[l] * C * sizeof(T);
  const bool fused = zero_init && s && s->plans && total < 0x7fffffffLL && K < 0x7fffffffLL &&
                     lv.batch <= 0x7fffffffLL && env_int("B200DET_BWD_FUSED", 1) != 0;
  if (fused) {
    // Phased mode (B200DET_BWD_PHASED=1): group = as many images as fit in a fraction of the L2
    // (B200DET_BWD_L2_PCT, default 30%).  Measured on B200 (profiles/r01_bwd_phased.md): with 37 MB
    // groups the reds do stay in L2 (DRAM traffic -26%), with 74 MB groups they do not, and in
    // both cases the kernel is NOT faster, because it is bound by the SM->L2 red/write path, not
    // by DRAM.  So the default is one group: zero fill and scatter fused in one launch.
    long long ipg = a.N;
    if (env_int("B200DET_BWD_PHASED", 0) != 0) {
      const double budget = (double)s->l2_bytes * env_int("B200DET_BWD_L2_PCT", 30) / 100.0;
      ipg = image_bytes ? (long long)(budget / (double)image_bytes) : a.N;
      if (ipg < 1) ipg = 1;
    }
    if (a.N > kMaxImages || ipg > a.N) ipg = a.N;
    const long long ngroups = (a.N + ipg - 1) / ipg;
    // one zero chunk per ~0.75 tasks so the fill hides behind the scatter of the previous group
    const double tasks_per_group = (double)total / (double)ngroups;
    double zc = (double)image_bytes * (double)ipg / (0.75 * (tasks_per_group > 1 ? tasks_per_group : 1));
    long long zero_chunk = ((long long)zc + 4095) / 4096 * 4096;
    if (zero_chunk < 16384) zero_chunk = 16384;
    if (zero_chunk > (1 << 20)) zero_chunk = 1 << 20;
    BwdPlan* plan = s->plans + (s->next++ % kPlanSlots);
    B200DET_CUDA_OK(cudaMemsetAsync(plan, 0, sizeof(BwdPlan), st));
    if (ipg < a.N) {
      const int th = 256;
      bwd_plan_kernel<<<(unsigned)((K + th - 1) / th), th, 0, st>>>((const float*)rois, (int)K, a.N, plan);
      B200DET_CUDA_OK(cudaGetLastError());
    }
    a.plan = plan;
    a.ipg = (int)ipg;
    a.zero_chunk = (int)zero_chunk;
  } else if (zero_init) {
    for (int l = 0; l < lv.num_levels; ++l) {
      const size_t bytes = (size_t)lv.batch * lv.channels * lv.height[l] * lv.width[l] * sizeof(T);
      if (bytes) B200DET_CUDA_OK(cudaMemsetAsync(lv.ptr[l], 0, bytes, st));
    }
  } else if (s && s->plans && total < 0x7fffffffLL) {
    // accumulate into what is there: dynamic tasks, no zero phases
    BwdPlan* plan = s->plans + (s->next++ % kPlanSlots);
    B200DET_CUDA_OK(cudaMemsetAsync(plan, 0, sizeof(BwdPlan), st));
    a.plan = plan;
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
  for (int d = 0; d < 64; ++d)
    if (g_scratch[d].plans) {
      int cur = 0;
      cudaGetDevice(&cur);
      cudaSetDevice(d);
      cudaFree(g_scratch[d].plans);
      cudaSetDevice(cur);
      g_scratch[d] = DeviceScratch();
    }
}

}  // namespace b200det
