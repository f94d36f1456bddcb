// roi_align_bwd_tile_sm100.cu -- tile-accumulating RoIAlign BACKWARD for sm_100a
// (channels_last, C % 128 == 0, sampling_ratio 1 or 2, contiguous-channel gradient).
//
// Why.  The scatter formulation (roi_align_bwd_sm100.cu, and the reference's
// atomics, cuda/roi_align_kernel.cu:303-327) sends every RoI's footprint through
// the SM->L2 red path: 4.1 GB of reds for cfg3 where only 1.2 GB of gradient
// exist, and that path (6 TB/s, tools/membw.cu) is what bounds it.  Here the
// gradient is produced tile by tile instead: a CTA owns a TH x TW pixel tile of
// one image/level with fp32 accumulators in shared memory, gathers the
// contributions of every RoI that overlaps the tile, and writes each gradient
// byte to HBM exactly once with a plain store.  No memset, no atomics, and the
// summation order is fixed (RoIs in ascending index), so the result is
// deterministic -- which the reference only offers through a slow pure-PyTorch
// path (ops/roi_align.py:35-200,251-255).
//
// Work split inside a CTA: warp w owns tile row w (no two warps ever touch the
// same accumulator, so shared memory needs no atomics); a lane owns CH channels.
// For a RoI that overlaps its row r the warp
//   * finds the pooled rows ph that tap r and their weights Wy[ph][r] (lane = ph),
//   * tabulates the x axis restricted to the tile's columns (sm100_common.cuh),
//   * sweeps the pooled columns: G[pw] = sum_ph Wy[ph][r] * grad[ph][pw] (register
//     double buffer, next column's loads in flight), spreads G over the two
//     feature columns under each x sample, and adds a finished column to its
//     shared-memory accumulator.
// A tiny plan kernel computes every RoI's footprint box once; a CTA finds the RoIs
// of its tile by scanning the boxes of its image (ordered compaction).
#include <stdlib.h>

#include <vector>

#include "sm100_common.cuh"

namespace b200det {
namespace {

constexpr int kTH = 8;             // tile rows = warps per CTA
constexpr int kTW = 16;            // tile columns
constexpr int kListCap = 1024;     // RoIs gathered per scan round
constexpr int kMaxImagesT = 4096;

struct TilePlan {
  int unsorted;
  int tile_ctr;
  int pad[2];
  int img_start[kMaxImagesT + 1];
};

// footprint of a RoI in its level: rows [ya, yb], columns [xa, xb] that receive taps
struct __align__(16) RoiBox { unsigned rows, cols; int level, batch; };

struct TileArgs {
  const float* rois;
  const int32_t* levels;
  const void* grad;
  long long gs_k, gs_h, gs_w;   // element strides of the upstream gradient (channel stride is 1)
  TilePlan* plan;
  const RoiBox* boxes;
  long long K;
  int N, C, PH, PW, CT;
  int num_tiles;
  int tile_base[B200DET_MAX_LEVELS + 1];  // first tile id of each level in processing order (coarse levels first)
  int tiles_x[B200DET_MAX_LEVELS], tiles_y[B200DET_MAX_LEVELS];
  bool aligned, accumulate;
};

// ---- plan: footprint boxes + first RoI of every image --------------------------------------
template <int SR>
__global__ void tile_plan_kernel(LevelTable<float> lv, const float* __restrict__ rois,
                                 const int32_t* __restrict__ levels, int K, int N, int PH, int PW, bool aligned,
                                 TilePlan* __restrict__ plan, RoiBox* __restrict__ boxes) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int l = levels ? __ldg(levels + k) : 0;
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(rois + (long long)k * 5, lv.scale[l], PH, PW, SR, aligned);
  int ya = 0x7fff, yb = -1, xa = 0x7fff, xb = -1;
  for (int s = 0; s < PH * SR; ++s) {
    const AxisTap<float> t = axis_tap<float>(sample_coord<float>(g.start_h, g.bin_h, s / SR, s % SR, SR), H);
    if (t.valid) { ya = min(ya, t.lo); yb = max(yb, t.hi); }
  }
  for (int s = 0; s < PW * SR; ++s) {
    const AxisTap<float> t = axis_tap<float>(sample_coord<float>(g.start_w, g.bin_w, s / SR, s % SR, SR), W);
    if (t.valid) { xa = min(xa, t.lo); xb = max(xb, t.hi); }
  }
  const int b = g.batch;
  const bool in_range = b >= 0 && b < N;
  if (!in_range || yb < 0 || xb < 0) { ya = 1; yb = 0; xa = 1; xb = 0; }  // never overlaps a tile
  RoiBox bx;
  bx.rows = (unsigned)ya | ((unsigned)yb << 16);
  bx.cols = (unsigned)xa | ((unsigned)xb << 16);
  bx.level = l;
  bx.batch = b;
  boxes[k] = bx;
  const int prev = k > 0 ? (int)__ldg(rois + (long long)(k - 1) * 5) : -1;
  if (!in_range || b < prev) {
    plan->unsorted = 1;
    return;
  }
  for (int i = prev + 1; i <= b; ++i) plan->img_start[i] = k;
  if (k == K - 1)
    for (int i = b + 1; i <= N; ++i) plan->img_start[i] = K;
}

// ---- x axis restricted to the tile's columns [x0, x1] -------------------------------------
// Same window / event scheme as build_x (sm100_common.cuh); a sample counts only
// if one of its two taps falls on a tile column, and columns are stored relative
// to x0 (-1 or tw for the neighbour just outside: the flush skips those).
struct __align__(16) TXTab {
  float4 samp[kMaxSX];          // (hx, lx, nnew as int bits, valid as int bits)
  int col[kMaxEvents];          // column (relative to x0) entering the window at shift e
  unsigned vmask;
};

template <int SR>
__device__ __forceinline__ void build_x_tile(const RoiGeom<float>& g, int W, int x0, int x1, int PW, int lane,
                                             TXTab& tb) {
  const int SX = PW * SR;
  const bool active = lane < SX;
  const int j = active ? lane : SX - 1;
  const float x = sample_coord<float>(g.start_w, g.bin_w, j / SR, j % SR, SR);
  const AxisTap<float> t = axis_tap<float>(x, W);
  const int xl = t.lo, xh = min(t.lo + 1, W - 1);
  const bool valid = t.valid && active && ((xl >= x0 && xl <= x1) || (xh >= x0 && xh <= x1));
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
  if (active) {
    tb.samp[lane] = make_float4(t.wl, t.wh, __int_as_float(nnew), __int_as_float(valid ? 1 : 0));
    if (nnew == 2) {
      tb.col[base] = xl - x0;
      tb.col[base + 1] = xh - x0;
    } else if (nnew == 1) {
      tb.col[base] = xh - x0;
    }
  }
  if (lane == 0) tb.vmask = vmask;
}

// ---- shared-memory accumulators ----------------------------------------------------------
// One pixel holds 32*CH floats in a lane-centric order: lane i keeps its first
// four channels at [4i, 4i+4) and (CH == 8) its other four at [128 + 4i, ...),
// whatever channels those are for T; every access is a conflict-free 16-byte one.
template <int CH>
__device__ __forceinline__ void smem_add(float* px, int lane, const f32x2 (&v)[CH / 2]) {
#pragma unroll
  for (int h = 0; h < CH / 4; ++h) {
    float4* p = reinterpret_cast<float4*>(px + h * 128 + lane * 4);
    float4 a = *p;
    float x, y, z, w;
    unpack2(v[2 * h], x, y);
    unpack2(v[2 * h + 1], z, w);
    a.x += x; a.y += y; a.z += z; a.w += w;
    *p = a;
  }
}

// ---- one (RoI, tile row): sweep of the pooled columns [pw_a, pw_b] -------------------------
template <typename T, int CH, int SR, int NR>
__device__ __forceinline__ void row_sweep(const TXTab& xt, float* __restrict__ acc_row, int tw,
                                          const T* const (&gp)[NR], const float (&wy)[NR], long long gs_w, int pw_a,
                                          int pw_b, int PW, int lane) {
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  f32x2 wv[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) wv[r] = pack2(wy[r], wy[r]);
  typename LV::raw A[NR], B[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) A[r] = LV::ld(gp[r] + pw_a * gs_w);
#pragma unroll
  for (int r = 0; r < NR; ++r) B[r] = LV::ld(gp[r] + min(pw_a + 1, PW - 1) * gs_w);
  f32x2 lo[H2], hi[H2];
#pragma unroll
  for (int i = 0; i < H2; ++i) { lo[i] = 0ull; hi[i] = 0ull; }
  int lo_col = -1, hi_col = -1, e = 0;
  auto flush = [&](const f32x2 (&a)[H2], int col) {
    if ((unsigned)col < (unsigned)tw) smem_add<CH>(acc_row + col * (32 * CH), lane, a);
  };
  // pooled column pw: G = sum_r wy_r * grad[ph_r][pw]; its registers are refilled with pw + 2
  auto bin = [&](typename LV::raw (&buf)[NR], int pw) {
    f32x2 G[H2], t[H2];
    const long long nxt = (long long)min(pw + 2, PW - 1) * gs_w;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      LV::pairs(buf[r], t);
#pragma unroll
      for (int i = 0; i < H2; ++i) G[i] = (r == 0) ? mul2(wv[0], t[i]) : fma2(wv[r], t[i], G[i]);
      buf[r] = LV::ld(gp[r] + nxt);
    }
#pragma unroll
    for (int s = 0; s < SR; ++s) {
      const float4 sm = xt.samp[pw * SR + s];
      if (__float_as_int(sm.w)) {
        for (int n = __float_as_int(sm.z); n > 0; --n) {  // the window shifts: the outgoing column is complete
          flush(lo, lo_col);
#pragma unroll
          for (int i = 0; i < H2; ++i) { lo[i] = hi[i]; hi[i] = 0ull; }
          lo_col = hi_col;
          hi_col = xt.col[e++];
        }
        const f32x2 hx = pack2(sm.x, sm.x), lx = pack2(sm.y, sm.y);
#pragma unroll
        for (int i = 0; i < H2; ++i) {
          lo[i] = fma2(hx, G[i], lo[i]);
          hi[i] = fma2(lx, G[i], hi[i]);
        }
      }
    }
  };
  for (int pw = pw_a; pw <= pw_b; pw += 2) {
    bin(A, pw);
    if (pw + 1 <= pw_b) bin(B, pw + 1);
  }
  flush(lo, lo_col);
  flush(hi, hi_col);
}

template <typename T, int CH, int SR>
__device__ __forceinline__ void roi_row(const LevelTable<T>& lv, const TileArgs& a, int l, long long k, int r, int x0,
                                        int tw, int ch, float* __restrict__ acc_row, TXTab& xt, int lane) {
  const int H = lv.height[l], W = lv.width[l];
  const RoiGeom<float> g = roi_geometry<float>(a.rois + k * 5, lv.scale[l], a.PH, a.PW, SR, a.aligned);
  // pooled rows that tap feature row r, and with which weight (lane = pooled row)
  float wph = 0.f;
  bool touched = false;
  if (lane < a.PH) {
#pragma unroll
    for (int iy = 0; iy < SR; ++iy) {
      const AxisTap<float> t = axis_tap<float>(sample_coord<float>(g.start_h, g.bin_h, lane, iy, SR), H);
      if (t.valid) {
        if (t.lo == r) { wph += t.wl; touched = true; }
        if (t.hi == r) { wph += t.wh; touched = true; }
      }
    }
  }
  wph *= 1.0f / (float)(SR * SR);
  unsigned mask = __ballot_sync(kFull, touched);
  if (!mask) return;
  __syncwarp();  // previous user of xt is done
  build_x_tile<SR>(g, W, x0, x0 + tw - 1, a.PW, lane, xt);
  __syncwarp();
  const unsigned vmask = xt.vmask;
  if (!vmask) return;
  const int pw_a = (__ffs(vmask) - 1) / SR, pw_b = (31 - __clz(vmask)) / SR;
  const T* gbase = reinterpret_cast<const T*>(a.grad) + k * a.gs_k + ch;
  while (mask) {
    const T* gp[4];
    float wy[4];
    int n = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (mask) {
        const int ph = __ffs(mask) - 1;
        mask &= mask - 1;
        gp[i] = gbase + ph * a.gs_h;
        wy[i] = __shfl_sync(kFull, wph, ph);
        n = i + 1;
      } else {
        gp[i] = gbase;
        wy[i] = 0.f;
      }
    }
    if (n == 1) {
      const T* const g1[1] = {gp[0]};
      const float w1[1] = {wy[0]};
      row_sweep<T, CH, SR, 1>(xt, acc_row, tw, g1, w1, a.gs_w, pw_a, pw_b, a.PW, lane);
    } else if (n == 2) {
      const T* const g2[2] = {gp[0], gp[1]};
      const float w2[2] = {wy[0], wy[1]};
      row_sweep<T, CH, SR, 2>(xt, acc_row, tw, g2, w2, a.gs_w, pw_a, pw_b, a.PW, lane);
    } else if (n == 3) {
      const T* const g3[3] = {gp[0], gp[1], gp[2]};
      const float w3[3] = {wy[0], wy[1], wy[2]};
      row_sweep<T, CH, SR, 3>(xt, acc_row, tw, g3, w3, a.gs_w, pw_a, pw_b, a.PW, lane);
    } else {
      const T* const g4[4] = {gp[0], gp[1], gp[2], gp[3]};
      const float w4[4] = {wy[0], wy[1], wy[2], wy[3]};
      row_sweep<T, CH, SR, 4>(xt, acc_row, tw, g4, w4, a.gs_w, pw_a, pw_b, a.PW, lane);
    }
  }
}

// ---- persistent tile kernel -----------------------------------------------------------------
template <typename T, int CH, int SR>
__global__ void __launch_bounds__(kTH * 32) roi_align_bwd_tile_kernel(const __grid_constant__ LevelTable<T> lv,
                                                                      const __grid_constant__ TileArgs a) {
  extern __shared__ __align__(16) float acc[];           // [kTH][kTW][32 * CH]
  __shared__ int2 list[kListCap];                        // (RoI index, rows ya | yb << 16), ascending RoI index
  __shared__ TXTab xtabs[kTH];
  __shared__ int s_tile, s_n, s_wsum[kTH];
  using LV = LaneVec<T, CH>;
  constexpr int H2 = CH / 2;
  constexpr int PX = 32 * CH;                            // floats per pixel
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float* acc_row = acc + warp * (kTW * PX);
  const bool single = a.plan->unsorted != 0;

  while (true) {
    __syncthreads();  // everybody is done with the previous tile's list
    if (tid == 0) s_tile = atomicAdd(&a.plan->tile_ctr, 1);
    __syncthreads();
    const int t = s_tile;
    if (t >= a.num_tiles) break;
    // decode: levels are processed coarse to fine (busy tiles first)
    int li = 0;
    while (t >= a.tile_base[li + 1]) ++li;
    const int l = lv.num_levels - 1 - li;
    int rem = t - a.tile_base[li];
    const int ct = rem % a.CT;
    rem /= a.CT;
    const int per_img = a.tiles_x[l] * a.tiles_y[l];
    const int b = rem / per_img;
    rem -= b * per_img;
    const int ty = rem / a.tiles_x[l], tx = rem - ty * a.tiles_x[l];
    const int H = lv.height[l], W = lv.width[l];
    const int y0 = ty * kTH, x0 = tx * kTW;
    const int th = min(kTH, H - y0), tw = min(kTW, W - x0);
    const int y1 = y0 + th - 1, x1 = x0 + tw - 1;
    const int r = y0 + warp;
    const bool have_row = warp < th;
    const int ch = ct * PX + lane_channel<T, CH>(lane);

    long long k0 = 0, k1 = a.K;
    if (!single) { k0 = a.plan->img_start[b]; k1 = a.plan->img_start[b + 1]; }
    bool dirty = false;  // this warp's accumulators hold something
    for (long long ks = k0; ks < k1; ks += kListCap) {
      if (ks != k0) __syncthreads();  // previous round's list fully consumed
      // ordered compaction of the RoIs whose footprint overlaps the tile: thread tid tests 4 consecutive RoIs
      unsigned hits = 0;
      int2 ent[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const long long k = ks + tid * 4 + i;
        ent[i] = make_int2(0, 0);
        if (k < k1) {
          const uint4 bx = __ldg(reinterpret_cast<const uint4*>(a.boxes + k));
          const int ya = bx.x & 0xffff, yb = bx.x >> 16, xa = bx.y & 0xffff, xb = bx.y >> 16;
          if ((int)bx.z == l && (int)bx.w == b && ya <= y1 && yb >= y0 && xa <= x1 && xb >= x0) {
            hits |= 1u << i;
            ent[i] = make_int2((int)(k - k0), (int)bx.x);
          }
        }
      }
      const int cnt = __popc(hits);
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_wsum[warp] = incl;
      __syncthreads();
      int pos = incl - cnt;
      for (int w = 0; w < warp; ++w) pos += s_wsum[w];
      if (tid == kTH * 32 - 1) s_n = pos + cnt;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (hits & (1u << i)) list[pos++] = ent[i];
      __syncthreads();
      const int n = s_n;
      if (have_row && n > 0) {
        for (int i = 0; i < n; ++i) {
          const int2 e = list[i];
          const int ya = e.y & 0xffff, yb = (unsigned)e.y >> 16;
          if (r < ya || r > yb) continue;
          if (!dirty) {  // first contribution: clear the row
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int o = lane; o < kTW * PX / 4; o += 32) reinterpret_cast<float4*>(acc_row)[o] = z;
            __syncwarp();
            dirty = true;
          }
          roi_row<T, CH, SR>(lv, a, l, k0 + e.x, r, x0, tw, ch, acc_row, xtabs[warp], lane);
        }
      }
    }
    // write the row out: every gradient byte goes to HBM once
    if (have_row) {
      __syncwarp();
      T* out = lv.ptr[l] + (long long)b * lv.sn[l] + ((long long)r * W + x0) * a.C + ch;
      for (int c = 0; c < tw; ++c) {
        f32x2 v[H2];
        if (dirty) {
#pragma unroll
          for (int h = 0; h < CH / 4; ++h) {
            const float4 f = *reinterpret_cast<const float4*>(acc_row + c * PX + h * 128 + lane * 4);
            v[2 * h] = pack2(f.x, f.y);
            v[2 * h + 1] = pack2(f.z, f.w);
          }
        } else {
#pragma unroll
          for (int i = 0; i < H2; ++i) v[i] = 0ull;
        }
        T* p = out + (long long)c * a.C;
        if (a.accumulate) {
          if (!dirty) continue;
          f32x2 old[H2];
          LV::pairs(LV::ld_stream(p), old);
#pragma unroll
          for (int i = 0; i < H2; ++i) {
            float x, y, ox, oy;
            unpack2(v[i], x, y);
            unpack2(old[i], ox, oy);
            v[i] = pack2(x + ox, y + oy);
          }
        }
        LV::st(p, v);
      }
    }
  }
}

// ---- host side ---------------------------------------------------------------------------------
struct TileScratch {
  char* buf = nullptr;
  size_t cap = 0;
  int sms = 0;
  bool tried = false;
};
TileScratch g_tscratch[64];
std::vector<void*>* g_garbage = nullptr;  // outgrown buffers, freed at release (kernels may still use them)

int tenv_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s ? atoi(s) : dflt;
}

// plan + boxes for K RoIs; nullptr when the memory cannot be had (e.g. inside a stream capture)
char* tile_scratch(size_t bytes, int* sms) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  TileScratch& s = g_tscratch[dev];
  if (!s.tried) {
    s.tried = true;
    cudaDeviceGetAttribute(&s.sms, cudaDevAttrMultiProcessorCount, dev);
  }
  *sms = s.sms > 0 ? s.sms : 148;
  if (s.cap < bytes) {
    void* p = nullptr;
    const size_t want = bytes + bytes / 2 + (1 << 20);
    if (cudaMalloc(&p, want) != cudaSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    if (s.buf) {
      if (!g_garbage) g_garbage = new std::vector<void*>();
      g_garbage->push_back(s.buf);
    }
    s.buf = (char*)p;
    s.cap = want;
  }
  return s.buf;
}

template <typename T, int CH, int SR>
int launch_tile_k(const LevelTable<T>& tab, const TileArgs& a, int sms, cudaStream_t st) {
  const int smem = kTH * kTW * 32 * CH * (int)sizeof(float);
  B200DET_CUDA_OK(cudaFuncSetAttribute(roi_align_bwd_tile_kernel<T, CH, SR>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int per_sm = 0;
  B200DET_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, roi_align_bwd_tile_kernel<T, CH, SR>,
                                                                kTH * 32, smem));
  if (per_sm < 1) per_sm = 1;
  long long blocks = (long long)sms * per_sm;
  if (blocks > a.num_tiles) blocks = a.num_tiles;
  roi_align_bwd_tile_kernel<T, CH, SR><<<(unsigned)blocks, kTH * 32, smem, st>>>(tab, a);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

template <typename T>
int launch_tile(const LevelArgs& lv, const void* grad, const int64_t* gs, const void* rois, const int32_t* levels,
                long long K, int C, int PH, int PW, int sr, bool aligned, bool zero_init, cudaStream_t st,
                int* taken) {
  const int CH = (C % 256 == 0 && tenv_int("B200DET_TILE_CH", 8) == 8) ? 8 : 4;
  TileArgs a;
  a.rois = (const float*)rois;
  a.levels = levels;
  a.grad = grad;
  a.gs_k = gs[0]; a.gs_h = gs[2]; a.gs_w = gs[3];
  a.K = K;
  a.N = (int)lv.batch; a.C = C; a.PH = PH; a.PW = PW;
  a.CT = C / (32 * CH);
  a.aligned = aligned;
  a.accumulate = !zero_init;
  long long tiles = 0;
  for (int li = 0; li < lv.num_levels; ++li) {
    const int l = lv.num_levels - 1 - li;
    a.tiles_x[l] = (int)((lv.width[l] + kTW - 1) / kTW);
    a.tiles_y[l] = (int)((lv.height[l] + kTH - 1) / kTH);
    a.tile_base[li] = (int)tiles;
    tiles += (long long)a.tiles_x[l] * a.tiles_y[l] * lv.batch * a.CT;
    if (tiles > 0x7fffffffLL) return 0;
  }
  for (int li = lv.num_levels; li <= B200DET_MAX_LEVELS; ++li) a.tile_base[li] = (int)tiles;
  a.num_tiles = (int)tiles;
  int sms = 148;
  const size_t plan_bytes = (sizeof(TilePlan) + 255) & ~(size_t)255;
  char* scratch = tile_scratch(plan_bytes + (size_t)K * sizeof(RoiBox), &sms);
  if (!scratch) return 0;
  a.plan = (TilePlan*)scratch;
  a.boxes = (const RoiBox*)(scratch + plan_bytes);
  B200DET_CUDA_OK(cudaMemsetAsync(a.plan, 0, sizeof(TilePlan), st));
  const LevelTable<float> ftab = make_nhwc_table<float>(lv);
  const int thr = 128, blk = (int)((K + thr - 1) / thr);
  if (sr == 1)
    tile_plan_kernel<1><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                             (RoiBox*)a.boxes);
  else
    tile_plan_kernel<2><<<blk, thr, 0, st>>>(ftab, (const float*)rois, levels, (int)K, a.N, PH, PW, aligned, a.plan,
                                             (RoiBox*)a.boxes);
  B200DET_CUDA_OK(cudaGetLastError());
  const LevelTable<T> tab = make_nhwc_table<T>(lv);
  int rc;
  if (CH == 8) rc = sr == 1 ? launch_tile_k<T, 8, 1>(tab, a, sms, st) : launch_tile_k<T, 8, 2>(tab, a, sms, st);
  else rc = sr == 1 ? launch_tile_k<T, 4, 1>(tab, a, sms, st) : launch_tile_k<T, 4, 2>(tab, a, sms, st);
  if (rc) return rc;
  *taken = 1;
  return 0;
}

}  // namespace

int roi_align_bwd_tile(const LevelArgs& lv, int dtype, const void* grad, const int64_t* gs, const void* rois,
                       const int32_t* levels, long long K, int PH, int PW, int sr, bool aligned, bool zero_init,
                       cudaStream_t st, int* taken) {
  *taken = 0;
  if (!sm100_eligible(lv, dtype, PH, PW, sr)) return 0;
  if (K == 0 || lv.batch == 0 || K > 0x7fffffffLL / 8 || lv.batch > kMaxImagesT) return 0;
  const size_t es = dtype_size(dtype);
  const long long per16 = 16 / (long long)es;
  const bool vec = gs[1] == 1 && gs[0] % per16 == 0 && gs[2] % per16 == 0 && gs[3] % per16 == 0 &&
                   ((uintptr_t)grad) % 16 == 0;
  if (!vec) return 0;  // strided gradients take the scatter kernel
  for (int l = 0; l < lv.num_levels; ++l)
    if (lv.height[l] > 0x7ffe || lv.width[l] > 0x7ffe) return 0;
  // RoIs that are not grouped by image make every tile scan all K boxes: fine up to a point
  const int C = (int)lv.channels;
  switch (dtype) {
    case B200DET_F32: return launch_tile<float>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
    case B200DET_F16: return launch_tile<__half>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
    case B200DET_BF16: return launch_tile<__nv_bfloat16>(lv, grad, gs, rois, levels, K, C, PH, PW, sr, aligned, zero_init, st, taken);
  }
  return 0;
}

void roi_align_bwd_tile_release() {
  for (int d = 0; d < 64; ++d)
    if (g_tscratch[d].buf) {
      int cur = 0;
      cudaGetDevice(&cur);
      cudaSetDevice(d);
      cudaFree(g_tscratch[d].buf);
      cudaSetDevice(cur);
      g_tscratch[d] = TileScratch();
    }
  if (g_garbage) {
    for (void* p : *g_garbage) cudaFree(p);
    g_garbage->clear();
  }
}

}  // namespace b200det
