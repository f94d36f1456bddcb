// nms_sm100.cu -- greedy NMS and class-aware NMS, entirely on the device.
//
// Pipeline (all on the caller's stream, no host round trip):
//   1-3. up to 13312 boxes: nms_front_rank_kernel -- every CTA stages all fp32 sort keys in shared memory and ranks
//        64 of them by counting (stable: ties keep ascending original index like scores.sort(stable=True,
//        descending=True), cuda/nms_kernel.cu:112-113), reduces boxes.max() for the coordinate trick
//        (boxes.py:91-93) and gathers its boxes into score order (B200DET_NMS_FRONT=block: the one-CTA block
//        radix sort nms_front_small_kernel).  Larger sets / fp64 scores: nms_prepare_kernel,
//        cub::DeviceRadixSort::SortPairsDescending, nms_max_coord_kernel, nms_gather_kernel.
//   4. nms_mask_kernel      upper-triangular 64x64 IoU>thr bit tiles
//                           (replaces nms_kernel_impl, cuda/nms_kernel.cu:31-78)
//   5b. nms_resolve_kernel  (sets of up to 13312 boxes) the greedy result by monotone propagation of finally kept /
//                           finally removed boxes on all SMs, a CTA per 64-box column block; decides everything
//                           in a few rounds unless suppression chains run deep across blocks
//   5. nms_scan_smem_kernel / nms_scan_kernel   greedy keep scan on the device, one CTA
//                           (replaces the D2H copy + host loop, :136-159);
//                           writes keep[] = order[kept] and the count.  Returns at once when 5b decided every box.
// Segmented (per-image) variant: see "segmented NMS" below.
//
// IoU follows devIoU (cuda/nms_kernel.cu:16-29) with every operation an
// explicit round-to-nearest intrinsic so no FMA contraction can change a
// comparison: inter / ((Sa + Sb) - inter) > (float)threshold.
#include <stdlib.h>

#include <mutex>
#include <type_traits>
#include <cub/block/block_radix_sort.cuh>
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "kernels.h"

namespace b200det {

namespace {

typedef unsigned long long u64;
constexpr int kTile = 64;

// Mask layout: upper-triangular, packed by row block.  Row block b keeps, for each of its 64 rows, only the
// words of column blocks j >= b: block base 64 * (b * cb - b (b - 1) / 2) words, row stride cb - b.  The rows of
// a block stay contiguous (one bulk copy stages them in the scan) and carry no words left of the diagonal,
// which halves what the scan reads.  Bases are multiples of 512 bytes.
__host__ __device__ __forceinline__ long long tri_base(int b, int cb) {
  return (long long)kTile * ((long long)b * cb - (long long)b * (b - 1) / 2);
}

// nms_resolve_kernel (section 5b): sets of up to kResolveMaxBlocks * 64 = 13312 boxes (110 KB of shared memory per
// CTA there: two CTAs per SM, so all 208 are resident at once), at most kResolveIters rounds
constexpr int kResolveMaxBlocks = 208;
constexpr int kResolveThreads = 256;
constexpr int kResolveIters = 24;
constexpr int kResolveCtlWords = 32;  // [1] CTAs done, [2] fail flag

// Class-grouped sets cut every block row at its last live tile jmax[b] (nothing right of it can hold a bit):
// row stride jmax[b] - b + 1, block bases from a prefix sum (wbase, in units of 64 words).  jmax == nullptr
// is the plain triangular layout.
struct MaskIdx {
  const int32_t* jmax;
  const int32_t* wbase;
  int cb;
  __device__ __forceinline__ int width(int b) const { return jmax ? jmax[b] - b + 1 : cb - b; }
  __device__ __forceinline__ long long base(int b) const {
    return jmax ? (long long)kTile * wbase[b] : tri_base(b, cb);
  }
  __device__ __forceinline__ long long word(int b, int i, int j) const {
    return base(b) + (long long)i * width(b) + (j - b);
  }
};

struct alignas(16) Box4f { float x1, y1, x2, y2; };
struct alignas(32) Box4d { double x1, y1, x2, y2; };

__device__ __forceinline__ bool iou_gt(const Box4f& a, const Box4f& b, float thr) {
  const float left = fmaxf(a.x1, b.x1), right = fminf(a.x2, b.x2);
  const float top = fmaxf(a.y1, b.y1), bottom = fminf(a.y2, b.y2);
  const float w = fmaxf(__fsub_rn(right, left), 0.f), h = fmaxf(__fsub_rn(bottom, top), 0.f);
  const float inter = __fmul_rn(w, h);
  const float sa = __fmul_rn(__fsub_rn(a.x2, a.x1), __fsub_rn(a.y2, a.y1));
  const float sb = __fmul_rn(__fsub_rn(b.x2, b.x1), __fsub_rn(b.y2, b.y1));
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(sa, sb), inter)) > thr;
}
__device__ __forceinline__ bool iou_gt(const Box4d& a, const Box4d& b, float thr) {
  const double left = fmax(a.x1, b.x1), right = fmin(a.x2, b.x2);
  const double top = fmax(a.y1, b.y1), bottom = fmin(a.y2, b.y2);
  const double w = fmax(__dsub_rn(right, left), 0.), h = fmax(__dsub_rn(bottom, top), 0.);
  const double inter = __dmul_rn(w, h);
  const double sa = __dmul_rn(__dsub_rn(a.x2, a.x1), __dsub_rn(a.y2, a.y1));
  const double sb = __dmul_rn(__dsub_rn(b.x2, b.x1), __dsub_rn(b.y2, b.y1));
  return __ddiv_rn(inter, __dsub_rn(__dadd_rn(sa, sb), inter)) > (double)thr;
}

template <typename T> struct BoxOf { using type = Box4f; };
template <> struct BoxOf<double> { using type = Box4d; };

// ---- 1. keys / iota / (optional) max coordinate ----------------------------
template <typename T>
__global__ void nms_prepare_kernel(const T* __restrict__ scores, long long n,
                                   float* __restrict__ keys, int32_t* __restrict__ idx) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = (float)Io<T>::ld(scores + i);
  idx[i] = (int32_t)i;
}
// fp64 scores cannot be ranked through fp32 keys without changing ties; they
// get a 64-bit key path.
__global__ void nms_prepare_kernel_f64(const double* __restrict__ scores, long long n,
                                       double* __restrict__ keys, int32_t* __restrict__ idx) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  keys[i] = scores[i];
  idx[i] = (int32_t)i;
}

// boxes.max() for the coordinate trick (boxes.py:91): one CTA, fp32 compare.
template <typename T>
__global__ void nms_max_coord_kernel(const T* __restrict__ boxes, long long n4,
                                     typename Io<T>::acc* __restrict__ out) {
  using A = typename Io<T>::acc;
  __shared__ A red[32];
  A m = -INFINITY;
  bool nan = false;
  for (long long i = threadIdx.x; i < n4; i += blockDim.x) {
    const A v = Io<T>::ld(boxes + i);
    nan |= (v != v);
    m = v > m ? v : m;
  }
  if (nan) m = NAN;  // torch.max propagates NaN
  for (int o = 16; o > 0; o >>= 1) {
    const A other = __shfl_xor_sync(0xffffffffu, m, o);
    m = (other != other || m != m) ? (A)NAN : (other > m ? other : m);
  }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : (A)-INFINITY;
    for (int o = 16; o > 0; o >>= 1) {
      const A other = __shfl_xor_sync(0xffffffffu, m, o);
      m = (other != other || m != m) ? (A)NAN : (other > m ? other : m);
    }
    if (threadIdx.x == 0) *out = m;
  }
}

// ---- 3. gather into score order ---------------------------------------------
// mode: 0 plain, 1 coordinate trick (adds idxs*(max+1) with the reference's
// fp32 rounding: boxes.py:92-93), 2 vanilla (keeps class id beside the box).
template <typename T>
__global__ void nms_gather_kernel(const T* __restrict__ boxes, const int64_t* __restrict__ idxs,
                                  const int32_t* __restrict__ order, long long n, int mode,
                                  const typename Io<T>::acc* __restrict__ max_coord,
                                  typename BoxOf<T>::type* __restrict__ sorted,
                                  int32_t* __restrict__ sorted_cls) {
  using A = typename Io<T>::acc;
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long src = order[i];
  A x1 = Io<T>::ld(boxes + src * 4 + 0), y1 = Io<T>::ld(boxes + src * 4 + 1);
  A x2 = Io<T>::ld(boxes + src * 4 + 2), y2 = Io<T>::ld(boxes + src * 4 + 3);
  if (mode == 1) {
    // offsets = idxs.to(boxes) * (max_coordinate + 1); boxes + offsets[:, None]
    // For fp16/bf16 boxes torch does this arithmetic in the boxes' dtype and
    // nms then runs in that dtype; we are only exact for fp32/fp64 here and
    // follow the autocast (fp32) semantics for the 16-bit types.
    const A off = Ar<A>::mul((A)idxs[src], Ar<A>::add(*max_coord, (A)1));
    x1 = Ar<A>::add(x1, off); y1 = Ar<A>::add(y1, off);
    x2 = Ar<A>::add(x2, off); y2 = Ar<A>::add(y2, off);
  }
  typename BoxOf<T>::type b;
  b.x1 = x1; b.y1 = y1; b.x2 = x2; b.y2 = y2;
  sorted[i] = b;
  if (sorted_cls) sorted_cls[i] = (int32_t)idxs[src];
}

// ---- 1-3 fused for detection-sized sets ----------------------------------------
// Up to kSortMax boxes: ONE CTA builds the keys, sorts them (block radix sort: LSD, stable, the same
// float ordering as the device-wide sort), reduces the maximum coordinate and gathers the boxes.  The
// device-wide radix sort costs 6 launches of 3-11 us for 5000 keys and prepare / max / gather three more;
// this is one launch.  Padding keys sort last (0xffffffff is the lowest float in radix order; a real key
// with that bit pattern still precedes the padding because the sort is stable).
// Programmatic dependent launch (sm_90+): the three launches of a small set (front end, IoU mask, scan) depend on
// one another, so the 5-10 us between them were launch latency, a third of the call.  A kernel that calls
// launch_dependents lets the next launch of the stream be scheduled at once; its CTAs sit in griddepcontrol.wait
// until the whole previous grid has finished and its writes are visible, then start without a launch gap.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }

template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

constexpr int kSortThreads = 1024, kSortItems = 8, kSortMax = kSortThreads * kSortItems;

template <typename T>
__global__ void __launch_bounds__(kSortThreads) nms_front_small_kernel(
    const T* __restrict__ boxes, const T* __restrict__ scores, const int64_t* __restrict__ idxs, int n, int mode,
    Box4f* __restrict__ sorted, int32_t* __restrict__ sorted_cls, int32_t* __restrict__ order) {
  using Sort = cub::BlockRadixSort<float, kSortThreads, kSortItems, int32_t>;
  __shared__ typename Sort::TempStorage tmp;
  __shared__ float red[32];
  __shared__ float s_max;
  pdl_launch_dependents();  // one CTA: the mask grid can take its places on the other SMs now
  const int tid = threadIdx.x;
  float k[kSortItems];
  int32_t v[kSortItems];
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int i = tid * kSortItems + j;
    k[j] = i < n ? (float)Io<T>::ld(scores + i) : __int_as_float(-1);
    v[j] = i < n ? i : 0x7fffffff;
  }
  Sort(tmp).SortDescending(k, v);
  if (mode == 1) {  // boxes.max() as nms_max_coord_kernel computes it (NaN propagates)
    float m = -INFINITY;
    bool nan = false;
    for (int i = tid; i < 4 * n; i += kSortThreads) {
      const float x = Io<T>::ld(boxes + i);
      nan |= (x != x);
      m = x > m ? x : m;
    }
    if (nan) m = NAN;
    for (int o = 16; o > 0; o >>= 1) {
      const float other = __shfl_xor_sync(0xffffffffu, m, o);
      m = (other != other || m != m) ? NAN : (other > m ? other : m);
    }
    if ((tid & 31) == 0) red[tid >> 5] = m;
    __syncthreads();
    if (tid < 32) {
      m = red[tid];
      for (int o = 16; o > 0; o >>= 1) {
        const float other = __shfl_xor_sync(0xffffffffu, m, o);
        m = (other != other || m != m) ? NAN : (other > m ? other : m);
      }
      if (tid == 0) s_max = m;
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < kSortItems; ++j) {
    const int p = tid * kSortItems + j;
    if (p >= n) continue;
    const int src = v[j];
    order[p] = src;
    float x1 = Io<T>::ld(boxes + (size_t)src * 4 + 0), y1 = Io<T>::ld(boxes + (size_t)src * 4 + 1);
    float x2 = Io<T>::ld(boxes + (size_t)src * 4 + 2), y2 = Io<T>::ld(boxes + (size_t)src * 4 + 3);
    if (mode == 1) {  // same arithmetic as nms_gather_kernel
      const float off = __fmul_rn((float)idxs[src], __fadd_rn(s_max, 1.f));
      x1 = __fadd_rn(x1, off); y1 = __fadd_rn(y1, off);
      x2 = __fadd_rn(x2, off); y2 = __fadd_rn(y2, off);
    }
    Box4f b;
    b.x1 = x1; b.y1 = y1; b.x2 = x2; b.y2 = y2;
    sorted[p] = b;
    if (sorted_cls) sorted_cls[p] = (int32_t)idxs[src];
  }
}

// ---- 1-3 fused, over all SMs: rank sort ------------------------------------------
// The block radix sort above runs on ONE SM (8 passes of ranking + exchange over 8192 slots, ~35 us for 5000
// boxes) while 147 SMs wait.  A detection-sized set fits every SM's shared memory, so each CTA stages ALL n keys
// (n * 4 bytes), and ranks kRankKeys of them by counting: rank(i) = #{j : key_j > key_i} + #{j < i : key_j == key_i},
// which is exactly the position a stable descending sort gives box i.  kRankSplit threads share the count of one
// key (16-byte shared loads, the same address for the threads of a warp that hold the same part: broadcasts).
// n^2 / 4 compares spread over n / 64 CTAs: 5000 boxes -> 79 CTAs x 1250 compares per thread.
// Keys are compared as cub compares floats (radix order: sign-flipped bit patterns, -0 ranks with +0), so ties,
// NaN scores and signed zeros land where the block sort and the device-wide sort put them.
constexpr int kRankKeys = 64, kRankSplit = 4, kRankThreads = kRankKeys * kRankSplit;
constexpr int kRankMax = 13312;  // 52 KB of keys per CTA; the sets the shared-memory scan and the resolve kernel take

__device__ __forceinline__ unsigned radix_order_key(float f) {
  unsigned b = __float_as_uint(f);
  if (b == 0x80000000u) b = 0u;
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

template <typename T>
__global__ void __launch_bounds__(kRankThreads) nms_front_rank_kernel(
    const T* __restrict__ boxes, const T* __restrict__ scores, const int64_t* __restrict__ idxs, int n, int mode,
    Box4f* __restrict__ sorted, int32_t* __restrict__ sorted_cls, int32_t* __restrict__ order) {
  extern __shared__ __align__(16) unsigned s_key[];  // n rounded up to 4, padding = 0 (below every real key's tie rule)
  __shared__ float red[kRankThreads / 32];
  __shared__ float s_max;
  pdl_launch_dependents();
  const int tid = threadIdx.x;
  const int n4 = (n + 3) >> 2;
  for (int i = tid; i < n4 * 4; i += kRankThreads) s_key[i] = i < n ? radix_order_key((float)Io<T>::ld(scores + i)) : 0u;
  if (mode == 1) {  // boxes.max() as nms_max_coord_kernel computes it (NaN propagates); every CTA needs it
    float m = -INFINITY;
    bool nan = false;
    for (int i = tid; i < 4 * n; i += kRankThreads) {
      const float x = Io<T>::ld(boxes + i);
      nan |= (x != x);
      m = x > m ? x : m;
    }
    if (nan) m = NAN;
    for (int o = 16; o > 0; o >>= 1) {
      const float other = __shfl_xor_sync(0xffffffffu, m, o);
      m = (other != other || m != m) ? NAN : (other > m ? other : m);
    }
    if ((tid & 31) == 0) red[tid >> 5] = m;
  }
  __syncthreads();
  if (mode == 1 && tid < 32) {
    float m = tid < kRankThreads / 32 ? red[tid] : -INFINITY;
    for (int o = 16; o > 0; o >>= 1) {
      const float other = __shfl_xor_sync(0xffffffffu, m, o);
      m = (other != other || m != m) ? NAN : (other > m ? other : m);
    }
    if (tid == 0) s_max = m;
  }
  const int part = tid % kRankSplit;
  const int i = blockIdx.x * kRankKeys + tid / kRankSplit;
  int rank = 0;
  if (i < n) {
    const unsigned ki = s_key[i];
    const uint4* k4 = reinterpret_cast<const uint4*>(s_key);
    for (int q = part; q < n4; q += kRankSplit) {
      const uint4 k = k4[q];
      const int j = q * 4;
      rank += (k.x > ki) | ((k.x == ki) & (j + 0 < i));
      rank += (k.y > ki) | ((k.y == ki) & (j + 1 < i));
      rank += (k.z > ki) | ((k.z == ki) & (j + 2 < i));
      rank += (k.w > ki) | ((k.w == ki) & (j + 3 < i));
    }
  }
  rank += __shfl_xor_sync(0xffffffffu, rank, 1);
  rank += __shfl_xor_sync(0xffffffffu, rank, 2);
  __syncthreads();  // s_max
  if (i >= n || part != 0) return;
  order[rank] = i;
  float x1 = Io<T>::ld(boxes + (size_t)i * 4 + 0), y1 = Io<T>::ld(boxes + (size_t)i * 4 + 1);
  float x2 = Io<T>::ld(boxes + (size_t)i * 4 + 2), y2 = Io<T>::ld(boxes + (size_t)i * 4 + 3);
  if (mode == 1) {  // same arithmetic as nms_gather_kernel
    const float off = __fmul_rn((float)idxs[i], __fadd_rn(s_max, 1.f));
    x1 = __fadd_rn(x1, off); y1 = __fadd_rn(y1, off);
    x2 = __fadd_rn(x2, off); y2 = __fadd_rn(y2, off);
  }
  Box4f b;
  b.x1 = x1; b.y1 = y1; b.x2 = x2; b.y2 = y2;
  sorted[rank] = b;
  if (sorted_cls) sorted_cls[rank] = (int32_t)idxs[i];
}

// ---- 4. IoU bit tiles ---------------------------------------------------------
// grid = (col_blocks, col_blocks); tiles below the diagonal exit at once and
// their words are never read (the scan starts at the diagonal word).
// One CTA = one 64-box row block x kColTiles consecutive column blocks (a CTA per single tile is
// launch-bound: 1.2 M CTAs of ~1 us for 64 images x 8663 boxes).  With a non-negative threshold a
// pair that does not intersect cannot pass (inter = 0 -> IoU is 0 or NaN), so it skips the division.
constexpr int kColTiles = 8;

template <typename BoxT>
__device__ __forceinline__ bool boxes_meet(const BoxT& a, const BoxT& b) {
  return fminf((float)a.x2, (float)b.x2) > fmaxf((float)a.x1, (float)b.x1) &&
         fminf((float)a.y2, (float)b.y2) > fmaxf((float)a.y1, (float)b.y1);
}
template <>
__device__ __forceinline__ bool boxes_meet<Box4d>(const Box4d& a, const Box4d& b) {
  return fmin(a.x2, b.x2) > fmax(a.x1, b.x1) && fmin(a.y2, b.y2) > fmax(a.y1, b.y1);
}

// fp32 fast path.  The reference decides fl(inter / union) > thr (nms_kernel.cu's devIoU); an IEEE division per
// pair is ~12 instructions plus a slow-path branch.  r = fma(-thr, union, inter) has the sign of
// inter / union - thr with ONE rounding, so whenever |r| > 2^-21 * union (and union > 0) the quotient is more
// than 4 ulp(thr) away from thr and its rounding cannot change the answer: such pairs are decided by the sign
// of r.  The rest (a few per million; also union <= 0, NaN, inf, where the comparisons come out false) are
// collected in a second bit word and decided by the exact division afterwards.  Valid for 2^-20 <= thr <= 1.
constexpr float kSureMargin = 4.76837158203125e-07f;  // 2^-21

__device__ __forceinline__ float box_area(const Box4f& b) {
  return __fmul_rn(__fsub_rn(b.x2, b.x1), __fsub_rn(b.y2, b.y1));
}

// packed pairs of fp32 (FADD2 / FMUL2 / FFMA2 on sm_100): each half is the IEEE .rn operation
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

__device__ __forceinline__ bool box_is_proper(const Box4f& b, float area) {
  return b.x2 >= b.x1 && b.y2 >= b.y1 && area > 0.f && area < INFINITY;
}

// column boxes of a tile, one array per field so that two neighbouring columns load as one packed pair
struct ColTile { float x1[kTile], y1[kTile], x2[kTile], y2[kTile], area[kTile]; };

// Two columns per step: min / max stay scalar (no packed form), the subtractions, products and the fma run packed.
template <bool HAS_CLS>
__device__ __forceinline__ u64 tile_bits_f32(const Box4f& a, float sa, int32_t ac, const ColTile& c,
                                             const int32_t* __restrict__ cc, float thr, u64 valid, u64 improper) {
  unsigned w[2] = {0u, 0u}, u[2] = {0u, 0u};
  const f32x2 sa2 = pk2(sa, sa), nthr2 = pk2(-thr, -thr), eps2 = pk2(kSureMargin, kSureMargin);
#pragma unroll
  for (int i = 0; i < kTile; i += 2) {
    const float2 bx1 = *reinterpret_cast<const float2*>(&c.x1[i]), by1 = *reinterpret_cast<const float2*>(&c.y1[i]);
    const float2 bx2 = *reinterpret_cast<const float2*>(&c.x2[i]), by2 = *reinterpret_cast<const float2*>(&c.y2[i]);
    const f32x2 sb2 = *reinterpret_cast<const f32x2*>(&c.area[i]);
    const f32x2 left = pk2(fmaxf(a.x1, bx1.x), fmaxf(a.x1, bx1.y)), right = pk2(fminf(a.x2, bx2.x), fminf(a.x2, bx2.y));
    const f32x2 top = pk2(fmaxf(a.y1, by1.x), fmaxf(a.y1, by1.y)), bottom = pk2(fminf(a.y2, by2.x), fminf(a.y2, by2.y));
    float w0, w1, h0, h1;
    upk2(sub2(right, left), w0, w1);
    upk2(sub2(bottom, top), h0, h1);
    const f32x2 inter = mul2(pk2(fmaxf(w0, 0.f), fmaxf(w1, 0.f)), pk2(fmaxf(h0, 0.f), fmaxf(h1, 0.f)));
    const f32x2 uni = sub2(add2(sa2, sb2), inter);
    float r0, r1, m0, m1;
    upk2(fma2(nthr2, uni, inter), r0, r1);
    upk2(mul2(uni, eps2), m0, m1);
    // union > 0 is guaranteed for proper boxes (x2 >= x1, y2 >= y1, finite area > 0: inter <= min(Sa, Sb), so
    // (Sa + Sb) - inter >= inter (1 - 2^-23) and > 0); improper rows / columns are sent to the exact path by `valid_fast`
    bool sure0 = fabsf(r0) > m0, sure1 = fabsf(r1) > m1;
    bool hit0 = sure0 && (r0 > 0.f), hit1 = sure1 && (r1 > 0.f);
    if (HAS_CLS) {
      const bool same0 = cc[i] == ac, same1 = cc[i + 1] == ac;
      hit0 = hit0 && same0; sure0 = sure0 || !same0;
      hit1 = hit1 && same1; sure1 = sure1 || !same1;
    }
    if (hit0) w[i >> 5] |= 1u << (i & 31);
    if (hit1) w[i >> 5] |= 2u << (i & 31);
    if (!sure0) u[i >> 5] |= 1u << (i & 31);
    if (!sure1) u[i >> 5] |= 2u << (i & 31);
  }
  u64 bits = (((u64)w[1] << 32) | w[0]) & valid & ~improper;
  u64 open = ((((u64)u[1] << 32) | u[0]) | improper) & valid;
  if (HAS_CLS) {  // an improper column of another class cannot set a bit
    for (u64 m = improper & valid; m; m &= m - 1) {
      const int i = __ffsll((long long)m) - 1;
      if (cc[i] != ac) open &= ~(1ULL << i);
    }
  }
  while (open) {  // rare: the exact quotient decides
    const int i = __ffsll((long long)open) - 1;
    open &= open - 1;
    Box4f b;
    b.x1 = c.x1[i]; b.y1 = c.y1[i]; b.x2 = c.x2[i]; b.y2 = c.y2[i];
    if (iou_gt(a, b, thr)) bits |= 1ULL << i;
  }
  return bits;
}

// Summary of one 64-box block of a class-grouped segment: the bounding box of its boxes (a NaN coordinate
// counts as unbounded on its side, as fmaxf / fminf treat it in devIoU) and the first / last class key.
struct BlockSum { float x1, y1, x2, y2; unsigned klo, khi; };

// No pair of the two blocks can set a bit: the bounding boxes do not meet (inter = 0 for every pair; only
// with thr >= 0) or, when bits need equal classes, the class-key ranges do not meet (keys ascend inside a segment).
__device__ __forceinline__ bool tile_is_empty(const BlockSum& r, const BlockSum& c, bool skip_apart, bool by_class) {
  if (by_class && c.klo > r.khi) return true;
  return skip_apart && !(fminf(r.x2, c.x2) > fmaxf(r.x1, c.x1) && fminf(r.y2, c.y2) > fmaxf(r.y1, c.y1));
}

template <typename BoxT>
__device__ __forceinline__ void mask_tiles(int n, float thr, const BoxT* __restrict__ boxes,
                                           const int32_t* __restrict__ cls, u64* __restrict__ mask,
                                           int row_start, int col_group, const BlockSum* __restrict__ blk = nullptr,
                                           const int32_t* __restrict__ jmax = nullptr,
                                           const int32_t* __restrict__ wbase = nullptr) {
  const MaskIdx mi{jmax, wbase, (n + kTile - 1) / kTile};
  const int col_blocks = (n + kTile - 1) / kTile;
  const int row_size = min(n - row_start * kTile, kTile);
  __shared__ BoxT cb[kTile];
  __shared__ int32_t cc[kTile];
  __shared__ __align__(16) ColTile ct;  // fp32 fast path only
  __shared__ unsigned s_improper[2];    // columns of the tile that must take the exact division
  const int cur = kTile * row_start + threadIdx.x;
  const bool have_row = (int)threadIdx.x < row_size;
  BoxT a = boxes[have_row ? cur : kTile * row_start];
  const int32_t ac = (cls && have_row) ? cls[cur] : 0;
  const bool skip_apart = thr >= 0.f;
  constexpr bool kF32 = std::is_same<BoxT, Box4f>::value;
  const bool fast = kF32 && thr >= 9.5367431640625e-07f && thr <= 1.f;
  float sa = 0.f;
  bool row_improper = false;
  if constexpr (kF32) {
    sa = box_area(a);
    row_improper = !box_is_proper(a, sa);
  }
  for (int t = 0; t < kColTiles; ++t) {
    const int col_start = col_group * kColTiles + t;
    if (col_start >= col_blocks) break;
    if (col_start < row_start) continue;
    const int col_size = min(n - col_start * kTile, kTile);
    if (jmax && col_start > jmax[row_start]) break;  // right of the block's last live tile: the scan never reads it
    if (blk && tile_is_empty(blk[row_start], blk[col_start], skip_apart, cls != nullptr)) {  // CTA-uniform
      if (have_row) mask[mi.word(row_start, threadIdx.x, col_start)] = 0ULL;
      continue;
    }
    __syncthreads();  // the previous tile's columns are no longer read
    bool col_improper = false;
    if ((int)threadIdx.x < col_size) {
      const BoxT b = boxes[kTile * col_start + threadIdx.x];
      cb[threadIdx.x] = b;
      cc[threadIdx.x] = cls ? cls[kTile * col_start + threadIdx.x] : 0;
      if constexpr (kF32) {
        const float area = box_area(b);
        ct.x1[threadIdx.x] = b.x1; ct.y1[threadIdx.x] = b.y1;
        ct.x2[threadIdx.x] = b.x2; ct.y2[threadIdx.x] = b.y2;
        ct.area[threadIdx.x] = area;
        col_improper = !box_is_proper(b, area);
      }
    }
    if constexpr (kF32) {
      const unsigned m = __ballot_sync(0xffffffffu, col_improper);
      if ((threadIdx.x & 31) == 0) s_improper[threadIdx.x >> 5] = m;
    }
    __syncthreads();
    if (!have_row) continue;
    const int start = (row_start == col_start) ? threadIdx.x + 1 : 0;
    u64 w = 0;
    if constexpr (kF32) {
      if (fast) {
        // entries past col_size hold stale boxes; their bits are masked off, like those up to the diagonal
        u64 valid = col_size == kTile ? ~0ULL : ((1ULL << col_size) - 1ULL);
        valid &= start >= kTile ? 0ULL : (~0ULL << start);
        const u64 improper = row_improper ? ~0ULL : (((u64)s_improper[1] << 32) | s_improper[0]);
        w = cls ? tile_bits_f32<true>(a, sa, ac, ct, cc, thr, valid, improper)
                : tile_bits_f32<false>(a, sa, ac, ct, cc, thr, valid, improper);
        mask[mi.word(row_start, threadIdx.x, col_start)] = w;
        continue;
      }
    }
    for (int i = start; i < col_size; ++i) {
      if (cc[i] != ac) continue;
      if (skip_apart && !boxes_meet(a, cb[i])) continue;
      if (iou_gt(a, cb[i], thr)) w |= 1ULL << i;
    }
    mask[mi.word(row_start, threadIdx.x, col_start)] = w;
  }
}

template <typename BoxT>
__global__ void __launch_bounds__(kTile) nms_mask_kernel(int n, float thr,
                                                         const BoxT* __restrict__ boxes,
                                                         const int32_t* __restrict__ cls,
                                                         u64* __restrict__ mask, unsigned* __restrict__ resolve_ctl,
                                                         u64* __restrict__ resolve_k) {
  pdl_launch_dependents();
  pdl_wait();  // (before the early exit too: the grid must not complete before its predecessor has)
  const int row_start = blockIdx.y, col_group = blockIdx.x;
  // nms_resolve_kernel starts from "nothing decided": its control words and its kept / removed words are zero
  if (resolve_ctl && row_start == 0 && col_group == 0) {
    if (threadIdx.x < kResolveCtlWords) resolve_ctl[threadIdx.x] = 0u;
    for (int w = threadIdx.x; w < 2 * ((n + kTile - 1) / kTile); w += kTile) resolve_k[w] = 0ULL;
  }
  if (row_start > col_group * kColTiles + kColTiles - 1) return;
  mask_tiles<BoxT>(n, thr, boxes, cls, mask, row_start, col_group);
}

// segmented: blockIdx.z = segment (image); boxes of segment s are sorted[off[s] .. off[s+1])
template <typename BoxT>
__global__ void __launch_bounds__(kTile) nms_seg_mask_kernel(const int32_t* __restrict__ off,
                                                             const long long* __restrict__ moff,
                                                             const int32_t* __restrict__ smode, float thr,
                                                             const BoxT* __restrict__ boxes,
                                                             const int32_t* __restrict__ cls,
                                                             u64* __restrict__ mask,
                                                             const int32_t* __restrict__ roff,
                                                             const BlockSum* __restrict__ blk,
                                                             const int32_t* __restrict__ jmax,
                                                             const int32_t* __restrict__ wbase) {
  const int s = blockIdx.z;
  const int o = off[s], n = off[s + 1] - o;
  const int row_start = blockIdx.y, col_group = blockIdx.x;
  if (row_start * kTile >= n || col_group * kColTiles * kTile >= n) return;
  if (row_start > col_group * kColTiles + kColTiles - 1) return;
  mask_tiles<BoxT>(n, thr, boxes + o, smode[s] == 2 ? cls + o : nullptr, mask + moff[s], row_start, col_group,
                   blk ? blk + roff[s] : nullptr, jmax ? jmax + roff[s] : nullptr, jmax ? wbase + roff[s] : nullptr);
}

// ---- 5. greedy scan -----------------------------------------------------------
// One CTA walks the 64-box blocks in score order; remv[] (one bit per box) lives in
// shared memory.  Per block: (a) one thread resolves the 64 in-block decisions from
// the diagonal words -- all 64 words are loaded up front, the dependent chain is a
// predicated AND per box; a box that is alive when visited is kept and is never
// cleared later (its row only has bits to its right), so kept == the final alive
// word; (b) all threads OR the kept rows into the words to the right.  Finally the
// kept bits are compacted into keep[].
//
// nms_scan_smem_kernel (col_blocks <= kSmemBlocks, i.e. N <= 13312 -- the detection
// sizes): the 64 mask rows of block b+1 are prefetched into shared memory with
// cp.async while block b is resolved, so no step waits on L2.
// nms_scan_kernel (any N): rows of kept boxes are read from global memory.
constexpr int kScanThreads = 1024;
constexpr int kSmemThreads = 512;
constexpr int kSmemBlocks = 208;  // 1064 bytes of shared memory per block + 2 KB of diagonal ring: 218 KB

__device__ __forceinline__ u64 resolve_block(u64 alive, const u64* __restrict__ diag, int stride) {
  u64 d[kTile];
#pragma unroll
  for (int i = 0; i < kTile; ++i) d[i] = diag[(size_t)i * stride];
#pragma unroll
  for (int i = 0; i < kTile; ++i)
    if ((alive >> i) & 1ULL) alive &= ~d[i];
  return alive;
}

// compaction of the kept bits: running offset over words, blockDim.x words per pass
__device__ __forceinline__ void compact_keep(const u64* keptw, int col_blocks, const int32_t* __restrict__ order,
                                             int64_t* __restrict__ keep, int64_t* __restrict__ num_keep,
                                             int* s_base, u64* s_total, int top_n = 0) {
  const int tid = threadIdx.x, nthreads = blockDim.x, nwarps = nthreads >> 5;
  int running = 0;
  for (int w0 = 0; w0 < col_blocks; w0 += nthreads) {
    const int w = w0 + tid;
    const u64 kw = w < col_blocks ? keptw[w] : 0ULL;
    const int cnt = __popcll(kw);
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) s_base[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
      const int v = tid < nwarps ? s_base[tid] : 0;
      int inc2 = v;
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, inc2, o);
        if (tid >= o) inc2 += u;
      }
      s_base[tid] = inc2 - v;  // exclusive
      if (tid == 31) *s_total = (u64)inc2;
    }
    __syncthreads();
    int pos = running + s_base[tid >> 5] + incl - cnt;
    u64 kk = kw;
    while (kk) {
      const int i = __ffsll((long long)kk) - 1;
      kk &= kk - 1;
      keep[pos++] = (int64_t)order[w * kTile + i];
    }
    running += (int)*s_total;
    __syncthreads();
  }
  if (tid == 0) *num_keep = (top_n > 0 && running > top_n) ? top_n : running;
}

// Class-grouped segments are scanned in (class, score) order (same kept SET: boxes of different classes never
// interact) but reported in score order: the kept bits move to their score rank first.  `scratch` holds
// col_blocks free words.
__device__ __forceinline__ void finish_keep(u64* keptw, u64* scratch, int col_blocks, const int32_t* __restrict__ rank,
                                            const int32_t* __restrict__ order, int64_t* __restrict__ keep,
                                            int64_t* __restrict__ num_keep, int* s_base, u64* s_total, int top_n) {
  if (rank) {
    const int tid = threadIdx.x, nthreads = blockDim.x;
    for (int j = tid; j < col_blocks; j += nthreads) scratch[j] = 0ULL;
    __syncthreads();
    for (int w = tid; w < col_blocks; w += nthreads) {
      u64 kk = keptw[w];
      while (kk) {
        const int i = __ffsll((long long)kk) - 1;
        kk &= kk - 1;
        const int r = rank[w * kTile + i];
        atomicOr(reinterpret_cast<unsigned*>(scratch) + (r >> 5), 1u << (r & 31));
      }
    }
    __syncthreads();
    keptw = scratch;
  }
  compact_keep(keptw, col_blocks, order, keep, num_keep, s_base, s_total, top_n);
}

// Pipelined: warp 0 is the serial resolver and runs one block AHEAD of the other warps.  In iteration b it
// completes remv[b + 1] with the kept rows of block b (only that one word) and resolves block b + 1 from its
// prefetched diagonal words, while warps 1.. OR the kept rows of block b into the words to the right of
// b + 1: one barrier per block, the dependent resolve overlaps the OR phase.
//   * the 64 rows of a block are contiguous in the mask, so ONE bulk copy (cp.async.bulk + mbarrier) stages
//     them; per-thread 8-byte cp.async made the LSU the bottleneck (1500 cycles per block).
//   * a block is resolved as a fixed point by the whole warp instead of a 64-step dependent chain:
//     K <- alive & ~OR{row j : j in K}, rows restricted to bits > j.  Bits below t are final after t rounds
//     (strictly upper-triangular dependence), so K' == K is reached in <= 64 rounds and is the greedy
//     result; typical boxes need a handful (the longest suppression chain inside the block).
constexpr int kDiagRing = 4;
#ifdef B200DET_SCAN_PROFILE
__device__ long long g_scan_prof[10];
#define SCAN_T(k) { const long long t_ = clock64(); prof_[k] += t_ - t0_; t0_ = t_; }
#else
#define SCAN_T(k)
#endif

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(u64* bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!ok);
}

// greedy keep bits of one 64-box block; lane l holds diagonal rows l and l + 32.  Whole warp, uniform result.
__device__ __forceinline__ u64 resolve_block_warp(u64 alive, u64 d0, u64 d1, int lane) {
  d0 &= ~((2ULL << lane) - 1ULL);         // bits > lane
  d1 &= ~((2ULL << (lane + 32)) - 1ULL);  // bits > lane + 32 (none for row 63)
  u64 K = alive;
  for (;;) {
    u64 v = ((K >> lane) & 1ULL) ? d0 : 0ULL;
    v |= ((K >> (lane + 32)) & 1ULL) ? d1 : 0ULL;
    const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
    const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
    const u64 Kn = alive & ~(((u64)hi << 32) | (u64)lo);
    if (Kn == K) return K;
    K = Kn;
  }
}

// OR phase of one block for worker warps 1..15.  remv is kept as kRowGroups partial words per column block:
// a unit of work is (row group g of 16 rows, 32 consecutive column blocks), its lanes read one row segment at
// a time (conflict-free) and OR into remv[g][j] with a plain read-modify-write -- each (g, j) has exactly one
// writer per block, so no atomics, and the resolver warp folds the partials of the one column it needs into
// the warp reduction it does anyway.  Measured alternatives: row groups meeting in shared atomics 670 cycles
// per block; one thread per column with 64 loads 1200-1500 (ptxas keeps two shared loads in flight at ~23
// cycles each whatever the source says, so warps have to hide the latency, not unrolling); row groups inside
// a warp combined by shuffles 1000 (the groups' row segments are 16 * stride words apart: 4-way bank conflicts).
constexpr int kRowGroups = 4, kGroupRows = kTile / kRowGroups;

__device__ __forceinline__ void or_rows(const u64* __restrict__ rowsb, u64* remv, u64 kept, int j0, int jend,
                                        int col_blocks, int stride, int warp, int lane, int b) {
  constexpr int kWorkers = kSmemThreads / 32 - 1;  // staged rows start at column block b, `stride` words apart
  const int units = ((jend - j0 + 31) / 32) * kRowGroups;
  for (int u = warp - 1; u < units; u += kWorkers) {  // warp-uniform
    const int g = u % kRowGroups, j = j0 + (u / kRowGroups) * 32 + lane;
    const unsigned kq = (unsigned)(kept >> (kGroupRows * g)) & ((1u << kGroupRows) - 1u);
    if (!kq || j >= jend) continue;
    const u64* col = rowsb + (size_t)(kGroupRows * g) * stride + (j - b);
    u64 acc = 0;
#pragma unroll
    for (int i = 0; i < kGroupRows; ++i) {
      const u64 v = col[(size_t)i * stride];
      acc |= ((kq >> i) & 1u) ? v : 0ULL;
    }
    if (acc) remv[g * col_blocks + j] |= acc;
  }
}

__device__ __forceinline__ void scan_smem_body(
    const u64* __restrict__ mask, int n, int col_blocks, const int32_t* __restrict__ order,
    int64_t* __restrict__ keep, int64_t* __restrict__ num_keep, const u64* __restrict__ init_remv, int top_n,
    const int32_t* __restrict__ rank = nullptr, const int32_t* __restrict__ jmax = nullptr,
    const int32_t* __restrict__ wbase = nullptr) {
  const MaskIdx mi{jmax, wbase, col_blocks};
  extern __shared__ __align__(16) u64 smem[];
  u64* remv = smem;                                    // [kRowGroups][col_blocks] partial words, OR = removed
  u64* keptw = smem + kRowGroups * col_blocks;         // [col_blocks]
  u64* ring = keptw + col_blocks;                      // [kDiagRing][kTile] diagonal words
  u64* stage = ring + kDiagRing * kTile + (col_blocks & 1);  // [2][kTile * col_blocks], on an even word (16 B)
  __shared__ __align__(8) u64 s_bar[2];
  __shared__ u64 s_kept;
  __shared__ int s_base[32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int j = tid; j < kRowGroups * col_blocks; j += kSmemThreads)
    remv[j] = (init_remv && j < col_blocks) ? init_remv[j] : 0ULL;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[0])) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&s_bar[1])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // rows of block b (all 64 are real for every block but the last, whose rows nobody reads): contiguous in the
  // packed-triangular mask and free of words left of the diagonal, so one bulk copy moves exactly what is read.
  // (With a row-major mask, a bulk copy per row for only the right part was measured at ~60 cycles of issue
  // per copy, 3900 cycles per block.)
  auto prefetch_rows = [&](int b) {  // one thread
    if (b + 1 >= col_blocks) return;
    const unsigned bytes = (unsigned)(kTile * mi.width(b) * sizeof(u64));
    const unsigned bar = smem_u32(&s_bar[b & 1]);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads of this stage are done
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(stage + (size_t)(b & 1) * kTile * col_blocks)),
                   "l"(mask + mi.base(b)), "r"(bytes), "r"(bar) : "memory");
  };
  auto prefetch_diag = [&](int b, int i) {  // mask[b*64 + i][b], i < rows
    if (b >= col_blocks || i >= min(n - b * kTile, kTile)) return;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;"
                 ::"r"(smem_u32(ring + (b & (kDiagRing - 1)) * kTile + i)),
                   "l"(mask + mi.word(b, i, b)) : "memory");
  };
  auto resolve = [&](int b, u64 removed) {  // warp 0, all lanes
    const int rows = min(n - b * kTile, kTile);
    u64 alive = ~removed;
    if (rows < kTile) alive &= (1ULL << rows) - 1ULL;  // stale ring entries belong to cleared bits
    const u64* d = ring + (b & (kDiagRing - 1)) * kTile;
    return resolve_block_warp(alive, d[lane], d[lane + 32], lane);
  };

  if (tid == kSmemThreads - 32) prefetch_rows(0);
  constexpr int kDiagT0 = kSmemThreads - 32 - kTile;  // the two warps before the last fetch diagonal words
  if (tid >= kDiagT0 && tid < kDiagT0 + kTile) {
    prefetch_diag(0, tid - kDiagT0);
    prefetch_diag(1, tid - kDiagT0);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    const u64 k0 = resolve(0, remv[0]);  // nothing has been ORed in yet: group 0 holds init_remv
    if (lane == 0) keptw[0] = k0;
  }
#ifdef B200DET_SCAN_PROFILE
  long long prof_[5] = {0, 0, 0, 0, 0};
  long long t0_ = clock64();
#endif
  for (int b = 0; b < col_blocks; ++b) {
    if (b + 1 < col_blocks) mbar_wait(&s_bar[b & 1], (unsigned)(b >> 1) & 1u);  // rows(b) landed
    asm volatile("cp.async.wait_group 0;" ::: "memory");                        // this thread's diag(b+1) words
    SCAN_T(0)
    __syncthreads();  // OR phase of block b-1 done; keptw[b] and diag(b+1) visible
    SCAN_T(1)
    if (tid == kSmemThreads - 32) prefetch_rows(b + 1);  // into the stage block b-1 has left (last warp: no OR work)
    if (tid >= kDiagT0 && tid < kDiagT0 + kTile) prefetch_diag(b + 2, tid - kDiagT0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    SCAN_T(2)
    const u64 kept = keptw[b];
    const u64* rowsb = stage + (size_t)(b & 1) * kTile * col_blocks;
    // class-grouped sets: words right of the block's last live tile were never written and are never read
    const int jend = jmax ? min(col_blocks, jmax[b] + 1) : col_blocks;
    const int stride = jend - b;  // row stride of the staged block = mi.width(b)
    if (warp == 0) {
      if (b + 1 < col_blocks) {
        // word b+1 of the kept rows of block b: two rows per lane, OR across the warp
        u64 v = 0;
        if (b + 1 < jend) {
          if ((kept >> lane) & 1ULL) v |= rowsb[(size_t)lane * stride + 1];
          if ((kept >> (lane + 32)) & 1ULL) v |= rowsb[(size_t)(lane + 32) * stride + 1];
        }
        if (lane < kRowGroups) v |= remv[lane * col_blocks + b + 1];  // the partial words of that column
        const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
        const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
        SCAN_T(3)
        const u64 k1 = resolve(b + 1, ((u64)hi << 32) | (u64)lo);
        if (lane == 0) keptw[b + 1] = k1;
        SCAN_T(4)
      }
    } else if (kept && b + 2 < jend) {
      or_rows(rowsb, remv, kept, b + 2, jend, col_blocks, stride, warp, lane, b);
    }
    if (warp != 0) { SCAN_T(3) }
  }
#ifdef B200DET_SCAN_PROFILE
  if (tid == 0 || tid == 40)
    for (int k = 0; k < 5; ++k) g_scan_prof[k + (tid ? 5 : 0)] = prof_[k];
#endif
  __syncthreads();
  finish_keep(keptw, remv, col_blocks, rank, order, keep, num_keep, s_base, &s_kept, top_n);
}

// ---- 5b. the greedy result by monotone propagation, on all SMs -------------------------------------
// A box is FINALLY REMOVED once one of its suppressors (a higher-ranked box with an IoU bit towards it) is finally
// kept, and FINALLY KEPT once all of its suppressors are finally removed; boxes without suppressors start out kept.
// Both sets only grow, their limit is the greedy result (the dependence is strictly upper-triangular), and a box
// is decided after as many rounds as its longest suppression chain is deep -- a handful for real detections.
// One CTA per 64-box column block j keeps ITS column of the bit mask in shared memory (the words (i, j) of all rows
// i < 64 (j + 1), diagonal words cut to the bits right of the row).  A round reads the published kept / removed
// words of the blocks left of it (plain L2 loads: the sets are monotone, so stale words only delay, and no barrier
// or fence orders the CTAs), ORs the rows that are finally kept (-> removed bits) and the rows that are not yet
// finally removed (-> still threatened bits), closes the chains INSIDE its own 64 boxes with a warp-level fixed
// point, and publishes its two words.  A CTA whose 64 boxes are decided waits for the blocks left of it, compacts
// its kept boxes into keep[] (prefix of popcounts) and counts itself done; the scan kernel behind returns at once
// when all CTAs are.  A CTA that is not decided after kResolveIters rounds (suppression chains deeper than
// that across blocks) raises the fail flag: everybody leaves and the serial scan kernel does the work as before --
// the result is the same either way.  Replaces the host-side scan of the reference (cuda/nms_kernel.cu:136-159).
// ctl: [1] CTAs done, [2] fail flag.
__device__ __forceinline__ u64 ld_l2_u64(const u64* p) {  // from L2, never hoisted or merged by the compiler
  u64 v;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_l2_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kResolveThreads) nms_resolve_kernel(
    const u64* __restrict__ mask, int n, int cb, const int32_t* __restrict__ order, int64_t* __restrict__ keep,
    int64_t* __restrict__ num_keep, u64* __restrict__ kw /* [2][cb]: kept, removed */, unsigned* __restrict__ ctl) {
  pdl_launch_dependents();
  pdl_wait();  // the mask is complete, the control words are zero
  extern __shared__ __align__(16) u64 smem[];
  u64* col = smem;                          // [(j + 1) * 64] words of this column
  u64* skf = smem + (size_t)cb * kTile;     // [cb] published kept words
  u64* srf = skf + cb;                      // [cb] published removed words
  __shared__ u64 s_red[2][kResolveThreads / 32];
  __shared__ u64 s_pub[2];
  __shared__ int s_state;                   // 0 go on, 1 decided, 2 give up
  const int j = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  u64* gkf = kw;
  u64* grf = kw + cb;
  const int tail = n - j * kTile;
  const u64 alive = tail >= kTile ? ~0ULL : ((1ULL << tail) - 1ULL);
  // boxes past the end count as removed; nothing is kept yet (the mask kernel left all published words zero:
  // "undecided", which a reader treats as a threat that has not gone away yet)
  if (tid == 0) __stcg(grf + j, ~alive);
  const int nent = (j + 1) * kTile;
  for (int e = tid; e < nent; e += kResolveThreads) {
    const int b = e >> 6, r = e & 63;
    // (the last block only has n - 64 b rows: nothing was written, and nothing is allocated, for the others)
    u64 w = r < n - b * kTile ? __ldcg(mask + tri_base(b, cb) + (long long)r * (cb - b) + (j - b)) : 0ULL;
    if (b == j) w &= ~((2ULL << r) - 1ULL);  // diagonal word: bits right of the row only (row 63: none)
    col[e] = w;
  }
  u64 kf = 0, rf = ~alive;   // this block's decided bits (uniform across the CTA after every round)
  __syncthreads();
  const int r = tid & 63;
  for (int round = 0;; ++round) {
    for (int b = tid; b < j; b += kResolveThreads) {
      skf[b] = ld_l2_u64(gkf + b);
      srf[b] = ld_l2_u64(grf + b);
    }
    __syncthreads();
    // rows of the blocks left of this one: thread tid owns row tid % 64 of blocks tid / 64, tid / 64 + 4, ...
    u64 a = 0, t = 0;  // a: bits removed by a finally kept row; t: bits still threatened by a row not yet removed
    for (int b = tid >> 6; b < j; b += kResolveThreads / kTile) {
      const u64 w = col[b * kTile + r];
      a |= ((skf[b] >> r) & 1ULL) ? w : 0ULL;
      t |= ((srf[b] >> r) & 1ULL) ? 0ULL : w;
    }
    {
      const unsigned alo = __reduce_or_sync(0xffffffffu, (unsigned)a), ahi = __reduce_or_sync(0xffffffffu, (unsigned)(a >> 32));
      const unsigned tlo = __reduce_or_sync(0xffffffffu, (unsigned)t), thi = __reduce_or_sync(0xffffffffu, (unsigned)(t >> 32));
      if (lane == 0) {
        s_red[0][warp] = ((u64)ahi << 32) | alo;
        s_red[1][warp] = ((u64)thi << 32) | tlo;
      }
    }
    __syncthreads();
    if (warp == 0) {
      u64 aoff = 0, toff = 0;
#pragma unroll
      for (int w = 0; w < kResolveThreads / 32; ++w) { aoff |= s_red[0][w]; toff |= s_red[1][w]; }
      // the chains inside the block: lane l holds diagonal rows l and l + 32
      const u64 d0 = col[j * kTile + lane], d1 = col[j * kTile + lane + 32];
      u64 nk = kf, nr = rf;
      for (;;) {
        u64 av = ((nk >> lane) & 1ULL) ? d0 : 0ULL, tv = ((nr >> lane) & 1ULL) ? 0ULL : d0;
        av |= ((nk >> (lane + 32)) & 1ULL) ? d1 : 0ULL;
        tv |= ((nr >> (lane + 32)) & 1ULL) ? 0ULL : d1;
        const u64 ad = ((u64)__reduce_or_sync(0xffffffffu, (unsigned)(av >> 32)) << 32) | __reduce_or_sync(0xffffffffu, (unsigned)av);
        const u64 td = ((u64)__reduce_or_sync(0xffffffffu, (unsigned)(tv >> 32)) << 32) | __reduce_or_sync(0xffffffffu, (unsigned)tv);
        const u64 r2 = nr | ((aoff | ad) & alive);
        const u64 k2 = nk | (alive & ~(toff | td) & ~r2);
        if (r2 == nr && k2 == nk) break;
        nr = r2;
        nk = k2;
      }
      if (lane == 0) {
        if (nk != kf) __stcg(gkf + j, nk);
        if (nr != rf) __stcg(grf + j, nr);
        s_pub[0] = nk;
        s_pub[1] = nr;
        int st = (nk | nr) == ~0ULL ? 1 : 0;
        if (!st && (round + 1 >= kResolveIters || ld_l2_u32(ctl + 2))) {
          st = 2;
          __stcg(ctl + 2, 1u);
        }
        s_state = st;
      }
    }
    __syncthreads();
    kf = s_pub[0];
    rf = s_pub[1];
    if (s_state == 2) return;  // the scan kernel takes over
    if (s_state == 1) break;
  }
  // decided: kf is this block's part of the greedy result.  Its position in keep[] needs the blocks left of it.
  for (int b = tid; b < j; b += kResolveThreads) {
    u64 k = 0;
    for (int spins = 0;; ++spins) {
      k = ld_l2_u64(gkf + b);
      if ((k | ld_l2_u64(grf + b)) == ~0ULL) break;
      if ((spins & 15) == 15 && ld_l2_u32(ctl + 2)) break;  // somebody gave up: the flag is checked below
      __nanosleep(100);
    }
    skf[b] = k;
  }
  __syncthreads();
  if (tid == 0) s_state = ld_l2_u32(ctl + 2) ? 2 : 1;  // one reader: the CTA leaves or stays as a whole
  __syncthreads();
  if (s_state == 2) return;
  int before = 0;
  for (int b = lane; b < j; b += 32) before += __popcll(skf[b]);
  before = __reduce_add_sync(0xffffffffu, before);  // every warp computes the same prefix
  if (tid < kTile && ((kf >> tid) & 1ULL))
    keep[before + __popcll(kf & ((1ULL << tid) - 1ULL))] = (int64_t)order[j * kTile + tid];
  if (tid == 0) {
    if (j == cb - 1) *num_keep = before + __popcll(kf);
    atomicAdd(ctl + 1, 1u);  // read by the scan kernel after this grid has completed
  }
}

__global__ void __launch_bounds__(kSmemThreads) nms_scan_smem_kernel(
    const u64* __restrict__ mask, int n, int col_blocks, const int32_t* __restrict__ order,
    int64_t* __restrict__ keep, int64_t* __restrict__ num_keep, const unsigned* __restrict__ resolve_ctl) {
  pdl_wait();
  // every CTA of nms_resolve_kernel decided its boxes and wrote its part of keep[]: nothing left to do
  if (resolve_ctl && __ldcg(resolve_ctl + 1) == (unsigned)col_blocks && !__ldcg(resolve_ctl + 2)) return;
  scan_smem_body(mask, n, col_blocks, order, keep, num_keep, nullptr, 0);
}

// segmented: one CTA per segment, all segments scan concurrently
__global__ void __launch_bounds__(kSmemThreads) nms_seg_scan_smem_kernel(
    const u64* __restrict__ mask, const int32_t* __restrict__ off, const long long* __restrict__ moff,
    const int32_t* __restrict__ roff, const int32_t* __restrict__ order, int64_t* __restrict__ keep,
    int64_t* __restrict__ num_keep, const u64* __restrict__ init_remv, int top_n,
    const int32_t* __restrict__ rank, const int32_t* __restrict__ jmax, const int32_t* __restrict__ wbase) {
  const int s = blockIdx.x;
  const int o = off[s], n = off[s + 1] - o;
  scan_smem_body(mask + moff[s], n, (n + kTile - 1) / kTile, order + o, keep + o, num_keep + s,
                 init_remv + roff[s], top_n, rank ? rank + o : nullptr, jmax ? jmax + roff[s] : nullptr,
                 jmax ? wbase + roff[s] : nullptr);
}

__device__ __forceinline__ void scan_body(
    const u64* __restrict__ mask, int n, int col_blocks, const int32_t* __restrict__ order,
    int64_t* __restrict__ keep, int64_t* __restrict__ num_keep, const u64* __restrict__ init_remv, int top_n,
    const int32_t* __restrict__ rank = nullptr, const int32_t* __restrict__ jmax = nullptr,
    const int32_t* __restrict__ wbase = nullptr) {
  const MaskIdx mi{jmax, wbase, col_blocks};
  extern __shared__ u64 smem[];
  u64* remv = smem;                 // [col_blocks]
  u64* keptw = smem + col_blocks;   // [col_blocks]
  __shared__ u64 diag[kTile];
  __shared__ u64 s_kept;
  __shared__ int s_base[kScanThreads / 32];
  const int tid = threadIdx.x;
  for (int j = tid; j < col_blocks; j += kScanThreads) remv[j] = init_remv ? init_remv[j] : 0ULL;
  __syncthreads();

  for (int b = 0; b < col_blocks; ++b) {
    const int rows = min(n - b * kTile, kTile);
    if (tid < kTile) diag[tid] = tid < rows ? mask[mi.word(b, tid, b)] : 0ULL;
    __syncthreads();
    if (tid == 0) {
      u64 alive = ~remv[b];
      if (rows < kTile) alive &= (1ULL << rows) - 1ULL;
      const u64 kept = resolve_block(alive, diag, 1);
      s_kept = kept;
      keptw[b] = kept;
    }
    __syncthreads();
    const u64 kept = s_kept;
    const int jend = jmax ? min(col_blocks, jmax[b] + 1) : col_blocks;
    for (int j = b + 1 + tid; j < jend; j += kScanThreads) {
      u64 acc = remv[j];
      u64 kk = kept;
      while (kk) {
        const int i = __ffsll((long long)kk) - 1;
        kk &= kk - 1;
        acc |= mask[mi.word(b, i, j)];
      }
      remv[j] = acc;
    }
    __syncthreads();
  }
  finish_keep(keptw, remv, col_blocks, rank, order, keep, num_keep, s_base, &s_kept, top_n);
}

__global__ void __launch_bounds__(kScanThreads) nms_scan_kernel(
    const u64* __restrict__ mask, int n, int col_blocks, const int32_t* __restrict__ order,
    int64_t* __restrict__ keep, int64_t* __restrict__ num_keep) {
  pdl_wait();
  scan_body(mask, n, col_blocks, order, keep, num_keep, nullptr, 0);
}

__global__ void __launch_bounds__(kScanThreads) nms_seg_scan_kernel(
    const u64* __restrict__ mask, const int32_t* __restrict__ off, const long long* __restrict__ moff,
    const int32_t* __restrict__ roff, const int32_t* __restrict__ order, int64_t* __restrict__ keep,
    int64_t* __restrict__ num_keep, const u64* __restrict__ init_remv, int top_n,
    const int32_t* __restrict__ rank, const int32_t* __restrict__ jmax, const int32_t* __restrict__ wbase) {
  const int s = blockIdx.x;
  const int o = off[s], n = off[s + 1] - o;
  scan_body(mask + moff[s], n, (n + kTile - 1) / kTile, order + o, keep + o, num_keep + s, init_remv + roff[s],
            top_n, rank ? rank + o : nullptr, jmax ? jmax + roff[s] : nullptr, jmax ? wbase + roff[s] : nullptr);
}

__global__ void nms_zero_count_kernel(int64_t* num_keep) { *num_keep = 0; }

// ---- workspace carving --------------------------------------------------------
struct Carve {
  char* p; size_t off, cap;
  void* take(size_t bytes) {
    off = (off + 255) & ~(size_t)255;
    void* r = p ? p + off : nullptr;
    off += bytes;
    return r;
  }
};

size_t cub_temp_bytes(long long n, bool f64) {
  size_t bytes = 0;
  if (f64)
    cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, (const double*)nullptr, (double*)nullptr,
                                              (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  else
    cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, (const float*)nullptr, (float*)nullptr,
                                              (const int32_t*)nullptr, (int32_t*)nullptr, (int)n);
  return bytes;
}

struct Layout {
  void *keys_in, *keys_out, *cub_tmp, *sorted, *maxc;
  int32_t *idx_in, *order, *cls;
  u64* mask;
  u64* resolve_k;          // [2][col_blocks] K words of nms_resolve_kernel
  unsigned* resolve_ctl;   // kResolveCtlWords control words
  size_t cub_bytes, total;
};

Layout carve(void* ws, long long n) {
  Layout L;
  Carve c{(char*)ws, 0, 0};
  const long long cb = (n + kTile - 1) / kTile;
  L.keys_in = c.take(n * 8);
  L.keys_out = c.take(n * 8);
  L.idx_in = (int32_t*)c.take(n * 4);
  L.order = (int32_t*)c.take(n * 4);
  L.cls = (int32_t*)c.take(n * 4);
  L.maxc = c.take(16);
  L.sorted = c.take(n * sizeof(Box4d));
  L.cub_bytes = cub_temp_bytes(n, true);
  const size_t f32b = cub_temp_bytes(n, false);
  if (f32b > L.cub_bytes) L.cub_bytes = f32b;
  L.cub_tmp = c.take(L.cub_bytes);
  L.mask = (u64*)c.take((size_t)n * cb * 8);
  L.resolve_k = (u64*)c.take((size_t)2 * cb * 8);
  L.resolve_ctl = (unsigned*)c.take(kResolveCtlWords * 4);
  L.total = c.off + 256;
  return L;
}

template <typename T>
int nms_t(const void* boxes, const void* scores, const int64_t* idxs, long long n, double thr,
          int mode, int64_t* keep, int64_t* num_keep, void* ws, cudaStream_t st) {
  using A = typename Io<T>::acc;
  using BoxT = typename BoxOf<T>::type;
  const Layout L = carve(ws, n);
  const int th = 256, bl = (int)((n + th - 1) / th);
  const int col_blocks = (int)((n + kTile - 1) / kTile);
  constexpr bool f64 = std::is_same<T, double>::value;
  // one-launch front ends: the rank sort over all SMs for sets of up to kRankMax boxes, the one-CTA block radix sort
  // (B200DET_NMS_FRONT=block, read per call) for up to kSortMax
  const char* front = getenv("B200DET_NMS_FRONT");
  const bool front_block = front && front[0] == 'b';
  const long long front_max = front_block ? kSortMax : kRankMax;
  if (f64) {
    nms_prepare_kernel_f64<<<bl, th, 0, st>>>((const double*)scores, n, (double*)L.keys_in, L.idx_in);
    B200DET_CUDA_OK(cub::DeviceRadixSort::SortPairsDescending(
        L.cub_tmp, const_cast<size_t&>(L.cub_bytes), (const double*)L.keys_in, (double*)L.keys_out,
        (const int32_t*)L.idx_in, L.order, (int)n, 0, 64, st));
  } else if (n <= front_max) {
    if constexpr (!f64) {
      if (front_block) {
        nms_front_small_kernel<T><<<1, kSortThreads, 0, st>>>((const T*)boxes, (const T*)scores, idxs, (int)n, mode,
                                                              (Box4f*)L.sorted, mode == 2 ? L.cls : nullptr, L.order);
      } else {
        const size_t ksmem = (size_t)((n + 3) / 4) * 16;  // all n keys in every CTA's shared memory
        if (ksmem > 48 * 1024)
          B200DET_CUDA_OK(cudaFuncSetAttribute(nms_front_rank_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)((size_t)kRankMax * 4)));
        nms_front_rank_kernel<T><<<(int)((n + kRankKeys - 1) / kRankKeys), kRankThreads, ksmem, st>>>(
            (const T*)boxes, (const T*)scores, idxs, (int)n, mode, (Box4f*)L.sorted, mode == 2 ? L.cls : nullptr, L.order);
      }
    }
  } else {
    nms_prepare_kernel<T><<<bl, th, 0, st>>>((const T*)scores, n, (float*)L.keys_in, L.idx_in);
    B200DET_CUDA_OK(cub::DeviceRadixSort::SortPairsDescending(
        L.cub_tmp, const_cast<size_t&>(L.cub_bytes), (const float*)L.keys_in, (float*)L.keys_out,
        (const int32_t*)L.idx_in, L.order, (int)n, 0, 32, st));
  }
  if (f64 || n > front_max) {
    if (mode == 1)
      nms_max_coord_kernel<T><<<1, 1024, 0, st>>>((const T*)boxes, n * 4, (A*)L.maxc);
    nms_gather_kernel<T><<<bl, th, 0, st>>>((const T*)boxes, idxs, L.order, n, mode, (const A*)L.maxc,
                                            (BoxT*)L.sorted, mode == 2 ? L.cls : nullptr);
  }
  dim3 grid((col_blocks + kColTiles - 1) / kColTiles, col_blocks);
  // B200DET_NMS_RESOLVE=0 (read per call) keeps the serial scan alone
  const char* res_s = getenv("B200DET_NMS_RESOLVE");
  const bool resolve = col_blocks >= 2 && col_blocks <= kResolveMaxBlocks && !(res_s && atoi(res_s) == 0);
  unsigned* ctl = resolve ? L.resolve_ctl : nullptr;
  B200DET_CUDA_OK(launch_pdl(nms_mask_kernel<BoxT>, grid, dim3(kTile), 0, st, (int)n, (float)thr, (const BoxT*)L.sorted,
                             (const int32_t*)(mode == 2 ? L.cls : nullptr), L.mask, ctl, L.resolve_k));
  if (resolve) {
    const size_t rsmem = ((size_t)col_blocks * kTile + 2 * (size_t)col_blocks) * sizeof(u64);
    if (rsmem > 48 * 1024)
      B200DET_CUDA_OK(cudaFuncSetAttribute(nms_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem));
    B200DET_CUDA_OK(launch_pdl(nms_resolve_kernel, dim3(col_blocks), dim3(kResolveThreads), rsmem, st, (const u64*)L.mask,
                               (int)n, col_blocks, (const int32_t*)L.order, keep, num_keep, L.resolve_k, ctl));
  }
  if (col_blocks <= kSmemBlocks) {
    const size_t smem = ((size_t)(kRowGroups + 1) * col_blocks + 1 + kDiagRing * kTile + (size_t)2 * kTile * col_blocks) * sizeof(u64);
    if (smem > 48 * 1024)
      B200DET_CUDA_OK(cudaFuncSetAttribute(nms_scan_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B200DET_CUDA_OK(launch_pdl(nms_scan_smem_kernel, dim3(1), dim3(kSmemThreads), smem, st, (const u64*)L.mask, (int)n,
                               col_blocks, (const int32_t*)L.order, keep, num_keep, (const unsigned*)ctl));
  } else {
    const size_t smem = (size_t)col_blocks * 16;
    if (smem > 48 * 1024)
      B200DET_CUDA_OK(cudaFuncSetAttribute(nms_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B200DET_CUDA_OK(launch_pdl(nms_scan_kernel, dim3(1), dim3(kScanThreads), smem, st, (const u64*)L.mask, (int)n,
                               col_blocks, (const int32_t*)L.order, keep, num_keep));
  }
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- segmented NMS: S independent box sets (images) in one launch sequence -------------------
// Replaces the per-image Python loops around batched_nms in RegionProposalNetwork.filter_proposals
// (models/detection/rpn.py:268-297) and RoIHeads.postprocess_detections (roi_heads.py:693-723).
constexpr int kSegBatch = 64;
struct SegBatch {
  int count, base;
  int off[kSegBatch + 1];          // first box of segment base + j
  long long moff[kSegBatch + 1];   // first mask word
  int roff[kSegBatch + 1];         // first remv word
  signed char mode[kSegBatch + 1]; // 0 plain, 1 coordinate trick, 2 vanilla (class compare)
};

__global__ void nms_seg_setup_kernel(const __grid_constant__ SegBatch b, int32_t* __restrict__ off,
                                     long long* __restrict__ moff, int32_t* __restrict__ roff,
                                     int32_t* __restrict__ smode) {
  const int j = threadIdx.x;
  if (j <= b.count) {
    off[b.base + j] = b.off[j];
    moff[b.base + j] = b.moff[j];
    roff[b.base + j] = b.roff[j];
    smode[b.base + j] = b.mode[j];
  }
}

// segment of global position i: largest s with off[s] <= i
__device__ __forceinline__ int seg_of(const int32_t* __restrict__ off, int S, int i) {
  int lo = 0, hi = S;  // off[lo] <= i < off[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(off + mid) <= i) lo = mid; else hi = mid;
  }
  return lo;
}

// 64-bit sort key: segment in the high word, the score mapped so that an ASCENDING sort gives
// descending scores in the low word; equal scores keep ascending index (stable radix sort), like
// scores.sort(stable=True, descending=True).  Boxes filtered out before NMS go last.
template <typename T>
__global__ void nms_seg_keys_kernel(const T* __restrict__ scores, const uint8_t* __restrict__ valid, int n,
                                    const int32_t* __restrict__ off, int S, u64* __restrict__ keys,
                                    int32_t* __restrict__ idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float f = (float)Io<T>::ld(scores + i);
  if (f == 0.f) f = 0.f;  // -0 sorts like +0
  const unsigned bits = __float_as_uint(f);
  const unsigned asc = bits ^ ((bits >> 31) ? 0xffffffffu : 0x80000000u);
  unsigned low = ~asc;
  if (valid && !valid[i]) low = 0xffffffffu;
  keys[i] = ((u64)(unsigned)seg_of(off, S, i) << 32) | (u64)low;
  idx[i] = i;
}

// boxes.max() of every segment's surviving boxes (coordinate trick, boxes.py:91); one CTA per segment
//
// It also decides whether the segment may be scanned grouped by class (sgroup).  Grouping changes the greedy
// order, so it needs classes that cannot interact.  The class compare of the vanilla strategy guarantees that;
// the coordinate trick only does when the offsets idx * (max + 1) really separate the classes: all surviving
// coordinates finite and >= 0 (so they lie in [0, max]) and |idx| * (max + 1) < 2^22, where consecutive offsets
// differ by at least max after rounding (ulp <= 1/2) and x + offset rounds monotonically: boxes of different
// classes then have inter = 0 exactly.  Otherwise the segment keeps the score order (all class keys 0).
template <typename T>
__global__ void nms_seg_max_kernel(const T* __restrict__ boxes, const int64_t* __restrict__ idxs,
                                   const uint8_t* __restrict__ valid, const int32_t* __restrict__ off,
                                   int32_t* __restrict__ smode, float* __restrict__ out,
                                   int32_t* __restrict__ sgroup, int thr_nonneg) {
  __shared__ float red[32];
  __shared__ int s_bad, s_cnt;
  const int s = blockIdx.x;
  if (smode[s] < 0) {
    // automatic strategy: the reference's rule for a CUDA tensor (boxes.py:72, numel() > 20000 -> vanilla) sees the
    // boxes that SURVIVED the filters (rpn.py:276-288, roi_heads.py:708-716 compact before batched_nms), so it is
    // decided here from the number of valid boxes of the set, not on the host from the set's size
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int c = 0;
    for (int i = off[s] + threadIdx.x; i < off[s + 1]; i += blockDim.x) c += (!valid || valid[i]) ? 1 : 0;
    if (c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) smode[s] = 4 * s_cnt > 20000 ? 2 : 1;
    __syncthreads();
  }
  if (smode[s] == 2) {
    if (threadIdx.x == 0) sgroup[s] = 1;
    return;
  }
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
  float m = -INFINITY, amax = 0.f;
  bool nan = false, bad = false;
  for (int i = off[s] + threadIdx.x; i < off[s + 1]; i += blockDim.x) {
    if (valid && !valid[i]) continue;
    amax = fmaxf(amax, fabsf((float)idxs[i]));
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float v = (float)Io<T>::ld(boxes + (long long)i * 4 + c);
      nan |= (v != v);
      bad |= !(v >= 0.f && v < INFINITY);
      m = v > m ? v : m;
    }
  }
  if (bad) atomicOr(&s_bad, 1);
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) amax = fmaxf(amax, red[w]);
    red[0] = amax;
  }
  __syncthreads();
  amax = red[0];
  __syncthreads();
  if (nan) m = NAN;
  for (int o = 16; o > 0; o >>= 1) {
    const float other = __shfl_xor_sync(0xffffffffu, m, o);
    m = (other != other || m != m) ? NAN : (other > m ? other : m);
  }
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : -INFINITY;
    for (int o = 16; o > 0; o >>= 1) {
      const float other = __shfl_xor_sync(0xffffffffu, m, o);
      m = (other != other || m != m) ? NAN : (other > m ? other : m);
    }
    if (threadIdx.x == 0) {
      out[s] = m;
      // a negative threshold lets disjoint boxes (IoU 0) suppress one another across classes
      sgroup[s] = (thr_nonneg && !s_bad && __fmul_rn(amax, __fadd_rn(m, 1.f)) < 4194304.f) ? 1 : 0;
    }
  }
}

// Class key of a box: any order-preserving-for-equality map of the class id; filtered boxes go last
__device__ __forceinline__ unsigned class_key(const int64_t* __restrict__ idxs, const uint8_t* __restrict__ valid,
                                              long long src, int grouped) {
  if (!grouped) return 0u;
  return (valid && !valid[src]) ? 0xffffffffu : ((unsigned)(int32_t)idxs[src] ^ 0x80000000u);
}

// second sort key of the class-grouped order: (segment, class key), applied stably to the score order
__global__ void nms_seg_class_keys_kernel(const int64_t* __restrict__ idxs, const uint8_t* __restrict__ valid,
                                          const int32_t* __restrict__ order, int n, const int32_t* __restrict__ off,
                                          int S, const int32_t* __restrict__ sgroup, u64* __restrict__ keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = seg_of(off, S, i);
  keys[i] = ((u64)(unsigned)s << 32) | (u64)class_key(idxs, valid, order[i], sgroup[s]);
}

// one warp per 64-box block of every segment: bounding box and class-key range
__global__ void nms_seg_blocks_kernel(const Box4f* __restrict__ sorted, const unsigned* __restrict__ ckey,
                                      const int32_t* __restrict__ off, const int32_t* __restrict__ roff, int S,
                                      int total_blocks, BlockSum* __restrict__ out) {
  const int g = (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (g >= total_blocks) return;
  const int s = seg_of(roff, S, g);
  const int first = off[s] + (g - roff[s]) * kTile;
  const int cnt = min(kTile, off[s + 1] - first);
  float x1 = INFINITY, y1 = INFINITY, x2 = -INFINITY, y2 = -INFINITY;
  for (int i = lane; i < cnt; i += 32) {
    const Box4f b = sorted[first + i];
    x1 = fminf(x1, b.x1 != b.x1 ? -INFINITY : b.x1);
    y1 = fminf(y1, b.y1 != b.y1 ? -INFINITY : b.y1);
    x2 = fmaxf(x2, b.x2 != b.x2 ? INFINITY : b.x2);
    y2 = fmaxf(y2, b.y2 != b.y2 ? INFINITY : b.y2);
  }
  for (int o = 16; o > 0; o >>= 1) {
    x1 = fminf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
    y1 = fminf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
    x2 = fmaxf(x2, __shfl_xor_sync(0xffffffffu, x2, o));
    y2 = fmaxf(y2, __shfl_xor_sync(0xffffffffu, y2, o));
  }
  if (lane == 0) {
    BlockSum b;
    b.x1 = x1; b.y1 = y1; b.x2 = x2; b.y2 = y2;
    b.klo = ckey[first];
    b.khi = ckey[first + cnt - 1];
    out[g] = b;
  }
}

// last live tile of every row block: jmax[b] = the largest column block j >= b whose tile (b, j) can hold a bit,
// and wbase[b] = sum of the widths jmax - b + 1 of the blocks before b (the packed layout's block bases).
// One CTA per segment; a warp finds jmax of a block by testing 32 column blocks at a time from the right.
constexpr int kJmaxThreads = 256;
__global__ void __launch_bounds__(kJmaxThreads) nms_seg_jmax_kernel(
    const BlockSum* __restrict__ blk, const int32_t* __restrict__ roff, const int32_t* __restrict__ smode,
    int skip_apart, int32_t* __restrict__ jmax, int32_t* __restrict__ wbase) {
  const int s = blockIdx.x;
  const int cb = roff[s + 1] - roff[s];
  const BlockSum* sb = blk + roff[s];
  int32_t* jm = jmax + roff[s];
  int32_t* wb = wbase + roff[s];
  const bool by_class = smode[s] == 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int b = warp; b < cb; b += kJmaxThreads / 32) {
    const BlockSum r = sb[b];
    int found = b;
    for (int hi = cb - 1; hi > b; hi -= 32) {  // columns hi - 31 .. hi
      const int j = hi - lane;
      const bool live = j > b && !tile_is_empty(r, sb[j], skip_apart != 0, by_class);
      const unsigned m = __ballot_sync(0xffffffffu, live);
      if (m) {
        found = hi - (__ffs(m) - 1);
        break;
      }
    }
    if (lane == 0) jm[b] = found;
  }
  __syncthreads();
  // exclusive prefix sum of the widths, a chunk of kJmaxThreads blocks at a time
  __shared__ int s_warp[kJmaxThreads / 32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < cb; b0 += kJmaxThreads) {
    const int b = b0 + threadIdx.x;
    const int w = b < cb ? jm[b] - b + 1 : 0;
    int incl = w;
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = s_carry;
    for (int k = 0; k < warp; ++k) before += s_warp[k];
    if (b < cb) wb[b] = before + incl - w;
    __syncthreads();
    if (threadIdx.x == kJmaxThreads - 1) s_carry = before + incl;
    __syncthreads();
  }
}

// boxes into (segment, score) order; mode as nms_gather_kernel; filtered boxes start out removed
template <typename T>
__global__ void nms_seg_gather_kernel(const T* __restrict__ boxes, const int64_t* __restrict__ idxs,
                                      const uint8_t* __restrict__ valid, const int32_t* __restrict__ order, int n,
                                      const int32_t* __restrict__ off, const int32_t* __restrict__ roff, int S,
                                      const int32_t* __restrict__ smode, const float* __restrict__ max_coord,
                                      Box4f* __restrict__ sorted, int32_t* __restrict__ sorted_cls,
                                      u64* __restrict__ init_remv, const int32_t* __restrict__ pos,
                                      const int32_t* __restrict__ sgroup, unsigned* __restrict__ ckey,
                                      int32_t* __restrict__ rank) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = seg_of(off, S, i);
  const int sp = pos ? pos[i] : i;  // class-grouped: position i holds the box of score position pos[i]
  const long long src = order[sp];
  if (pos) {
    rank[i] = sp - off[s];
    ckey[i] = class_key(idxs, valid, src, sgroup[s]);
  }
  float x1 = (float)Io<T>::ld(boxes + src * 4 + 0), y1 = (float)Io<T>::ld(boxes + src * 4 + 1);
  float x2 = (float)Io<T>::ld(boxes + src * 4 + 2), y2 = (float)Io<T>::ld(boxes + src * 4 + 3);
  const int mode = smode[s];
  if (mode == 1) {
    const float o = __fmul_rn((float)idxs[src], __fadd_rn(max_coord[s], 1.f));
    x1 = __fadd_rn(x1, o); y1 = __fadd_rn(y1, o);
    x2 = __fadd_rn(x2, o); y2 = __fadd_rn(y2, o);
  }
  Box4f b;
  b.x1 = x1; b.y1 = y1; b.x2 = x2; b.y2 = y2;
  sorted[i] = b;
  if (idxs) sorted_cls[i] = (int32_t)idxs[src];
  if (valid && !valid[src]) {
    const int p = i - off[s];
    atomicOr(init_remv + roff[s] + (p >> 6), 1ULL << (p & 63));
  }
}

struct SegLayout {
  u64 *keys_in, *keys_out, *mask, *init_remv;
  int32_t *idx_in, *order, *cls, *off, *roff, *smode, *pos, *rank, *sgroup, *jmax, *wbase;
  unsigned* ckey;
  BlockSum* blk;
  long long* moff;
  float* maxc;
  Box4f* sorted;
  void* cub_tmp;
  size_t cub_bytes, total;
};

size_t cub_temp_bytes_u64(long long n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const u64*)nullptr, (u64*)nullptr, (const int32_t*)nullptr,
                                  (int32_t*)nullptr, (int)n);
  return bytes;
}

SegLayout seg_carve(void* ws, long long n, long long S, long long mask_words, long long remv_words) {
  SegLayout L;
  Carve c{(char*)ws, 0, 0};
  L.keys_in = (u64*)c.take(n * 8);
  L.keys_out = (u64*)c.take(n * 8);
  L.idx_in = (int32_t*)c.take(n * 4);
  L.order = (int32_t*)c.take(n * 4);
  L.cls = (int32_t*)c.take(n * 4);
  L.pos = (int32_t*)c.take(n * 4);
  L.rank = (int32_t*)c.take(n * 4);
  L.ckey = (unsigned*)c.take(n * 4);
  L.blk = (BlockSum*)c.take((remv_words + 1) * sizeof(BlockSum));
  L.jmax = (int32_t*)c.take((remv_words + 1) * 4);
  L.wbase = (int32_t*)c.take((remv_words + 1) * 4);
  L.off = (int32_t*)c.take((S + 1) * 4);
  L.roff = (int32_t*)c.take((S + 1) * 4);
  L.smode = (int32_t*)c.take((S + 1) * 4);
  L.sgroup = (int32_t*)c.take((S + 1) * 4);
  L.moff = (long long*)c.take((S + 1) * 8);
  L.maxc = (float*)c.take((S + 1) * 4);
  L.sorted = (Box4f*)c.take(n * sizeof(Box4f));
  L.init_remv = (u64*)c.take((remv_words + 1) * 8);
  L.cub_bytes = cub_temp_bytes_u64(n);
  L.cub_tmp = c.take(L.cub_bytes);
  L.mask = (u64*)c.take((size_t)(mask_words + 1) * 8);
  L.total = c.off + 256;
  return L;
}

// host-side shape of the segmentation: returns false when the offsets are not a partition of [0, N)
bool seg_shape(const int64_t* seg_offsets, long long S, long long N, long long* mask_words, long long* remv_words,
               int* cb_max) {
  *mask_words = 0; *remv_words = 0; *cb_max = 0;
  if (S < 0 || (S == 0 && N != 0)) return false;
  if (S == 0) return true;
  if (seg_offsets[0] != 0 || seg_offsets[S] != N) return false;
  for (long long s = 0; s < S; ++s) {
    const long long n = seg_offsets[s + 1] - seg_offsets[s];
    if (n < 0) return false;
    const long long cb = (n + kTile - 1) / kTile;
    *mask_words += (n * cb + 1) & ~1LL;
    *remv_words += cb;
    if (cb > *cb_max) *cb_max = (int)cb;
  }
  return true;
}

template <typename T>
int nms_seg_t(const void* boxes, const void* scores, const int64_t* idxs, const uint8_t* valid, long long n,
              const int64_t* seg_offsets, long long S, double thr, int mode, int top_n, int64_t* keep,
              int64_t* counts, void* ws, cudaStream_t st) {
  long long mask_words, remv_words;
  int cb_max;
  seg_shape(seg_offsets, S, n, &mask_words, &remv_words, &cb_max);
  const SegLayout L = seg_carve(ws, n, S, mask_words, remv_words);
  // segment tables: host offsets travel as kernel parameters, 64 segments per launch
  long long mo = 0;
  int ro = 0;
  for (long long base = 0; base < S; base += kSegBatch) {
    SegBatch b;
    b.base = (int)base;
    b.count = (int)((S - base) < kSegBatch ? (S - base) : kSegBatch);
    for (int j = 0; j <= b.count; ++j) {
      b.off[j] = (int)seg_offsets[base + j];
      b.moff[j] = mo;
      b.roff[j] = ro;
      b.mode[j] = 0;
      if (j < b.count) {
        const long long ns = seg_offsets[base + j + 1] - seg_offsets[base + j];
        // mode < 0: the reference's rule for a CUDA tensor, set by set, decided on the device from the number of
        // boxes that pass the filters (nms_seg_max_kernel)
        b.mode[j] = (signed char)(!idxs ? 0 : (mode >= 0 ? mode : -1));
        const long long cb = (ns + kTile - 1) / kTile;
        mo += (ns * cb + 1) & ~1LL;  // 16-byte aligned segments: the scan stages rows with bulk copies
        ro += (int)cb;
      }
    }
    nms_seg_setup_kernel<<<1, kSegBatch + 1, 0, st>>>(b, L.off, L.moff, L.roff, L.smode);
  }
  const int th = 256, bl = (int)((n + th - 1) / th);
  nms_seg_keys_kernel<T><<<bl, th, 0, st>>>((const T*)scores, valid, (int)n, L.off, (int)S, L.keys_in, L.idx_in);
  int seg_bits = 1;
  while ((1LL << seg_bits) < S) ++seg_bits;
  B200DET_CUDA_OK(cub::DeviceRadixSort::SortPairs(L.cub_tmp, const_cast<size_t&>(L.cub_bytes), (const u64*)L.keys_in,
                                                  L.keys_out, (const int32_t*)L.idx_in, L.order, (int)n, 0,
                                                  32 + seg_bits, st));
  B200DET_CUDA_OK(cudaMemsetAsync(L.init_remv, 0, (size_t)(remv_words + 1) * 8, st));
  // Class-aware sets are scanned grouped by class: a second, stable sort of the score order by (segment, class)
  // makes the IoU mask block-diagonal, and whole tiles between different classes are skipped from the block
  // summaries (5 FPN levels: ~4.6x fewer pairs; 80 classes: ~80x).  The kept set is the same -- boxes of
  // different classes never suppress one another in either reference strategy, and where the coordinate trick
  // fails to separate them (negative or non-finite coordinates) the bounding boxes meet and the tile is computed.
  const bool grouped = idxs != nullptr && remv_words > 0;
  if (idxs)
    nms_seg_max_kernel<T><<<(unsigned)S, 256, 0, st>>>((const T*)boxes, idxs, valid, L.off, L.smode, L.maxc, L.sgroup,
                                               (float)thr >= 0.f ? 1 : 0);
  if (grouped) {
    nms_seg_class_keys_kernel<<<bl, th, 0, st>>>(idxs, valid, L.order, (int)n, L.off, (int)S, L.sgroup, L.keys_in);
    B200DET_CUDA_OK(cub::DeviceRadixSort::SortPairs(L.cub_tmp, const_cast<size_t&>(L.cub_bytes),
                                                    (const u64*)L.keys_in, L.keys_out, (const int32_t*)L.idx_in, L.pos,
                                                    (int)n, 0, 32 + seg_bits, st));
  }
  nms_seg_gather_kernel<T><<<bl, th, 0, st>>>((const T*)boxes, idxs, valid, L.order, (int)n, L.off, L.roff, (int)S,
                                              L.smode, L.maxc, L.sorted, L.cls, L.init_remv,
                                              grouped ? L.pos : nullptr, L.sgroup, L.ckey, L.rank);
  if (grouped) {
    const int warps = (int)remv_words;
    nms_seg_blocks_kernel<<<(warps * 32 + th - 1) / th, th, 0, st>>>(L.sorted, L.ckey, L.off, L.roff, (int)S, warps,
                                                                    L.blk);
    nms_seg_jmax_kernel<<<(unsigned)S, kJmaxThreads, 0, st>>>(L.blk, L.roff, L.smode, (float)thr >= 0.f ? 1 : 0, L.jmax,
                                                              L.wbase);
  }
  const int32_t* rank = grouped ? L.rank : nullptr;
  const int32_t* jmax = grouped ? L.jmax : nullptr;
  if (cb_max > 0) {
    dim3 grid((cb_max + kColTiles - 1) / kColTiles, cb_max, (unsigned)S);
    nms_seg_mask_kernel<Box4f><<<grid, kTile, 0, st>>>(L.off, L.moff, L.smode, (float)thr, L.sorted, L.cls, L.mask,
                                                       L.roff, grouped ? L.blk : nullptr, jmax, L.wbase);
  }
  if (cb_max <= kSmemBlocks) {
    const size_t smem = ((size_t)(kRowGroups + 1) * cb_max + 1 + kDiagRing * kTile + (size_t)2 * kTile * cb_max) * sizeof(u64) + 16;
    if (smem > 48 * 1024)
      B200DET_CUDA_OK(cudaFuncSetAttribute(nms_seg_scan_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nms_seg_scan_smem_kernel<<<(unsigned)S, kSmemThreads, smem, st>>>(L.mask, L.off, L.moff, L.roff, L.order, keep,
                                                                      counts, L.init_remv, top_n, rank, jmax, L.wbase);
  } else {
    const size_t smem = (size_t)cb_max * 16 + 16;
    if (smem > 48 * 1024)
      B200DET_CUDA_OK(cudaFuncSetAttribute(nms_seg_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    nms_seg_scan_kernel<<<(unsigned)S, kScanThreads, smem, st>>>(L.mask, L.off, L.moff, L.roff, L.order, keep, counts,
                                                                 L.init_remv, top_n, rank, jmax, L.wbase);
  }
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

size_t nms_workspace_bytes(long long N) {
  if (N <= 0) return 256;
  return carve(nullptr, N).total;
}

int nms_device(const void* boxes, const void* scores, const int64_t* idxs, int dtype, long long N,
               double thr, int mode, int64_t* keep, int64_t* num_keep, void* ws, size_t ws_bytes,
               cudaStream_t st) {
  if (N == 0) {
    nms_zero_count_kernel<<<1, 1, 0, st>>>(num_keep);
    B200DET_CUDA_OK(cudaGetLastError());
    return 0;
  }
  if (N > 100000000LL / 1) return fail(B200DET_EINVAL, "nms: N=%lld too large", N);
  if (ws_bytes < nms_workspace_bytes(N))
    return fail(B200DET_EWORKSPACE, "nms: workspace %zu < required %zu", ws_bytes, nms_workspace_bytes(N));
  const long long cb = (N + kTile - 1) / kTile;
  if (cb * 16 > 200 * 1024) return fail(B200DET_ENOSUP, "nms: N=%lld exceeds the single-CTA scan capacity", N);
  switch (dtype) {
    case B200DET_F32: return nms_t<float>(boxes, scores, idxs, N, thr, mode, keep, num_keep, ws, st);
    case B200DET_F16: return nms_t<__half>(boxes, scores, idxs, N, thr, mode, keep, num_keep, ws, st);
    case B200DET_BF16: return nms_t<__nv_bfloat16>(boxes, scores, idxs, N, thr, mode, keep, num_keep, ws, st);
    case B200DET_F64: return nms_t<double>(boxes, scores, idxs, N, thr, mode, keep, num_keep, ws, st);
  }
  return fail(B200DET_ENOSUP, "nms: unknown dtype %d", dtype);
}

size_t nms_segmented_workspace_bytes(long long N, const int64_t* seg_offsets, long long S) {
  long long mw, rw;
  int cbm;
  if (!seg_shape(seg_offsets, S, N, &mw, &rw, &cbm)) return 0;
  return seg_carve(nullptr, N > 0 ? N : 1, S, mw, rw).total;
}

int nms_segmented_device(const void* boxes, const void* scores, const int64_t* idxs, const uint8_t* valid, int dtype,
                         long long N, const int64_t* seg_offsets, long long S, double thr, int mode, int top_n,
                         int64_t* keep, int64_t* counts, void* ws, size_t ws_bytes, cudaStream_t st) {
  long long mw, rw;
  int cbm;
  if (!seg_offsets || !seg_shape(seg_offsets, S, N, &mw, &rw, &cbm))
    return fail(B200DET_EINVAL, "nms_segmented: seg_offsets must be a non-decreasing partition of [0, N)");
  if (S == 0) return 0;
  if (N >= 0x7fffffffLL || S > 65535) return fail(B200DET_EINVAL, "nms_segmented: N=%lld S=%lld too large", N, S);
  if ((long long)cbm * 16 > 200 * 1024)
    return fail(B200DET_ENOSUP, "nms_segmented: a segment exceeds the single-CTA scan capacity");
  if (N == 0) {
    B200DET_CUDA_OK(cudaMemsetAsync(counts, 0, (size_t)S * 8, st));
    return 0;
  }
  const size_t need = nms_segmented_workspace_bytes(N, seg_offsets, S);
  if (ws_bytes < need) return fail(B200DET_EWORKSPACE, "nms_segmented: workspace %zu < required %zu", ws_bytes, need);
  switch (dtype) {
    case B200DET_F32: return nms_seg_t<float>(boxes, scores, idxs, valid, N, seg_offsets, S, thr, mode, top_n, keep, counts, ws, st);
    case B200DET_F16: return nms_seg_t<__half>(boxes, scores, idxs, valid, N, seg_offsets, S, thr, mode, top_n, keep, counts, ws, st);
    case B200DET_BF16: return nms_seg_t<__nv_bfloat16>(boxes, scores, idxs, valid, N, seg_offsets, S, thr, mode, top_n, keep, counts, ws, st);
  }
  return fail(B200DET_ENOSUP, "nms_segmented: dtype %d not supported (fp32 / fp16 / bf16 boxes)", dtype);
}

}  // namespace b200det

// ---- allocation-free NMS: b200det_nms_static ---------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) and do
// not want to carry a workspace either: the same launch sequence as b200det_nms, with the scratch (sort keys,
// order, score-ordered boxes, IoU bit mask) in a FIXED array that lives in the library's device image.  Every GPU
// that loads the library holds B200DET_NMS_STATIC_BYTES (10 MiB) of it until the process exits; it fits sets of
// up to B200DET_NMS_STATIC_MAX_BOXES (8192) boxes (8.9 MB: the bit mask is 8 MiB of it), larger sets are refused.
// The array's name below is a host-side stand-in in host code, not a device address: the device address is taken
// once per device, with that device current, through cudaGetSymbolAddress, and that pointer goes to the kernels.
// Nothing in the scratch has to start at zero (every word the scan reads was written by this call's front end
// or mask kernel), so it is not cleared; it keeps the last call's contents.  ALL calls on a device share the one
// array: calls that overlap on the GPU (different streams, or graphs replayed concurrently) race; serialise them.
static __device__ __align__(256) unsigned char g_nms_static_scratch[B200DET_NMS_STATIC_BYTES];

namespace b200det {
namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
void* g_static_ptr[kStaticMaxDevices] = {};

int nms_static_scratch(void** out) {
  int dev = -1;
  B200DET_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) return fail(B200DET_ENOSUP, "nms_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    B200DET_CUDA_OK(cudaGetSymbolAddress(&p, g_nms_static_scratch));
    if (!p) return fail(B200DET_EINVAL, "nms_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return 0;
}
}  // namespace
}  // namespace b200det

extern "C" int b200det_nms_static(const void* boxes, const void* scores, int dtype, int64_t N, double iou_threshold,
                                  int64_t* keep, int64_t* num_keep, void* stream) {
  using namespace b200det;
  if (N < 0) return fail(B200DET_EINVAL, "nms_static: negative N");
  if (!num_keep) return fail(B200DET_EINVAL, "nms_static: null num_keep");
  if (N > 0 && (!boxes || !scores || !keep)) return fail(B200DET_EINVAL, "nms_static: null pointer");
  if (dtype != B200DET_F32 && dtype != B200DET_F16 && dtype != B200DET_BF16 && dtype != B200DET_F64)
    return fail(B200DET_ENOSUP, "nms_static: unknown dtype %d", dtype);
  if (N > B200DET_NMS_STATIC_MAX_BOXES)
    return fail(B200DET_EINVAL, "nms_static: N=%lld exceeds the built-in scratch (at most %d boxes); "
                                "use b200det_nms with a caller-owned workspace", (long long)N, B200DET_NMS_STATIC_MAX_BOXES);
  const size_t need = nms_workspace_bytes(N);
  if (need > (size_t)B200DET_NMS_STATIC_BYTES)
    return fail(B200DET_EWORKSPACE, "nms_static: N=%lld needs %zu bytes of scratch, the built-in array has %d",
                (long long)N, need, B200DET_NMS_STATIC_BYTES);
  void* scratch = nullptr;
  const int rc = nms_static_scratch(&scratch);
  if (rc != 0) return rc;
  return nms_device(boxes, scores, nullptr, dtype, N, iou_threshold, 0, keep, num_keep, scratch,
                    (size_t)B200DET_NMS_STATIC_BYTES, (cudaStream_t)stream);
}
