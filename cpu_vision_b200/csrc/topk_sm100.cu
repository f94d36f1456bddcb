// topk_sm100.cu -- per-level top-k of the RPN objectness, all images and pyramid levels in ONE launch.
//
// Replaces RegionProposalNetwork._get_top_n_idx (torchvision/models/detection/rpn.py:231-240): a Python loop over
// the levels with `ob.topk(min(pre_nms_top_n, num_anchors), dim=1)` on a strided slice of the [B, A] objectness
// (torch clones the slice, then runs its multi-block radix top-k and a sort: 5 levels x ~8 kernels, 0.50 ms of the
// 1.97 ms filter_proposals takes at the cfg5 shape) followed by `+ offset` and `cat`.
//
// One CTA per (image, level).  (0) a floor under the k-th largest key from a sorted sample, so that the histograms only
// see a few percent of the keys; (1) radix select of the k-th largest key in three passes of 11 / 11 / 10 bits over
// the level's scores (shared-memory histogram, suffix scan from the top; keys ordered like torch's top-k orders
// floats: sign-flipped bit patterns, every NaN on top); (2) one more pass collects the keys above the threshold and
// the first `need` keys equal to it IN INDEX ORDER (ordered compaction: ballot + running offset), as 64-bit
// (key, ~index) composites in shared memory; (3) a block radix sort puts them in descending score order, equal
// scores by ascending index -- the order a stable descending sort gives; torch.topk leaves the order of ties
// unspecified -- and the indices are written at the level's offset inside the image's row, already shifted by
// the level's first anchor.  The scores are read four times: once from HBM, three times from L2.
#include <cub/block/block_radix_sort.cuh>
#include <cub/block/block_scan.cuh>

#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kTopkThreads = 1024, kTopkItems = 2, kTopkMax = kTopkThreads * kTopkItems;  // k <= 2048
constexpr int kBins = 2048;
constexpr int kTopkUnroll = 8;  // global loads a thread keeps in flight in the passes over the scores

struct TopkLevels {
  int end[B200DET_MAX_LEVELS];      // cumulative anchors per level
  int out_off[B200DET_MAX_LEVELS];  // first output column of the level
};

// torch's TopKTypeConfig<float>::convert: larger float -> larger key, every NaN the largest key
__device__ __forceinline__ unsigned topk_key(float v) {
  const unsigned x = __float_as_uint(v);
  const unsigned m = (x & 0x80000000u) ? 0xffffffffu : 0x80000000u;
  return (v == v) ? (x ^ m) : 0xffffffffu;
}

__global__ void __launch_bounds__(kTopkThreads) segmented_topk_kernel(const float* __restrict__ scores, long long A,
                                                                      const __grid_constant__ TopkLevels lv, int k,
                                                                      long long P, int64_t* __restrict__ out_idx) {
  using Scan = cub::BlockScan<int, kTopkThreads>;
  using Sort = cub::BlockRadixSort<unsigned long long, kTopkThreads, kTopkItems>;
  using SortS = cub::BlockRadixSort<unsigned, kTopkThreads, kTopkItems>;
  using SortP = cub::BlockRadixSort<unsigned, kTopkThreads, kTopkItems, unsigned>;
  __shared__ union {  // the phases never overlap in time
    typename SortS::TempStorage sample;    // (0) sort of the sample
    struct {
      int hist[kBins];
      typename Scan::TempStorage scan;
    } sel;                                 // (1) radix select
    unsigned long long buf[kTopkMax];      // (2) collected (key, ~index) composites
    typename SortP::TempStorage sortp;     // (3) block radix sort of (key, index) pairs
    typename Sort::TempStorage sort;       // (3') of 64-bit composites when equal keys came out of index order
  } tmp;
  unsigned long long* buf = tmp.buf;
  __shared__ unsigned s_prefix, s_need, s_floor;
  __shared__ int s_cnt, s_run, s_warp[kTopkThreads / 32];
  const int l = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long b = blockIdx.y;
  const int lo = l ? lv.end[l - 1] : 0, n = lv.end[l] - lo;
  const int kk = min(k, n);
  if (kk <= 0) return;
  const float* s = scores + b * A + lo;

  unsigned T = 0u, need = (unsigned)n;  // n <= k: every element is taken (all keys >= 0)
  if (n > kk) {
    // (0) a floor under the k-th largest key from a sample of 2048 keys: the histograms below then only see the
    // few percent of the keys above it (shared-memory atomics on the handful of bins a bell-shaped input fills
    // would serialise: 0.32 ms per call at the cfg5 shape without this).  The floor is the sample's key of rank
    // m + 4 sqrt(m) + 8, m = the expected number of sample keys above the k-th largest; should fewer than k keys
    // lie above it after all (probability ~1e-8), the first histogram notices and the floor is dropped.
    unsigned floor_key = 0u;
    if (n >= 4 * kTopkMax) {
      unsigned smp[kTopkItems];
#pragma unroll
      for (int q = 0; q < kTopkItems; ++q)
        smp[q] = topk_key(__ldg(s + (long long)(tid * kTopkItems + q) * n / kTopkMax));
      SortS(tmp.sample).SortDescending(smp, 16, 32);  // by the upper 16 bits: 4 passes instead of 8
      const float m = (float)kTopkMax * (float)kk / (float)n;
      const int r = (int)(m + 4.f * sqrtf(m) + 8.f);
      if (r < kTopkMax) {
#pragma unroll
        for (int q = 0; q < kTopkItems; ++q)
          if (tid * kTopkItems + q == r) s_floor = smp[q] & 0xffff0000u;  // (a lower bound of the sample's r-th key)
      } else if (tid == 0) {
        s_floor = 0u;
      }
      __syncthreads();
      floor_key = s_floor;
    }
    unsigned prefix = 0u, mask = 0u;
    need = (unsigned)kk;
    const int shifts[3] = {21, 10, 0}, bits[3] = {11, 11, 10};
#pragma unroll 1
    for (int p = 0; p < 3; ++p) {
      const int shift = shifts[p];
      const unsigned binmask = (1u << bits[p]) - 1u;
      for (int i = tid; i < kBins; i += kTopkThreads) tmp.sel.hist[i] = 0;
      __syncthreads();
      for (int i0 = tid; i0 < n; i0 += kTopkUnroll * kTopkThreads) {  // kTopkUnroll loads in flight per thread
        float v[kTopkUnroll];
#pragma unroll
        for (int u = 0; u < kTopkUnroll; ++u) {
          const int i = i0 + u * kTopkThreads;
          v[u] = i < n ? __ldg(s + i) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < kTopkUnroll; ++u) {
          const unsigned key = topk_key(v[u]);
          if (i0 + u * kTopkThreads < n && key >= floor_key && (key & mask) == prefix)
            atomicAdd(&tmp.sel.hist[(key >> shift) & binmask], 1);
        }
      }
      __syncthreads();
      // suffix sums from the top bin: thread t owns bins 2047 - 2t and 2046 - 2t
      int c[2] = {tmp.sel.hist[kBins - 1 - 2 * tid], tmp.sel.hist[kBins - 2 - 2 * tid]};
      int incl[2];
      Scan(tmp.sel.scan).InclusiveSum(c, incl);
      if (tid == kTopkThreads - 1) s_cnt = incl[1];  // keys counted in this pass
      if (tid == 0) s_prefix = 0xffffffffu;
      __syncthreads();
      if (p == 0 && (unsigned)s_cnt < need) {  // the floor was too high (block-uniform): once more without it
        floor_key = 0u;
        p = -1;
        __syncthreads();
        continue;
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const unsigned excl = (unsigned)(incl[q] - c[q]);
        if (excl < need && need <= (unsigned)incl[q]) {  // exactly one bin: the one that holds the need-th largest
          s_prefix = prefix | ((unsigned)(kBins - 1 - 2 * tid - q) << shift);
          s_need = need - excl;
        }
      }
      __syncthreads();
      prefix = s_prefix;
      need = s_need;
      mask |= binmask << shift;
      __syncthreads();
    }
    T = prefix;  // the kk-th largest key; `need` of the keys equal to it are taken
  }

  // collect: keys above T in any order, the first `need` keys equal to T in index order (a chunk is scanned in
  // order only when it holds such a key and some are still wanted)
  if (tid == 0) { s_cnt = 0; s_run = 0; }
  __syncthreads();
  const bool all = n <= kk;
  for (int c0 = 0; c0 < n; c0 += kTopkUnroll * kTopkThreads) {  // kTopkUnroll chunks of 1024 keys, loads in flight together
    float v[kTopkUnroll];
#pragma unroll
    for (int u = 0; u < kTopkUnroll; ++u) {
      const int i = c0 + u * kTopkThreads + tid;
      v[u] = i < n ? __ldg(s + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < kTopkUnroll; ++u) {
      const int i = c0 + u * kTopkThreads + tid;
      if (c0 + u * kTopkThreads < n) {  // block-uniform
        const unsigned key = topk_key(v[u]);
        const bool gt = i < n && !all && key > T;
        const bool eq = i < n && (all || key == T);
        if (gt) buf[atomicAdd(&s_cnt, 1)] = ((unsigned long long)key << 32) | (0xffffffffu - (unsigned)i);
        if (__syncthreads_or(eq && (unsigned)s_run < need)) {
          const unsigned bal = __ballot_sync(0xffffffffu, eq);
          if (lane == 0) s_warp[warp] = __popc(bal);
          __syncthreads();
          int before = s_run;  // equal keys of the chunks before this one
          for (int w = 0; w < warp; ++w) before += s_warp[w];
          const int my = before + __popc(bal & ((1u << lane) - 1u));
          if (eq && (unsigned)my < need)
            buf[atomicAdd(&s_cnt, 1)] = ((unsigned long long)key << 32) | (0xffffffffu - (unsigned)i);
          __syncthreads();
          if (tid == 0) {
            int tot = 0;
            for (int w = 0; w < kTopkThreads / 32; ++w) tot += s_warp[w];
            s_run += tot;
          }
          __syncthreads();
        }
      }
    }
  }
  __syncthreads();
  // s_cnt == kk.  Descending keys, equal keys by ascending index.  The pairs are sorted by key alone (8 passes of
  // a stable 32-bit sort instead of 16 of a 64-bit one); equal keys keep the order they were collected in, which is
  // arbitrary -- ties among the top k are rare, so they are looked for afterwards and only a block that finds a
  // run out of index order sorts the 64-bit (key, ~index) composites.
  unsigned long long items[kTopkItems];
  unsigned keys[kTopkItems], vals[kTopkItems];
#pragma unroll
  for (int q = 0; q < kTopkItems; ++q) {
    const int j = tid * kTopkItems + q;
    items[q] = j < kk ? buf[j] : 0ULL;   // (0 is below every real key)
    keys[q] = (unsigned)(items[q] >> 32);
    vals[q] = 0xffffffffu - (unsigned)(items[q] & 0xffffffffULL);
  }
  __syncthreads();
  SortP(tmp.sortp).SortDescending(keys, vals);
  __syncthreads();
#pragma unroll
  for (int q = 0; q < kTopkItems; ++q) buf[tid * kTopkItems + q] = ((unsigned long long)keys[q] << 32) | vals[q];
  __syncthreads();
  bool bad = false;
#pragma unroll
  for (int q = 0; q < kTopkItems; ++q) {
    const int j = tid * kTopkItems + q;
    if (j + 1 < kk) {
      const unsigned long long nx = buf[j + 1];
      bad |= (unsigned)(nx >> 32) == keys[q] && (unsigned)(nx & 0xffffffffULL) < vals[q];
    }
  }
  if (__syncthreads_or(bad)) {
    Sort(tmp.sort).SortDescending(items);
#pragma unroll
    for (int q = 0; q < kTopkItems; ++q) vals[q] = 0xffffffffu - (unsigned)(items[q] & 0xffffffffULL);
  }
  int64_t* o = out_idx + b * P + lv.out_off[l];
#pragma unroll
  for (int q = 0; q < kTopkItems; ++q) {
    const int j = tid * kTopkItems + q;
    if (j < kk) o[j] = (int64_t)lo + (int64_t)vals[q];
  }
}

}  // namespace

int segmented_topk(const float* scores, long long B, long long A, const int32_t* level_ends, int num_levels,
                   long long k, int64_t* out_idx, long long P, cudaStream_t st) {
  if (num_levels < 1 || num_levels > B200DET_MAX_LEVELS)
    return fail(B200DET_EINVAL, "segmented_topk: num_levels=%d outside [1,%d]", num_levels, B200DET_MAX_LEVELS);
  if (k < 1 || k > kTopkMax) return fail(B200DET_ENOSUP, "segmented_topk: k=%lld outside [1,%d]", k, kTopkMax);
  if (B < 0 || B > 65535) return fail(B200DET_EINVAL, "segmented_topk: batch %lld outside [0,65535]", B);
  TopkLevels lv;
  long long prev = 0, off = 0;
  for (int l = 0; l < B200DET_MAX_LEVELS; ++l) {
    const long long e = l < num_levels ? level_ends[l] : prev;
    if (e < prev || e > 0x7fffffffLL) return fail(B200DET_EINVAL, "segmented_topk: level ends must ascend");
    lv.end[l] = (int)e;
    lv.out_off[l] = (int)off;
    off += (e - prev) < k ? (e - prev) : k;
    prev = e;
  }
  if (prev != A) return fail(B200DET_EINVAL, "segmented_topk: the last level end (%lld) must be A (%lld)", prev, A);
  if (off != P) return fail(B200DET_EINVAL, "segmented_topk: P must be the sum of min(k, anchors of a level) = %lld", off);
  if (B == 0 || P == 0) return 0;
  if (!scores || !out_idx) return fail(B200DET_EINVAL, "segmented_topk: null pointer");
  segmented_topk_kernel<<<dim3((unsigned)num_levels, (unsigned)B), kTopkThreads, 0, st>>>(scores, A, lv, (int)k, P, out_idx);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace b200det
