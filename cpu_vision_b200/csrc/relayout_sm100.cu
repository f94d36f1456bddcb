// relayout_sm100.cu -- batched 2-D transposition dst[b][c][r] = src[b][r][c], the NCHW <-> NHWC
// re-layout around the channels_last RoIAlign kernels.  torchvision's detection models hand NCHW-
// contiguous feature maps to roi_align; the generic strided kernels serve those at the reference's
// parallelisation (cfg3 fp32: 9.4 ms forward, 11.5 ms backward), the channels_last kernels at 0.59 /
// 0.97 ms.  Two passes of this kernel (features in, pooled output back) cost about 1 ms, so the NCHW
// caller is served by relayout -> fast kernel -> relayout.
//
// 64 x 64 element tiles through shared memory (padded rows: conflict-free both ways); a warp reads and
// writes 64 consecutive elements, every DRAM sector is used in full.  HBM-bound: 2 x bytes per pass.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace b200det {
namespace {

constexpr int kRT = 64;

template <typename T>
__global__ void __launch_bounds__(256) relayout_kernel(const T* __restrict__ src, T* __restrict__ dst, long long R,
                                                       long long S, int tiles_s) {
  __shared__ T tile[kRT][kRT + (sizeof(T) == 2 ? 2 : 1)];
  const long long b = blockIdx.x / tiles_s;
  const long long s0 = (long long)(blockIdx.x % tiles_s) * kRT, r0 = (long long)blockIdx.y * kRT;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
  const T* sp = src + b * R * S;
  T* dp = dst + b * R * S;
  if (s0 + tx < S)
#pragma unroll 4
    for (int j = ty; j < kRT; j += 4)
      if (r0 + j < R) tile[j][tx] = sp[(r0 + j) * S + s0 + tx];
  __syncthreads();
  if (r0 + tx < R)
#pragma unroll 4
    for (int j = ty; j < kRT; j += 4)
      if (s0 + j < S) dp[(s0 + j) * R + r0 + tx] = tile[tx][j];
}

// Tall variant for R <= kTallRows (channels, or pooled bins: 256, 196, 49 ...): a CTA takes ALL rows of a
// 64-column strip.  Its destination -- rows s0 .. s0+63 of dst[b], R elements each -- is one contiguous
// range of 64 * R elements, so the write phase is a linear, fully coalesced copy out of shared memory
// (odd pitch: the strided fill is conflict-free).
constexpr int kTallRows = 256;

template <typename T>
__global__ void __launch_bounds__(256) relayout_tall_kernel(const T* __restrict__ src, T* __restrict__ dst, int R,
                                                            long long S, int tiles_s) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  const int Rp = R | 1;
  const long long b = blockIdx.x / tiles_s;
  const long long s0 = (long long)(blockIdx.x % tiles_s) * kRT;
  const int ncol = (int)min((long long)kRT, S - s0);
  const T* sp = src + b * R * S + s0;
  T* dp = dst + (b * S + s0) * R;
  {
    const int j = threadIdx.x & 63, rr = threadIdx.x >> 6;
    if (j < ncol) {
#pragma unroll 8
      for (int r = rr; r < R; r += 4) tile[j * Rp + r] = sp[(long long)r * S + j];
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = warp; j < ncol; j += 8) {
    const T* t = tile + j * Rp;
    T* d = dp + (long long)j * R;
#pragma unroll 8
    for (int r = lane; r < R; r += 32) d[r] = t[r];
  }
}

// Wide variant for S <= kTallRows (NHWC -> NCHW of a feature map: S = C): the mirror image.  A CTA takes
// 64 source rows, which are one contiguous range of 64 * S elements (linear coalesced read), and writes
// 64-element runs of every destination row.
template <typename T>
__global__ void __launch_bounds__(256) relayout_wide_kernel(const T* __restrict__ src, T* __restrict__ dst,
                                                            long long R, int S, int tiles_r) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);
  const int Sp = S | 1;
  const long long b = blockIdx.x / tiles_r;
  const long long r0 = (long long)(blockIdx.x % tiles_r) * kRT;
  const int nrow = (int)min((long long)kRT, R - r0);
  const T* sp = src + (b * R + r0) * S;
  T* dp = dst + b * R * S + r0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = warp; j < nrow; j += 8) {
    const T* g = sp + (long long)j * S;
    T* t = tile + j * Sp;
#pragma unroll 8
    for (int c = lane; c < S; c += 32) t[c] = g[c];
  }
  __syncthreads();
  const int j = threadIdx.x & 63, ss = threadIdx.x >> 6;
  if (j < nrow) {
#pragma unroll 8
    for (int c = ss; c < S; c += 4) dp[(long long)c * R + j] = tile[j * Sp + c];
  }
}


// ---- pipelined variants (round 2) ----------------------------------------------------------------------
// The kernels above load a tile, barrier, store it: inside a CTA the two phases never overlap, and ncu showed the
// re-layout at 46-59 % of DRAM throughput.  Here a persistent CTA keeps TWO tiles in shared memory: while the
// threads store tile i (transposed reads from shared memory, coalesced stores), the 16-byte cp.async copies of
// tile i + 1 are in flight (no registers, no instruction waits on them until the next barrier).  Rows keep a
// 16-byte aligned pitch (row bytes + 16), so the transposed side reads with 4-way bank conflicts; that costs
// ~2600 shared-memory cycles per 64 KB tile against ~6000 cycles of HBM time for its 128 KB, and overlaps them.
// Measured (tools/kbench_relayout.py, profiles/r02_kbench_relayout.log): cfg3 features NCHW -> NHWC 0.40 -> 0.30 ms
// (5.9 TB/s), pooled 14x14 outputs 0.86 -> 0.57 ms / back 0.67 -> 0.59 ms, pooled 7x7 back 0.77 -> 0.65 ms.  The
// mirror-image kernel for the wide case (NHWC -> NCHW of a feature map: 256-byte runs scattered over C rows) was
// slower than the load / barrier / store kernel (0.45 -> 0.48 ms: the scattered stores are the long pole and want
// more resident warps, not overlapped loads) and is not in the tree.
constexpr int kPipeThreads = 256;
constexpr int kRunBytes = 256;  // source row piece of one tile

__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(g)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// R <= 256 rows, any S with S * sizeof(T) % 16 == 0: tile = all R rows x (256 / sizeof(T)) columns
template <typename T>
__global__ void __launch_bounds__(kPipeThreads) relayout_tall_pipe_kernel(const T* __restrict__ src, T* __restrict__ dst,
                                                                         int R, long long S, int tiles_s,
                                                                         long long tiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int EPC = 16 / (int)sizeof(T);         // elements per 16-byte chunk
  constexpr int W = kRunBytes / (int)sizeof(T);    // columns per tile
  constexpr int kPitch = kRunBytes + 16;
  const int tile_bytes = R * kPitch;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto issue = [&](long long t, int buf) {
    const long long b = t / tiles_s, s0 = (t % tiles_s) * W;
    const int nchunk = (int)min((long long)W, S - s0) / EPC;
    const T* sp = src + b * R * S + s0;
    unsigned char* tb = smem_raw + buf * tile_bytes;
    for (int c = threadIdx.x; c < R * 16; c += kPipeThreads) {
      const int r = c >> 4, k = c & 15;
      if (k < nchunk) cp_async16(tb + r * kPitch + k * 16, sp + (long long)r * S + k * EPC);
    }
    cp_async_commit();
  };
  long long t = blockIdx.x;
  int buf = 0;
  if (t < tiles) issue(t, 0);
  for (; t < tiles; t += gridDim.x) {
    const long long tn = t + gridDim.x;
    if (tn < tiles) {
      issue(tn, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const long long b = t / tiles_s, s0 = (t % tiles_s) * W;
    const int ncol = (int)min((long long)W, S - s0);
    T* dp = dst + (b * S + s0) * R;
    const unsigned char* tb = smem_raw + buf * tile_bytes;
    for (int j = warp; j < ncol; j += kPipeThreads / 32) {
      T* d = dp + (long long)j * R;
      const unsigned char* col = tb + j * (int)sizeof(T);
#pragma unroll 4
      for (int r = lane; r < R; r += 32) d[r] = *reinterpret_cast<const T*>(col + r * kPitch);
    }
    __syncthreads();  // the next iteration's copies overwrite the other buffer, read one iteration ago
    buf ^= 1;
  }
}

// persistent grid: as many CTAs as fit at once (the occupancy of the last shared-memory size is remembered per kernel
// and device: a re-layout is called several times per NCHW roi_align call)
template <typename K>
int pipe_grid(K kernel, int smem, long long tiles, unsigned* grid) {
  static int c_dev = -1, c_smem = -1;
  static long long c_cap = 0;
  static int c_max_set = 0;
  int dev = 0;
  B200DET_CUDA_OK(cudaGetDevice(&dev));
  if (dev != c_dev || smem != c_smem) {
    int sms = 0, per_sm = 0;
    B200DET_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (dev != c_dev || smem > c_max_set) {
      B200DET_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      c_max_set = smem;
    }
    B200DET_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kPipeThreads, smem));
    if (per_sm < 1) return fail(B200DET_EINVAL, "relayout: tile does not fit shared memory");
    c_dev = dev;
    c_smem = smem;
    c_cap = (long long)sms * per_sm;
  }
  *grid = (unsigned)(tiles < c_cap ? tiles : c_cap);
  return 0;
}

template <typename T>
int launch_relayout(const void* src, void* dst, long long B, long long R, long long S, cudaStream_t st) {
  const long long ts = (S + kRT - 1) / kRT, tr = (R + kRT - 1) / kRT;
  if (B * ts > 0x7fffffffLL) return fail(B200DET_EINVAL, "relayout: shape too large");
  // pipelined variants: 16-byte copies need 16-byte aligned rows.  B200DET_RELAYOUT_PIPE=0 (read per call) keeps
  // the load / barrier / store kernels below.
  const char* pipe_s = getenv("B200DET_RELAYOUT_PIPE");
  const bool pipe_ok = !(pipe_s && atoi(pipe_s) == 0) && sizeof(T) <= 8 && (S * sizeof(T)) % 16 == 0 &&
                       ((uintptr_t)src) % 16 == 0;
  if (pipe_ok && R <= kTallRows && R * S >= 4096) {
    constexpr int W = kRunBytes / (int)sizeof(T);
    const long long tps = (S + W - 1) / W, tiles = B * tps;
    const int smem = 2 * (int)R * (kRunBytes + 16);
    unsigned grid = 0;
    if (tps <= 0x7fffffffLL) {
      int rc = pipe_grid(relayout_tall_pipe_kernel<T>, smem, tiles, &grid);
      if (rc) return rc;
      relayout_tall_pipe_kernel<T><<<grid, kPipeThreads, smem, st>>>((const T*)src, (T*)dst, (int)R, S, (int)tps, tiles);
      B200DET_CUDA_OK(cudaGetLastError());
      return 0;
    }
  }
  if (R <= kTallRows && sizeof(T) <= 4) {
    const int smem = kRT * (int)(R | 1) * (int)sizeof(T);
    static bool attr_set = false;  // per T instantiation
    if (!attr_set) {
      B200DET_CUDA_OK(cudaFuncSetAttribute(relayout_tall_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           kRT * (kTallRows | 1) * (int)sizeof(T)));
      attr_set = true;
    }
    relayout_tall_kernel<T><<<(unsigned)(B * ts), 256, smem, st>>>((const T*)src, (T*)dst, (int)R, S, (int)ts);
    B200DET_CUDA_OK(cudaGetLastError());
    return 0;
  }
  if (S <= kTallRows && sizeof(T) <= 4 && B * tr <= 0x7fffffffLL) {
    const int smem = kRT * (int)(S | 1) * (int)sizeof(T);
    static bool attr_set = false;
    if (!attr_set) {
      B200DET_CUDA_OK(cudaFuncSetAttribute(relayout_wide_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           kRT * (kTallRows | 1) * (int)sizeof(T)));
      attr_set = true;
    }
    relayout_wide_kernel<T><<<(unsigned)(B * tr), 256, smem, st>>>((const T*)src, (T*)dst, R, (int)S, (int)tr);
    B200DET_CUDA_OK(cudaGetLastError());
    return 0;
  }
  if (tr > 65535) return fail(B200DET_EINVAL, "relayout: shape too large");
  dim3 grid((unsigned)(B * ts), (unsigned)tr);
  relayout_kernel<T><<<grid, 256, 0, st>>>((const T*)src, (T*)dst, R, S, (int)ts);
  B200DET_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int relayout(const void* src, void* dst, int dtype, long long B, long long R, long long S, cudaStream_t st) {
  if (B == 0 || R == 0 || S == 0) return 0;
  switch (dtype_size(dtype)) {
    case 2: return launch_relayout<unsigned short>(src, dst, B, R, S, st);
    case 4: return launch_relayout<unsigned int>(src, dst, B, R, S, st);
    case 8: return launch_relayout<unsigned long long>(src, dst, B, R, S, st);
  }
  return fail(B200DET_ENOSUP, "relayout: unknown dtype %d", dtype);
}

}  // namespace b200det
