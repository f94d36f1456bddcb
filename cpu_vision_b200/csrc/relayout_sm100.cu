// relayout_sm100.cu -- batched 2-D transposition dst[b][c][r] = src[b][r][c], the NCHW <-> NHWC
// re-layout around the channels_last RoIAlign kernels.  torchvision's detection models hand NCHW-
// contiguous feature maps to roi_align; the generic strided kernels serve those at the reference's
// parallelisation (cfg3 fp32: 9.4 ms forward, 11.5 ms backward), the channels_last kernels at 0.59 /
// 0.97 ms.  Two passes of this kernel (features in, pooled output back) cost about 1 ms, so the NCHW
// caller is served by relayout -> fast kernel -> relayout.
//
// 64 x 64 element tiles through shared memory (padded rows: conflict-free both ways); a warp reads and
// writes 64 consecutive elements, every DRAM sector is used in full.  HBM-bound: 2 x bytes per pass.
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

template <typename T>
int launch_relayout(const void* src, void* dst, long long B, long long R, long long S, cudaStream_t st) {
  const long long ts = (S + kRT - 1) / kRT, tr = (R + kRT - 1) / kRT;
  if (B * ts > 0x7fffffffLL) return fail(B200DET_EINVAL, "relayout: shape too large");
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
