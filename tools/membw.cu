// membw.cu -- what the memory system of THIS B200 gives a gather/scatter kernel
// (development aid; numbers quoted in DESIGN.md "Bounds").
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/membw tools/membw.cu && tools/membw
// Legs: 16-byte loads / stores / vector reds over a working set that fits L2
// (48 MB) and one that does not (4 GB); bulk-async (TMA 1D) loads into shared
// memory over the same sets.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define OK(x)                                                                       \
  do {                                                                              \
    cudaError_t e = (x);                                                            \
    if (e != cudaSuccess) {                                                         \
      printf("%s: %s\n", #x, cudaGetErrorString(e));                                \
      return 1;                                                                     \
    }                                                                               \
  } while (0)

__global__ void k_read(const float4* __restrict__ p, size_t n, int reps, float* sink) {
  float4 a = make_float4(0, 0, 0, 0);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride) {
      const float4 v = __ldg(p + i);
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
  if (a.x + a.y + a.z + a.w == 123.456f) *sink = a.x;
}

__global__ void k_write(float4* __restrict__ p, size_t n, int reps) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride)
      p[i] = make_float4(r, 1, 2, 3);
}

__global__ void k_red(float* __restrict__ p, size_t n4, int reps) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += stride)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %1, %1, %1};" ::"l"(p + 4 * i), "f"(1.0f) : "memory");
}

__global__ void k_red_scalar(float* __restrict__ p, size_t n, int reps) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride)
      atomicAdd(p + i, 1.0f);
}

// bulk-async loads: each CTA pulls `chunk`-byte pieces into a ring of `stages`
// shared buffers and does nothing with them (pure L2/HBM -> SM transport).
__global__ void k_bulk(const char* __restrict__ p, size_t bytes, int chunk, int stages, int reps) {
  extern __shared__ __align__(128) char sm[];
  __shared__ __align__(8) unsigned long long bar[8];
  const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const size_t nchunks = bytes / chunk;
  unsigned phase[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long issued = 0, waited = 0;
  for (int r = 0; r < reps; ++r)
    for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
      const int s = (int)(issued % stages);
      if (issued >= stages) {  // wait for the previous use of this stage
        const unsigned b = (unsigned)__cvta_generic_to_shared(&bar[s]);
        unsigned done = 0;
        while (!done)
          asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }"
                       : "=r"(done) : "r"(b), "r"(phase[s]));
        phase[s] ^= 1;
        ++waited;
      }
      const unsigned b = (unsigned)__cvta_generic_to_shared(&bar[s]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(chunk));
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(sbase + s * chunk), "l"(p + c * (size_t)chunk), "r"(chunk), "r"(b) : "memory");
      ++issued;
    }
  for (; waited < issued; ++waited) {
    const int s = (int)(waited % stages);
    const unsigned b = (unsigned)__cvta_generic_to_shared(&bar[s]);
    unsigned done = 0;
    while (!done)
      asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2; selp.u32 %0, 1, 0, q; }"
                   : "=r"(done) : "r"(b), "r"(phase[s]));
    phase[s] ^= 1;
  }
}


// TMA bulk store / bulk reduce-add (fp32) from shared memory: does the SM -> L2 port move more bytes per cycle with
// bulk requests than with st.128 / red.v4.f32 (round 2: that port bounds the scatter backward)?
// One thread per CTA issues cp.async.bulk / cp.reduce.async.bulk of `chunk` bytes, `depth` groups in flight.
template <bool RED>
__global__ void k_bulk_out(char* __restrict__ p, size_t bytes, int chunk, int depth, int reps) {
  extern __shared__ __align__(128) char sm[];
  for (int i = threadIdx.x; i < chunk / 4; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = RED ? 1.0f : 0.0f;
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x != 0) return;
  const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
  const size_t nchunks = bytes / chunk;
  int inflight = 0;
  for (int r = 0; r < reps; ++r)
    for (size_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
      char* g = p + c * (size_t)chunk;
      if (RED)
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                     ::"l"(g), "r"(sbase), "r"(chunk) : "memory");
      else
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(sbase), "r"(chunk) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (++inflight >= depth) {
        asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory");
      }
    }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
float timeit(F f, int iters) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < iters; ++i) f();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  return ms / iters;
}

int main() {
  const size_t big = 4ull << 30, small = 48ull << 20;
  char* buf;
  float* sink;
  OK(cudaMalloc(&buf, big));
  OK(cudaMalloc(&sink, 4));
  OK(cudaMemset(buf, 0, big));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int l2 = 0;
  cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0);
  printf("SMs %d, L2 %.1f MB\n", sms, l2 / 1048576.0);
  for (int pass = 0; pass < 2; ++pass) {
    const size_t bytes = pass ? big : small;
    const int reps = pass ? 1 : 40;
    const char* tag = pass ? "HBM 4GB " : "L2 48MB ";
    for (int occ = 1; occ <= 2; ++occ) {
      const int grid = sms * 4 * occ, th = 512;
      float ms = timeit([&] { k_read<<<grid, th>>>((const float4*)buf, bytes / 16, reps, sink); }, 5);
      printf("%s read  ld.128   grid %5d: %8.1f GB/s\n", tag, grid, bytes * (double)reps / ms / 1e6);
      ms = timeit([&] { k_write<<<grid, th>>>((float4*)buf, bytes / 16, reps); }, 5);
      printf("%s write st.128   grid %5d: %8.1f GB/s\n", tag, grid, bytes * (double)reps / ms / 1e6);
      ms = timeit([&] { k_red<<<grid, th>>>((float*)buf, bytes / 16, reps); }, 5);
      printf("%s red.v4.f32     grid %5d: %8.1f GB/s\n", tag, grid, bytes * (double)reps / ms / 1e6);
    }
    {
      const int grid = sms * 8, th = 512;
      const size_t b4 = pass ? (1ull << 30) : small;
      float ms = timeit([&] { k_red_scalar<<<grid, th>>>((float*)buf, b4 / 4, reps); }, 3);
      printf("%s atomicAdd.f32  grid %5d: %8.1f GB/s\n", tag, grid, b4 * (double)reps / ms / 1e6);
    }
    for (int chunk = 4096; chunk <= 32768; chunk *= 2)
      for (int stages = 2; stages <= 4; stages += 2) {
        const int smem = chunk * stages;
        cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        const int per_sm = smem <= 64 * 1024 ? 3 : 1;
        const int grid = sms * per_sm;
        float ms = timeit([&] { k_bulk<<<grid, 32, smem>>>(buf, bytes, chunk, stages, reps); }, 3);
        printf("%s bulk %5d B x%d stages, %d CTA/SM: %8.1f GB/s\n", tag, chunk, stages, per_sm,
               bytes * (double)reps / ms / 1e6);
      }
    for (int chunk = 1024; chunk <= 16384; chunk *= 4)
      for (int per_sm = 4; per_sm <= 16; per_sm *= 2) {
        const int grid = sms * per_sm;
        float ms = timeit([&] { k_bulk_out<false><<<grid, 128, chunk>>>(buf, bytes, chunk, 8, reps); }, 3);
        printf("%s bulk STORE  %5d B, %2d CTA/SM: %8.1f GB/s\n", tag, chunk, per_sm, bytes * (double)reps / ms / 1e6);
        ms = timeit([&] { k_bulk_out<true><<<grid, 128, chunk>>>(buf, bytes, chunk, 8, reps); }, 3);
        printf("%s bulk RED.f32 %5d B, %2d CTA/SM: %8.1f GB/s\n", tag, chunk, per_sm, bytes * (double)reps / ms / 1e6);
      }
  }
  cudaMemset(buf, 0, 1024);
  printf("done\n");
  return 0;
}
