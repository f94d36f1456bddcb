// Where does the single-CTA NMS scan spend its cycles?  clock64() counters per phase of scan_smem_body.
// Build: nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -DB200DET_SCAN_PROFILE \
//          -o tools/bin/scan_prof tools/scan_prof.cu      (tools/bin/ is git-ignored; run it under gpurun)
#include "../cpu_vision_b200/csrc/nms_sm100.cu"
#include <cstdio>
#include <vector>
#include <random>
int main() {
  const int n = 5000;
  std::mt19937 g(1);
  std::uniform_real_distribution<float> u(0.f, 1.f);
  std::vector<float> b(n * 4), s(n);
  for (int i = 0; i < n; ++i) {
    const float x = u(g) * 1000, y = u(g) * 1000;
    b[4 * i] = x; b[4 * i + 1] = y; b[4 * i + 2] = x + u(g) * 200 + 10; b[4 * i + 3] = y + u(g) * 200 + 10;
    s[i] = u(g);
  }
  float *db, *ds; int64_t *keep, *nk; void* ws;
  cudaMalloc(&db, n * 16); cudaMalloc(&ds, n * 4); cudaMalloc(&keep, n * 8); cudaMalloc(&nk, 8);
  const size_t wb = b200det::nms_workspace_bytes(n);
  cudaMalloc(&ws, wb);
  cudaMemcpy(db, b.data(), n * 16, cudaMemcpyHostToDevice);
  cudaMemcpy(ds, s.data(), n * 4, cudaMemcpyHostToDevice);
  for (int it = 0; it < 3; ++it) {
    long long z[10] = {0};
    cudaMemcpyToSymbol(b200det::g_scan_prof, z, sizeof(z));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    int rc = b200det::nms_device(db, ds, nullptr, 0, n, 0.5, 0, keep, nk, ws, wb, 0);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpyFromSymbol(z, b200det::g_scan_prof, sizeof(z));
    long long k; cudaMemcpy(&k, nk, 8, cudaMemcpyDeviceToHost);
    printf("rc %d kept %lld total %.1f us | warp0: wait %lld barrier %lld prefetch %lld word %lld resolve %lld | warp1: wait %lld barrier %lld prefetch %lld or %lld (cycles, 79 blocks)\n",
           rc, k, ms * 1000, z[0], z[1], z[2], z[3], z[4], z[5], z[6], z[7], z[8]);
  }
  return 0;
}
namespace b200det { int fail(int code, const char* fmt, ...) { fprintf(stderr, "fail %d: %s\n", code, fmt); return code; } }
