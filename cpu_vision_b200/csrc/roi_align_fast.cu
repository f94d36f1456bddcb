// roi_align_fast.cu -- placeholder until the specialised kernels land.
#include "common.cuh"
#include "kernels.h"
namespace b200det {
int roi_align_fwd_fast(const LevelArgs&, int, const void*, const int32_t*, void*, int, long long, int,
                       int, int, bool, cudaStream_t, int* taken) { *taken = 0; return 0; }
int roi_align_bwd_fast(const LevelArgs&, int, const void*, const int64_t*, const void*,
                       const int32_t*, long long, int, int, int, bool, cudaStream_t, int* taken) { *taken = 0; return 0; }
}
