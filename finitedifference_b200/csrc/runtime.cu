// Library-wide state: launch counter and last CUDA error text.
#include "common.cuh"

#include <string.h>

namespace bfd {
uint64_t g_launches = 0;
char g_cuda_error[256] = "";

int record_cuda_error(cudaError_t e, const char *where) {
    snprintf(g_cuda_error, sizeof(g_cuda_error), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
    return BFD_ECUDA;
}
}  // namespace bfd

extern "C" int bfd_abi_version(void) { return 2; }
extern "C" const char *bfd_last_cuda_error(void) { return bfd::g_cuda_error; }
extern "C" uint64_t bfd_launch_count(void) { return bfd::g_launches; }
