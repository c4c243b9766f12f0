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

extern "C" int bfd_abi_version(void) { return 5; }
extern "C" const char *bfd_last_cuda_error(void) { return bfd::g_cuda_error; }
extern "C" uint64_t bfd_launch_count(void) { return bfd::g_launches; }

// Device memory shareable between the processes of a slab run (include/burgers_b200.h).
#ifdef BFD_EMULATE
#include <stdlib.h>
extern "C" int bfd_mem_alloc(size_t bytes, void **ptr) {
    if (!ptr || !bytes) return BFD_EINVAL;
    *ptr = calloc(1, bytes);
    return *ptr ? BFD_OK : BFD_ECUDA;
}
extern "C" int bfd_mem_free(void *ptr) { free(ptr); return BFD_OK; }
extern "C" int bfd_ipc_export(void *ptr, unsigned char *handle) {   // one process: the handle is the pointer
    if (!ptr || !handle) return BFD_EINVAL;
    memset(handle, 0, BFD_IPC_HANDLE_BYTES);
    memcpy(handle, &ptr, sizeof(ptr));
    return BFD_OK;
}
extern "C" int bfd_ipc_open(const unsigned char *handle, void **ptr) {
    if (!ptr || !handle) return BFD_EINVAL;
    memcpy(ptr, handle, sizeof(*ptr));
    return BFD_OK;
}
extern "C" int bfd_ipc_close(void *) { return BFD_OK; }
#else
static_assert(sizeof(cudaIpcMemHandle_t) == BFD_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
extern "C" int bfd_mem_alloc(size_t bytes, void **ptr) {
    if (!ptr || !bytes) return BFD_EINVAL;
    BFD_CUDA(cudaMalloc(ptr, bytes));
    return BFD_OK;
}
extern "C" int bfd_mem_free(void *ptr) {
    if (ptr) BFD_CUDA(cudaFree(ptr));
    return BFD_OK;
}
extern "C" int bfd_ipc_export(void *ptr, unsigned char *handle) {
    if (!ptr || !handle) return BFD_EINVAL;
    cudaIpcMemHandle_t h;
    BFD_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle, &h, sizeof(h));
    return BFD_OK;
}
extern "C" int bfd_ipc_open(const unsigned char *handle, void **ptr) {
    if (!ptr || !handle) return BFD_EINVAL;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    BFD_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return BFD_OK;
}
extern "C" int bfd_ipc_close(void *ptr) {
    if (ptr) BFD_CUDA(cudaIpcCloseMemHandle(ptr));
    return BFD_OK;
}
#endif
