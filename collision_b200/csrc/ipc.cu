// ipc.cu -- device buffers shared between the ranks of ONE box (one process per GPU).
//
// The sharded build of collision_b200/sharded.py keeps every rank's part of the traversal data
// in that rank's own HBM and lets the other ranks read the pieces they need straight over
// NVLink (peer loads from a kernel on the reading rank), instead of broadcasting the whole tree.
// That needs buffers whose addresses can be opened in another process: plain cudaMalloc
// allocations exported with CUDA IPC.  These are the only entry points of the library that
// allocate; the reference has no counterpart (it is single-device).
#include <cstring>

#include "common.cuh"

COL_API int col_ipc_alloc(void** ptr, uint64_t bytes) {
    if (!ptr) return COL_ERR_BAD_ARG;
    COL_CUDA(cudaMalloc(ptr, bytes ? bytes : 1));
    return 0;
}

COL_API int col_ipc_free(void* ptr) {
    COL_CUDA(cudaFree(ptr));
    return 0;
}

// handle64: 64 bytes (cudaIpcMemHandle_t) to be sent to the peers (host memory)
COL_API int col_ipc_export(void* ptr, void* h_handle64) {
    if (!ptr || !h_handle64) return COL_ERR_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    COL_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(h_handle64), ptr));
    return 0;
}

// opens a peer's allocation in this process, on the CURRENT device (peer access is enabled lazily)
COL_API int col_ipc_open(void** ptr, const void* h_handle64) {
    if (!ptr || !h_handle64) return COL_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, h_handle64, sizeof(h));
    COL_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

COL_API int col_ipc_close(void* ptr) {
    COL_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}
