// Symmetric receive-arena plumbing: device allocation outside the torch caching allocator (so the
// whole allocation can be exported), CUDA IPC handle export / import, and peer-access checks.
//
// Reference behaviour being replaced: one big registered memory region per process plus an rkey
// exchanged in the RDMA-CM private data (src/DataNet/RDMAComm.cc:67-154,178-218,
// src/DataNet/RDMAComm.h:35-39 connreq_data). Here the "registration" is a cudaMalloc'd arena whose IPC
// handle is all-gathered through the torch process group once at start-up; after that every rank holds
// a plain device pointer into every peer's arena and the transport is st.global over NVLink.
#include "common.cuh"
#include "../../include/uda_b200.h"
#include <string.h>

extern "C" int uda_dev_malloc(void** ptr, long long bytes) {
    UDA_CUDA(cudaMalloc(ptr, (size_t)bytes));
    return 0;
}
extern "C" int uda_dev_free(void* ptr) {
    UDA_CUDA(cudaFree(ptr));
    return 0;
}
extern "C" int uda_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
extern "C" int uda_ipc_get_handle(void* ptr, void* handle_out) {
    cudaIpcMemHandle_t h;
    UDA_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}
extern "C" int uda_ipc_open_handle(const void* handle_in, void** ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_in, sizeof(h));
    UDA_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int uda_ipc_close_handle(void* ptr) {
    UDA_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}
extern "C" int uda_can_access_peer(int dev, int peer, int* out) {
    UDA_CUDA(cudaDeviceCanAccessPeer(out, dev, peer));
    return 0;
}
extern "C" int uda_memset_async(void* ptr, int value, long long bytes, cudaStream_t stream) {
    UDA_CUDA(cudaMemsetAsync(ptr, value, (size_t)bytes, stream));
    return 0;
}
extern "C" int uda_host_alloc_pinned(void** ptr, long long bytes) {
    UDA_CUDA(cudaHostAlloc(ptr, (size_t)bytes, cudaHostAllocPortable));
    return 0;
}
extern "C" int uda_host_free_pinned(void* ptr) {
    UDA_CUDA(cudaFreeHost(ptr));
    return 0;
}

// Plain runtime services for hosts that drive the kernels through this C ABI without linking a CUDA runtime of their
// own (the C++ consumer of libuda.so dlopens this library for its device merge).
extern "C" int uda_device_count(int* out) {
    cudaError_t e = cudaGetDeviceCount(out);
    if (e != cudaSuccess) { *out = 0; cudaGetLastError(); return (int)e; }
    return 0;
}
extern "C" int uda_set_device(int dev) {
    UDA_CUDA(cudaSetDevice(dev));
    return 0;
}
extern "C" int uda_mem_info(long long* free_bytes, long long* total_bytes) {
    size_t f = 0, t = 0;
    UDA_CUDA(cudaMemGetInfo(&f, &t));
    *free_bytes = (long long)f;
    *total_bytes = (long long)t;
    return 0;
}
extern "C" int uda_memcpy_h2d_async(void* dst_dev, const void* src_host, long long bytes, cudaStream_t stream) {
    UDA_CUDA(cudaMemcpyAsync(dst_dev, src_host, (size_t)bytes, cudaMemcpyHostToDevice, stream));
    return 0;
}
extern "C" int uda_memcpy_d2h_async(void* dst_host, const void* src_dev, long long bytes, cudaStream_t stream) {
    UDA_CUDA(cudaMemcpyAsync(dst_host, src_dev, (size_t)bytes, cudaMemcpyDeviceToHost, stream));
    return 0;
}
extern "C" int uda_stream_synchronize(cudaStream_t stream) {
    UDA_CUDA(cudaStreamSynchronize(stream));
    return 0;
}

// Human-readable text for a return code of this library (negative: UDA_E_*, positive: cudaError_t).
extern "C" const char* uda_error_string(int rc) {
    if (rc == 0) return "success";
    if (rc == UDA_E_INVALID) return "uda: invalid argument";
    if (rc == UDA_E_TOO_LARGE) return "uda: size exceeds the documented maximum of this entry point";
    if (rc > 0) return cudaGetErrorString((cudaError_t)rc);
    return "uda: unknown error";
}
