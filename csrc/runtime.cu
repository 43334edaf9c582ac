// Device runtime: peer-mapped slab allocation (cudaMalloc + CUDA IPC), device queries, error strings.
//
// The parameter slab of every rank is allocated here (not by torch's caching allocator) so that it
// can be exported with cudaIpcGetMemHandle and mapped by the other ranks of the node; kernels then
// read peers' parameters with ordinary loads that travel over NVLink 5 / NVSwitch.  The reference has
// no counterpart: its "network" is a Python list of objects in one process.
#include <cuda_runtime.h>
#include <string.h>
#include "../include/fedmse_b200.h"

extern "C" int fm_slab_alloc(uint64_t bytes, void** ptr) {
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaMemset(*ptr, 0, bytes);
}
extern "C" int fm_slab_free(void* ptr) { return (int)cudaFree(ptr); }

extern "C" int fm_ipc_get_handle(void* ptr, uint8_t handle[64]) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
    if (e != cudaSuccess) return (int)e;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle, &h, 64);
    return 0;
}
extern "C" int fm_ipc_open_handle(const uint8_t handle[64], void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}
extern "C" int fm_ipc_close_handle(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }

extern "C" int fm_enable_peer_access(int32_t peer_device) {
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
    return (int)e;
}
extern "C" const char* fm_error_string(int code) { return cudaGetErrorString((cudaError_t)code); }

extern "C" int fm_device_info(int32_t* sm_count, int32_t* max_smem_optin, int32_t* cc_major, int32_t* cc_minor) {
    int dev;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev);
    return 0;
}
