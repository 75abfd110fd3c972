// Peer-mapped ("symmetric") device buffers for the one-process-per-GPU mode.
//
// The reference moves every payload through pickled CPU tensors over multiprocessing
// pipes (simulation_lib/algorithm_factory.py:26-28).  Here each rank cudaMallocs its
// upload arena / global-model buffers / flag words, publishes a cudaIpcMemHandle, and
// maps every peer's allocation into its own address space, so a kernel on GPU a
// dereferences GPU b's memory directly over NVLink (reduce.cu does exactly that).
#include "common.cuh"

using namespace dls;

extern "C" {

int dls_ipc_alloc(long long bytes, void** out_ptr) {
    if (bytes <= 0) return fail_msg("dls_ipc_alloc: bytes must be positive");
    void* p = nullptr;
    DLS_CHECK(cudaMalloc(&p, (size_t)bytes));
    DLS_CHECK(cudaMemset(p, 0, (size_t)bytes));
    DLS_CHECK(cudaDeviceSynchronize());
    *out_ptr = p;
    return 0;
}

int dls_ipc_free(void* ptr) {
    DLS_CHECK(cudaFree(ptr));
    return 0;
}

// handle_out: 64 bytes (CUDA_IPC_HANDLE_SIZE)
int dls_ipc_get_handle(void* ptr, unsigned char* handle_out) {
    cudaIpcMemHandle_t h;
    DLS_CHECK(cudaIpcGetMemHandle(&h, ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out, &h, sizeof(h));
    return 0;
}

int dls_ipc_open_handle(const unsigned char* handle, void** out_ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    DLS_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *out_ptr = p;
    return 0;
}

int dls_ipc_close_handle(void* ptr) {
    DLS_CHECK(cudaIpcCloseMemHandle(ptr));
    return 0;
}

// same-process multi-GPU: let `device` dereference `peer`'s allocations
int dls_enable_peer_access(int device, int peer) {
    int can = 0;
    DLS_CHECK(cudaDeviceCanAccessPeer(&can, device, peer));
    if (!can) return fail_msg("dls_enable_peer_access: devices are not peer capable");
    int cur = 0;
    DLS_CHECK(cudaGetDevice(&cur));
    DLS_CHECK(cudaSetDevice(device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    cudaSetDevice(cur);
    if (e != cudaSuccess) return fail("cudaDeviceEnablePeerAccess", e);
    return 0;
}

}  // extern "C"

// plain device-to-device copy whose source may be a peer-mapped pointer
extern "C" int dls_copy_device(void* dst, const void* src, long long bytes, void* stream) {
    DLS_CHECK(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}
