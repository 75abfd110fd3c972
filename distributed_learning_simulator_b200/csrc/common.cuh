// Shared helpers for the hand-written sm_100a kernels of this framework.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>

namespace dls {

// number of kernels this library has launched (bench.py reports it as "gpu_launches")
extern std::atomic<long long> g_launch_count;
extern thread_local char g_last_error[512];

inline int fail(const char* what, cudaError_t err) {
    snprintf(g_last_error, sizeof(g_last_error), "%s: %s", what, cudaGetErrorString(err));
    return (int)err ? (int)err : -1;
}
inline int fail_msg(const char* what) {
    snprintf(g_last_error, sizeof(g_last_error), "%s", what);
    return -1;
}

#define DLS_CHECK(expr)                                         \
    do {                                                        \
        cudaError_t _e = (expr);                                \
        if (_e != cudaSuccess) return ::dls::fail(#expr, _e);   \
    } while (0)

#define DLS_LAUNCHED(name)                                      \
    do {                                                        \
        ::dls::g_launch_count.fetch_add(1);                     \
        cudaError_t _e = cudaGetLastError();                    \
        if (_e != cudaSuccess) return ::dls::fail(name, _e);    \
    } while (0)

constexpr int kNumSMs = 148;  // B200

// ---- programmatic dependent launch ---------------------------------------------------------------
// Inside the captured training step ~150 small kernels run back to back; the gap between two dependent
// kernels (drain, launch, prologue) is a visible share of the step.  Kernels launched through launch_pdl()
// may start while their predecessor in the stream is still running: every CTA first lets ITS successor
// launch (pdl_launch_dependents), does the work that does not touch the predecessor's data (barrier
// init, tensor-memory allocation, descriptor prefetch, index maths) and only then waits for the
// predecessor grid to complete and flush (pdl_wait).  Without the launch attribute both are no-ops.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

inline bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("DLS_PDL"); return e == nullptr || e[0] != '0'; }();
    return on;
}

// cluster == 0: no cluster attribute
template <typename K, typename... Args>
inline cudaError_t launch_pdl(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster,
                              Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int n = 0;
    if (cluster > 0) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = (unsigned)cluster;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

inline int grid_for(long long work_items, int threads, int blocks_per_sm) {
    long long blocks = (work_items + threads - 1) / threads;
    long long cap = (long long)kNumSMs * blocks_per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ---- device helpers ---------------------------------------------------------------
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
    // streaming 128-bit load that does not allocate in L1 (data is touched once)
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}
// peer / volatile-safe variant (no .nc: the row may live in another GPU's memory and is
// written by that GPU shortly before; ordering comes from the acquire on the flag)
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// upper_bound over a small sorted table: returns t with table[t] <= x < table[t+1]
__device__ __forceinline__ int find_segment(const long long* __restrict__ table, int n_segments, long long x) {
    int lo = 0, hi = n_segments;  // table has n_segments + 1 entries
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (table[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// Philox4x32-10 counter-based RNG (stateless: stream = (seed, element index))
__device__ __forceinline__ uint4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                            uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ float u32_to_unit(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }

}  // namespace dls
