// Issue-rate microbenchmark of tcgen05.mma (kind::f16, bf16 operands from shared memory, M = 128) on sm_100a.
// One elected thread per CTA issues `n` UMMAs back to back on operands that already sit in shared memory and
// commits once; the clock64 delta up to the commit's mbarrier completion divided by n is the cost per UMMA.
// Variants: N of the instruction, a wait + fence + commit every 4 UMMAs (the pipeline hand-shake of the
// convolution kernels), A groups 2 048 B apart and off the swizzle period (the halo-patch descriptors).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o umma_rate tools/umma_rate.cu && ./umma_rate
#include <cstdio>
#include <cstdlib>

#include "../distributed_learning_simulator_b200/csrc/tcgen05.cuh"

using namespace dls;

namespace dls {
thread_local char g_last_error[512];
}  // namespace dls

__device__ __forceinline__ uint64_t desc_sbo(uint32_t addr, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(sbo >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

// issue mode: 0 = `if (threadIdx.x == 0)` (divergent single lane), 1 = elect.sync once around the whole loop,
// 2 = whole warp walks the loop, elect.sync around each group of 4 UMMAs
template <int kN, int kMode>
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int n, int handshake, int halo, int n_acc, int accumulate, int issuers, long long* out) {
    extern __shared__ unsigned char raw[];
    unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar_done, bar_ready, bar_sink;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u + i;
    if (threadIdx.x == 0) {
        mbar_init(&bar_done, 1);
        mbar_init(&bar_ready, 1);
        mbar_init(&bar_sink, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async_smem();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) mbar_arrive(&bar_ready);  // phase 0 complete: waits on parity 0 return at once
    __syncthreads();
    if (threadIdx.x == 32 && issuers == 2) {  // a second issuer on its own accumulator columns, not timed
        const uint32_t idesc2 = umma_instr_desc<kN>();
        const uint32_t a0 = smem_u32(sm) + 16384, b0 = smem_u32(sm + 64 * 1024);
        for (int i = 0; i < n; ++i)
            umma_bf16(tmem + 256, umma_smem_desc(a0 + (i & 3) * 32), umma_smem_desc(b0 + (i & 3) * 32), idesc2, 1);
    }
    if (kMode == 2) {
        // warp-uniform control flow and operands (warp index and the tensor-memory base broadcast from lane 0), one
        // elect.sync around each instruction: the compiler can then keep the operands in uniform registers
        const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
        if (warp_u == 0) {
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t idesc = umma_instr_desc<kN>();
            const uint32_t a0 = smem_u32(sm), b0 = smem_u32(sm + 48 * 1024);
            const long long t0 = clock64();
            for (int i = 0; i < n; i += 4) {
                if (handshake) {
                    mbar_wait(&bar_ready, 0);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                const uint32_t a = a0 + ((i >> 2) & 1) * 16384;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    if (elect_one())
                        umma_bf16(tmem_u + (uint32_t)(((i >> 2) % n_acc) * kN), umma_smem_desc(a + kk * 32),
                                  umma_smem_desc(b0 + kk * 32), idesc, (uint32_t)accumulate);
                if (handshake && elect_one()) umma_commit(&bar_sink);
            }
            const long long t1 = clock64();
            if (elect_one()) umma_commit(&bar_done);
            mbar_wait(&bar_done, 0);
            const long long t2 = clock64();
            if (threadIdx.x == 0) {
                out[blockIdx.x * 2] = t1 - t0;
                out[blockIdx.x * 2 + 1] = t2 - t0;
            }
        }
    } else if (kMode == 3) {
        // straight-line issue: one elect.sync around an unrolled run of 16 UMMAs whose descriptors differ from a
        // precomputed base by constants -- the shortest instruction stream between two UMMAs
        const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
        if (warp_u == 0) {
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
            const uint32_t idesc = umma_instr_desc<kN>();
            const uint64_t da0 = umma_smem_desc(smem_u32(sm)), db0 = umma_smem_desc(smem_u32(sm + 48 * 1024));
            long long t0 = 0, t1 = 0;
            if (elect_one()) {
                t0 = clock64();
                for (int i = 0; i < n; i += 16) {
#pragma unroll
                    for (int u = 0; u < 16; ++u)
                        umma_bf16(tmem_u + (uint32_t)((u & (n_acc - 1)) * kN), da0 + (uint64_t)((u & 3) * 2 + (u >> 2) * 512),
                                  db0 + (uint64_t)((u & 3) * 2), idesc, (uint32_t)accumulate);
                }
                t1 = clock64();
                umma_commit(&bar_done);
            }
            __syncwarp();
            mbar_wait(&bar_done, 0);
            const long long t2 = clock64();
            if (t0 != 0) {
                out[blockIdx.x * 2] = t1 - t0;
                out[blockIdx.x * 2 + 1] = t2 - t0;
            }
        }
    } else if (kMode == 1 ? (warp == 0 && elect_one()) : threadIdx.x == 0) {
        uint32_t idesc = umma_instr_desc<kN>();
        const uint32_t a0 = smem_u32(sm), b0 = smem_u32(sm + 48 * 1024);
        const long long t0 = clock64();
        for (int i = 0; i < n; i += 4) {
            if (handshake) {
                mbar_wait(&bar_ready, 0);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const uint32_t a = halo ? a0 + (uint32_t)(((i >> 2) % 9 / 3) * 16 + ((i >> 2) % 3)) * 128 : a0 + ((i >> 2) & 1) * 16384;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const uint64_t da = halo ? desc_sbo(a + kk * 32, 2048) : umma_smem_desc(a + kk * 32);
                const uint64_t db = umma_smem_desc(b0 + kk * 32);
                umma_bf16(tmem + (uint32_t)(((i >> 2) % n_acc) * kN), da, db, idesc, (uint32_t)accumulate);
            }
            if (handshake) umma_commit(&bar_sink);
        }
        const long long t1 = clock64();
        umma_commit(&bar_done);
        mbar_wait(&bar_done, 0);
        const long long t2 = clock64();
        out[blockIdx.x * 2] = t1 - t0;
        out[blockIdx.x * 2 + 1] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

template <int kN, int kMode = 0>
static void run(int grid, int n, int handshake, int halo, int n_acc = 1, int accumulate = 1, int issuers = 1) {
    long long* out;
    cudaMalloc(&out, sizeof(long long) * 2 * grid);
    const size_t smem = 100 * 1024;
    cudaFuncSetAttribute(umma_rate_kernel<kN, kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int rep = 0; rep < 3; ++rep) umma_rate_kernel<kN, kMode><<<grid, 128, smem>>>(n, handshake, halo, n_acc, accumulate, issuers, out);
    if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("N=%d failed: %s\n", kN, cudaGetErrorString(cudaGetLastError()));
        exit(1);
    }
    long long h[2];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("{\"issue_mode\": %d, \"n_dim\": %d, \"grid\": %d, \"ummas\": %d, \"handshake\": %d, \"halo_desc\": %d, \"accumulators\": %d, \"accumulate\": %d, \"issuers\": %d, \"issue_clk_per_umma\": %.1f, "
           "\"done_clk_per_umma\": %.1f}\n",
           kMode, kN, grid, n, handshake, halo, n_acc, accumulate, issuers, (double)h[0] / n, (double)h[1] / n);
    cudaFree(out);
}

int main() {
    for (int hs : {0, 1})
        for (int halo : {0, 1}) {
            run<64>(148, 144, hs, halo);
            run<128>(148, 144, hs, halo);
            run<256>(148, 144, hs, halo);
        }
    run<64>(148, 144, 0, 0, 2);
    run<64>(148, 144, 0, 0, 4);
    run<128>(148, 144, 0, 0, 2);
    run<64>(148, 144, 0, 0, 1, 0);
    run<64>(148, 144, 0, 0, 4, 0);
    run<64>(148, 144, 0, 0, 1, 1, 2);
    run<128>(148, 144, 0, 0, 1, 1, 2);
    for (int hs : {0, 1}) {
        run<64, 1>(148, 144, hs, 0);
        run<128, 1>(148, 144, hs, 0);
        run<256, 1>(148, 144, hs, 0);
        run<64, 2>(148, 144, hs, 0);
        run<128, 2>(148, 144, hs, 0);
        run<256, 2>(148, 144, hs, 0);
    }
    run<64, 2>(148, 144, 1, 0, 2);
    run<64, 3>(148, 144, 0, 0);
    run<128, 3>(148, 144, 0, 0);
    run<256, 3>(148, 144, 0, 0);
    run<64, 3>(148, 144, 0, 0, 4);
    run<64, 3>(148, 1440, 0, 0, 2);
    run<64, 3>(1, 1440, 0, 0, 1);
    return 0;
}
