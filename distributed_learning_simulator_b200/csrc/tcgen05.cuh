// tcgen05 / TMA / mbarrier building blocks shared by the hand-written tensor-core kernels
// (gemm_tcgen05.cu, subset_gemm.cu).  sm_100a only.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace dls {

constexpr int kGemmBlockM = 128;
// BLOCK_N is a template parameter: 64, 128 or 256 columns per tile (= TMEM columns)
constexpr int kGemmBlockK = 64;   // 64 bf16 = 128 bytes = one swizzle atom row
constexpr int kGemmUmmaK = 16;
constexpr int kGemmStages = 4;
constexpr int kGemmThreads = 192;
constexpr uint32_t kBytesA = kGemmBlockM * kGemmBlockK * 2;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Single-thread roles (TMA producer, MMA issuer).  tcgen05.mma / cp.async.bulk.tensor take their operands from
// uniform registers.  Under `if (lane == 0)` the compiler cannot prove the operands warp-uniform and wraps EVERY
// such instruction in a per-lane loop (ELECT ... BRA.U.ANY) that waits on the instruction's scoreboard: measured on
// B200 (tools/umma_rate.cu) that costs 143 clk per UMMA whatever its N, against 48 / 64 / 126 clk for N = 64 /
// 128 / 256 when the UMMAs are issued back to back.  So: the warp index and everything read from shared memory
// are broadcast from lane 0 (warp-uniform by construction) and the role's thread is picked with elect.sync.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ int uniform_warp_index() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }
__device__ __forceinline__ uint32_t warp_uniform(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug must end in a trap (error reported to the host), not a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spins = 0;; ++spins) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (spins > 200000000u) {  // tens of seconds: a protocol bug, not a busy GPU
            printf("dls: mbarrier wait timed out (block %d thread %d barrier 0x%x parity %u)\n", (int)blockIdx.x,
                   (int)threadIdx.x, addr, parity);
            asm volatile("trap;");
        }
    }
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

// K-major operand tile in shared memory, 128-byte swizzle: rows are 128 bytes apart, groups of
// 8 rows 1024 bytes apart (SBO), LBO = 1 (unused for swizzled K-major), descriptor version 1.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);   // start address, bits [0,14)
    desc |= (uint64_t)1 << 16;                          // leading byte offset (>>4), bits [16,30)
    desc |= (uint64_t)(1024 >> 4) << 32;                // stride byte offset (>>4), bits [32,46)
    desc |= (uint64_t)1 << 46;                          // descriptor version (Blackwell)
    desc |= (uint64_t)2 << 61;                          // layout type: SWIZZLE_128B
    return desc;
}

// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, M = 128, N = kBlockN
template <int kBlockN>
__device__ __forceinline__ uint32_t umma_instr_desc() {
    uint32_t d = 0;
    d |= 1u << 4;                         // c_format: F32
    d |= 1u << 7;                         // a_format: BF16
    d |= 1u << 10;                        // b_format: BF16
    d |= (uint32_t)(kBlockN >> 3) << 17;  // n_dim
    d |= (uint32_t)(kGemmBlockM >> 4) << 24;  // m_dim
    return d;
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// generic-proxy writes to shared memory (st.shared by CUDA threads) must be made visible to the async
// proxy (tcgen05.mma operand reads) before the MMA is issued
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// [batch][rows][k] bf16, k contiguous; box = 64 (k) x box_rows x 1, 128-byte swizzle, zero fill
static int make_map(CUtensorMap* map, const void* base, int rows, int k, int batch, int box_rows) {
    EncodeTiledFn encode = get_encode_fn();
    if (encode == nullptr) return fail_msg("cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, (cuuint64_t)batch};
    const cuuint64_t strides[2] = {(cuuint64_t)k * 2, (cuuint64_t)rows * (cuuint64_t)k * 2};
    const cuuint32_t box[3] = {(cuuint32_t)kGemmBlockK, (cuuint32_t)box_rows, 1};
    const cuuint32_t elem_strides[3] = {1, 1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box,
                              elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_last_error, sizeof(g_last_error), "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return -1;
    }
    return 0;
}

}  // namespace dls
