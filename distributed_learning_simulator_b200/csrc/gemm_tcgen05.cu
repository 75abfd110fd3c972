// bf16 GEMM on the 5th-generation tensor cores: C[b] = act(A[b] * B[b]^T + bias[b]).
//
// Used for the evaluation GEMMs (fully connected layers of the server's test pass) and for
// the batched GTG-Shapley coalition evaluator, where the batch index b is the coalition
// (reference path: one CPU model average + one full eager test inference per coalition,
// simulation_lib/method/shapley_value/shapley_value_algorithm.py:67-76).
//
// Structure (sm_100a, one CTA per 128 x {64,128,256} output tile, no library GEMM):
//   warp 0      TMA producer: cp.async.bulk.tensor loads of the A (128 x 64) and B (BLOCK_N x 64)
//               K-slices into a 4-stage 128B-swizzled shared-memory ring, mbarrier complete_tx
//   warp 1      allocates TMEM, then one lane issues tcgen05.mma (M128 N64/128/256 K16, bf16 -> fp32
//               accumulators in tensor memory); tcgen05.commit releases ring slots and finally
//               signals the epilogue
//   warps 2-5   epilogue: tcgen05.ld (32 lanes x 32 columns per instruction) -> bias / ReLU ->
//               128-bit global stores (fp32 or bf16)
// The CTAs are persistent (one per SM, static tile schedule) and the accumulator is double
// buffered in TMEM, so the epilogue of one tile overlaps the MMAs of the next.
// Out-of-range rows/columns/K are zero-filled by the TMA unit, so any M, N and any K % 8 == 0
// work without padding copies.
#include <stdlib.h>

#include "tcgen05.cuh"
#include "../../include/dls_b200.h"

namespace dls {

template <int kBlockN>
struct __align__(1024) GemmSmem {
    __nv_bfloat16 a[kGemmStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kGemmStages][kBlockN * kGemmBlockK];
    uint64_t full[kGemmStages];
    uint64_t empty[kGemmStages];
    uint64_t tmem_full[2];   // accumulator stage ready for the epilogue
    uint64_t tmem_empty[2];  // accumulator stage drained, the MMA warp may overwrite it
    uint32_t tmem_base;
};

template <int kBlockN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                         const float* __restrict__ bias, long long bias_batch_stride, void* __restrict__ c,
                         int out_bf16, int m, int n, int k, int batch_count, int b_shared, int relu) {
    // Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ... ; two accumulator stages in TMEM
    // so the epilogue of tile i overlaps the MMAs of tile i + 1.
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kBytesB = kBlockN * kGemmBlockK * 2;
    constexpr uint32_t kTmemCols = 2 * kBlockN;  // 128 / 256 / 512: a power of two
    using Smem = GemmSmem<kBlockN>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int num_kb = (k + kGemmBlockK - 1) / kGemmBlockK;
    const int m_tiles = (m + kGemmBlockM - 1) / kGemmBlockM;
    const int n_tiles = (n + kBlockN - 1) / kBlockN;
    const long long num_tiles = (long long)m_tiles * n_tiles * batch_count;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < kGemmStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], 128);  // every epilogue thread arrives
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // whole warp: allocate the accumulator columns, then give the permit back
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer: the ring runs straight through tile boundaries
            uint32_t it = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int mt = (int)(tile % m_tiles), nt = (int)((tile / m_tiles) % n_tiles);
                const int batch = (int)(tile / ((long long)m_tiles * n_tiles));
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t phase = (it / kGemmStages) & 1;
                    mbar_wait(&sm.empty[s], phase ^ 1);
                    mbar_expect_tx(&sm.full[s], kBytesA + kBytesB);
                    tma_load_3d(sm.a[s], &map_a, &sm.full[s], kb * kGemmBlockK, mt * kGemmBlockM, batch);
                    tma_load_3d(sm.b[s], &map_b, &sm.full[s], kb * kGemmBlockK, nt * kBlockN, b_shared ? 0 : batch);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = umma_instr_desc<kBlockN>();
            uint32_t it = 0, local_tile = 0;
            for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);  // epilogue has drained this stage
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t phase = (it / kGemmStages) & 1;
                    mbar_wait(&sm.full[s], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                    for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                        // advancing 16 bf16 = 32 bytes along K stays inside the 128-byte swizzle row
                        umma_bf16(tmem_acc, umma_smem_desc(a_addr + kk * kGemmUmmaK * 2),
                                  umma_smem_desc(b_addr + kk * kGemmUmmaK * 2), idesc, (kb | kk) != 0);
                    }
                    umma_commit(&sm.empty[s]);  // frees the ring slot when these MMAs have read it
                }
                umma_commit(&sm.tmem_full[acc]);  // accumulator stage complete
            }
        }
    } else {
        // ---- epilogue: warp w may touch TMEM lanes [32 * (w % 4), +32)
        const int quarter = warp & 3;
        uint32_t local_tile = 0;
        for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            const int mt = (int)(tile % m_tiles), nt = (int)((tile / m_tiles) % n_tiles);
            const int batch = (int)(tile / ((long long)m_tiles * n_tiles));
            const int m0 = mt * kGemmBlockM, n0 = nt * kBlockN;
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            const int row = m0 + quarter * 32 + lane;
            mbar_wait(&sm.tmem_full[acc], use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_acc = tmem_base + acc * kBlockN;
            const float* bias_b = bias != nullptr ? bias + (long long)batch * bias_batch_stride : nullptr;
#pragma unroll 1
            for (int chunk = 0; chunk < kBlockN / 32; ++chunk) {
                if (n0 + chunk * 32 >= n) break;  // columns past N hold zeros: nothing to store
                uint32_t r[32];
                const uint32_t taddr = tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                      "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
                      "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
                      "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
                      "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                    : "r"(taddr)
                    : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row < m) {
                    const int col0 = n0 + chunk * 32;
                    const long long out_off = ((long long)batch * m + row) * n + col0;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        v[j] = __uint_as_float(r[j]);
                        if (bias_b != nullptr && col0 + j < n) v[j] += bias_b[col0 + j];
                        if (relu) v[j] = fmaxf(v[j], 0.f);
                    }
                    const bool full = col0 + 32 <= n;
                    if (out_bf16) {
                        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(c) + out_off;
                        if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                            // 32 columns = 64 bytes: four 128-bit stores of packed bf16 pairs
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint4 pk;
                                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                                for (int e = 0; e < 4; ++e)
                                    h[e] = __floats2bfloat162_rn(v[8 * q + 2 * e], v[8 * q + 2 * e + 1]);
                                reinterpret_cast<uint4*>(dst)[q] = pk;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < n) dst[j] = __float2bfloat16(v[j]);
                        }
                    } else {
                        float* dst = reinterpret_cast<float*>(c) + out_off;
                        if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                reinterpret_cast<float4*>(dst)[q] =
                                    make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < n) dst[j] = v[j];
                        }
                    }
                }
            }
            // this thread's TMEM reads of the stage are complete: hand the stage back to the MMA warp
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&sm.tmem_empty[acc]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}


// ------------------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2): two CTAs on neighbouring SMs work on ONE 256 x kBlockN tile.
// CTA r holds rows [128 r, 128 r + 128) of the A tile and rows [r kBlockN/2, ...) of the B tile in its own
// shared memory; the leader CTA issues M = 256 MMAs that read both halves, each CTA's accumulator
// (its 128 rows x kBlockN columns) lives in its own TMEM.  Every B byte is fetched once per pair
// instead of once per CTA and each instruction does twice the work.
//   * both CTAs' TMA loads complete on the LEADER's full barrier (cp.async.bulk.tensor ... cta_group::2,
//     barrier address with the peer bit cleared), only the leader posts the expected byte count;
//   * the leader releases ring slots and publishes accumulators with multicast tcgen05.commit
//     (cta_group::2) to both CTAs;
//   * the peer's epilogue threads hand accumulator stages back with remote mbarrier arrives.
constexpr int kPairStages = 6;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // shared::cluster address of the same offset in the pair's CTA 0

template <int kBlockN>
struct __align__(1024) PairSmem {
    __nv_bfloat16 a[kPairStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kPairStages][(kBlockN / 2) * kGemmBlockK];
    uint64_t full[kPairStages];
    uint64_t empty[kPairStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ void tma_load_3d_pair(void* smem_dst, const CUtensorMap* map, uint64_t* leader_bar, int c0,
                                                 int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {  // arrives on the barrier of BOTH CTAs
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(smem_u32(bar)), "h"((uint16_t)3)
        : "memory");
}

template <int kBlockN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                              const float* __restrict__ bias, long long bias_batch_stride, void* __restrict__ c,
                              int out_bf16, int m, int n, int k, int batch_count, int b_shared, int relu) {
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kBytesBHalf = (kBlockN / 2) * kGemmBlockK * 2;
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    constexpr int kPairM = 2 * kGemmBlockM;
    using Smem = PairSmem<kBlockN>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    uint32_t rank;
    rank = blockIdx.x & 1;  // == %cluster_ctarank for the (2, 1, 1) cluster, and provably warp-uniform
    const bool leader = rank == 0;
    const int num_kb = (k + kGemmBlockK - 1) / kGemmBlockK;
    const int m_tiles = (m + kPairM - 1) / kPairM;
    const int n_tiles = (n + kBlockN - 1) / kBlockN;
    const long long num_tiles = (long long)m_tiles * n_tiles * batch_count;
    const long long first_tile = blockIdx.x >> 1, tile_step = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < kPairStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], 256);  // the epilogue threads of both CTAs (used on the leader only)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // collective over the pair: the same warp of both CTAs allocates the accumulator columns
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer, both CTAs: my half of A and my half of B
            uint32_t it = 0;
            for (long long tile = first_tile; tile < num_tiles; tile += tile_step) {
                const int mt = (int)(tile % m_tiles), nt = (int)((tile / m_tiles) % n_tiles);
                const int batch = (int)(tile / ((long long)m_tiles * n_tiles));
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kPairStages;
                    const uint32_t phase = (it / kPairStages) & 1;
                    mbar_wait(&sm.empty[s], phase ^ 1);
                    if (leader) mbar_expect_tx(&sm.full[s], 2 * (kBytesA + kBytesBHalf));
                    tma_load_3d_pair(sm.a[s], &map_a, &sm.full[s], kb * kGemmBlockK,
                                     mt * kPairM + (int)rank * kGemmBlockM, batch);
                    tma_load_3d_pair(sm.b[s], &map_b, &sm.full[s], kb * kGemmBlockK,
                                     nt * kBlockN + (int)rank * (kBlockN / 2), b_shared ? 0 : batch);
                }
            }
        }
    } else if (warp == 1) {
        if (!leader) {
            // the peer CTA's tensor core is driven by the leader's cta_group::2 instructions
        } else if (elect_one()) {  // ---- MMA issuer: the leader CTA only
            // instruction descriptor as in the single-CTA kernel but M = 256
            uint32_t idesc = 0;
            idesc |= 1u << 4;
            idesc |= 1u << 7;
            idesc |= 1u << 10;
            idesc |= (uint32_t)(kBlockN >> 3) << 17;
            idesc |= (uint32_t)(kPairM >> 4) << 24;
            uint32_t it = 0, local_tile = 0;
            for (long long tile = first_tile; tile < num_tiles; tile += tile_step, ++local_tile) {
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kPairStages;
                    const uint32_t phase = (it / kPairStages) & 1;
                    mbar_wait(&sm.full[s], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                    for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                        umma_bf16_pair(tmem_acc, umma_smem_desc(a_addr + kk * kGemmUmmaK * 2),
                                       umma_smem_desc(b_addr + kk * kGemmUmmaK * 2), idesc, (kb | kk) != 0);
                    }
                    umma_commit_pair(&sm.empty[s]);
                }
                umma_commit_pair(&sm.tmem_full[acc]);
            }
        }
    } else {
        // ---- epilogue, both CTAs: my 128 rows of the pair tile
        const int quarter = warp & 3;
        uint32_t leader_empty[2];
#pragma unroll
        for (int a = 0; a < 2; ++a)
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(leader_empty[a]) : "r"(smem_u32(&sm.tmem_empty[a])), "r"(0));
        uint32_t local_tile = 0;
        for (long long tile = first_tile; tile < num_tiles; tile += tile_step, ++local_tile) {
            const int mt = (int)(tile % m_tiles), nt = (int)((tile / m_tiles) % n_tiles);
            const int batch = (int)(tile / ((long long)m_tiles * n_tiles));
            const int m0 = mt * kPairM + (int)rank * kGemmBlockM, n0 = nt * kBlockN;
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            const int row = m0 + quarter * 32 + lane;
            mbar_wait(&sm.tmem_full[acc], use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_acc = tmem_base + acc * kBlockN;
            const float* bias_b = bias != nullptr ? bias + (long long)batch * bias_batch_stride : nullptr;
#pragma unroll 1
            for (int chunk = 0; chunk < kBlockN / 32; ++chunk) {
                if (n0 + chunk * 32 >= n) break;
                uint32_t r[32];
                const uint32_t taddr = tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                      "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
                      "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
                      "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
                      "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                    : "r"(taddr)
                    : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row < m) {
                    const int col0 = n0 + chunk * 32;
                    const long long out_off = ((long long)batch * m + row) * n + col0;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        v[j] = __uint_as_float(r[j]);
                        if (bias_b != nullptr && col0 + j < n) v[j] += bias_b[col0 + j];
                        if (relu) v[j] = fmaxf(v[j], 0.f);
                    }
                    const bool full = col0 + 32 <= n;
                    if (out_bf16) {
                        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(c) + out_off;
                        if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint4 pk;
                                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                                for (int e = 0; e < 4; ++e)
                                    h[e] = __floats2bfloat162_rn(v[8 * q + 2 * e], v[8 * q + 2 * e + 1]);
                                reinterpret_cast<uint4*>(dst)[q] = pk;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < n) dst[j] = __float2bfloat16(v[j]);
                        }
                    } else {
                        float* dst = reinterpret_cast<float*>(c) + out_off;
                        if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                reinterpret_cast<float4*>(dst)[q] =
                                    make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < n) dst[j] = v[j];
                        }
                    }
                }
            }
            // hand the accumulator stage back to the leader's MMA warp (remote arrive from the peer CTA)
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(leader_empty[acc]) : "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

}  // namespace dls

using namespace dls;

extern "C" int dls_gemm_bf16_tn(const void* a, const void* b, const float* bias, void* c, int out_dtype, int m, int n,
                                int k, int batch, int b_shared, int relu, void* stream) {
    if (m < 1 || n < 1 || k < 8 || batch < 1) return fail_msg("dls_gemm_bf16_tn: empty problem");
    if (k % 8) return fail_msg("dls_gemm_bf16_tn: K must be a multiple of 8 (16-byte rows for TMA)");
    if (((uintptr_t)a | (uintptr_t)b) & 15) return fail_msg("dls_gemm_bf16_tn: operands must be 16-byte aligned");
    if (out_dtype != 0 && out_dtype != 1) return fail_msg("dls_gemm_bf16_tn: out_dtype must be 0 (fp32) or 1 (bf16)");
    CUtensorMap map_a, map_b;
    if (int rc = make_map(&map_a, a, m, k, batch, kGemmBlockM)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    {
        // CTA pairs (cta_group::2) for problems with at least one wave of 256 x 256 (or 256 x 128) tiles
        const char* pair_env = getenv("DLS_GEMM_2CTA");
        const int pair_n = n > 128 ? 256 : 128;
        const long long pair_tiles = (long long)((m + 255) / 256) * ((n + pair_n - 1) / pair_n) * batch;
        const bool want_pair = pair_env != nullptr ? pair_env[0] != '0' : (m >= 512 && n >= 128 && pair_tiles >= kNumSMs / 2);
        if (want_pair && m >= 256 && n >= 64) {
            if (int rc = make_map(&map_b, b, n, k, b_shared ? 1 : batch, pair_n / 2)) return rc;
            const long long clusters = pair_tiles < kNumSMs / 2 ? pair_tiles : kNumSMs / 2;
            const long long bias_stride_p = b_shared ? 0 : (long long)n;
#define DLS_GEMM_PAIR_LAUNCH(BN)                                                                               \
    do {                                                                                                       \
        const size_t smem = sizeof(PairSmem<BN>) + 1024;                                                       \
        static thread_local bool configured = false;                                                           \
        if (!configured) {                                                                                     \
            DLS_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_pair_kernel<BN>,                                  \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
            configured = true;                                                                                 \
        }                                                                                                      \
        cudaLaunchConfig_t cfg = {};                                                                           \
        cfg.gridDim = dim3((unsigned)(clusters * 2));                                                          \
        cfg.blockDim = dim3(kGemmThreads);                                                                     \
        cfg.dynamicSmemBytes = smem;                                                                           \
        cfg.stream = s;                                                                                        \
        cudaLaunchAttribute attr[1];                                                                           \
        attr[0].id = cudaLaunchAttributeClusterDimension;                                                      \
        attr[0].val.clusterDim.x = 2;                                                                          \
        attr[0].val.clusterDim.y = 1;                                                                          \
        attr[0].val.clusterDim.z = 1;                                                                          \
        cfg.attrs = attr;                                                                                      \
        cfg.numAttrs = 1;                                                                                      \
        DLS_CHECK(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_pair_kernel<BN>, map_a, map_b, bias,              \
                                     bias_stride_p, c, out_dtype, m, n, k, batch, b_shared, relu));            \
    } while (0)
            if (pair_n == 256) DLS_GEMM_PAIR_LAUNCH(256);
            else DLS_GEMM_PAIR_LAUNCH(128);
#undef DLS_GEMM_PAIR_LAUNCH
            DLS_LAUNCHED("gemm_bf16_tcgen05_pair_kernel");
            return 0;
        }
    }
    // tile width: narrow outputs keep a narrow tile; wide ones take 256 columns when that still
    // leaves at least one wave of CTAs
    const long long m_tiles = (m + kGemmBlockM - 1) / kGemmBlockM;
    int block_n = 64;
    if (n > 128 && m_tiles * ((n + 255) / 256) * batch >= kNumSMs) block_n = 256;
    else if (n > 64) block_n = 128;
    static const int forced = [] { const char* e = getenv("DLS_GEMM_BN"); return e ? atoi(e) : 0; }();  // tuning knob
    if ((forced == 64 || forced == 128 || forced == 256) && n > forced / 2) block_n = forced;
    if (int rc = make_map(&map_b, b, n, k, b_shared ? 1 : batch, block_n)) return rc;
    const long long num_tiles = m_tiles * ((n + block_n - 1) / block_n) * batch;
    dim3 grid((unsigned)(num_tiles < kNumSMs ? num_tiles : kNumSMs));  // persistent: one CTA per SM
    const long long bias_stride = b_shared ? 0 : (long long)n;
#define DLS_GEMM_LAUNCH(BN)                                                                                   \
    do {                                                                                                      \
        const size_t smem = sizeof(GemmSmem<BN>) + 1024;                                                      \
        static thread_local bool configured = false;                                                          \
        if (!configured) {                                                                                    \
            DLS_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           (int)smem));                                                       \
            configured = true;                                                                                \
        }                                                                                                     \
        gemm_bf16_tcgen05_kernel<BN><<<grid, kGemmThreads, smem, s>>>(map_a, map_b, bias, bias_stride, c, out_dtype, \
                                                                      m, n, k, batch, b_shared, relu);        \
    } while (0)
    if (block_n == 256) DLS_GEMM_LAUNCH(256);
    else if (block_n == 128) DLS_GEMM_LAUNCH(128);
    else DLS_GEMM_LAUNCH(64);
#undef DLS_GEMM_LAUNCH
    DLS_LAUNCHED("gemm_bf16_tcgen05_kernel");
    return 0;
}
