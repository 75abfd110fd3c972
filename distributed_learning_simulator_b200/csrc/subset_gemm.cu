// Coalition ("subset") GEMM for the Shapley engines: the coalition model is never materialised.
//
//   C[s] = act( A[s] * W_s^T + b_s ),   W_s = sum_i r[s][i] * W_i,   b_s = sum_i r[s][i] * b_i
//
// W_i / b_i are layer slices of the N workers' uploaded fp32 parameter arenas (one base pointer per
// worker plus the layer's offset), r[s][i] the FedAvg weight of worker i inside coalition s (zero
// for non-members).  The reference builds W_s on the CPU for one coalition at a time, loads it
// into the tester and runs a whole eager inference
// (/root/reference/simulation_lib/method/shapley_value/shapley_value_algorithm.py:67-76,
// algorithm/fed_avg_algorithm.py:83-97).  Here the construction is the B-operand producer of the
// tcgen05 GEMM itself:
//
//   warp 0      TMA producer of the activation tiles A (128 x 64 bf16, 128B swizzle)
//   warp 1      TMEM allocation + tcgen05.mma issue (M128, N64/128, K16; fp32 accumulators in TMEM)
//   warps 2-5   epilogue: per-tile coalition bias built in shared memory, tcgen05.ld -> bias / ReLU ->
//               128-bit stores
//   warps 6-9   B builders: read the workers' fp32 weight rows (128-bit loads, L2 resident), form
//               the weighted sum in registers, round to bf16 and write the tile into the SAME
//               128B-swizzled shared-memory slot layout a TMA load would have produced, then
//               fence.proxy.async + mbarrier arrive
//
// Cluster variant (kCluster = 2): two CTAs that work on neighbouring M tiles of the same coalition and
// N tile need the SAME W_s tile.  Each builds half of its rows and writes them into BOTH CTAs'
// shared-memory slots (st.shared::cluster through DSMEM), arrives on both CTAs' full barriers, and the
// MMA warps release a slot with a multicast tcgen05.commit to both CTAs - the workers' base tiles are
// read and combined once per cluster instead of once per CTA.
//
// Two row mappings: "batched" (batch index = coalition; every coalition has its own activations)
// and "stacked" (one GEMM whose N dimension stacks the rows of all coalitions - the first layer,
// where all coalitions see the same input patches).
#include <stdlib.h>

#include "tcgen05.cuh"
#include "../../include/dls_b200.h"

namespace dls {

constexpr int kSubsetThreads = 320;
constexpr int kSubsetBuilders = 128;
constexpr int kSubsetMaxWorkers = 64;

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait with cluster-scope acquire: the data behind the barrier was written by the peer CTA
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spins = 0;; ++spins) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (spins > 20000000u) asm volatile("trap;");
    }
}
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t cta_mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(smem_u32(bar)), "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// One 64-wide K block of the coalition weight tile: rows [rank * kBlockN / kCluster, +kBlockN / kCluster) of
// the tile, written into `slot` (128B-swizzle layout) of this CTA and of every peer CTA of the cluster.
template <int kBlockN, int kCluster>
__device__ __forceinline__ void build_b_block(unsigned char* slot, uint32_t rank, int bt, int n0, int k0, int batch,
                                              const float* const* wptr, int n_workers, long long w_offset,
                                              const float* __restrict__ coefs, int n_layer, int stacked, int n_total,
                                              int k_w, int ptrs_aligned) {
    constexpr int kRowsMine = kBlockN / kCluster;
#pragma unroll 2
    for (int chunk = bt; chunk < kRowsMine * 8; chunk += kSubsetBuilders) {
        const int r = (int)rank * kRowsMine + (chunk >> 3), cc = chunk & 7;
        const int j = n0 + r;
        const int kpos = k0 + cc * 8;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        int valid = k_w - kpos;
        valid = valid > 8 ? 8 : valid;
        if (j < n_total && valid > 0) {
            const int sidx = stacked ? j / n_layer : batch;
            const int row = stacked ? j - sidx * n_layer : j;
            const float* coef = coefs + (long long)sidx * n_workers;
            const long long off = w_offset + (long long)row * k_w + kpos;
            const bool vec = ptrs_aligned && valid == 8 && (off & 3) == 0;
            for (int i = 0; i < n_workers; ++i) {
                const float ri = __ldg(coef + i);
                if (ri == 0.f) continue;  // not a member of this coalition: no traffic at all
                const float* src = wptr[i] + off;
                if (vec) {
                    const float4 x = __ldg(reinterpret_cast<const float4*>(src));
                    const float4 y = __ldg(reinterpret_cast<const float4*>(src) + 1);
                    acc[0] = fmaf(ri, x.x, acc[0]); acc[1] = fmaf(ri, x.y, acc[1]);
                    acc[2] = fmaf(ri, x.z, acc[2]); acc[3] = fmaf(ri, x.w, acc[3]);
                    acc[4] = fmaf(ri, y.x, acc[4]); acc[5] = fmaf(ri, y.y, acc[5]);
                    acc[6] = fmaf(ri, y.z, acc[6]); acc[7] = fmaf(ri, y.w, acc[7]);
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e)
                        if (e < valid) acc[e] = fmaf(ri, __ldg(src + e), acc[e]);
                }
            }
        }
        uint4 pk;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(acc[2 * e], acc[2 * e + 1]);
        // Swizzle<3,4,3>: 16-byte chunk index XOR (row mod 8) inside each 128-byte row
        unsigned char* dst = slot + r * 128 + ((cc ^ (r & 7)) << 4);
        *reinterpret_cast<uint4*>(dst) = pk;
        if (kCluster > 1) {
#pragma unroll
            for (int peer = 0; peer < kCluster; ++peer)
                if (peer != (int)rank) st_cluster_v4(map_to_cta(smem_u32(dst), peer), pk);
        }
    }
}

// Epilogue of one output tile: coalition bias for the tile's columns (built in shared memory while the
// MMAs run), then TMEM -> registers -> bias / ReLU -> global.
template <int kBlockN>
__device__ __forceinline__ void epilogue_tile(float* bias_smem, uint64_t* tmem_full_bar, uint32_t use, uint32_t tmem_base,
                                              uint32_t acc, int quarter, int lane, int et, int m0, int n0, int batch,
                                              const float* const* wptr, int n_workers, long long bias_offset,
                                              const float* __restrict__ coefs, int n_layer, int stacked, int n_total,
                                              int m, void* __restrict__ c, int out_bf16, int relu) {
    // coalition bias of this tile's columns, built while the MMAs run
    for (int col = et; col < kBlockN; col += 128) {
        const int j = n0 + col;
        float b = 0.f;
        if (bias_offset >= 0 && j < n_total) {
            const int sidx = stacked ? j / n_layer : batch;
            const int row = stacked ? j - sidx * n_layer : j;
            const float* coef = coefs + (long long)sidx * n_workers;
            for (int i = 0; i < n_workers; ++i) {
                const float ri = __ldg(coef + i);
                if (ri != 0.f) b = fmaf(ri, __ldg(wptr[i] + bias_offset + row), b);
            }
        }
        bias_smem[col] = b;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
    const int row = m0 + quarter * 32 + lane;
    mbar_wait(tmem_full_bar, use & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_acc = tmem_base + acc * kBlockN;
#pragma unroll 1
    for (int chunk = 0; chunk < kBlockN / 32; ++chunk) {
        if (n0 + chunk * 32 >= n_total) break;
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
            const long long out_off = ((long long)batch * m + row) * n_total + col0;
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                v[j] = __uint_as_float(r[j]) + bias_smem[chunk * 32 + j];
                if (relu) v[j] = fmaxf(v[j], 0.f);
            }
            const bool full = col0 + 32 <= n_total;
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
                        if (col0 + j < n_total) dst[j] = __float2bfloat16(v[j]);
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
                        if (col0 + j < n_total) dst[j] = v[j];
                }
            }
        }
    }
}

template <int kBlockN>
struct __align__(1024) SubsetSmem {
    __nv_bfloat16 a[kGemmStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kGemmStages][kBlockN * kGemmBlockK];
    float bias[2][kBlockN];
    const float* wptr[kSubsetMaxWorkers];
    uint64_t full_a[kGemmStages];
    uint64_t full_b[kGemmStages];
    uint64_t empty[kGemmStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
};

template <int kBlockN, int kCluster>
__global__ void __launch_bounds__(kSubsetThreads, 1)
subset_gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, int a_shared,
                           const float* const* __restrict__ worker_ptrs, int n_workers, long long w_offset,
                           long long bias_offset, const float* __restrict__ coefs, int n_layer, int stacked,
                           void* __restrict__ c, int out_bf16, int m, int n_total, int k_a, int k_w,
                           int batch_count, int relu, int ptrs_aligned) {
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    using Smem = SubsetSmem<kBlockN>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int num_kb = (k_a + kGemmBlockK - 1) / kGemmBlockK;
    const int m_tiles = (m + kGemmBlockM - 1) / kGemmBlockM;
    const int n_tiles = (n_total + kBlockN - 1) / kBlockN;
    // The schedule walks "M groups" of kCluster neighbouring M tiles; CTA rank r of a cluster takes the
    // r-th tile of the group (a tile past the last row is computed on zero-filled A and stores nothing).
    const int m_groups = (m_tiles + kCluster - 1) / kCluster;
    const long long num_tiles = (long long)m_groups * n_tiles * batch_count;
    const uint32_t rank = kCluster > 1 ? cluster_ctarank() : 0;
    const long long first_tile = blockIdx.x / kCluster;
    const long long tile_step = gridDim.x / kCluster;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        for (int s = 0; s < kGemmStages; ++s) {
            mbar_init(&sm.full_a[s], 1);
            mbar_init(&sm.full_b[s], kCluster * kSubsetBuilders / 32);  // one arrive per builder warp of the cluster
            mbar_init(&sm.empty[s], kCluster);                            // every CTA's MMA warp releases the slot
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < n_workers; i += kSubsetThreads) sm.wptr[i] = worker_ptrs[i];
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (kCluster > 1) cluster_sync_all();  // the peer's barriers are initialised before anyone signals them
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer (activations only)
            uint32_t it = 0;
            for (long long tile = first_tile; tile < num_tiles; tile += tile_step) {
                const int mt = (int)(tile % m_groups) * kCluster + (int)rank;
                const int batch = (int)(tile / ((long long)m_groups * n_tiles));
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t phase = (it / kGemmStages) & 1;
                    mbar_wait(&sm.empty[s], phase ^ 1);
                    mbar_expect_tx(&sm.full_a[s], kBytesA);
                    tma_load_3d(sm.a[s], &map_a, &sm.full_a[s], kb * kGemmBlockK, mt * kGemmBlockM,
                                a_shared ? 0 : batch);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = umma_instr_desc<kBlockN>();
            uint32_t it = 0, local_tile = 0;
            for (long long tile = first_tile; tile < num_tiles; tile += tile_step, ++local_tile) {
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kGemmStages;
                    const uint32_t phase = (it / kGemmStages) & 1;
                    mbar_wait(&sm.full_a[s], phase);
                    if (kCluster > 1) mbar_wait_cluster(&sm.full_b[s], phase);
                    else mbar_wait(&sm.full_b[s], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                    for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                        umma_bf16(tmem_acc, umma_smem_desc(a_addr + kk * kGemmUmmaK * 2),
                                  umma_smem_desc(b_addr + kk * kGemmUmmaK * 2), idesc, (kb | kk) != 0);
                    }
                    if (kCluster > 1) umma_commit_multicast(&sm.empty[s], (uint16_t)((1u << kCluster) - 1));
                    else umma_commit(&sm.empty[s]);
                }
                umma_commit(&sm.tmem_full[acc]);
            }
        }
    } else if (warp >= 6) {
        // ---- B builders: W_s tile = sum_i r[s][i] * W_i tile, written in the 128B-swizzle slot layout
        const int bt = threadIdx.x - 6 * 32;  // 0 .. 127
        uint32_t it = 0;
        for (long long tile = first_tile; tile < num_tiles; tile += tile_step) {
            const int nt = (int)((tile / m_groups) % n_tiles);
            const int batch = (int)(tile / ((long long)m_groups * n_tiles));
            const int n0 = nt * kBlockN;
            for (int kb = 0; kb < num_kb; ++kb, ++it) {
                const int s = it % kGemmStages;
                const uint32_t phase = (it / kGemmStages) & 1;
                mbar_wait(&sm.empty[s], phase ^ 1);
                build_b_block<kBlockN, kCluster>(reinterpret_cast<unsigned char*>(sm.b[s]), rank, bt, n0,
                                                 kb * kGemmBlockK, batch, sm.wptr, n_workers, w_offset, coefs,
                                                 n_layer, stacked, n_total, k_w, ptrs_aligned);
                if (kCluster > 1) {
                    // generic-proxy stores (local and DSMEM) -> visible to the tensor cores of both CTAs
                    asm volatile("fence.proxy.async;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) {
#pragma unroll
                        for (int peer = 0; peer < kCluster; ++peer)
                            mbar_arrive_cluster(map_to_cta(smem_u32(&sm.full_b[s]), peer));
                    }
                } else {
                    fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's reads
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.full_b[s]);
                }
            }
        }
    } else {
        // ---- epilogue warps 2-5: TMEM lane quarter = warp % 4
        const int quarter = warp & 3;
        const int et = (warp - 2) * 32 + lane;  // 0 .. 127
        uint32_t local_tile = 0;
        for (long long tile = first_tile; tile < num_tiles; tile += tile_step, ++local_tile) {
            const int mt = (int)(tile % m_groups) * kCluster + (int)rank, nt = (int)((tile / m_groups) % n_tiles);
            const int batch = (int)(tile / ((long long)m_groups * n_tiles));
            const int m0 = mt * kGemmBlockM, n0 = nt * kBlockN;
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            epilogue_tile<kBlockN>(sm.bias[acc], &sm.tmem_full[acc], use, tmem_base, acc, quarter, lane, et, m0, n0,
                                   batch, sm.wptr, n_workers, bias_offset, coefs, n_layer, stacked, n_total, m, c,
                                   out_bf16, relu);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&sm.tmem_empty[acc]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (kCluster > 1) cluster_sync_all();  // no CTA leaves while its peer may still write or signal into it
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}


// ------------------------------------------------------------------------------------------------------
// Weight-stationary variant (K <= 512): the built W_s tile stays resident in shared memory while the CTA
// (pair) sweeps a run of `m_run` neighbouring M groups of the same coalition and N tile, so the workers'
// slices are read and combined once per run instead of once per output tile.  The ring only carries A.
constexpr int kStationaryMaxKb = 8;

template <int kBlockN>
struct __align__(1024) StationarySmem {
    __nv_bfloat16 a[kGemmStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kStationaryMaxKb][kBlockN * kGemmBlockK];
    float bias[2][kBlockN];
    const float* wptr[kSubsetMaxWorkers];
    uint64_t full_a[kGemmStages];
    uint64_t empty[kGemmStages];
    uint64_t b_ready[kStationaryMaxKb];  // K block kb of the resident tile is built (all CTAs of the cluster)
    uint64_t b_free;                     // every CTA's MMAs of the previous run have finished reading it
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
};

template <int kBlockN, int kCluster>
__global__ void __launch_bounds__(kSubsetThreads, 1)
subset_gemm_stationary_kernel(const __grid_constant__ CUtensorMap map_a, int a_shared,
                              const float* const* __restrict__ worker_ptrs, int n_workers, long long w_offset,
                              long long bias_offset, const float* __restrict__ coefs, int n_layer, int stacked,
                              void* __restrict__ c, int out_bf16, int m, int n_total, int k_a, int k_w,
                              int batch_count, int relu, int ptrs_aligned, int m_run) {
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    using Smem = StationarySmem<kBlockN>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int num_kb = (k_a + kGemmBlockK - 1) / kGemmBlockK;  // <= kStationaryMaxKb (host checks)
    const int m_tiles = (m + kGemmBlockM - 1) / kGemmBlockM;
    const int n_tiles = (n_total + kBlockN - 1) / kBlockN;
    const int m_groups = (m_tiles + kCluster - 1) / kCluster;
    const int m_runs = (m_groups + m_run - 1) / m_run;
    const long long num_units = (long long)m_runs * n_tiles * batch_count;  // unit = one resident W_s tile
    const uint32_t rank = kCluster > 1 ? cluster_ctarank() : 0;
    const long long first_unit = blockIdx.x / kCluster;
    const long long unit_step = gridDim.x / kCluster;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        for (int s = 0; s < kGemmStages; ++s) {
            mbar_init(&sm.full_a[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int kb = 0; kb < kStationaryMaxKb; ++kb) mbar_init(&sm.b_ready[kb], kCluster * kSubsetBuilders / 32);
        mbar_init(&sm.b_free, kCluster);
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < n_workers; i += kSubsetThreads) sm.wptr[i] = worker_ptrs[i];
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (kCluster > 1) cluster_sync_all();
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);

    // unit -> (first M group, number of M groups, N tile, coalition)
#define DLS_UNIT_DECODE(unit)                                                          \
    const int run_idx = (int)((unit) % m_runs);                                        \
    const int nt = (int)(((unit) / m_runs) % n_tiles);                                 \
    const int batch = (int)((unit) / ((long long)m_runs * n_tiles));                   \
    const int mg0 = run_idx * m_run;                                                   \
    const int mg1 = mg0 + m_run < m_groups ? mg0 + m_run : m_groups;

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer (activations)
            uint32_t it = 0;
            for (long long unit = first_unit; unit < num_units; unit += unit_step) {
                DLS_UNIT_DECODE(unit)
                (void)nt;
                for (int mg = mg0; mg < mg1; ++mg) {
                    const int mt = mg * kCluster + (int)rank;
                    for (int kb = 0; kb < num_kb; ++kb, ++it) {
                        const int s = it % kGemmStages;
                        const uint32_t phase = (it / kGemmStages) & 1;
                        mbar_wait(&sm.empty[s], phase ^ 1);
                        mbar_expect_tx(&sm.full_a[s], kBytesA);
                        tma_load_3d(sm.a[s], &map_a, &sm.full_a[s], kb * kGemmBlockK, mt * kGemmBlockM,
                                    a_shared ? 0 : batch);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = umma_instr_desc<kBlockN>();
            uint32_t it = 0, local_tile = 0, local_unit = 0;
            for (long long unit = first_unit; unit < num_units; unit += unit_step, ++local_unit) {
                DLS_UNIT_DECODE(unit)
                (void)nt; (void)batch;
                for (int mg = mg0; mg < mg1; ++mg, ++local_tile) {
                    const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                    mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                    for (int kb = 0; kb < num_kb; ++kb, ++it) {
                        const int s = it % kGemmStages;
                        const uint32_t phase = (it / kGemmStages) & 1;
                        mbar_wait(&sm.full_a[s], phase);
                        if (mg == mg0) {  // first tile of the run: the resident block may still be under construction
                            if (kCluster > 1) mbar_wait_cluster(&sm.b_ready[kb], local_unit & 1);
                            else mbar_wait(&sm.b_ready[kb], local_unit & 1);
                        }
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[kb]);
#pragma unroll
                        for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                            umma_bf16(tmem_acc, umma_smem_desc(a_addr + kk * kGemmUmmaK * 2),
                                      umma_smem_desc(b_addr + kk * kGemmUmmaK * 2), idesc, (kb | kk) != 0);
                        }
                        umma_commit(&sm.empty[s]);
                    }
                    umma_commit(&sm.tmem_full[acc]);
                }
                // the resident tile may be rebuilt once every CTA of the cluster has finished reading it
                if (kCluster > 1) umma_commit_multicast(&sm.b_free, (uint16_t)((1u << kCluster) - 1));
                else umma_commit(&sm.b_free);
            }
        }
    } else if (warp >= 6) {
        // ---- B builders: one resident W_s tile (all K blocks) per unit
        const int bt = threadIdx.x - 6 * 32;
        uint32_t local_unit = 0;
        for (long long unit = first_unit; unit < num_units; unit += unit_step, ++local_unit) {
            DLS_UNIT_DECODE(unit)
            (void)mg1;
            mbar_wait(&sm.b_free, (local_unit & 1) ^ 1);
            for (int kb = 0; kb < num_kb; ++kb) {
                build_b_block<kBlockN, kCluster>(reinterpret_cast<unsigned char*>(sm.b[kb]), rank, bt, nt * kBlockN,
                                                 kb * kGemmBlockK, batch, sm.wptr, n_workers, w_offset, coefs,
                                                 n_layer, stacked, n_total, k_w, ptrs_aligned);
                if (kCluster > 1) {
                    asm volatile("fence.proxy.async;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) {
#pragma unroll
                        for (int peer = 0; peer < kCluster; ++peer)
                            mbar_arrive_cluster(map_to_cta(smem_u32(&sm.b_ready[kb]), peer));
                    }
                } else {
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.b_ready[kb]);
                }
            }
        }
    } else {
        // ---- epilogue warps 2-5
        const int quarter = warp & 3;
        const int et = (warp - 2) * 32 + lane;
        uint32_t local_tile = 0;
        for (long long unit = first_unit; unit < num_units; unit += unit_step) {
            DLS_UNIT_DECODE(unit)
            for (int mg = mg0; mg < mg1; ++mg, ++local_tile) {
                const int mt = mg * kCluster + (int)rank;
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                epilogue_tile<kBlockN>(sm.bias[acc], &sm.tmem_full[acc], use, tmem_base, acc, quarter, lane, et,
                                       mt * kGemmBlockM, nt * kBlockN, batch, sm.wptr, n_workers, bias_offset, coefs,
                                       n_layer, stacked, n_total, m, c, out_bf16, relu);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                mbar_arrive(&sm.tmem_empty[acc]);
            }
        }
    }
#undef DLS_UNIT_DECODE
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (kCluster > 1) cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

}  // namespace dls

using namespace dls;

extern "C" int dls_subset_gemm_bf16(const void* a, int a_shared, const float* const* worker_ptrs, int n_workers,
                                    long long w_offset, long long bias_offset, const float* coefs, int n_sets,
                                    int n_layer, int stacked, void* c, int out_dtype, int m, int k_a, int k_w,
                                    int relu, int ptrs_aligned, void* stream) {
    if (m < 1 || n_layer < 1 || n_sets < 1 || k_w < 1) return fail_msg("dls_subset_gemm_bf16: empty problem");
    if (k_a % 8 || k_a < k_w) return fail_msg("dls_subset_gemm_bf16: A's row length must be a multiple of 8 and >= K");
    if ((uintptr_t)a & 15) return fail_msg("dls_subset_gemm_bf16: A must be 16-byte aligned");
    if (n_workers < 1 || n_workers > kSubsetMaxWorkers) return fail_msg("dls_subset_gemm_bf16: 1..64 workers");
    if (out_dtype != 0 && out_dtype != 1) return fail_msg("dls_subset_gemm_bf16: out_dtype must be 0 (fp32) or 1 (bf16)");
    const int batch_count = stacked ? 1 : n_sets;
    const long long n_total_ll = stacked ? (long long)n_sets * n_layer : n_layer;
    if (n_total_ll > 0x7fffffff) return fail_msg("dls_subset_gemm_bf16: N too large");
    const int n_total = (int)n_total_ll;
    CUtensorMap map_a;
    if (int rc = make_map(&map_a, a, m, k_a, (a_shared || stacked) ? 1 : batch_count, kGemmBlockM)) return rc;
    const long long m_tiles = (m + kGemmBlockM - 1) / kGemmBlockM;
    const int block_n = n_total > 64 ? 128 : 64;
    // pairs of neighbouring M tiles share one built W_s tile through a 2-CTA cluster
    const char* cluster_env = getenv("DLS_SUBSET_CLUSTER");  // "1" selects the single-CTA variant (A/B runs, tests)
    // a pair only helps when both halves of the tile hold real rows (narrow layers keep one CTA per tile)
    const bool pair_pays = m_tiles >= 2 && n_total > block_n / 2;
    const int cluster = (cluster_env != nullptr ? atoi(cluster_env) >= 2 && m_tiles >= 2 : pair_pays) ? 2 : 1;
    const long long m_groups = (m_tiles + cluster - 1) / cluster;
    const long long n_tiles = (n_total + block_n - 1) / block_n;
    const long long max_clusters = kNumSMs / cluster;
    cudaStream_t s = (cudaStream_t)stream;
    const int a_one = (a_shared || stacked) ? 1 : 0;
    // weight-stationary schedule when the whole K extent of a tile fits in shared memory: runs of M groups
    // reuse one built W_s tile; about two runs per cluster keep the tail short
    const char* stat_env = getenv("DLS_SUBSET_STATIONARY");
    const bool stationary = (k_a + kGemmBlockK - 1) / kGemmBlockK <= kStationaryMaxKb &&
                            !(stat_env != nullptr && stat_env[0] == '0');
    long long m_run = 1, num_units = m_groups * n_tiles * batch_count;
    if (stationary) {
        const long long pairs = n_tiles * batch_count;
        long long runs_per_pair = (2 * max_clusters + pairs - 1) / pairs;
        if (runs_per_pair < 1) runs_per_pair = 1;
        if (runs_per_pair > m_groups) runs_per_pair = m_groups;
        m_run = (m_groups + runs_per_pair - 1) / runs_per_pair;
        num_units = ((m_groups + m_run - 1) / m_run) * pairs;
    }
    const unsigned grid = (unsigned)((num_units < max_clusters ? num_units : max_clusters) * cluster);
#define DLS_SUBSET_LAUNCH(KERNEL, SMEM, BN, CL, ...)                                                           \
    do {                                                                                                       \
        const size_t smem = sizeof(SMEM<BN>) + 1024;                                                           \
        static thread_local bool configured = false;                                                           \
        if (!configured) {                                                                                     \
            DLS_CHECK(cudaFuncSetAttribute(KERNEL<BN, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                           (int)smem));                                                        \
            configured = true;                                                                                 \
        }                                                                                                      \
        cudaLaunchConfig_t cfg = {};                                                                           \
        cfg.gridDim = dim3(grid);                                                                              \
        cfg.blockDim = dim3(kSubsetThreads);                                                                   \
        cfg.dynamicSmemBytes = smem;                                                                           \
        cfg.stream = s;                                                                                        \
        cudaLaunchAttribute attr[1];                                                                           \
        attr[0].id = cudaLaunchAttributeClusterDimension;                                                      \
        attr[0].val.clusterDim.x = CL;                                                                         \
        attr[0].val.clusterDim.y = 1;                                                                          \
        attr[0].val.clusterDim.z = 1;                                                                          \
        cfg.attrs = attr;                                                                                      \
        cfg.numAttrs = 1;                                                                                      \
        DLS_CHECK(cudaLaunchKernelEx(&cfg, KERNEL<BN, CL>, map_a, a_one, worker_ptrs, n_workers, w_offset,     \
                                     bias_offset, coefs, n_layer, stacked, c, out_dtype, m, n_total, k_a, k_w, \
                                     batch_count, relu, ptrs_aligned, ##__VA_ARGS__));                         \
    } while (0)
    if (stationary) {
        const int run = (int)m_run;
        if (block_n == 128 && cluster == 2) DLS_SUBSET_LAUNCH(subset_gemm_stationary_kernel, StationarySmem, 128, 2, run);
        else if (block_n == 128) DLS_SUBSET_LAUNCH(subset_gemm_stationary_kernel, StationarySmem, 128, 1, run);
        else if (cluster == 2) DLS_SUBSET_LAUNCH(subset_gemm_stationary_kernel, StationarySmem, 64, 2, run);
        else DLS_SUBSET_LAUNCH(subset_gemm_stationary_kernel, StationarySmem, 64, 1, run);
    } else {
        if (block_n == 128 && cluster == 2) DLS_SUBSET_LAUNCH(subset_gemm_tcgen05_kernel, SubsetSmem, 128, 2);
        else if (block_n == 128) DLS_SUBSET_LAUNCH(subset_gemm_tcgen05_kernel, SubsetSmem, 128, 1);
        else if (cluster == 2) DLS_SUBSET_LAUNCH(subset_gemm_tcgen05_kernel, SubsetSmem, 64, 2);
        else DLS_SUBSET_LAUNCH(subset_gemm_tcgen05_kernel, SubsetSmem, 64, 1);
    }
#undef DLS_SUBSET_LAUNCH
    DLS_LAUNCHED("subset_gemm_tcgen05_kernel");
    return 0;
}
