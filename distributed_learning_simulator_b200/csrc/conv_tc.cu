// Convolutions of the training / evaluation hot path as implicit GEMMs on the 5th-generation tensor
// cores (NHWC bf16 activations, [Cout][R][S][Cin] bf16 filters = the channels-last segments of the
// parameter arena, fp32 accumulation in tensor memory).
//
// Reference counterpart: none of its own -- every convolution of every local step runs inside the
// external eager ``trainer.train()`` (/root/reference/simulation_lib/worker/worker.py:86-88) and of
// every server evaluation inside ``tester.inference()`` (simulation_lib/server/server.py:40-55),
// i.e. cuDNN.  Here they are three hand-written sm_100a kernels with no library on the path:
//
//   conv_igemm_kernel<BN, false>   fprop   Y[p][k]  = sum_{tap,c} X[p + tap][c] * W[k][tap][c]
//   conv_igemm_kernel<BN, true>    dgrad   dX[p][c] = sum_{tap,k} dY[p - tap][k] * W[k][tap][c]
//   conv_wgrad_kernel<BN>          wgrad   dW[k][tap][c] = sum_p dY[p][k] * X[p + tap][c]
//
// No im2col buffer exists: the A operand of a K-block is ONE TMA box load per filter tap -- a
// 5-d tensor map over the activation ({C*s, W/s, s, H/s, N}: the stride-s parity view, s = 1 or 2)
// whose box {64 channels, bw, 1, bh, bn} covers the 128 output pixels of the tile shifted by the
// tap; padding is the TMA unit's out-of-bounds zero fill (negative coordinates included).  Filters
// come through a 3-d map {Cin, taps, Cout}.
//   * fprop reads filters K-major (rows = output channels), the layout of the plain GEMM;
//   * dgrad reads the SAME filter memory as an MN-major B operand (rows = reduction index k,
//     64 contiguous input channels per row): no transposed filter copy is ever made;
//   * stride-2 dgrad is decomposed into the four output-parity classes, each a stride-1 problem
//     over a subset of the taps, scheduled as extra tiles of the same launch;
//   * wgrad reduces over pixels, so BOTH operands are MN-major tiles ([128 pixels][64 channels],
//     exactly what the pixel-box load produces); the pixel range is split over CTAs and partial
//     tiles are combined with red.global.add.v4.f32 when split > 1.
// Pipeline per CTA as in gemm_tcgen05.cu: warp 0 = TMA producer over an mbarrier ring, warp 1 =
// tcgen05.mma issuer, warps 2-5 = epilogue (tcgen05.ld -> bias / BatchNorm statistics -> stores),
// persistent tile loop, double-buffered TMEM accumulator.
//
// TMA loads here read only this GPU's memory and use no multicast.
#include <stdlib.h>

#include "tcgen05.cuh"
#include "../../include/dls_b200.h"

namespace dls {

constexpr int kConvMaxTaps = 36;
constexpr int kConvMaxClasses = 4;
constexpr int kConvThreads = 192;
// fprop / dgrad kernels: 8 epilogue warps (two per tensor-memory lane quarter, each taking every other 32-column
// chunk) -- a lone warp per scheduler cannot hide the latencies of the convert / store / statistics chain
constexpr int kConvWideThreads = 320;

struct ConvTap {
    int16_t dh, dw;     // shift of the box origin in the (parity-view) h / w coordinates
    uint8_t ph, pw;     // parity plane of the strided view (0 for stride 1)
    uint16_t widx;      // filter tap index r * S + s
};

struct ConvGemmParams {
    int pn, ph, pw;               // pixel space of one class (tile space)
    int bw, bh, bn;               // box: bw * bh * bn == 128 pixels
    int tiles_w, tiles_h, tiles_n;
    int a_channels;               // reduction channels per tap
    int a_cview;                  // channel count of the A tensor (parity plane offset = pw * a_cview)
    int n_out;                    // GEMM N (output channels of this launch)
    int n_classes;
    int class_ntaps[kConvMaxClasses];
    int class_tap0[kConvMaxClasses];
    int class_oph[kConvMaxClasses];
    int class_opw[kConvMaxClasses];
    int out_h, out_w, out_pitch;  // output tensor [pn][out_h][out_w][out_pitch] (elements)
    int os;                       // output pixel = (n, h * os + oph, w * os + opw)
    ConvTap taps[kConvMaxTaps];
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], "
        "[%2];"
        ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

// MN-major operand tile, 128-byte swizzle: each K row holds 64 contiguous MN elements (128 bytes),
// 8 K rows form one 1024-byte swizzle atom (SBO = distance between 8-row groups along K),
// the next 64 MN elements start `lbo_bytes` further (LBO).
__device__ __forceinline__ uint64_t umma_smem_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    desc |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    desc |= (uint64_t)(1024 >> 4) << 32;
    desc |= (uint64_t)1 << 46;
    desc |= (uint64_t)2 << 61;
    return desc;
}

template <int kBlockN, bool kAMn, bool kBMn>
__device__ __forceinline__ uint32_t conv_instr_desc() {
    uint32_t d = umma_instr_desc<kBlockN>();
    if (kAMn) d |= 1u << 15;
    if (kBMn) d |= 1u << 16;
    return d;
}

__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// column sums over the warp's 32 rows: butterfly that halves the value count per step, lane l ends with the
// totals of column l
__device__ __forceinline__ void conv_chunk_stats(const float (&v)[32], bool row_ok, int col0, int n_out,
                                                 float* sum_dst, float* sq_dst, int lane) {
    float s1[32], s2[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const float x = row_ok ? v[j] : 0.f;
        s1[j] = x;
        s2[j] = x * x;
    }
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        const bool upper = (lane & step) != 0;
#pragma unroll
        for (int j = 0; j < step; ++j) {
            // keep columns [j + (upper ? step : 0)], send the other half to the partner lane
            const float keep1 = upper ? s1[j + step] : s1[j];
            const float send1 = upper ? s1[j] : s1[j + step];
            const float keep2 = upper ? s2[j + step] : s2[j];
            const float send2 = upper ? s2[j] : s2[j + step];
            s1[j] = keep1 + __shfl_xor_sync(0xffffffffu, send1, step);
            s2[j] = keep2 + __shfl_xor_sync(0xffffffffu, send2, step);
        }
    }
    if (col0 + lane < n_out) {
        atomicAdd(sum_dst + col0 + lane, s1[0]);
        atomicAdd(sq_dst + col0 + lane, s2[0]);
    }
}

// Epilogue of one 32-column chunk of an output row (one thread = one pixel row of the tile): bias, rounding to
// the stored bf16, 128-bit stores, and -- when `stats` is given -- per-channel sum / sum of squares over the
// warp's 32 rows added to stats[0][c], stats[1][c] (the BatchNorm that follows needs no reduction pass).
__device__ __forceinline__ void conv_store_chunk(float (&v)[32], bool row_ok, __nv_bfloat16* dst_row, int col0,
                                                 int n_out, const float* __restrict__ bias, float* sum_dst, float* sq_dst,
                                                 int lane) {
    __nv_bfloat16* dst = dst_row + col0;
    if (col0 + 32 <= n_out && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
        // whole chunk, 16-byte aligned row: the common case, kept free of per-column predicates
        if (bias != nullptr) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col0) + q);
                v[4 * q] += b.x; v[4 * q + 1] += b.y; v[4 * q + 2] += b.z; v[4 * q + 3] += b.w;
            }
        }
        uint32_t packed[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            packed[e] = *reinterpret_cast<const uint32_t*>(&h);
        }
        if (row_ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                reinterpret_cast<uint4*>(dst)[q] =
                    make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
        }
        if (sum_dst != nullptr) {
            // statistics and consumers see the value that is stored: widen the packed bf16 back (a shift / a mask)
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                v[2 * e] = __uint_as_float(packed[e] << 16);
                v[2 * e + 1] = __uint_as_float(packed[e] & 0xffff0000u);
            }
            conv_chunk_stats(v, row_ok, col0, n_out, sum_dst, sq_dst, lane);
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        if (bias != nullptr && col0 + j < n_out) v[j] += bias[col0 + j];
        v[j] = __bfloat162float(__float2bfloat16(v[j]));
    }
    if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (col0 + j < n_out) dst[j] = __float2bfloat16(v[j]);
    }
    if (sum_dst != nullptr) conv_chunk_stats(v, row_ok, col0, n_out, sum_dst, sq_dst, lane);
}

// A persistent CTA adds the column sums of all its tiles into shared memory and flushes them once: with one global
// atomic per warp, chunk and tile the ResNet stage-1 layer sent 2 048 atomics to each of its 128 addresses, 4 us
// of L2 serialisation.
constexpr int kConvStatChannels = 512;
struct ConvStatAcc {
    float sum[kConvStatChannels];
    float sq[kConvStatChannels];
};
__device__ __forceinline__ void conv_stat_acc_zero(ConvStatAcc& acc) {
    for (int i = threadIdx.x; i < kConvStatChannels; i += blockDim.x) {
        acc.sum[i] = 0.f;
        acc.sq[i] = 0.f;
    }
}
__device__ __forceinline__ void conv_stat_acc_flush(const ConvStatAcc& acc, float* __restrict__ stats, int n_out) {
    for (int i = threadIdx.x; i < n_out; i += blockDim.x) {
        const float a = acc.sum[i], b = acc.sq[i];
        if (b != 0.f) {  // a channel this CTA never produced, or one that is all zero, adds nothing
            atomicAdd(stats + i, a);
            atomicAdd(stats + n_out + i, b);
        }
    }
}

template <int kBlockN>
constexpr int conv_stages() { return kBlockN == 64 ? 6 : 4; }

template <int kBlockN>
struct __align__(1024) ConvSmem {
    static constexpr int kStages = conv_stages<kBlockN>();
    __nv_bfloat16 a[kStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kStages][kBlockN * kGemmBlockK];
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
    ConvStatAcc stat;
};

// ---- fprop / dgrad ------------------------------------------------------------------------------
// stats (optional, fp32 [2][n_out]): per-output-channel sum and sum of squares of the STORED (bf16
// rounded) outputs, accumulated with atomics -- the BatchNorm that follows needs no reduction pass.
template <int kBlockN, bool kDgrad>
__global__ void __launch_bounds__(kConvWideThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const __grid_constant__ ConvGemmParams p, __nv_bfloat16* __restrict__ out,
                  const float* __restrict__ bias, float* __restrict__ stats) {
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kBytesB = kBlockN * kGemmBlockK * 2;
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    constexpr int kStages = conv_stages<kBlockN>();
    using Smem = ConvSmem<kBlockN>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int cblocks = (p.a_channels + kGemmBlockK - 1) / kGemmBlockK;
    const int n_tiles = (p.n_out + kBlockN - 1) / kBlockN;
    const int p_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int num_tiles = n_tiles * p_tiles * p.n_classes;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], kConvWideThreads - 64);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const bool stat_local = stats != nullptr && p.n_out <= kConvStatChannels;
    if (stat_local) conv_stat_acc_zero(sm.stat);
    float* const stat_sum = stats == nullptr ? nullptr : (stat_local ? sm.stat.sum : stats);
    float* const stat_sq = stats == nullptr ? nullptr : (stat_local ? sm.stat.sq : stats + p.n_out);
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);
    // Let the next kernel of the stream launch only now that this CTA owns its tensor memory: a dependent that
    // were resident earlier could take the columns this CTA is still waiting for and then wait for this grid.
    pdl_launch_dependents();
    pdl_wait();  // everything above is independent of the previous kernel's output

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int nt = tile % n_tiles;
                int rest = tile / n_tiles;
                const int pt = rest % p_tiles, cls = rest / p_tiles;
                const int w0 = (pt % p.tiles_w) * p.bw;
                const int h0 = ((pt / p.tiles_w) % p.tiles_h) * p.bh;
                const int n0 = (pt / (p.tiles_w * p.tiles_h)) * p.bn;
                const int ntaps = p.class_ntaps[cls], tap0 = p.class_tap0[cls];
                for (int t = 0; t < ntaps; ++t) {
                    const ConvTap tap = p.taps[tap0 + t];
                    for (int cb = 0; cb < cblocks; ++cb, ++it) {
                        const int s = it % kStages;
                        const uint32_t phase = (it / kStages) & 1;
                        mbar_wait(&sm.empty[s], phase ^ 1);
                        mbar_expect_tx(&sm.full[s], kBytesA + kBytesB);
                        tma_load_5d(sm.a[s], &map_a, &sm.full[s], tap.pw * p.a_cview + cb * kGemmBlockK, w0 + tap.dw,
                                    tap.ph, h0 + tap.dh, n0);
                        if (!kDgrad) {
                            // [kBlockN output channels][64 reduction channels] of this tap: K-major
                            tma_load_3d(sm.b[s], &map_b, &sm.full[s], cb * kGemmBlockK, tap.widx, nt * kBlockN);
                        } else {
                            // [64 reduction rows k][64 output (input-image) channels] chunks: MN-major
#pragma unroll
                            for (int j = 0; j < kBlockN / 64; ++j)
                                tma_load_3d(reinterpret_cast<unsigned char*>(sm.b[s]) + j * 8192, &map_b, &sm.full[s],
                                            nt * kBlockN + j * 64, tap.widx, cb * kGemmBlockK);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = conv_instr_desc<kBlockN, false, kDgrad>();
            uint32_t it = 0, local_tile = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
                const int cls = (tile / n_tiles) / p_tiles;
                const int num_kb = p.class_ntaps[cls] * cblocks;
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % kStages;
                    const uint32_t phase = (it / kStages) & 1;
                    mbar_wait(&sm.full[s], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                    for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                        const uint64_t da = umma_smem_desc(a_addr + kk * kGemmUmmaK * 2);
                        const uint64_t db = kDgrad ? umma_smem_desc_mn(b_addr + kk * 2048, 8192)
                                                   : umma_smem_desc(b_addr + kk * kGemmUmmaK * 2);
                        umma_bf16(tmem_acc, da, db, idesc, (kb | kk) != 0);
                    }
                    umma_commit(&sm.empty[s]);
                }
                umma_commit(&sm.tmem_full[acc]);
            }
        }
    } else {
        // ---- epilogue
        const int quarter = warp & 3, cgroup = (warp - 2) >> 2;
        const int m_local = quarter * 32 + lane;
        const int lw = m_local % p.bw, lh = (m_local / p.bw) % p.bh, ln = m_local / (p.bw * p.bh);
        uint32_t local_tile = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            const int nt = tile % n_tiles;
            int rest = tile / n_tiles;
            const int pt = rest % p_tiles, cls = rest / p_tiles;
            const int w = (pt % p.tiles_w) * p.bw + lw;
            const int h = ((pt / p.tiles_w) % p.tiles_h) * p.bh + lh;
            const int n = (pt / (p.tiles_w * p.tiles_h)) * p.bn + ln;
            const bool row_ok = w < p.pw && h < p.ph && n < p.pn;
            const bool has_acc = p.class_ntaps[cls] > 0;
            const int n0 = nt * kBlockN;
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            mbar_wait(&sm.tmem_full[acc], use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_acc = tmem_base + acc * kBlockN;
            __nv_bfloat16* dst_row = out +
                (((long long)n * p.out_h + (h * p.os + p.class_oph[cls])) * p.out_w + (w * p.os + p.class_opw[cls])) *
                    (long long)p.out_pitch;
#pragma unroll 1
            for (int chunk = cgroup; chunk < kBlockN / 32; chunk += 2) {
                const int col0 = n0 + chunk * 32;
                if (col0 >= p.n_out) break;
                uint32_t r[32];
                if (has_acc) {
                    tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32), r);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = 0u;
                }
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                conv_store_chunk(v, row_ok, dst_row, col0, p.n_out, bias, stat_sum, stat_sq, lane);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&sm.tmem_empty[acc]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (stat_local) conv_stat_acc_flush(sm.stat, stats, p.n_out);
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// ---- 3 x 3, stride 1 with a shared-memory halo patch --------------------------------------------------------
// The tap-by-tap kernel above brings the A tile in once PER TAP: 9 x 16 KB per 64 reduction channels, and the
// ~46 B/clk an SM takes in from L2 is what bounds it on the wide feature maps (ResNet stages 1 / 2).  Here a CTA
// owns an 8 (w) x 16 (h) patch of output pixels of one image and loads, per 64 reduction channels, ONE box of
// 16 x 18 input pixels -- the tile plus its halo, 36 KB instead of 144 KB.  All nine taps read that box in place:
// with the 128-byte swizzle one pixel is one 128-byte row, 8 rows (8 consecutive w) form a swizzle atom, and the
// UMMA descriptor of tap (r, s) simply starts (r * 16 + s) rows further with the stride between 8-row groups
// set to one patch line (2 048 B).  The start is then s rows off the 1 024-byte swizzle period; measured on
// B200 the tensor core derives the swizzle phase from the absolute shared-memory address bits [7, 10), which
// is exactly how TMA wrote the patch, so the base-offset field of the descriptor stays 0 (setting it to s, or
// to 8 - s, gives wrong results).  Filters: nine 64 x 64 tiles per 64 reduction channels; when the layer has a
// single such set (Cin, Cout <= 64) they are loaded once per CTA and stay resident.
constexpr int kHaloW = 8, kHaloH = 16;              // output pixels of a tile
constexpr int kHaloPatchW = 16, kHaloPatchH = 18;   // input pixels loaded for it
constexpr int kHaloAStages = 3;
constexpr uint32_t kHaloBytesA = kHaloPatchW * kHaloPatchH * 128;
constexpr uint32_t kHaloBytesB = 64 * 64 * 2;

struct HaloParams {
    int n, h, w;
    int k_channels;   // reduction channels
    int n_out;        // output channels
    int out_pitch;
    int tiles_w, tiles_h;
    int resident;     // one filter set for the whole launch
};

#ifdef DLS_CONV_TIMELINE
// per-CTA clock64 stamps of the kernel's phases (development aid, tools/conv_timeline.py)
__device__ long long g_conv_timeline[kNumSMs][16];
#define DLS_TL(slot) (g_conv_timeline[blockIdx.x % kNumSMs][slot] = clock64())
#else
#define DLS_TL(slot) ((void)0)
#endif

struct __align__(1024) HaloSmem {
    unsigned char a[kHaloAStages][kHaloBytesA];
    unsigned char b[9][kHaloBytesB];
    uint64_t a_full[kHaloAStages];
    uint64_t a_empty[kHaloAStages];
    uint64_t b_full[9];
    uint64_t b_empty[9];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
    ConvStatAcc stat;
};

__device__ __forceinline__ uint64_t umma_smem_desc_halo(uint32_t smem_addr) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    desc |= (uint64_t)1 << 16;
    desc |= (uint64_t)((kHaloPatchW * 128) >> 4) << 32;  // next 8-pixel group = next patch line
    desc |= (uint64_t)1 << 46;
    desc |= (uint64_t)2 << 61;
    return desc;
}

template <bool kDgrad>
__global__ void __launch_bounds__(kConvWideThreads, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ HaloParams p, __nv_bfloat16* __restrict__ out,
                 const float* __restrict__ bias, float* __restrict__ stats) {
    extern __shared__ unsigned char smem_raw[];
    constexpr int kBlockN = 64;
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    HaloSmem& sm = *reinterpret_cast<HaloSmem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    if (threadIdx.x == 0) DLS_TL(0);
    const int cblocks = (p.k_channels + 63) / 64;
    const int n_tiles = (p.n_out + kBlockN - 1) / kBlockN;
    const int p_tiles = p.tiles_w * p.tiles_h * p.n;
    const int num_tiles = n_tiles * p_tiles;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < kHaloAStages; ++s) {
            mbar_init(&sm.a_full[s], 1);
            mbar_init(&sm.a_empty[s], 1);
        }
        for (int s = 0; s < 9; ++s) {
            mbar_init(&sm.b_full[s], 1);
            mbar_init(&sm.b_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], kConvWideThreads - 64);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const bool stat_local = stats != nullptr && p.n_out <= kConvStatChannels;
    if (stat_local) conv_stat_acc_zero(sm.stat);
    float* const stat_sum = stats == nullptr ? nullptr : (stat_local ? sm.stat.sum : stats);
    float* const stat_sq = stats == nullptr ? nullptr : (stat_local ? sm.stat.sq : stats + p.n_out);
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);
    pdl_launch_dependents();  // after the tensor-memory allocation, see conv_igemm_kernel
    if (threadIdx.x == 0) DLS_TL(1);

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer
            pdl_wait();
            DLS_TL(2);
            if (p.resident) {
                for (int t = 0; t < 9; ++t) {
                    mbar_expect_tx(&sm.b_full[t], kHaloBytesB);
                    if (!kDgrad) tma_load_3d(sm.b[t], &map_b, &sm.b_full[t], 0, t, 0);
                    else tma_load_3d(sm.b[t], &map_b, &sm.b_full[t], 0, 8 - t, 0);
                }
            }
            uint32_t ia = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int nt = tile % n_tiles, pt = tile / n_tiles;
                const int w0 = (pt % p.tiles_w) * kHaloW;
                const int h0 = ((pt / p.tiles_w) % p.tiles_h) * kHaloH;
                const int img = pt / (p.tiles_w * p.tiles_h);
                for (int cb = 0; cb < cblocks; ++cb, ++ia) {
                    const int s = ia % kHaloAStages;
                    mbar_wait(&sm.a_empty[s], ((ia / kHaloAStages) & 1) ^ 1);
                    mbar_expect_tx(&sm.a_full[s], kHaloBytesA);
                    tma_load_5d(sm.a[s], &map_a, &sm.a_full[s], cb * 64, w0 - 1, 0, h0 - 1, img);
                    if (!p.resident) {
                        for (int t = 0; t < 9; ++t) {
                            // slot t holds the filter tap that patch position t (row t / 3, column t % 3) meets
                            mbar_wait(&sm.b_empty[t], (ia & 1) ^ 1);
                            mbar_expect_tx(&sm.b_full[t], kHaloBytesB);
                            if (!kDgrad) tma_load_3d(sm.b[t], &map_b, &sm.b_full[t], cb * 64, t, nt * kBlockN);
                            else tma_load_3d(sm.b[t], &map_b, &sm.b_full[t], nt * kBlockN, 8 - t, cb * 64);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = conv_instr_desc<kBlockN, false, kDgrad>();
            uint32_t ia = 0, local_tile = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int cb = 0; cb < cblocks; ++cb, ++ia) {
                    const int s = ia % kHaloAStages;
                    mbar_wait(&sm.a_full[s], (ia / kHaloAStages) & 1);
                    if (ia == 0) DLS_TL(3);
                    const uint32_t a_addr = smem_u32(sm.a[s]);
#pragma unroll 1
                    for (int t = 0; t < 9; ++t) {
                        mbar_wait(&sm.b_full[t], p.resident ? 0u : (ia & 1));
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const int pr = t / 3, pc = t - 3 * pr;
                        const uint32_t a_tap = a_addr + (uint32_t)(pr * kHaloPatchW + pc) * 128;
                        const uint32_t b_addr = smem_u32(sm.b[t]);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            const uint64_t da = umma_smem_desc_halo(a_tap + kk * 32);
                            const uint64_t db = kDgrad ? umma_smem_desc_mn(b_addr + kk * 2048, 8192)
                                                       : umma_smem_desc(b_addr + kk * 32);
                            umma_bf16(tmem_acc, da, db, idesc, (cb | t | kk) != 0);
                        }
                        if (!p.resident) umma_commit(&sm.b_empty[t]);
                    }
                    umma_commit(&sm.a_empty[s]);
                }
                umma_commit(&sm.tmem_full[acc]);
            }
        }
    } else {
        // ---- epilogue
        pdl_wait();
        const int quarter = warp & 3, cgroup = (warp - 2) >> 2;
        const int m_local = quarter * 32 + lane;
        const int lw = m_local & (kHaloW - 1), lh = m_local / kHaloW;
        uint32_t local_tile = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            const int nt = tile % n_tiles, pt = tile / n_tiles;
            const int w = (pt % p.tiles_w) * kHaloW + lw;
            const int h = ((pt / p.tiles_w) % p.tiles_h) * kHaloH + lh;
            const int img = pt / (p.tiles_w * p.tiles_h);
            const bool row_ok = w < p.w && h < p.h;
            const int n0 = nt * kBlockN;
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            mbar_wait(&sm.tmem_full[acc], use & 1);
            if (threadIdx.x == 64 && local_tile < 4) DLS_TL(4 + 2 * local_tile);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_acc = tmem_base + acc * kBlockN;
            __nv_bfloat16* dst_row = out + (((long long)img * p.h + h) * p.w + w) * (long long)p.out_pitch;
#pragma unroll 1
            for (int chunk = cgroup; chunk < kBlockN / 32; chunk += 2) {
                const int col0 = n0 + chunk * 32;
                if (col0 >= p.n_out) break;
                uint32_t r[32];
                tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32), r);
                if (threadIdx.x == 64 && local_tile == 0 && chunk == 0) DLS_TL(12);
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                conv_store_chunk(v, row_ok, dst_row, col0, p.n_out, bias, stat_sum, stat_sq, lane);
                if (threadIdx.x == 64 && local_tile == 0 && chunk == 0) DLS_TL(13);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&sm.tmem_empty[acc]);
            if (threadIdx.x == 64 && local_tile < 4) DLS_TL(5 + 2 * local_tile);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) DLS_TL(14);
    if (stat_local) conv_stat_acc_flush(sm.stat, stats, p.n_out);
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// ---- split-K variant for layers with few output tiles and a long reduction ---------------------------------
// (ResNet stages 3 / 4: 4 096 or 1 024 output pixels, K = 9 x 256 .. 9 x 512).  An SM takes in ~46 B/clk from
// L2, so a lone CTA walking 72 K blocks of one 128 x 64 tile needs ~27 us while most SMs idle.  Here the
// kSplit CTAs of a thread-block cluster share ONE 128 x 128 tile: CTA r accumulates K blocks
// [r * per, (r + 1) * per) in its own tensor memory, then the partial tiles are combined by a reduce-scatter
// through distributed shared memory -- CTA j owns columns [j * 128 / kSplit, +128 / kSplit): every other CTA
// writes its partial of those columns into CTA j's receive buffer (st.shared::cluster) and arrives on CTA j's
// mbarrier (release at cluster scope); CTA j adds them to its own partial and runs the normal epilogue for
// its columns.  One tile per cluster (no persistent loop): these layers have at most ~64 tiles.
template <int kSplit>
struct __align__(1024) SplitSmem {
    static constexpr int kStages = 4;
    __nv_bfloat16 a[kStages][kGemmBlockM * kGemmBlockK];
    __nv_bfloat16 b[kStages][128 * kGemmBlockK];
    float recv[kSplit - 1][128 / kSplit][kGemmBlockM];  // [source slot][column of my slice][tile row]
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t tmem_full;
    uint64_t recv_full;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
    return r;
}

template <bool kDgrad, int kSplit>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_igemm_splitk_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                         const __grid_constant__ ConvGemmParams p, __nv_bfloat16* __restrict__ out,
                         const float* __restrict__ bias, float* __restrict__ stats) {
    extern __shared__ unsigned char smem_raw[];
    constexpr int kBlockN = 128;
    constexpr int kSlice = kBlockN / kSplit;          // columns finalised by one CTA
    constexpr uint32_t kBytesB = kBlockN * kGemmBlockK * 2;
    constexpr uint32_t kTmemCols = kBlockN;
    using Smem = SplitSmem<kSplit>;
    constexpr int kStages = Smem::kStages;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int crank = (int)cluster_rank();
    const int cblocks = (p.a_channels + kGemmBlockK - 1) / kGemmBlockK;
    const int n_tiles = (p.n_out + kBlockN - 1) / kBlockN;
    const int p_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int tile = blockIdx.x / kSplit;             // grid = tiles * kSplit exactly
    const int nt = tile % n_tiles;
    const int pt = (tile / n_tiles) % p_tiles, cls = (tile / n_tiles) / p_tiles;
    const int num_kb = p.class_ntaps[cls] * cblocks;
    const int per = (num_kb + kSplit - 1) / kSplit;
    const int kb_begin = min(crank * per, num_kb), kb_end = min(kb_begin + per, num_kb);

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        mbar_init(&sm.tmem_full, 1);
        mbar_init(&sm.recv_full, (kSplit - 1) * 128);  // every epilogue thread of every other CTA
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    // every CTA's receive barrier exists before a peer writes to it
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);
    // Let the next kernel of the stream launch only now that this CTA owns its tensor memory: a dependent that
    // were resident earlier could take the columns this CTA is still waiting for and then wait for this grid.
    pdl_launch_dependents();
    pdl_wait();  // everything above is independent of the previous kernel's output
    const int w0 = (pt % p.tiles_w) * p.bw;
    const int h0 = ((pt / p.tiles_w) % p.tiles_h) * p.bh;
    const int n0 = (pt / (p.tiles_w * p.tiles_h)) * p.bn;

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer over this CTA's share of the K blocks
            const int tap0 = p.class_tap0[cls];
            uint32_t it = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
                const ConvTap tap = p.taps[tap0 + kb / cblocks];
                const int cb = kb % cblocks;
                const int s = it % kStages;
                const uint32_t phase = (it / kStages) & 1;
                mbar_wait(&sm.empty[s], phase ^ 1);
                mbar_expect_tx(&sm.full[s], kBytesA + kBytesB);
                tma_load_5d(sm.a[s], &map_a, &sm.full[s], tap.pw * p.a_cview + cb * kGemmBlockK, w0 + tap.dw, tap.ph,
                            h0 + tap.dh, n0);
                if (!kDgrad) {
                    tma_load_3d(sm.b[s], &map_b, &sm.full[s], cb * kGemmBlockK, tap.widx, nt * kBlockN);
                } else {
#pragma unroll
                    for (int j = 0; j < kBlockN / 64; ++j)
                        tma_load_3d(reinterpret_cast<unsigned char*>(sm.b[s]) + j * 8192, &map_b, &sm.full[s],
                                    nt * kBlockN + j * 64, tap.widx, cb * kGemmBlockK);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = conv_instr_desc<kBlockN, false, kDgrad>();
            uint32_t it = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
                const int s = it % kStages;
                const uint32_t phase = (it / kStages) & 1;
                mbar_wait(&sm.full[s], phase);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                for (int kk = 0; kk < kGemmBlockK / kGemmUmmaK; ++kk) {
                    const uint64_t da = umma_smem_desc(a_addr + kk * kGemmUmmaK * 2);
                    const uint64_t db = kDgrad ? umma_smem_desc_mn(b_addr + kk * 2048, 8192)
                                               : umma_smem_desc(b_addr + kk * kGemmUmmaK * 2);
                    umma_bf16(tmem_base, da, db, idesc, (kb > kb_begin) || kk != 0);
                }
                umma_commit(&sm.empty[s]);
            }
            umma_commit(&sm.tmem_full);
        }
    } else {
        // ---- epilogue: reduce-scatter of the partial tiles, then the usual store of my column slice
        const int quarter = warp & 3;
        const int m_local = quarter * 32 + lane;
        const bool has_acc = kb_end > kb_begin;
        mbar_wait(&sm.tmem_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t recv_base = smem_u32(&sm.recv[0][0][0]);
        const uint32_t recv_bar = smem_u32(&sm.recv_full);
#pragma unroll 1
        for (int d = 1; d < kSplit; ++d) {
            const int dest = (crank + d) % kSplit;
            const int slot = crank > dest ? crank - 1 : crank;  // my index among dest's kSplit - 1 sources
            const uint32_t remote = map_to_cta(recv_base, (uint32_t)dest) + (uint32_t)((slot * kSlice) * kGemmBlockM + m_local) * 4u;
#pragma unroll 1
            for (int chunk = 0; chunk < kSlice / 32; ++chunk) {
                uint32_t r[32];
                if (has_acc) {
                    tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(dest * kSlice + chunk * 32), r);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = 0u;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(remote + (uint32_t)((chunk * 32 + j) * kGemmBlockM) * 4u),
                                 "r"(r[j])
                                 : "memory");
            }
            // release my stores to the destination CTA and count this thread in
            asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(map_to_cta(recv_bar, (uint32_t)dest))
                         : "memory");
        }
        // my slice: own partial + the kSplit - 1 received ones
        {
            uint32_t done = 0;
            for (uint32_t spins = 0; !done; ++spins) {
                asm volatile(
                    "{\n\t.reg .pred p;\n\t"
                    "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
                    "selp.u32 %0, 1, 0, p;\n\t}"
                    : "=r"(done)
                    : "r"(recv_bar), "r"(0u)
                    : "memory");
                if (spins > 200000000u) {
                    printf("dls: split-K receive barrier timed out (block %d thread %d)\n", (int)blockIdx.x, (int)threadIdx.x);
                    asm volatile("trap;");
                }
            }
        }
        const int lw = m_local % p.bw, lh = (m_local / p.bw) % p.bh, ln = m_local / (p.bw * p.bh);
        const int w = w0 + lw, h = h0 + lh, n = n0 + ln;
        const bool row_ok = w < p.pw && h < p.ph && n < p.pn;
        __nv_bfloat16* dst_row = out +
            (((long long)n * p.out_h + (h * p.os + p.class_oph[cls])) * p.out_w + (w * p.os + p.class_opw[cls])) *
                (long long)p.out_pitch;
#pragma unroll 1
        for (int chunk = 0; chunk < kSlice / 32; ++chunk) {
            const int col0 = nt * kBlockN + crank * kSlice + chunk * 32;
            if (col0 >= p.n_out) break;
            float v[32];
            if (has_acc) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(crank * kSlice + chunk * 32), r);
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.f;
            }
#pragma unroll
            for (int src = 0; src < kSplit - 1; ++src)
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += sm.recv[src][chunk * 32 + j][m_local];
            conv_store_chunk(v, row_ok, dst_row, col0, p.n_out, bias, stats, stats ? stats + p.n_out : nullptr, lane);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    // no CTA leaves while a peer may still write into its receive buffer or arrive on its barrier
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// ---- wgrad ----------------------------------------------------------------------------------------
constexpr int kWgradPixels = 128;  // pixels per K block (8 MMAs of K = 16)

template <int kBlockN, int kStages>
struct __align__(1024) WgradSmem {
    __nv_bfloat16 a[kStages][2 * kWgradPixels * 64];              // dY: two 64-channel chunks (M = 128)
    __nv_bfloat16 b[kStages][(kBlockN / 64) * kWgradPixels * 64];  // X: kBlockN / 64 chunks
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
};

struct WgradParams {
    int bw, bh, bn;
    int tiles_w, tiles_h, tiles_n;   // pixel tiles of the output-pixel space (N, OH, OW)
    int c_out, c_in, x_cview;
    int n_taps, taps_total;          // taps handled here / taps of the filter (row pitch of dW)
    int split, tiles_per_split;
    int atomic;
    int out_bf16;                    // store mode only: dw is bf16
    int dw_pitch, dw_cols;           // elements per (k, tap) row of dw, columns actually written
    ConvTap taps[kConvMaxTaps];
};

template <int kBlockN, int kStages>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x,
                  const __grid_constant__ WgradParams p, float* __restrict__ dw) {
    extern __shared__ unsigned char smem_raw[];
    constexpr uint32_t kChunkBytes = kWgradPixels * 64 * 2;  // 16 KiB
    constexpr uint32_t kStageBytes = (2 + kBlockN / 64) * kChunkBytes;
    constexpr uint32_t kTmemCols = 2 * kBlockN;
    using Smem = WgradSmem<kBlockN, kStages>;
    Smem& sm = *reinterpret_cast<Smem*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int warp = uniform_warp_index(), lane = threadIdx.x & 31;
    const int m_tiles = (p.c_out + 127) / 128;
    const int n_tiles = (p.c_in + kBlockN - 1) / kBlockN;
    const int p_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int num_tiles = m_tiles * n_tiles * p.n_taps * p.split;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_dy) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sm.tmem_full[a], 1);
            mbar_init(&sm.tmem_empty[a], 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = warp_uniform(sm.tmem_base);
    // Let the next kernel of the stream launch only now that this CTA owns its tensor memory: a dependent that
    // were resident earlier could take the columns this CTA is still waiting for and then wait for this grid.
    pdl_launch_dependents();
    pdl_wait();  // everything above is independent of the previous kernel's output

    // tile -> (nt, mt, tap, split): the split index is slowest, so concurrently running CTAs share pixels
    auto decode = [&](int tile, int& nt, int& mt, int& tap, int& sp) {
        nt = tile % n_tiles;
        int rest = tile / n_tiles;
        mt = rest % m_tiles;
        rest /= m_tiles;
        tap = rest % p.n_taps;
        sp = rest / p.n_taps;
    };

    if (warp == 0) {
        if (elect_one()) {  // ---- TMA producer
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int nt, mt, ti, sp;
                decode(tile, nt, mt, ti, sp);
                const ConvTap tap = p.taps[ti];
                const int pt_begin = sp * p.tiles_per_split;
                const int pt_end = min(pt_begin + p.tiles_per_split, p_tiles);
                for (int pt = pt_begin; pt < pt_end; ++pt, ++it) {
                    const int w0 = (pt % p.tiles_w) * p.bw;
                    const int h0 = ((pt / p.tiles_w) % p.tiles_h) * p.bh;
                    const int n0 = (pt / (p.tiles_w * p.tiles_h)) * p.bn;
                    const int s = it % kStages;
                    const uint32_t phase = (it / kStages) & 1;
                    mbar_wait(&sm.empty[s], phase ^ 1);
                    mbar_expect_tx(&sm.full[s], kStageBytes);
                    unsigned char* a = reinterpret_cast<unsigned char*>(sm.a[s]);
                    unsigned char* b = reinterpret_cast<unsigned char*>(sm.b[s]);
#pragma unroll
                    for (int j = 0; j < 2; ++j)  // channels past c_out are zero filled
                        tma_load_5d(a + j * kChunkBytes, &map_dy, &sm.full[s], mt * 128 + j * 64, w0, 0, h0, n0);
#pragma unroll
                    for (int j = 0; j < kBlockN / 64; ++j)
                        tma_load_5d(b + j * kChunkBytes, &map_x, &sm.full[s],
                                    tap.pw * p.x_cview + nt * kBlockN + j * 64, w0 + tap.dw, tap.ph, h0 + tap.dh, n0);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {  // ---- MMA issuer
            const uint32_t idesc = conv_instr_desc<kBlockN, true, true>();
            uint32_t it = 0, local_tile = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
                int nt, mt, ti, sp;
                decode(tile, nt, mt, ti, sp);
                const int pt_begin = sp * p.tiles_per_split;
                const int pt_end = min(pt_begin + p.tiles_per_split, p_tiles);
                const uint32_t acc = local_tile & 1, use = local_tile >> 1;
                mbar_wait(&sm.tmem_empty[acc], (use & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_acc = tmem_base + acc * kBlockN;
                for (int pt = pt_begin; pt < pt_end; ++pt, ++it) {
                    const int s = it % kStages;
                    const uint32_t phase = (it / kStages) & 1;
                    mbar_wait(&sm.full[s], phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_addr = smem_u32(sm.a[s]), b_addr = smem_u32(sm.b[s]);
#pragma unroll
                    for (int kk = 0; kk < kWgradPixels / kGemmUmmaK; ++kk) {
                        umma_bf16(tmem_acc, umma_smem_desc_mn(a_addr + kk * 2048, kChunkBytes),
                                  umma_smem_desc_mn(b_addr + kk * 2048, kChunkBytes), idesc, (pt > pt_begin) || kk != 0);
                    }
                    umma_commit(&sm.empty[s]);
                }
                umma_commit(&sm.tmem_full[acc]);
            }
        }
    } else {
        const int quarter = warp & 3;
        uint32_t local_tile = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
            int nt, mt, ti, sp;
            decode(tile, nt, mt, ti, sp);
            const bool has_acc = sp * p.tiles_per_split < p_tiles;
            const int widx = p.taps[ti].widx;
            const int row = mt * 128 + quarter * 32 + lane;  // output channel k
            const uint32_t acc = local_tile & 1, use = local_tile >> 1;
            mbar_wait(&sm.tmem_full[acc], use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_acc = tmem_base + acc * kBlockN;
            const long long row_off = ((long long)row * p.taps_total + widx) * p.dw_pitch;
            float* dst_row = dw + row_off;
            __nv_bfloat16* dst_row_h = reinterpret_cast<__nv_bfloat16*>(dw) + row_off;
#pragma unroll 1
            for (int chunk = 0; chunk < kBlockN / 32; ++chunk) {
                const int col0 = nt * kBlockN + chunk * 32;
                if (col0 >= p.dw_cols) break;
                uint32_t r[32];
                tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32), r);
                if (row < p.c_out && has_acc) {
                    float* dst = dst_row + col0;
                    const bool vec = col0 + 32 <= p.dw_cols && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
                    if (p.out_bf16) {
                        __nv_bfloat16* dh = dst_row_h + col0;
                        if (col0 + 32 <= p.dw_cols && ((reinterpret_cast<uintptr_t>(dh) & 15) == 0)) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint4 pk;
                                __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                                for (int e = 0; e < 4; ++e)
                                    hh[e] = __floats2bfloat162_rn(__uint_as_float(r[8 * q + 2 * e]),
                                                                  __uint_as_float(r[8 * q + 2 * e + 1]));
                                reinterpret_cast<uint4*>(dh)[q] = pk;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < p.dw_cols) dh[j] = __float2bfloat16(__uint_as_float(r[j]));
                        }
                    } else if (p.atomic) {
                        if (vec) {
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * q),
                                             "f"(__uint_as_float(r[4 * q])), "f"(__uint_as_float(r[4 * q + 1])),
                                             "f"(__uint_as_float(r[4 * q + 2])), "f"(__uint_as_float(r[4 * q + 3]))
                                             : "memory");
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (col0 + j < p.dw_cols) atomicAdd(dst + j, __uint_as_float(r[j]));
                        }
                    } else if (vec) {
#pragma unroll
                        for (int q = 0; q < 8; ++q)
                            reinterpret_cast<float4*>(dst)[q] =
                                make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                            __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (col0 + j < p.dw_cols) dst[j] = __uint_as_float(r[j]);
                    }
                }
            }
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

// ---- host side ------------------------------------------------------------------------------------
static int pow2_ceil(int v) {
    int r = 1;
    while (r < v) r <<= 1;
    return r;
}

static void choose_box(int pw, int ph, int pn, int& bw, int& bh, int& bn) {
    bw = pow2_ceil(pw);
    if (bw > 128) bw = 128;
    bh = pow2_ceil(ph);
    if (bh > 128 / bw) bh = 128 / bw;
    bn = 128 / (bw * bh);
}

// 5-d parity view of an NHWC bf16 activation: {C * s, W / s, s, H / s, N}; box {64, bw, 1, bh, bn}
static int make_act_map(CUtensorMap* map, const void* base, int n, int h, int w, int c, int pitch, int s, int bw, int bh,
                        int bn) {
    EncodeTiledFn encode = get_encode_fn();
    if (encode == nullptr) return fail_msg("cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t pix = (cuuint64_t)pitch * 2;  // bytes per pixel row
    const cuuint64_t dims[5] = {(cuuint64_t)(s == 1 ? c : pitch * (s - 1) + c), (cuuint64_t)(w / s), (cuuint64_t)s,
                                (cuuint64_t)(h / s), (cuuint64_t)n};
    const cuuint64_t strides[4] = {pix * s, pix * w, pix * w * s, pix * w * h};
    const cuuint32_t box[5] = {64, (cuuint32_t)bw, 1, (cuuint32_t)bh, (cuuint32_t)bn};
    const cuuint32_t es[5] = {1, 1, 1, 1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_last_error, sizeof(g_last_error), "activation tensor map: CUresult %d (n %d h %d w %d c %d pitch %d s %d)",
                 (int)r, n, h, w, c, pitch, s);
        return -1;
    }
    return 0;
}

// filters [Cout][taps][Cin] bf16: dims {Cin, taps, Cout}; box {64, 1, rows}
static int make_filter_map(CUtensorMap* map, const void* base, int c_out, int taps, int c_in, int rows) {
    EncodeTiledFn encode = get_encode_fn();
    if (encode == nullptr) return fail_msg("cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[3] = {(cuuint64_t)c_in, (cuuint64_t)taps, (cuuint64_t)c_out};
    const cuuint64_t strides[2] = {(cuuint64_t)c_in * 2, (cuuint64_t)c_in * 2 * taps};
    const cuuint32_t box[3] = {64, 1, (cuuint32_t)rows};
    const cuuint32_t es[3] = {1, 1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_last_error, sizeof(g_last_error), "filter tensor map: CUresult %d (cout %d taps %d cin %d)", (int)r,
                 c_out, taps, c_in);
        return -1;
    }
    return 0;
}

static int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

static int check_conv(const DlsConvShape& cs, const char* who) {
    if (cs.n < 1 || cs.h < 1 || cs.w < 1 || cs.c_in < 8 || cs.c_out < 8 || cs.r < 1 || cs.s < 1)
        return fail_msg("conv: empty problem");
    if (cs.c_in % 8 || cs.c_out % 8) return fail_msg("conv: channel counts must be multiples of 8 (16-byte TMA rows)");
    if (cs.stride != 1 && cs.stride != 2) return fail_msg("conv: stride must be 1 or 2");
    if (cs.stride == 2 && (cs.h % 2 || cs.w % 2 || cs.c_in % 64))
        return fail_msg("conv: stride 2 needs even H, W and Cin % 64 == 0");
    if (cs.r * cs.s > kConvMaxTaps) return fail_msg("conv: too many filter taps");
    (void)who;
    return 0;
}

template <typename Kernel>
static int set_smem(Kernel kernel, size_t smem) {
    DLS_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return 0;
}

template <bool kDgrad, int kSplit>
static int launch_splitk(const CUtensorMap& map_a, const CUtensorMap& map_b, const ConvGemmParams& p, void* out,
                         const float* bias, float* stats, long long tiles, cudaStream_t st) {
    auto kernel = conv_igemm_splitk_kernel<kDgrad, kSplit>;
    const size_t smem = sizeof(SplitSmem<kSplit>) + 1024;
    static thread_local bool configured = false;
    if (!configured) {
        if (int rc = set_smem(kernel, smem)) return rc;
        configured = true;
    }
    DLS_CHECK(launch_pdl(kernel, dim3((unsigned)(tiles * kSplit)), dim3(kConvThreads), smem, st, kSplit, map_a, map_b, p,
                         (__nv_bfloat16*)out, bias, stats));
    return 0;
}

// split-K factor (0 = use the persistent kernel): only when 128 x 128 tiles leave most SMs idle and every CTA of
// the cluster still gets a few K blocks
static int pick_split(int n_out, long long tiles128, int min_kb) {
    static const int forced = [] { const char* e = getenv("DLS_CONV_SPLITK"); return e ? atoi(e) : -1; }();
    if (forced == 0) return 0;
    if (n_out < 128 || n_out % 32) return 0;
    int split = 0;
    if (tiles128 * 4 <= kNumSMs + 12 && min_kb >= 8) split = 4;
    else if (tiles128 * 2 <= kNumSMs + 12 && min_kb >= 4) split = 2;
    if (forced == 2 || forced == 4) split = (min_kb >= forced) ? forced : split;
    return split;
}

static int pick_block_n(int n_out, long long p_tiles) {
    static const int forced = [] { const char* e = getenv("DLS_CONV_BN"); return e ? atoi(e) : 0; }();
    if (forced == 64 || forced == 128) return n_out > 64 ? forced : 64;
    if (n_out <= 64) return 64;
    // wide tiles only when they still give every SM a tile
    return (p_tiles * ((n_out + 127) / 128) >= 100) ? 128 : 64;
}

// 3 x 3 / stride 1 / pad 1 layers wide enough for 8 x 16 pixel tiles take the halo-patch kernel
static bool halo_eligible(const DlsConvShape& cs) {
    static const int enabled = [] { const char* e = getenv("DLS_CONV_HALO"); return e ? atoi(e) : 1; }();
    return enabled && cs.r == 3 && cs.s == 3 && cs.stride == 1 && cs.pad_h == 1 && cs.pad_w == 1 && cs.w % kHaloW == 0 &&
           cs.w >= kHaloPatchW && cs.h >= kHaloH;
}

template <bool kDgrad>
static int launch_halo(const void* a, const void* w, void* out, const float* bias, float* stats, const DlsConvShape& cs,
                       cudaStream_t st) {
    HaloParams p = {};
    p.n = cs.n; p.h = cs.h; p.w = cs.w;
    p.k_channels = kDgrad ? cs.c_out : cs.c_in;
    p.n_out = kDgrad ? cs.c_in : cs.c_out;
    const int a_pitch = kDgrad ? (cs.y_pitch ? cs.y_pitch : cs.c_out) : (cs.x_pitch ? cs.x_pitch : cs.c_in);
    p.out_pitch = kDgrad ? (cs.x_pitch ? cs.x_pitch : cs.c_in) : (cs.y_pitch ? cs.y_pitch : cs.c_out);
    p.tiles_w = cs.w / kHaloW; p.tiles_h = (cs.h + kHaloH - 1) / kHaloH;
    p.resident = (p.k_channels <= 64 && p.n_out <= 64) ? 1 : 0;
    CUtensorMap map_a, map_b;
    if (int rc = make_act_map(&map_a, a, cs.n, cs.h, cs.w, p.k_channels, a_pitch, 1, kHaloPatchW, kHaloPatchH, 1)) return rc;
    if (int rc = make_filter_map(&map_b, w, cs.c_out, 9, cs.c_in, 64)) return rc;
    const long long tiles = (long long)p.tiles_w * p.tiles_h * cs.n * ((p.n_out + 63) / 64);
    auto kernel = conv_halo_kernel<kDgrad>;
    const size_t smem = sizeof(HaloSmem) + 1024;
    static thread_local bool configured = false;
    if (!configured) {
        if (int rc = set_smem(kernel, smem)) return rc;
        configured = true;
    }
    DLS_CHECK(launch_pdl(kernel, dim3((unsigned)(tiles < kNumSMs ? tiles : kNumSMs)), dim3(kConvWideThreads), smem, st, 0,
                         map_a, map_b, p, (__nv_bfloat16*)out, bias, stats));
    return 0;
}

}  // namespace dls

using namespace dls;

// debugging aid: the CUresult of cuTensorMapEncodeTiled for a bf16 tensor described by the caller
extern "C" int dls_tensor_map_probe(int rank, const unsigned long long* dims, const unsigned long long* strides,
                                    const unsigned int* box, const unsigned int* elem_strides, int swizzle,
                                    const void* base) {
    EncodeTiledFn encode = get_encode_fn();
    if (encode == nullptr) return -1;
    CUtensorMap map;
    return (int)encode(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                       (const cuuint64_t*)dims, (const cuuint64_t*)strides, (const cuuint32_t*)box,
                       (const cuuint32_t*)elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle,
                       CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

#ifdef DLS_CONV_TIMELINE
extern "C" int dls_conv_timeline(long long* host_out, int reset) {
    if (reset) {
        static long long zeros[kNumSMs][16];
        DLS_CHECK(cudaMemcpyToSymbol(g_conv_timeline, zeros, sizeof(zeros)));
        return 0;
    }
    DLS_CHECK(cudaDeviceSynchronize());
    DLS_CHECK(cudaMemcpyFromSymbol(host_out, g_conv_timeline, sizeof(long long) * kNumSMs * 16));
    return 0;
}
#endif

extern "C" int dls_conv_supported(const DlsConvShape* cs) { return check_conv(*cs, "probe") == 0 ? 1 : 0; }

extern "C" int dls_conv_fprop(const void* x, const void* w, void* y, const float* bias, float* stats,
                              const DlsConvShape* shape, void* stream) {
    const DlsConvShape cs = *shape;
    if (int rc = check_conv(cs, "fprop")) return rc;
    if (((uintptr_t)x | (uintptr_t)w | (uintptr_t)y) & 15) return fail_msg("dls_conv_fprop: operands must be 16-byte aligned");
    const int oh = (cs.h + 2 * cs.pad_h - cs.r) / cs.stride + 1, ow = (cs.w + 2 * cs.pad_w - cs.s) / cs.stride + 1;
    ConvGemmParams p = {};
    p.pn = cs.n; p.ph = oh; p.pw = ow;
    choose_box(ow, oh, cs.n, p.bw, p.bh, p.bn);
    p.tiles_w = (ow + p.bw - 1) / p.bw; p.tiles_h = (oh + p.bh - 1) / p.bh; p.tiles_n = (cs.n + p.bn - 1) / p.bn;
    p.a_channels = cs.c_in; p.a_cview = cs.x_pitch ? cs.x_pitch : cs.c_in; p.n_out = cs.c_out;
    p.n_classes = 1;
    p.class_ntaps[0] = cs.r * cs.s; p.class_tap0[0] = 0;
    p.out_h = oh; p.out_w = ow; p.out_pitch = cs.y_pitch ? cs.y_pitch : cs.c_out; p.os = 1;
    for (int r = 0; r < cs.r; ++r)
        for (int s = 0; s < cs.s; ++s) {
            ConvTap& t = p.taps[r * cs.s + s];
            const int qh = r - cs.pad_h, qw = s - cs.pad_w;
            t.dh = (int16_t)floor_div(qh, cs.stride); t.ph = (uint8_t)(qh - t.dh * cs.stride);
            t.dw = (int16_t)floor_div(qw, cs.stride); t.pw = (uint8_t)(qw - t.dw * cs.stride);
            t.widx = (uint16_t)(r * cs.s + s);
        }
    if (halo_eligible(cs)) {
        if (int rc = launch_halo<false>(x, w, y, bias, stats, cs, (cudaStream_t)stream)) return rc;
        DLS_LAUNCHED("conv_halo_kernel<fprop>");
        return 0;
    }
    const long long p_tiles = (long long)p.tiles_w * p.tiles_h * p.tiles_n;
    const int cblocks = (cs.c_in + 63) / 64;
    if (const int split = pick_split(cs.c_out, p_tiles * ((cs.c_out + 127) / 128), cs.r * cs.s * cblocks)) {
        CUtensorMap ma, mb;
        if (int rc = make_act_map(&ma, x, cs.n, cs.h, cs.w, cs.c_in, p.a_cview, cs.stride, p.bw, p.bh, p.bn)) return rc;
        if (int rc = make_filter_map(&mb, w, cs.c_out, cs.r * cs.s, cs.c_in, 128)) return rc;
        const long long tiles = p_tiles * ((cs.c_out + 127) / 128);
        const int rc = split == 4 ? launch_splitk<false, 4>(ma, mb, p, y, bias, stats, tiles, (cudaStream_t)stream)
                                  : launch_splitk<false, 2>(ma, mb, p, y, bias, stats, tiles, (cudaStream_t)stream);
        if (rc) return rc;
        DLS_LAUNCHED("conv_igemm_splitk_kernel<fprop>");
        return 0;
    }
    const int block_n = pick_block_n(cs.c_out, p_tiles);
    CUtensorMap map_a, map_b;
    if (int rc = make_act_map(&map_a, x, cs.n, cs.h, cs.w, cs.c_in, p.a_cview, cs.stride, p.bw, p.bh, p.bn)) return rc;
    if (int rc = make_filter_map(&map_b, w, cs.c_out, cs.r * cs.s, cs.c_in, block_n)) return rc;
    const long long tiles = p_tiles * ((cs.c_out + block_n - 1) / block_n);
    dim3 grid((unsigned)(tiles < kNumSMs ? tiles : kNumSMs));
    cudaStream_t st = (cudaStream_t)stream;
#define DLS_CONV_LAUNCH(BN)                                                                                   \
    do {                                                                                                      \
        const size_t smem = sizeof(ConvSmem<BN>) + 1024;                                                      \
        static thread_local bool configured = false;                                                          \
        if (!configured) {                                                                                    \
            if (int rc = set_smem(conv_igemm_kernel<BN, false>, smem)) return rc;                             \
            configured = true;                                                                                \
        }                                                                                                     \
        DLS_CHECK(launch_pdl(conv_igemm_kernel<BN, false>, grid, dim3(kConvWideThreads), smem, st, 0, map_a, map_b, p, \
                             (__nv_bfloat16*)y, bias, stats));                                                \
    } while (0)
    if (block_n == 128) DLS_CONV_LAUNCH(128);
    else DLS_CONV_LAUNCH(64);
#undef DLS_CONV_LAUNCH
    DLS_LAUNCHED("conv_igemm_kernel<fprop>");
    return 0;
}

extern "C" int dls_conv_dgrad(const void* dy, const void* w, void* dx, const DlsConvShape* shape, void* stream) {
    const DlsConvShape cs = *shape;
    if (int rc = check_conv(cs, "dgrad")) return rc;
    if (((uintptr_t)dy | (uintptr_t)w | (uintptr_t)dx) & 15) return fail_msg("dls_conv_dgrad: operands must be 16-byte aligned");
    const int oh = (cs.h + 2 * cs.pad_h - cs.r) / cs.stride + 1, ow = (cs.w + 2 * cs.pad_w - cs.s) / cs.stride + 1;
    const int st_ = cs.stride;
    if (st_ == 2 && (oh != cs.h / 2 || ow != cs.w / 2)) return fail_msg("dls_conv_dgrad: stride 2 needs OH == H/2, OW == W/2");
    if (halo_eligible(cs)) {
        if (int rc = launch_halo<true>(dy, w, dx, nullptr, nullptr, cs, (cudaStream_t)stream)) return rc;
        DLS_LAUNCHED("conv_halo_kernel<dgrad>");
        return 0;
    }
    ConvGemmParams p = {};
    p.pn = cs.n; p.ph = cs.h / st_; p.pw = cs.w / st_;
    choose_box(p.pw, p.ph, cs.n, p.bw, p.bh, p.bn);
    p.tiles_w = (p.pw + p.bw - 1) / p.bw; p.tiles_h = (p.ph + p.bh - 1) / p.bh; p.tiles_n = (cs.n + p.bn - 1) / p.bn;
    p.a_channels = cs.c_out; p.a_cview = cs.y_pitch ? cs.y_pitch : cs.c_out; p.n_out = cs.c_in;
    p.out_h = cs.h; p.out_w = cs.w; p.out_pitch = cs.x_pitch ? cs.x_pitch : cs.c_in; p.os = st_;
    p.n_classes = st_ * st_;
    int n_taps = 0;
    for (int cph = 0; cph < st_; ++cph)
        for (int cpw = 0; cpw < st_; ++cpw) {
            const int cls = cph * st_ + cpw;
            p.class_tap0[cls] = n_taps; p.class_oph[cls] = cph; p.class_opw[cls] = cpw;
            for (int r = 0; r < cs.r; ++r)
                for (int s = 0; s < cs.s; ++s) {
                    const int qh = cph + cs.pad_h - r, qw = cpw + cs.pad_w - s;  // = stride * (dY coordinate - h2)
                    if (((qh % st_) + st_) % st_ != 0 || ((qw % st_) + st_) % st_ != 0) continue;
                    ConvTap& t = p.taps[n_taps++];
                    t.dh = (int16_t)floor_div(qh, st_); t.dw = (int16_t)floor_div(qw, st_);
                    t.ph = 0; t.pw = 0; t.widx = (uint16_t)(r * cs.s + s);
                }
            p.class_ntaps[cls] = n_taps - p.class_tap0[cls];
        }
    const long long p_tiles = (long long)p.tiles_w * p.tiles_h * p.tiles_n * p.n_classes;
    const int kblocks = (cs.c_out + 63) / 64;
    int min_kb = 1 << 30;
    for (int c = 0; c < p.n_classes; ++c) min_kb = p.class_ntaps[c] * kblocks < min_kb ? p.class_ntaps[c] * kblocks : min_kb;
    if (const int split = pick_split(cs.c_in, p_tiles * ((cs.c_in + 127) / 128), min_kb)) {
        CUtensorMap ma, mb;
        if (int rc = make_act_map(&ma, dy, cs.n, oh, ow, cs.c_out, p.a_cview, 1, p.bw, p.bh, p.bn)) return rc;
        if (int rc = make_filter_map(&mb, w, cs.c_out, cs.r * cs.s, cs.c_in, 64)) return rc;
        const long long tiles = p_tiles * ((cs.c_in + 127) / 128);
        const int rc = split == 4 ? launch_splitk<true, 4>(ma, mb, p, dx, nullptr, nullptr, tiles, (cudaStream_t)stream)
                                  : launch_splitk<true, 2>(ma, mb, p, dx, nullptr, nullptr, tiles, (cudaStream_t)stream);
        if (rc) return rc;
        DLS_LAUNCHED("conv_igemm_splitk_kernel<dgrad>");
        return 0;
    }
    const int block_n = pick_block_n(cs.c_in, p_tiles);
    CUtensorMap map_a, map_b;
    if (int rc = make_act_map(&map_a, dy, cs.n, oh, ow, cs.c_out, p.a_cview, 1, p.bw, p.bh, p.bn)) return rc;
    if (int rc = make_filter_map(&map_b, w, cs.c_out, cs.r * cs.s, cs.c_in, 64)) return rc;
    const long long tiles = p_tiles * ((cs.c_in + block_n - 1) / block_n);
    dim3 grid((unsigned)(tiles < kNumSMs ? tiles : kNumSMs));
    cudaStream_t st = (cudaStream_t)stream;
#define DLS_CONV_LAUNCH(BN)                                                                                   \
    do {                                                                                                      \
        const size_t smem = sizeof(ConvSmem<BN>) + 1024;                                                      \
        static thread_local bool configured = false;                                                          \
        if (!configured) {                                                                                    \
            if (int rc = set_smem(conv_igemm_kernel<BN, true>, smem)) return rc;                              \
            configured = true;                                                                                \
        }                                                                                                     \
        DLS_CHECK(launch_pdl(conv_igemm_kernel<BN, true>, grid, dim3(kConvWideThreads), smem, st, 0, map_a, map_b, p, \
                             (__nv_bfloat16*)dx, (const float*)nullptr, (float*)nullptr));                    \
    } while (0)
    if (block_n == 128) DLS_CONV_LAUNCH(128);
    else DLS_CONV_LAUNCH(64);
#undef DLS_CONV_LAUNCH
    DLS_LAUNCHED("conv_igemm_kernel<dgrad>");
    return 0;
}

extern "C" int dls_conv_wgrad(const void* dy, const void* x, void* dw_out, const DlsConvShape* shape, int mode,
                              void* stream) {
    // mode 0: store fp32, 1: accumulate into fp32 (may split the pixel range), 2: store bf16
    if (mode < 0 || mode > 2) return fail_msg("dls_conv_wgrad: mode must be 0, 1 or 2");
    float* dw = reinterpret_cast<float*>(dw_out);
    const int zeroed = mode == 1;
    const DlsConvShape cs = *shape;
    if (int rc = check_conv(cs, "wgrad")) return rc;
    if (((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dw) & 15) return fail_msg("dls_conv_wgrad: operands must be 16-byte aligned");
    const int oh = (cs.h + 2 * cs.pad_h - cs.r) / cs.stride + 1, ow = (cs.w + 2 * cs.pad_w - cs.s) / cs.stride + 1;
    WgradParams p = {};
    choose_box(ow, oh, cs.n, p.bw, p.bh, p.bn);
    p.tiles_w = (ow + p.bw - 1) / p.bw; p.tiles_h = (oh + p.bh - 1) / p.bh; p.tiles_n = (cs.n + p.bn - 1) / p.bn;
    p.c_out = cs.c_out; p.c_in = cs.c_in; p.x_cview = cs.x_pitch ? cs.x_pitch : cs.c_in;
    p.n_taps = p.taps_total = cs.r * cs.s;
    p.dw_cols = cs.dw_cols > 0 ? cs.dw_cols : cs.c_in;
    p.dw_pitch = cs.dw_pitch > 0 ? cs.dw_pitch : p.dw_cols;
    for (int r = 0; r < cs.r; ++r)
        for (int s = 0; s < cs.s; ++s) {
            ConvTap& t = p.taps[r * cs.s + s];
            const int qh = r - cs.pad_h, qw = s - cs.pad_w;
            t.dh = (int16_t)floor_div(qh, cs.stride); t.ph = (uint8_t)(qh - t.dh * cs.stride);
            t.dw = (int16_t)floor_div(qw, cs.stride); t.pw = (uint8_t)(qw - t.dw * cs.stride);
            t.widx = (uint16_t)(r * cs.s + s);
        }
    const int p_tiles = p.tiles_w * p.tiles_h * p.tiles_n;
    const int block_n = cs.c_in > 64 ? 128 : 64;
    const int out_tiles = ((cs.c_out + 127) / 128) * ((cs.c_in + block_n - 1) / block_n) * p.n_taps;
    // split the pixel range so that roughly two waves of CTAs exist, at least 4 pixel tiles per CTA
    int split = 1;
    if (zeroed) {
        split = (2 * kNumSMs + out_tiles - 1) / out_tiles;
        const int max_split = (p_tiles + 3) / 4;
        if (split > max_split) split = max_split;
        if (split < 1) split = 1;
    }
    p.tiles_per_split = (p_tiles + split - 1) / split;
    split = (p_tiles + p.tiles_per_split - 1) / p.tiles_per_split;
    p.split = split;
    p.atomic = zeroed ? 1 : 0;  // accumulate mode always adds (dw may hold a running sum)
    p.out_bf16 = mode == 2 ? 1 : 0;
    CUtensorMap map_dy, map_x;
    if (int rc = make_act_map(&map_dy, dy, cs.n, oh, ow, cs.c_out, cs.y_pitch ? cs.y_pitch : cs.c_out, 1, p.bw, p.bh, p.bn))
        return rc;
    if (int rc = make_act_map(&map_x, x, cs.n, cs.h, cs.w, cs.c_in, p.x_cview, cs.stride, p.bw, p.bh, p.bn)) return rc;
    const long long tiles = (long long)out_tiles * split;
    dim3 grid((unsigned)(tiles < kNumSMs ? tiles : kNumSMs));
    cudaStream_t st = (cudaStream_t)stream;
#define DLS_WGRAD_LAUNCH(BN, STAGES)                                                                          \
    do {                                                                                                      \
        const size_t smem = sizeof(WgradSmem<BN, STAGES>) + 1024;                                             \
        static thread_local bool configured = false;                                                          \
        if (!configured) {                                                                                    \
            if (int rc = set_smem(conv_wgrad_kernel<BN, STAGES>, smem)) return rc;                            \
            configured = true;                                                                                \
        }                                                                                                     \
        DLS_CHECK(launch_pdl(conv_wgrad_kernel<BN, STAGES>, grid, dim3(kConvThreads), smem, st, 0, map_dy, map_x, \
                             p, dw));                                                                         \
    } while (0)
    if (block_n == 128) DLS_WGRAD_LAUNCH(128, 3);
    else DLS_WGRAD_LAUNCH(64, 4);
#undef DLS_WGRAD_LAUNCH
    DLS_LAUNCHED("conv_wgrad_kernel");
    return 0;
}
