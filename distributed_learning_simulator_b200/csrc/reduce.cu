// Fused FedAvg reduction (+ decode at load, + broadcast in the epilogue).
//
// What the reference does per round (simulation_lib/algorithm/fed_avg_algorithm.py:39-81,
// message.py:41-49, server/server.py:114-120): every upload is pickled to the server, the
// delta is restored onto a deep copy of the old model, cast to fp64 on the CPU, multiplied
// by the dataset size and added tensor by tensor in a Python loop; the sum is divided,
// cast back, NaN-checked and then sent to every worker one pickle at a time.
//
// Here an upload is a [P] row in HBM (this GPU's or a peer's, mapped over NVLink).  One
// launch streams all K rows once, accumulates in fp64 registers, applies
// old + sum(w_k * row_k)/W, raises a NaN flag and stores the result into every
// destination buffer (local and peer: the broadcast).  With n_ranks > 1 each rank owns a
// [lo, hi) slice: the reduce-scatter + all-gather shape, with a device-side flag barrier
// (st.release.sys / ld.acquire.sys on peer-mapped words) on both sides, no NCCL.
#include "common.cuh"
#include "../../include/dls_b200.h"

namespace dls {

std::atomic<long long> g_launch_count{0};
thread_local char g_last_error[512] = "";

// ---- device-side barrier over ranks ------------------------------------------------
// flag block layout per rank (32 uint32 words): arrive[2][8] (phase 0 = "my uploads are
// ready", phase 1 = "my slice is written everywhere"), nan[8] at word 16, timeout at word 24; rank r writes word [phase][r] of EVERY
// rank's block, and waits until its own block shows `epoch` in all n_ranks words.
__device__ __forceinline__ void barrier_signal(uint32_t* const* peer_flags, int n_ranks, int rank, int phase,
                                               uint32_t epoch) {
    __threadfence_system();
    for (int r = 0; r < n_ranks; ++r) st_release_sys(peer_flags[r] + phase * 8 + rank, epoch);
}
__device__ __forceinline__ void barrier_wait(uint32_t* const* peer_flags, int n_ranks, int rank, int phase,
                                             uint32_t epoch) {
    const uint32_t* mine = peer_flags[rank] + phase * 8;
    // bounded wait (~20 s at 2 GHz): a dead peer must surface as an error, not hang the GPU
    const long long t0 = clock64();
    for (int r = 0; r < n_ranks; ++r) {
        while ((int32_t)(ld_acquire_sys(mine + r) - epoch) < 0) {
            __nanosleep(64);
            if (clock64() - t0 > 40000000000LL) {
                atomicOr(peer_flags[rank] + 24, 1u);  // word 24: "barrier timed out"
                return;
            }
        }
    }
}

// NVLS: one store to a multicast address lands in the same offset of every GPU bound to the multicast
// object (the NVSwitch replicates it), so the broadcast costs one NVLink write instead of one per peer.
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_st_f1(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

constexpr int kReduceThreads = 256;
constexpr int kMaxDsts = 8;
// "last block out" counters of the cross-GPU reduce, one slot per launch in flight (module globals exist once per
// device, need no allocation and are therefore safe under stream capture)
constexpr int kDoneSlots = 64;
__device__ unsigned int g_done_counters[kDoneSlots];

struct DstTable { float* p[kMaxDsts]; };

template <bool kPeer>
__global__ void __launch_bounds__(kReduceThreads, 2)
fedavg_reduce_f32_kernel(const float* const* __restrict__ rows, const float* __restrict__ weights, int n_rows,
                         const float* __restrict__ old, float old_scale, double inv_total, DstTable dsts,
                         int n_dsts, float* __restrict__ mc_dst, long long lo, long long hi, uint32_t* const* peer_flags, int n_ranks, int rank,
                         uint32_t epoch, uint32_t* __restrict__ nan_flag, int done_slot) {
    extern __shared__ unsigned char smem_raw[];
    const float** s_rows = reinterpret_cast<const float**>(smem_raw);
    float* s_w = reinterpret_cast<float*>(s_rows + n_rows);
    for (int k = threadIdx.x; k < n_rows; k += blockDim.x) {
        s_rows[k] = rows[k];
        s_w[k] = weights[k];
    }
    if (kPeer) {
        if (blockIdx.x == 0 && threadIdx.x == 0) barrier_signal(peer_flags, n_ranks, rank, 0, epoch);
        if (threadIdx.x == 0) barrier_wait(peer_flags, n_ranks, rank, 0, epoch);
    }
    __syncthreads();

    bool saw_nan = false;
    const long long n = hi - lo;
    const long long n_vec = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    // [lo, hi) is 4-aligned on both ends except for the global tail, handled below
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n_rows <= 4) {
        // few rows (e.g. one worker per GPU on 2-4 GPUs): keep 4 positions x n_rows 128-bit loads
        // in flight per thread -- peer loads over NVLink have ~2 us latency
        for (; v + 3 * stride < n_vec; v += 4 * stride) {
            double acc[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { acc[u][0] = acc[u][1] = acc[u][2] = acc[u][3] = 0.0; }
            for (int k = 0; k < n_rows; ++k) {
                const float* row = s_rows[k] + lo;
                float4 x[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    x[u] = kPeer ? ld_peer_f4(row + ((v + u * stride) << 2)) : ld_stream_f4(row + ((v + u * stride) << 2));
                const double w = s_w[k];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    acc[u][0] += w * x[u].x; acc[u][1] += w * x[u].y; acc[u][2] += w * x[u].z; acc[u][3] += w * x[u].w;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long i = lo + ((v + u * stride) << 2);
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                if (old != nullptr) o = ld_stream_f4(old + i);
                float4 r;
                r.x = (float)((double)old_scale * o.x + acc[u][0] * inv_total);
                r.y = (float)((double)old_scale * o.y + acc[u][1] * inv_total);
                r.z = (float)((double)old_scale * o.z + acc[u][2] * inv_total);
                r.w = (float)((double)old_scale * o.w + acc[u][3] * inv_total);
                saw_nan |= (r.x != r.x) | (r.y != r.y) | (r.z != r.z) | (r.w != r.w);
                if (kPeer && mc_dst != nullptr) {
                    multimem_st_f4(mc_dst + i, r);
                } else {
#pragma unroll
                    for (int d = 0; d < kMaxDsts; ++d)
                        if (d < n_dsts) st_stream_f4(dsts.p[d] + i, r);
                }
            }
        }
    }
    for (; v < n_vec; v += stride) {
        const long long i = lo + (v << 2);
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int k = 0;
        for (; k + 4 <= n_rows; k += 4) {  // 4 independent 128-bit loads in flight per thread
            float4 x0 = kPeer ? ld_peer_f4(s_rows[k] + i) : ld_stream_f4(s_rows[k] + i);
            float4 x1 = kPeer ? ld_peer_f4(s_rows[k + 1] + i) : ld_stream_f4(s_rows[k + 1] + i);
            float4 x2 = kPeer ? ld_peer_f4(s_rows[k + 2] + i) : ld_stream_f4(s_rows[k + 2] + i);
            float4 x3 = kPeer ? ld_peer_f4(s_rows[k + 3] + i) : ld_stream_f4(s_rows[k + 3] + i);
            const double w0 = s_w[k], w1 = s_w[k + 1], w2 = s_w[k + 2], w3 = s_w[k + 3];
            a0 += w0 * x0.x + w1 * x1.x + w2 * x2.x + w3 * x3.x;
            a1 += w0 * x0.y + w1 * x1.y + w2 * x2.y + w3 * x3.y;
            a2 += w0 * x0.z + w1 * x1.z + w2 * x2.z + w3 * x3.z;
            a3 += w0 * x0.w + w1 * x1.w + w2 * x2.w + w3 * x3.w;
        }
        for (; k < n_rows; ++k) {
            float4 x = kPeer ? ld_peer_f4(s_rows[k] + i) : ld_stream_f4(s_rows[k] + i);
            const double w = s_w[k];
            a0 += w * x.x; a1 += w * x.y; a2 += w * x.z; a3 += w * x.w;
        }
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
        if (old != nullptr) o = ld_stream_f4(old + i);
        float4 r;
        r.x = (float)((double)old_scale * o.x + a0 * inv_total);
        r.y = (float)((double)old_scale * o.y + a1 * inv_total);
        r.z = (float)((double)old_scale * o.z + a2 * inv_total);
        r.w = (float)((double)old_scale * o.w + a3 * inv_total);
        saw_nan |= (r.x != r.x) | (r.y != r.y) | (r.z != r.z) | (r.w != r.w);
        if (kPeer && mc_dst != nullptr) {
            multimem_st_f4(mc_dst + i, r);
        } else {
#pragma unroll
            for (int d = 0; d < kMaxDsts; ++d)
                if (d < n_dsts) st_stream_f4(dsts.p[d] + i, r);
        }
    }
    // scalar tail (at most 3 elements, only in the slice that ends at P)
    for (long long i = lo + (n_vec << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
        double a = 0.0;
        // peer rows were written by other GPUs: read them past the (non-coherent) L1 like the vector path does
        for (int k = 0; k < n_rows; ++k)
            a += (double)s_w[k] * (kPeer ? *reinterpret_cast<const volatile float*>(s_rows[k] + i) : s_rows[k][i]);
        float r = (float)((double)old_scale * (old ? old[i] : 0.f) + a * inv_total);
        saw_nan |= (r != r);
        if (kPeer && mc_dst != nullptr) multimem_st_f1(mc_dst + i, r);
        else
            for (int d = 0; d < n_dsts; ++d) dsts.p[d][i] = r;
    }
    if (__syncthreads_or(saw_nan) && threadIdx.x == 0) atomicOr(nan_flag, 1u);

    if (kPeer) {
        // last block out tells every rank "my slice has landed", then waits for theirs so
        // that stream order makes the complete model visible to whatever is enqueued next
        __shared__ bool is_last;
        __threadfence_system();
        __syncthreads();
        unsigned int* done_counter = &g_done_counters[done_slot];
        if (threadIdx.x == 0) is_last = (atomicAdd(done_counter, 1u) == gridDim.x - 1);
        __syncthreads();
        if (is_last && threadIdx.x == 0) {
            *done_counter = 0;
            // publish this slice's NaN verdict to every rank (words 16..23) so all replicas agree
            const uint32_t verdict = *reinterpret_cast<volatile uint32_t*>(nan_flag);
            for (int r = 0; r < n_ranks; ++r) peer_flags[r][16 + rank] = verdict;
            barrier_signal(peer_flags, n_ranks, rank, 1, epoch);
            barrier_wait(peer_flags, n_ranks, rank, 1, epoch);
        }
    }
}

// ---- packed / masked variant -------------------------------------------------------
__device__ __forceinline__ void decode_group(const DlsPackedRow& row, int t, long long t_off, long long j0, int cnt,
                                             float (&x)[8]) {
    if (row.kind == 0) {
        const float* p = reinterpret_cast<const float*>(row.payload) + t_off + j0;
#pragma unroll
        for (int e = 0; e < 8; ++e) x[e] = e < cnt ? p[e] : 0.f;
    } else if (row.kind == 1) {
        const uint8_t* q = reinterpret_cast<const uint8_t*>(row.payload) + t_off + j0;
        const float s = row.scale[t];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (e < cnt) {
                const long long g = t_off + j0 + e;
                const bool neg = (row.aux[g >> 3] >> (g & 7)) & 1;
                const float m = (float)q[e] * s;
                x[e] = neg ? -m : m;
            } else x[e] = 0.f;
        }
    } else {
        const int b = row.bits[t];
        const uint8_t* p = reinterpret_cast<const uint8_t*>(row.payload) + row.byte_offsets[t] + (j0 >> 3) * b;
        // b bytes hold 8 codes of b bits; assemble as two 64-bit words (b <= 16)
        unsigned long long w0 = 0, w1 = 0;
        for (int i = 0; i < b; ++i) {
            const unsigned long long byte = p[i];
            if (i < 8) w0 |= byte << (8 * i); else w1 |= byte << (8 * (i - 8));
        }
        const float s = row.scale[t], o = row.offset[t];
        const unsigned int mask = (1u << b) - 1u;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int bit = e * b;
            unsigned long long v;
            if (bit >= 64) v = w1 >> (bit - 64);
            else if (bit + b <= 64) v = w0 >> bit;
            else v = (w0 >> bit) | (w1 << (64 - bit));
            x[e] = e < cnt ? (float)((unsigned int)v & mask) * s + o : 0.f;
        }
    }
}

__global__ void __launch_bounds__(256, 2)
fedavg_reduce_packed_kernel(const DlsPackedRow* __restrict__ rows, int n_rows, const float* __restrict__ old,
                            const long long* __restrict__ tensor_offsets,
                            const long long* __restrict__ group_offsets, int n_tensors, int nonzero_denominator,
                            float* __restrict__ out, uint32_t* __restrict__ nan_flag) {
    extern __shared__ long long s_tab[];  // group_offsets[T+1] then tensor_offsets[T+1]
    long long* s_goff = s_tab;
    long long* s_toff = s_tab + n_tensors + 1;
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) {
        s_goff[i] = group_offsets[i];
        s_toff[i] = tensor_offsets[i];
    }
    __syncthreads();
    const long long n_groups = s_goff[n_tensors];
    bool saw_nan = false;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
         g += (long long)gridDim.x * blockDim.x) {
        const int t = find_segment(s_goff, n_tensors, g);
        const long long t_off = s_toff[t];
        const long long j0 = (g - s_goff[t]) << 3;
        const long long n_t = s_toff[t + 1] - t_off;
        const int cnt = (int)min((long long)8, n_t - j0);
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (old != nullptr && e < cnt) ? old[t_off + j0 + e] : 0.f;
        double acc[8], den[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { acc[e] = 0.0; den[e] = 0.0; }
        for (int k = 0; k < n_rows; ++k) {
            const DlsPackedRow row = rows[k];
            const double w = row.weight;
            float x[8];
            if (row.present != nullptr && !row.present[t]) {
                // FedOBD: a dropped block contributes the old value (ParameterMessage.complete)
#pragma unroll
                for (int e = 0; e < 8; ++e) { acc[e] += w * o[e]; den[e] += w; }
                continue;
            }
            decode_group(row, t, t_off, j0, cnt, x);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float v = row.is_delta ? o[e] + x[e] : x[e];
                if (nonzero_denominator) {
                    // FedDropoutAvg: an element weighs in only where the upload is non-zero
                    const double we = (v != 0.f) ? w : 0.0;
                    acc[e] += we * v; den[e] += we;
                } else {
                    acc[e] += w * v; den[e] += w;
                }
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (e < cnt) {
                const double d = den[e] == 0.0 ? 1.0 : den[e];
                const float r = (float)(acc[e] / d);
                saw_nan |= (r != r);
                out[t_off + j0 + e] = r;
            }
        }
    }
    if (__syncthreads_or(saw_nan) && threadIdx.x == 0) atomicOr(nan_flag, 1u);
}

}  // namespace dls

using namespace dls;

extern "C" const char* dls_last_error(void) { return g_last_error; }
extern "C" long long dls_launch_count(void) { return g_launch_count.load(); }
extern "C" void dls_reset_launch_count(void) { g_launch_count.store(0); }


static int fedavg_reduce_f32_impl(const float* const* rows, const float* weights, int n_rows, const float* old,
                                  float old_scale, double inv_total, float* const* dsts, int n_dsts, float* mc_dst,
                                  long long lo, long long hi, uint32_t* const* peer_flags, int n_ranks, int rank,
                                  uint32_t epoch, uint32_t* nan_flag, void* stream) {
    if (n_dsts < 1 || n_dsts > kMaxDsts) return fail_msg("dls_fedavg_reduce_f32: n_dsts must be in [1, 8]");
    if (n_rows < 1 || n_rows > 4096) return fail_msg("dls_fedavg_reduce_f32: n_rows must be in [1, 4096]");
    if (hi < lo || (lo & 3)) return fail_msg("dls_fedavg_reduce_f32: bad [lo, hi) (lo must be 4-aligned)");
    DstTable table;
    for (int d = 0; d < kMaxDsts; ++d) table.p[d] = d < n_dsts ? dsts[d] : nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem = (size_t)n_rows * (sizeof(float*) + sizeof(float));
    const long long n_vec = ((hi - lo) >> 2) + 1;
    const bool peer = peer_flags != nullptr && n_ranks > 1;
    if (peer) {
        // last-block counter of this launch: a slot of the per-device pool, so reductions that overlap on one GPU
        // (two tasks, two streams) never share one; the kernel leaves its slot at zero
        static std::atomic<unsigned int> next_slot{0};
        const int done_slot = (int)(next_slot.fetch_add(1) % kDoneSlots);
        // every block must be co-resident: it spins on the peers' flags
        int per_sm = 0;
        DLS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fedavg_reduce_f32_kernel<true>,
                                                                kReduceThreads, smem));
        if (per_sm < 1) return fail_msg("dls_fedavg_reduce_f32: kernel does not fit on an SM");
        const int grid = grid_for(n_vec, kReduceThreads, per_sm > 2 ? 2 : per_sm);
        fedavg_reduce_f32_kernel<true><<<grid, kReduceThreads, smem, s>>>(
            rows, weights, n_rows, old, old_scale, inv_total, table, n_dsts, mc_dst, lo, hi, peer_flags, n_ranks, rank,
            epoch, nan_flag, done_slot);
    } else {
        const int grid = grid_for(n_vec, kReduceThreads, 4);
        fedavg_reduce_f32_kernel<false><<<grid, kReduceThreads, smem, s>>>(
            rows, weights, n_rows, old, old_scale, inv_total, table, n_dsts, nullptr, lo, hi, nullptr, 1, 0, 0, nan_flag,
            0);
    }
    DLS_LAUNCHED("fedavg_reduce_f32_kernel");
    return 0;
}

extern "C" int dls_fedavg_reduce_f32(const float* const* rows, const float* weights, int n_rows, const float* old,
                                     float old_scale, double inv_total, float* const* dsts, int n_dsts,
                                     long long lo, long long hi, uint32_t* const* peer_flags, int n_ranks, int rank,
                                     uint32_t epoch, uint32_t* nan_flag, void* stream) {
    return fedavg_reduce_f32_impl(rows, weights, n_rows, old, old_scale, inv_total, dsts, n_dsts, nullptr, lo, hi,
                                  peer_flags, n_ranks, rank, epoch, nan_flag, stream);
}

extern "C" int dls_fedavg_reduce_f32_mc(const float* const* rows, const float* weights, int n_rows, const float* old,
                                        float old_scale, double inv_total, float* mc_dst, long long lo, long long hi,
                                        uint32_t* const* peer_flags, int n_ranks, int rank, uint32_t epoch,
                                        uint32_t* nan_flag, void* stream) {
    if (mc_dst == nullptr || peer_flags == nullptr || n_ranks < 2)
        return fail_msg("dls_fedavg_reduce_f32_mc: needs a multicast destination and the peers' flag blocks");
    if ((uintptr_t)mc_dst & 15) return fail_msg("dls_fedavg_reduce_f32_mc: destination must be 16-byte aligned");
    float* none[1] = {mc_dst};  // the table is unused on the multicast path; keep n_dsts valid
    return fedavg_reduce_f32_impl(rows, weights, n_rows, old, old_scale, inv_total, none, 1, mc_dst, lo, hi, peer_flags,
                                  n_ranks, rank, epoch, nan_flag, stream);
}

// K9 on its own: copy n floats from src to the same offset of every GPU bound to the multicast address
namespace dls {
__global__ void __launch_bounds__(256)
multimem_bcast_kernel(const float* __restrict__ src, float* __restrict__ mc_dst, long long n) {
    const long long n_vec = n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n_vec; v += stride)
        multimem_st_f4(mc_dst + (v << 2), ld_stream_f4(src + (v << 2)));
    for (long long i = (n_vec << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        multimem_st_f1(mc_dst + i, src[i]);
    __threadfence_system();
}
}  // namespace dls

extern "C" int dls_multimem_bcast_f32(const float* src, float* mc_dst, long long n, void* stream) {
    if (n < 1 || mc_dst == nullptr) return fail_msg("dls_multimem_bcast_f32: empty");
    if (((uintptr_t)src | (uintptr_t)mc_dst) & 15) return fail_msg("dls_multimem_bcast_f32: 16-byte alignment");
    multimem_bcast_kernel<<<grid_for((n >> 2) + 1, 256, 4), 256, 0, (cudaStream_t)stream>>>(src, mc_dst, n);
    DLS_LAUNCHED("multimem_bcast_kernel");
    return 0;
}

extern "C" int dls_fedavg_reduce_packed(const DlsPackedRow* rows, int n_rows, const float* old,
                                        const long long* tensor_offsets, const long long* group_offsets,
                                        int n_tensors, int nonzero_denominator, float* out, uint32_t* nan_flag,
                                        void* stream) {
    if (n_rows < 1) return fail_msg("dls_fedavg_reduce_packed: no rows");
    if (n_tensors < 1 || n_tensors > 2900) return fail_msg("dls_fedavg_reduce_packed: n_tensors must be in [1, 2900]");
    const size_t smem = 2 * (size_t)(n_tensors + 1) * sizeof(long long);
    const int grid = kNumSMs * 4;
    fedavg_reduce_packed_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(
        rows, n_rows, old, tensor_offsets, group_offsets, n_tensors, nonzero_denominator, out, nan_flag);
    DLS_LAUNCHED("fedavg_reduce_packed_kernel");
    return 0;
}
