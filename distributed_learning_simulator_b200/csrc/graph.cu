// Graph-family kernels (fed_gnn / fed_gcn / fed_aas): SURVEY K13 (edge / node mask + ordered
// compaction, optional Bernoulli edge drop) and K14 (boundary-embedding gather, halo scatter-merge),
// plus the normalised GCN aggregation that the message-passing layers spend their time in.
//
// Reference behaviour: /root/reference/simulation_lib/worker/graph_worker.py:205-216 (edge masks
// per mini-batch), :252-269 (boundary rows picked with python loops over n_id.tolist()), :271-321
// (merge of received embeddings with torch.where + per-row assignment).  Here every one of those is
// one or two launches over index maps that never leave the device.
#include <cuda_bf16.h>

#include "common.cuh"

namespace dls {

constexpr int kCompactThreads = 256;
constexpr int kCompactItems = 8;  // per thread
constexpr int kCompactTile = kCompactThreads * kCompactItems;

// keep[i] for pair i: mask[map(a[i])] && (b == nullptr || mask[map(b[i])]) && Bernoulli(keep_prob)
__device__ __forceinline__ bool compact_keep(const long long* __restrict__ a, const long long* __restrict__ b,
                                             const long long* __restrict__ n_id, const uint8_t* __restrict__ mask,
                                             long long i, float keep_prob, unsigned long long seed) {
    long long va = a[i];
    if (n_id != nullptr) va = n_id[va];
    bool keep = mask[va] != 0;
    if (keep && b != nullptr) {
        long long vb = b[i];
        if (n_id != nullptr) vb = n_id[vb];
        keep = mask[vb] != 0;
    }
    if (keep && keep_prob < 1.f) {
        const uint4 r = philox4x32((uint32_t)i, (uint32_t)(i >> 32), 13u, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
        keep = u32_to_unit(r.x) < keep_prob;
    }
    return keep;
}

__global__ void __launch_bounds__(kCompactThreads)
compact_count_kernel(const long long* __restrict__ a, const long long* __restrict__ b,
                     const long long* __restrict__ n_id, const uint8_t* __restrict__ mask, long long n,
                     float keep_prob, unsigned long long seed, int* __restrict__ block_counts) {
    const long long base = (long long)blockIdx.x * kCompactTile;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
        const long long i = base + (long long)k * kCompactThreads + threadIdx.x;
        if (i < n) mine += compact_keep(a, b, n_id, mask, i, keep_prob, seed) ? 1 : 0;
    }
    __shared__ int s_warp[kCompactThreads / 32];
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        int total = 0;
        for (int w = 0; w < kCompactThreads / 32; ++w) total += s_warp[w];
        block_counts[blockIdx.x] = total;
    }
}

// Ordered scatter: block b starts at sum(block_counts[0..b)); inside the tile, thread t owns the
// kCompactItems CONSECUTIVE pairs base + t*kCompactItems .. so a warp/block exclusive scan of the
// per-thread counts gives a stable order.
__global__ void __launch_bounds__(kCompactThreads)
compact_scatter_kernel(const long long* __restrict__ a, const long long* __restrict__ b,
                       const long long* __restrict__ n_id, const uint8_t* __restrict__ mask, long long n,
                       float keep_prob, unsigned long long seed, const int* __restrict__ block_counts,
                       long long* __restrict__ out_a, long long* __restrict__ out_b,
                       long long* __restrict__ out_pos, long long* __restrict__ count) {
    __shared__ long long s_red[kCompactThreads / 32];
    __shared__ int s_scan[kCompactThreads / 32];
    __shared__ long long s_base;
    // (1) this block's global offset = sum of the counts of the blocks before it
    long long before = 0;
    for (int j = threadIdx.x; j < (int)blockIdx.x; j += kCompactThreads) before += block_counts[j];
    for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = before;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long t = 0;
        for (int w = 0; w < kCompactThreads / 32; ++w) t += s_red[w];
        s_base = t;
        if (blockIdx.x == gridDim.x - 1) *count = t + block_counts[blockIdx.x];
    }
    // (2) flags of this thread's consecutive items
    const long long first = (long long)blockIdx.x * kCompactTile + (long long)threadIdx.x * kCompactItems;
    unsigned flags = 0;
    int mine = 0;
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
        const long long i = first + k;
        if (i < n && compact_keep(a, b, n_id, mask, i, keep_prob, seed)) {
            flags |= 1u << k;
            ++mine;
        }
    }
    // (3) block exclusive scan of per-thread counts
    int incl = mine;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    int warp_off = 0;
    for (int w = 0; w < warp; ++w) warp_off += s_scan[w];
    long long dst = s_base + warp_off + (incl - mine);
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
        if (flags & (1u << k)) {
            const long long i = first + k;
            out_a[dst] = a[i];
            if (out_b != nullptr) out_b[dst] = b[i];
            if (out_pos != nullptr) out_pos[dst] = i;
            ++dst;
        }
    }
}

// ---- row gather (K14a): out[k, :] = x[rows[k], :], rows of row_words 32-bit words -------------------
__global__ void __launch_bounds__(256)
row_gather_kernel(const uint32_t* __restrict__ x, const long long* __restrict__ rows, long long n_rows,
                  int row_words, uint32_t* __restrict__ out) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (long long k = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); k < n_rows;
         k += (long long)gridDim.x * warps_per_block) {
        const uint32_t* src = x + rows[k] * row_words;
        uint32_t* dst = out + k * row_words;
        if ((row_words & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
            for (int j = lane; j < row_words / 4; j += 32)
                reinterpret_cast<uint4*>(dst)[j] = reinterpret_cast<const uint4*>(src)[j];
        } else {
            for (int j = lane; j < row_words; j += 32) dst[j] = src[j];
        }
    }
}

// ---- halo merge (K14b) -----------------------------------------------------------------------------
// For row r of a mini-batch (global node id g = n_id[r]):
//   local training node  -> out[r] = x[r]
//   foreign node         -> out[r] = emb[perm[j]] if sorted_ids[j] == g exists, else 0
// foreign[r] is written for the backward pass (gradient flows through local rows only).
// With emb == nullptr and foreign_in != nullptr the kernel is the backward: out[r] = foreign ? 0 : x[r].
__global__ void __launch_bounds__(256)
halo_merge_kernel(const uint32_t* __restrict__ x, const long long* __restrict__ n_id, long long n_rows,
                  int row_words, const uint8_t* __restrict__ train_mask, const long long* __restrict__ sorted_ids,
                  const long long* __restrict__ perm, long long n_ids, const uint32_t* __restrict__ emb,
                  const uint8_t* __restrict__ foreign_in, uint8_t* __restrict__ foreign_out,
                  uint32_t* __restrict__ out) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (long long r = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); r < n_rows;
         r += (long long)gridDim.x * warps_per_block) {
        const uint32_t* src = x + r * row_words;
        bool zero = false;
        if (foreign_in != nullptr) {
            zero = foreign_in[r] != 0;
        } else {
            const long long g = n_id[r];
            const bool foreign = train_mask[g] == 0;
            if (lane == 0 && foreign_out != nullptr) foreign_out[r] = foreign ? 1 : 0;
            if (foreign) {
                zero = true;
                if (emb != nullptr && n_ids > 0) {
                    long long lo = 0, hi = n_ids;  // lower_bound
                    while (lo < hi) {
                        const long long mid = (lo + hi) >> 1;
                        if (sorted_ids[mid] < g) lo = mid + 1; else hi = mid;
                    }
                    if (lo < n_ids && sorted_ids[lo] == g) {
                        src = emb + perm[lo] * row_words;
                        zero = false;
                    }
                }
            }
        }
        uint32_t* dst = out + r * row_words;
        for (int j = lane; j < row_words; j += 32) dst[j] = zero ? 0u : src[j];
    }
}

// ---- GCN aggregation ---------------------------------------------------------------------------------
// out = D^-1/2 (A + I) D^-1/2 h (+ bias), deg counted over destination nodes incl. the self loop.
__global__ void gcn_degree_kernel(const long long* __restrict__ dst, long long n_edges, int* __restrict__ deg) {
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_edges;
         e += (long long)gridDim.x * blockDim.x)
        atomicAdd(deg + dst[e], 1);
}
__global__ void gcn_dinv_kernel(const int* __restrict__ deg, long long n_nodes, float* __restrict__ dinv) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes;
         i += (long long)gridDim.x * blockDim.x)
        dinv[i] = rsqrtf((float)(deg[i] + 1));
}
// self-loop term initialises the output (so no memset): out[i] = h[i] * dinv[i]^2 (+ bias)
__global__ void __launch_bounds__(256)
gcn_self_kernel(const float* __restrict__ h, const float* __restrict__ dinv, const float* __restrict__ bias,
                long long n_nodes, int channels, float* __restrict__ out) {
    const long long total = n_nodes * channels;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (long long)gridDim.x * blockDim.x) {
        const long long i = k / channels;
        const float d = dinv[i];
        out[k] = h[k] * d * d + (bias != nullptr ? bias[k - i * channels] : 0.f);
    }
}
// one warp per edge: out[to] += h[from] * dinv[from] * dinv[to]; 128-bit vector reductions when aligned
__global__ void __launch_bounds__(256)
gcn_edge_kernel(const float* __restrict__ h, const long long* __restrict__ from, const long long* __restrict__ to,
                long long n_edges, int channels, const float* __restrict__ dinv, float* __restrict__ out) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    const bool vec = (channels & 3) == 0 &&
                     ((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    for (long long e = (long long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); e < n_edges;
         e += (long long)gridDim.x * warps_per_block) {
        const long long s = from[e], d = to[e];
        const float w = dinv[s] * dinv[d];
        const float* src = h + s * channels;
        float* dst = out + d * channels;
        if (vec) {
            for (int j = lane; j < channels / 4; j += 32) {
                const float4 v = reinterpret_cast<const float4*>(src)[j];
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 4 * j), "f"(v.x * w),
                             "f"(v.y * w), "f"(v.z * w), "f"(v.w * w)
                             : "memory");
            }
        } else {
            for (int j = lane; j < channels; j += 32) atomicAdd(dst + j, src[j] * w);
        }
    }
}

}  // namespace dls

using namespace dls;

extern "C" long long dls_index_compact_scratch_ints(long long n) { return (n + kCompactTile - 1) / kCompactTile + 1; }

extern "C" int dls_index_compact(const long long* a, const long long* b, long long n, const long long* n_id,
                                 const uint8_t* mask, float keep_prob, unsigned long long seed, long long* out_a,
                                 long long* out_b, long long* out_pos, long long* count, int* scratch, void* stream) {
    if (n < 0 || (b != nullptr) != (out_b != nullptr)) return fail_msg("dls_index_compact: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        DLS_CHECK(cudaMemsetAsync(count, 0, sizeof(long long), s));
        return 0;
    }
    const long long blocks = (n + kCompactTile - 1) / kCompactTile;
    if (blocks > 0x7fffffffLL) return fail_msg("dls_index_compact: too many items");
    compact_count_kernel<<<(unsigned)blocks, kCompactThreads, 0, s>>>(a, b, n_id, mask, n, keep_prob, seed, scratch);
    DLS_LAUNCHED("compact_count_kernel");
    compact_scatter_kernel<<<(unsigned)blocks, kCompactThreads, 0, s>>>(a, b, n_id, mask, n, keep_prob, seed, scratch,
                                                                       out_a, out_b, out_pos, count);
    DLS_LAUNCHED("compact_scatter_kernel");
    return 0;
}

extern "C" int dls_row_gather(const void* x, const long long* rows, long long n_rows, long long row_bytes, void* out,
                              void* stream) {
    if (n_rows < 0 || row_bytes < 4 || row_bytes % 4) return fail_msg("dls_row_gather: rows must be whole 32-bit words");
    if (n_rows == 0) return 0;
    const int grid = grid_for(n_rows * 32, 256, 8);
    row_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)x, rows, n_rows, (int)(row_bytes / 4),
                                                              (uint32_t*)out);
    DLS_LAUNCHED("row_gather_kernel");
    return 0;
}

extern "C" int dls_halo_merge(const void* x, const long long* n_id, long long n_rows, long long row_bytes,
                              const uint8_t* train_mask, const long long* sorted_ids, const long long* perm,
                              long long n_ids, const void* emb, const uint8_t* foreign_in, uint8_t* foreign_out,
                              void* out, void* stream) {
    if (n_rows < 0 || row_bytes < 4 || row_bytes % 4) return fail_msg("dls_halo_merge: rows must be whole 32-bit words");
    if ((foreign_in == nullptr) && (n_id == nullptr || train_mask == nullptr))
        return fail_msg("dls_halo_merge: forward needs n_id and the training-node mask");
    if (n_rows == 0) return 0;
    const int grid = grid_for(n_rows * 32, 256, 8);
    halo_merge_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)x, n_id, n_rows, (int)(row_bytes / 4),
                                                              train_mask, sorted_ids, perm, n_ids,
                                                              (const uint32_t*)emb, foreign_in, foreign_out,
                                                              (uint32_t*)out);
    DLS_LAUNCHED("halo_merge_kernel");
    return 0;
}

extern "C" int dls_gcn_dinv(const long long* dst, long long n_edges, long long n_nodes, int* deg_scratch, float* dinv,
                            void* stream) {
    if (n_nodes < 1 || n_edges < 0) return fail_msg("dls_gcn_dinv: empty graph");
    cudaStream_t s = (cudaStream_t)stream;
    DLS_CHECK(cudaMemsetAsync(deg_scratch, 0, sizeof(int) * (size_t)n_nodes, s));
    if (n_edges > 0) {
        gcn_degree_kernel<<<grid_for(n_edges, 256, 8), 256, 0, s>>>(dst, n_edges, deg_scratch);
        DLS_LAUNCHED("gcn_degree_kernel");
    }
    gcn_dinv_kernel<<<grid_for(n_nodes, 256, 8), 256, 0, s>>>(deg_scratch, n_nodes, dinv);
    DLS_LAUNCHED("gcn_dinv_kernel");
    return 0;
}

extern "C" int dls_gcn_aggregate(const float* h, const long long* from, const long long* to, long long n_edges,
                                 long long n_nodes, int channels, const float* dinv, const float* bias, float* out,
                                 void* stream) {
    if (n_nodes < 1 || channels < 1 || n_edges < 0) return fail_msg("dls_gcn_aggregate: empty input");
    cudaStream_t s = (cudaStream_t)stream;
    gcn_self_kernel<<<grid_for(n_nodes * channels, 256, 8), 256, 0, s>>>(h, dinv, bias, n_nodes, channels, out);
    DLS_LAUNCHED("gcn_self_kernel");
    if (n_edges > 0) {
        gcn_edge_kernel<<<grid_for(n_edges * 32, 256, 8), 256, 0, s>>>(h, from, to, n_edges, channels, dinv, out);
        DLS_LAUNCHED("gcn_edge_kernel");
    }
    return 0;
}
