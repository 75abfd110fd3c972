// Parameter quantizers over the flat arena (FedPAQ / fed_obd_sq: 255-level stochastic;
// FedOBD: NNADQ adaptive deterministic).
//
// The reference calls external Python/ATen quantizers once per tensor inside the endpoint
// (simulation_lib/topology/quantized_endpoint.py:29-34,58-68,76,82,89).  Here a model is
// quantized in a fixed number of launches regardless of how many tensors it has:
// a segmented statistics pass (per-tensor abs-max, or min/max/sum/sum-of-squares), a
// one-block pass that turns statistics into per-tensor scale / bit width / byte offsets,
// and a pack pass.  Bit layout is defined in algorithm/quantization/packed.py and is what
// fedavg_reduce_packed decodes at load time.
#include "common.cuh"
#include "../../include/dls_b200.h"

namespace dls {

__device__ __forceinline__ int float_to_ordered(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_to_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

// ---------------------------------------------------------------- stochastic (sq)
// Segmented reductions over the flat arena share one access pattern: a warp walks a contiguous
// span of kSpanIters * 32 elements, every lane keeps a private partial for the tensor it is
// currently inside (flushing it with an atomic only when it crosses a tensor boundary), and at
// the end of the span the lanes combine with shuffles and issue ONE atomic when they all sit in
// the same tensor -- a few thousand atomics per model instead of one per 32 elements.
constexpr int kSpanIters = 64;

__global__ void __launch_bounds__(256)
sq_absmax_kernel(const float* __restrict__ flat, long long n, const long long* __restrict__ tensor_offsets,
                 int n_tensors, uint32_t* __restrict__ absmax_bits) {
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long base = warp_id * (kSpanIters * 32); base < n; base += n_warps * (kSpanIters * 32)) {
        int t_cur = -1;
        long long t_end = -1;
        float acc = 0.f;
        for (int j = 0; j < kSpanIters; ++j) {
            const long long i = base + (long long)j * 32 + lane;
            if (i >= n) break;
            if (i >= t_end) {
                if (t_cur >= 0) atomicMax(absmax_bits + t_cur, __float_as_uint(acc));
                t_cur = find_segment(s_off, n_tensors, i);
                t_end = s_off[t_cur + 1];
                acc = 0.f;
            }
            acc = fmaxf(acc, fabsf(flat[i]));
        }
        const int t0 = __shfl_sync(0xffffffffu, t_cur, 0);
        if (__all_sync(0xffffffffu, t_cur == t0) && t0 >= 0) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc = fmaxf(acc, __shfl_xor_sync(0xffffffffu, acc, o));
            if (lane == 0) atomicMax(absmax_bits + t0, __float_as_uint(acc));
        } else if (t_cur >= 0) {
            atomicMax(absmax_bits + t_cur, __float_as_uint(acc));
        }
    }
}

__global__ void __launch_bounds__(256)
sq_pack_kernel(const float* __restrict__ flat, long long n, const long long* __restrict__ tensor_offsets,
               int n_tensors, int level, uint64_t seed, const uint32_t* __restrict__ absmax_bits,
               uint8_t* __restrict__ payload, uint8_t* __restrict__ sign_plane, float* __restrict__ scale) {
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    if (blockIdx.x == 0)
        for (int t = threadIdx.x; t < n_tensors; t += blockDim.x)
            scale[t] = __uint_as_float(absmax_bits[t]) / (float)level;
    const long long n_groups = (n + 7) >> 3;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
         g += (long long)gridDim.x * blockDim.x) {
        const long long i0 = g << 3;
        // one Philox call per group: eight 16-bit uniforms (the rounding probability is resolved to 2^-16)
        const uint4 r0 = philox4x32((uint32_t)g, (uint32_t)(g >> 32), 0u, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
        const uint32_t raw[4] = {r0.x, r0.y, r0.z, r0.w};
        int t = find_segment(s_off, n_tensors, i0);
        uint32_t sign = 0;
        uint8_t q[8];
        if (i0 + 8 <= s_off[t + 1]) {
            // the whole group lies inside one tensor (all but a handful do): two 128-bit loads, one scale
            const float4 a = ld_stream_f4(flat + i0), b = ld_stream_f4(flat + i0 + 4);
            const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            const float s = __uint_as_float(absmax_bits[t]);
            const float inv = s > 0.f ? (float)level / s : 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float y = fminf(fabsf(x[e]) * inv, (float)level);
                const float lo = floorf(y);
                const float u = (float)((raw[e >> 1] >> ((e & 1) * 16)) & 0xFFFFu) * (1.0f / 65536.0f);
                q[e] = (uint8_t)fminf(lo + (u < (y - lo) ? 1.f : 0.f), (float)level);
                sign |= (x[e] < 0.f ? 1u : 0u) << e;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const long long i = i0 + e;
                q[e] = 0;
                if (i < n) {
                    while (i >= s_off[t + 1]) ++t;
                    const float xv = flat[i];
                    const float s = __uint_as_float(absmax_bits[t]);
                    const float y = s > 0.f ? fminf(__fdiv_rn(fabsf(xv), s) * (float)level, (float)level) : 0.f;
                    const float lo = floorf(y);
                    const float u = (float)((raw[e >> 1] >> ((e & 1) * 16)) & 0xFFFFu) * (1.0f / 65536.0f);
                    q[e] = (uint8_t)fminf(lo + (u < (y - lo) ? 1.f : 0.f), (float)level);
                    sign |= (xv < 0.f ? 1u : 0u) << e;
                }
            }
        }
        if (i0 + 8 <= n) {
            uint2 w;
            w.x = q[0] | (q[1] << 8) | (q[2] << 16) | ((uint32_t)q[3] << 24);
            w.y = q[4] | (q[5] << 8) | (q[6] << 16) | ((uint32_t)q[7] << 24);
            *reinterpret_cast<uint2*>(payload + i0) = w;
        } else {
            for (int e = 0; e < 8 && i0 + e < n; ++e) payload[i0 + e] = q[e];
        }
        sign_plane[g] = (uint8_t)sign;
    }
}

__global__ void __launch_bounds__(256)
sq_dequant_kernel(const uint8_t* __restrict__ payload, const uint8_t* __restrict__ sign_plane,
                  const float* __restrict__ scale, const long long* __restrict__ tensor_offsets, int n_tensors,
                  long long n, float* __restrict__ out) {
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    const long long n_groups = (n + 7) >> 3;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
         g += (long long)gridDim.x * blockDim.x) {
        const long long i0 = g << 3;
        const uint32_t sign = sign_plane[g];
        int t = find_segment(s_off, n_tensors, i0);
        if (i0 + 8 <= s_off[t + 1]) {
            // one 64-bit load of the eight codes, two 128-bit stores
            const uint2 w = *reinterpret_cast<const uint2*>(payload + i0);
            const float sc = scale[t];
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float m = (float)(((e < 4 ? w.x : w.y) >> ((e & 3) * 8)) & 0xFFu) * sc;
                v[e] = ((sign >> e) & 1u) ? -m : m;
            }
            st_stream_f4(out + i0, make_float4(v[0], v[1], v[2], v[3]));
            st_stream_f4(out + i0 + 4, make_float4(v[4], v[5], v[6], v[7]));
            continue;
        }
        for (int e = 0; e < 8 && i0 + e < n; ++e) {
            const long long i = i0 + e;
            while (i >= s_off[t + 1]) ++t;
            const float m = (float)payload[i] * scale[t];
            out[i] = ((sign >> e) & 1u) ? -m : m;
        }
    }
}

// ---------------------------------------------------------------- NNADQ
struct NnadqStats {  // 32 bytes per tensor, lives in the caller's scratch
    int min_key;
    int max_key;
    double sum;
    double sumsq;
    long long pad;
};

__global__ void nnadq_init_kernel(NnadqStats* stats, int n_tensors) {
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_tensors; t += gridDim.x * blockDim.x) {
        stats[t].min_key = 0x7fffffff;
        stats[t].max_key = (int)0x80000000;
        stats[t].sum = 0.0;
        stats[t].sumsq = 0.0;
    }
}

__device__ __forceinline__ void nnadq_flush(NnadqStats* stats, int t, float mn, float mx, double s, double ss) {
    atomicMin(&stats[t].min_key, float_to_ordered(mn));
    atomicMax(&stats[t].max_key, float_to_ordered(mx));
    atomicAdd(&stats[t].sum, s);
    atomicAdd(&stats[t].sumsq, ss);
}

__global__ void __launch_bounds__(256)
nnadq_stats_kernel(const float* __restrict__ flat, const long long* __restrict__ tensor_offsets, int n_tensors,
                   NnadqStats* __restrict__ stats) {
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    const long long n = s_off[n_tensors];
    const int lane = threadIdx.x & 31;
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long base = warp_id * (kSpanIters * 32); base < n; base += n_warps * (kSpanIters * 32)) {
        int t_cur = -1;
        long long t_end = -1;
        float mn = INFINITY, mx = -INFINITY;
        double sum = 0.0, sq = 0.0;
        for (int j = 0; j < kSpanIters; ++j) {
            const long long i = base + (long long)j * 32 + lane;
            if (i >= n) break;
            if (i >= t_end) {
                if (t_cur >= 0) nnadq_flush(stats, t_cur, mn, mx, sum, sq);
                t_cur = find_segment(s_off, n_tensors, i);
                t_end = s_off[t_cur + 1];
                mn = INFINITY; mx = -INFINITY; sum = 0.0; sq = 0.0;
            }
            const float x = flat[i];
            mn = fminf(mn, x);
            mx = fmaxf(mx, x);
            sum += x;
            sq += (double)x * x;
        }
        const int t0 = __shfl_sync(0xffffffffu, t_cur, 0);
        if (__all_sync(0xffffffffu, t_cur == t0) && t0 >= 0) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                sum += __shfl_xor_sync(0xffffffffu, sum, o);
                sq += __shfl_xor_sync(0xffffffffu, sq, o);
            }
            if (lane == 0) nnadq_flush(stats, t0, mn, mx, sum, sq);
        } else if (t_cur >= 0) {
            nnadq_flush(stats, t_cur, mn, mx, sum, sq);
        }
    }
}

__global__ void nnadq_plan_kernel(const NnadqStats* __restrict__ stats, const long long* __restrict__ tensor_offsets,
                                  const long long* __restrict__ group_offsets, int n_tensors, float weight,
                                  float* __restrict__ scale, float* __restrict__ offset, int32_t* __restrict__ bits,
                                  long long* __restrict__ byte_offsets) {
    for (int t = threadIdx.x; t < n_tensors; t += blockDim.x) {
        const long long n = tensor_offsets[t + 1] - tensor_offsets[t];
        float mn = 0.f, rng = 0.f, sd = 0.f;
        if (n > 0) {
            mn = ordered_to_float(stats[t].min_key);
            rng = ordered_to_float(stats[t].max_key) - mn;
            const double mean = stats[t].sum / (double)n;
            double var = stats[t].sumsq / (double)n - mean * mean;
            sd = (float)sqrt(var > 0.0 ? var : 0.0);
        }
        int b = 1;
        if (rng > 0.f) {
            const float c = (0.6931471805599453f * rng) / ((3.4641016151377544f * fmaxf(sd, 1e-30f)) * weight);
            b = (int)rintf(log2f(fmaxf(c, 1e-30f)));
            b = max(1, min(16, b));
        }
        const float levels = (float)((1 << b) - 1);
        bits[t] = b;
        offset[t] = mn;
        scale[t] = rng > 0.f ? __fdiv_rn(rng, levels) : 1.f;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        long long acc = 0;
        for (int t = 0; t < n_tensors; ++t) {
            byte_offsets[t] = acc;
            acc += (group_offsets[t + 1] - group_offsets[t]) * bits[t];
        }
        byte_offsets[n_tensors] = acc;
    }
}

__global__ void __launch_bounds__(256)
nnadq_pack_kernel(const float* __restrict__ flat, const long long* __restrict__ tensor_offsets,
                  const long long* __restrict__ group_offsets, int n_tensors, const float* __restrict__ scale,
                  const float* __restrict__ offset, const int32_t* __restrict__ bits,
                  const long long* __restrict__ byte_offsets, uint8_t* __restrict__ payload,
                  long long payload_capacity) {
    extern __shared__ long long s_tab[];
    long long* s_goff = s_tab;
    long long* s_toff = s_tab + n_tensors + 1;
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) {
        s_goff[i] = group_offsets[i];
        s_toff[i] = tensor_offsets[i];
    }
    __syncthreads();
    const long long n_groups = s_goff[n_tensors];
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
         g += (long long)gridDim.x * blockDim.x) {
        const int t = find_segment(s_goff, n_tensors, g);
        const long long gl = g - s_goff[t];
        const long long j0 = gl << 3;
        const long long n_t = s_toff[t + 1] - s_toff[t];
        const int b = bits[t];
        const float s = scale[t], o = offset[t];
        const float top = (float)((1 << b) - 1);
        unsigned long long w0 = 0, w1 = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            unsigned long long code = 0;
            if (j0 + e < n_t) {
                const float x = flat[s_toff[t] + j0 + e];
                float c = rintf(__fdiv_rn(x - o, s));
                c = fminf(fmaxf(c, 0.f), top);
                code = (unsigned long long)c;
            }
            const int bit = e * b;
            if (bit >= 64) w1 |= code << (bit - 64);
            else {
                w0 |= code << bit;
                if (bit + b > 64) w1 |= code >> (64 - bit);
            }
        }
        const long long base = byte_offsets[t] + gl * b;
        if (base + b > payload_capacity) continue;  // never with a correctly sized buffer
        for (int i = 0; i < b; ++i)
            payload[base + i] = (uint8_t)(i < 8 ? (w0 >> (8 * i)) : (w1 >> (8 * (i - 8))));
    }
}

__global__ void __launch_bounds__(256)
nnadq_dequant_kernel(const uint8_t* __restrict__ payload, const float* __restrict__ scale,
                     const float* __restrict__ offset, const int32_t* __restrict__ bits,
                     const long long* __restrict__ byte_offsets, const long long* __restrict__ tensor_offsets,
                     const long long* __restrict__ group_offsets, int n_tensors, float* __restrict__ out) {
    extern __shared__ long long s_tab[];
    long long* s_goff = s_tab;
    long long* s_toff = s_tab + n_tensors + 1;
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) {
        s_goff[i] = group_offsets[i];
        s_toff[i] = tensor_offsets[i];
    }
    __syncthreads();
    const long long n_groups = s_goff[n_tensors];
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups;
         g += (long long)gridDim.x * blockDim.x) {
        const int t = find_segment(s_goff, n_tensors, g);
        const long long gl = g - s_goff[t];
        const long long j0 = gl << 3;
        const long long n_t = s_toff[t + 1] - s_toff[t];
        const int b = bits[t];
        const uint8_t* p = payload + byte_offsets[t] + gl * b;
        unsigned long long w0 = 0, w1 = 0;
        for (int i = 0; i < b; ++i) {
            const unsigned long long byte = p[i];
            if (i < 8) w0 |= byte << (8 * i); else w1 |= byte << (8 * (i - 8));
        }
        const unsigned int mask = (1u << b) - 1u;
        const float s = scale[t], o = offset[t];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int bit = e * b;
            unsigned long long v;
            if (bit >= 64) v = w1 >> (bit - 64);
            else if (bit + b <= 64) v = w0 >> bit;
            else v = (w0 >> bit) | (w1 << (64 - bit));
            if (j0 + e < n_t) out[s_toff[t] + j0 + e] = (float)((unsigned int)v & mask) * s + o;
        }
    }
}

}  // namespace dls

using namespace dls;

extern "C" int dls_sq_quant(const float* flat, long long n, const long long* tensor_offsets, int n_tensors, int level,
                            uint64_t seed, uint8_t* payload, uint8_t* sign_plane, float* scale,
                            uint32_t* scratch_absmax, void* stream) {
    if (n < 1 || n_tensors < 1 || n_tensors > 5000) return fail_msg("dls_sq_quant: bad sizes");
    if (level < 1 || level > 255) return fail_msg("dls_sq_quant: level must be in [1, 255]");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem = (size_t)(n_tensors + 1) * sizeof(long long);
    DLS_CHECK(cudaMemsetAsync(scratch_absmax, 0, sizeof(uint32_t) * n_tensors, s));
    sq_absmax_kernel<<<grid_for(n, 256, 8), 256, smem, s>>>(flat, n, tensor_offsets, n_tensors, scratch_absmax);
    DLS_LAUNCHED("sq_absmax_kernel");
    sq_pack_kernel<<<grid_for((n + 7) / 8, 256, 8), 256, smem, s>>>(flat, n, tensor_offsets, n_tensors, level, seed,
                                                                     scratch_absmax, payload, sign_plane, scale);
    DLS_LAUNCHED("sq_pack_kernel");
    return 0;
}

extern "C" int dls_sq_dequant(const uint8_t* payload, const uint8_t* sign_plane, const float* scale,
                              const long long* tensor_offsets, int n_tensors, long long n, float* out, void* stream) {
    if (n < 1 || n_tensors < 1 || n_tensors > 5000) return fail_msg("dls_sq_dequant: bad sizes");
    const size_t smem = (size_t)(n_tensors + 1) * sizeof(long long);
    sq_dequant_kernel<<<grid_for((n + 7) / 8, 256, 8), 256, smem, (cudaStream_t)stream>>>(
        payload, sign_plane, scale, tensor_offsets, n_tensors, n, out);
    DLS_LAUNCHED("sq_dequant_kernel");
    return 0;
}

extern "C" int dls_nnadq_quant(const float* flat, const long long* tensor_offsets, const long long* group_offsets,
                               int n_tensors, float weight, double* scratch_stats, float* scale, float* offset,
                               int32_t* bits, long long* byte_offsets, uint8_t* payload, long long payload_capacity,
                               void* stream) {
    if (n_tensors < 1 || n_tensors > 2900) return fail_msg("dls_nnadq_quant: n_tensors must be in [1, 2900]");
    if (!(weight > 0.f)) return fail_msg("dls_nnadq_quant: weight must be positive");
    cudaStream_t s = (cudaStream_t)stream;
    NnadqStats* stats = reinterpret_cast<NnadqStats*>(scratch_stats);
    static_assert(sizeof(NnadqStats) == 4 * sizeof(double), "scratch layout");
    const size_t smem1 = (size_t)(n_tensors + 1) * sizeof(long long);
    nnadq_init_kernel<<<(n_tensors + 255) / 256, 256, 0, s>>>(stats, n_tensors);
    DLS_LAUNCHED("nnadq_init_kernel");
    nnadq_stats_kernel<<<kNumSMs * 8, 256, smem1, s>>>(flat, tensor_offsets, n_tensors, stats);
    DLS_LAUNCHED("nnadq_stats_kernel");
    nnadq_plan_kernel<<<1, 256, 0, s>>>(stats, tensor_offsets, group_offsets, n_tensors, weight, scale, offset, bits,
                                        byte_offsets);
    DLS_LAUNCHED("nnadq_plan_kernel");
    nnadq_pack_kernel<<<kNumSMs * 8, 256, 2 * smem1, s>>>(flat, tensor_offsets, group_offsets, n_tensors, scale,
                                                          offset, bits, byte_offsets, payload, payload_capacity);
    DLS_LAUNCHED("nnadq_pack_kernel");
    return 0;
}

extern "C" int dls_nnadq_dequant(const uint8_t* payload, const float* scale, const float* offset, const int32_t* bits,
                                 const long long* byte_offsets, const long long* tensor_offsets,
                                 const long long* group_offsets, int n_tensors, float* out, void* stream) {
    if (n_tensors < 1 || n_tensors > 2900) return fail_msg("dls_nnadq_dequant: n_tensors must be in [1, 2900]");
    const size_t smem = 2 * (size_t)(n_tensors + 1) * sizeof(long long);
    nnadq_dequant_kernel<<<kNumSMs * 8, 256, smem, (cudaStream_t)stream>>>(
        payload, scale, offset, bits, byte_offsets, tensor_offsets, group_offsets, n_tensors, out);
    DLS_LAUNCHED("nnadq_dequant_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// dls_sq_quant_noalloc: the stochastic quantizer for callers that cannot allocate.
//
// dls_sq_quant needs one uint32 of scratch per tensor (the abs-max reduction target) and
// leaves it to the caller to provide that memory.  Some callers have no way to get it:
//   * a host that drives this library through the C ABI alone (no torch, no pooling
//     allocator of its own) and runs with every byte of HBM already handed out to the
//     packed workers' arenas -- a late cudaMalloc there can simply fail;
//   * a worker thread on a GPU it shares with many other workers' streams: cudaMalloc and
//     cudaFree are device-wide operations that stall every one of those streams, so the
//     per-round upload path must not contain them.
// For them the scratch lives in the library image itself: one file-scope __device__ array
// of DLS_SQ_NOALLOC_MAX_TENSORS words (16 KiB).  Every GPU that loads the library carries
// its own copy of that array until the process exits.
//
// Rules that follow from sharing one array:
//   * n_tensors above DLS_SQ_NOALLOC_MAX_TENSORS is refused with an error, nothing runs;
//   * the array keeps whatever the previous call left in it, so each call first zeroes the
//     n_tensors words it is going to use (on `stream`, ahead of its kernels);
//   * two calls whose GPU work overlaps (different streams, or different host threads on
//     the same device) would reduce into the same words and corrupt each other's scales:
//     callers must order them, e.g. by issuing them on one stream;
//   * the first call on a device looks the array's device address up with
//     cudaGetSymbolAddress (in host code the array's name is only a host-side handle, not
//     something a kernel may dereference) and remembers it per device.  Make that first
//     call, on every device you use, before you begin any stream capture.
// The entry point itself never calls cudaMalloc, cudaFree or any allocating routine.
// It has no counterpart in the reference (which has no native code at all).
#define DLS_SQ_NOALLOC_MAX_TENSORS 4096

static __device__ __align__(256) uint32_t g_sq_noalloc_scratch[DLS_SQ_NOALLOC_MAX_TENSORS];

#include <mutex>
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
uint32_t* g_noalloc_scratch_addr[kMaxDevices] = {nullptr};  // per device: address of that device's copy
}  // namespace

extern "C" int dls_sq_quant_noalloc_max_tensors(void) { return DLS_SQ_NOALLOC_MAX_TENSORS; }

extern "C" int dls_sq_quant_noalloc(const float* flat, long long n, const long long* tensor_offsets, int n_tensors,
                                    int level, uint64_t seed, uint8_t* payload, uint8_t* sign_plane, float* scale,
                                    void* stream) {
    if (n < 1 || n_tensors < 1) return fail_msg("dls_sq_quant_noalloc: bad sizes");
    if (n_tensors > DLS_SQ_NOALLOC_MAX_TENSORS) {
        snprintf(g_last_error, sizeof(g_last_error),
                 "dls_sq_quant_noalloc: %d tensors exceed the built-in scratch of %d tensors; "
                 "use dls_sq_quant with caller-provided scratch",
                 n_tensors, DLS_SQ_NOALLOC_MAX_TENSORS);
        return -1;
    }
    if (level < 1 || level > 255) return fail_msg("dls_sq_quant_noalloc: level must be in [1, 255]");
    int dev = -1;
    DLS_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail_msg("dls_sq_quant_noalloc: device index out of range");
    uint32_t* scratch = nullptr;
    {
        // one-time set-up per device, with that device current (it is: we just asked for it)
        std::lock_guard<std::mutex> guard(g_noalloc_mutex);
        if (g_noalloc_scratch_addr[dev] == nullptr) {
            uint32_t* addr = nullptr;
            DLS_CHECK(cudaGetSymbolAddress((void**)&addr, g_sq_noalloc_scratch));
            if (addr == nullptr) return fail_msg("dls_sq_quant_noalloc: cudaGetSymbolAddress returned NULL");
            g_noalloc_scratch_addr[dev] = addr;
        }
        scratch = g_noalloc_scratch_addr[dev];
    }
    cudaStream_t s = (cudaStream_t)stream;
    const size_t smem = (size_t)(n_tensors + 1) * sizeof(long long);
    // the array still holds the previous call's maxima: clear the words this call reduces into
    DLS_CHECK(cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * n_tensors, s));
    sq_absmax_kernel<<<grid_for(n, 256, 8), 256, smem, s>>>(flat, n, tensor_offsets, n_tensors, scratch);
    DLS_LAUNCHED("sq_absmax_kernel");
    sq_pack_kernel<<<grid_for((n + 7) / 8, 256, 8), 256, smem, s>>>(flat, n, tensor_offsets, n_tensors, level, seed,
                                                                     scratch, payload, sign_plane, scale);
    DLS_LAUNCHED("sq_pack_kernel");
    return 0;
}
