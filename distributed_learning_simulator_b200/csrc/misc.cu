// sign_SGD packing / majority vote, FedDropoutAvg masking, FedOBD block scores.
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/dls_b200.h"

namespace dls {

// 32 signs per word: bit e of word w is 1 when grad[32*w + e] < 0 (reference scaffold:
// simulation_lib/worker/gradient_worker.py:47-48 `_process_gradient` override point)
__global__ void __launch_bounds__(256)
sign_pack_kernel(const float* __restrict__ grad, long long n, uint32_t* __restrict__ packed) {
    const int lane = threadIdx.x & 31;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i - lane < n;
         i += (long long)gridDim.x * blockDim.x) {
        const bool neg = i < n && grad[i] < 0.f;
        const uint32_t word = __ballot_sync(0xffffffffu, neg);
        if (lane == 0) packed[i >> 5] = word;
    }
}

// element-wise majority vote over K workers' packed signs: sign(sum_k s_k[i]), s in {-1,+1}
__global__ void __launch_bounds__(256)
sign_majority_kernel(const uint32_t* const* __restrict__ rows, int n_rows, long long n, float* __restrict__ vote) {
    const int lane = threadIdx.x & 31;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i - lane < n;
         i += (long long)gridDim.x * blockDim.x) {
        const long long w = i >> 5;
        int neg = 0;
        for (int k = 0; k < n_rows; ++k) neg += (rows[k][w] >> lane) & 1u;  // one broadcast load per warp
        const int s = n_rows - 2 * neg;
        if (i < n) vote[i] = s > 0 ? 1.f : (s < 0 ? -1.f : 0.f);
    }
}

// signSGD step in one pass: majority vote over the K workers' packed sign words (rows may live in peer GPUs'
// memory: plain loads over NVLink) fused with the optimizer update of THIS worker's parameters -- momentum,
// dampening, nesterov, p -= lr * d, and the bf16 compute shadow when there is one.  Replaces the server-side
// vote launch, the fp32 vote tensor (4 bytes per parameter written once and read by every worker) and the
// separate update launch of the scaffold (reference: worker/gradient_worker.py:96-116 applies the update with
// a Python loop over parameter tensors).
__global__ void __launch_bounds__(256)
sign_vote_sgd_kernel(const uint32_t* const* __restrict__ rows, int n_rows, long long n, float* __restrict__ params,
                     float* __restrict__ mom, float lr, float momentum, float dampening, int nesterov, int mom_ready,
                     __nv_bfloat16* __restrict__ shadow, const uint8_t* __restrict__ shadow_mask,
                     const long long* __restrict__ tensor_offsets, int n_tensors) {
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i - lane < n;
         i += (long long)gridDim.x * blockDim.x) {
        const long long w = i >> 5;
        int neg = 0;
        for (int k = 0; k < n_rows; ++k) neg += (rows[k][w] >> lane) & 1u;  // one broadcast load per warp
        if (i >= n) continue;
        const int s = n_rows - 2 * neg;
        float d = s > 0 ? 1.f : (s < 0 ? -1.f : 0.f);
        if (momentum != 0.f) {
            const float m = mom_ready ? momentum * mom[i] + (1.f - dampening) * d : d;
            mom[i] = m;
            d = nesterov ? d + momentum * m : m;
        }
        const float p = params[i] - lr * d;
        params[i] = p;
        if (shadow != nullptr) {
            bool shadowed = true;
            if (shadow_mask != nullptr) shadowed = shadow_mask[find_segment(tensor_offsets, n_tensors, i)] != 0;
            if (shadowed) shadow[i] = __float2bfloat16(p);
        }
    }
}

// upload masking of FedDropoutAvg (reference: method/fed_dropout_avg/worker.py:23-27 draws a
// Bernoulli tensor, multiplies, then count_nonzero().item() per tensor); one pass here
__global__ void __launch_bounds__(256)
dropout_mask_kernel(const float* __restrict__ in, long long n, float keep_prob, uint64_t seed, float* __restrict__ out,
                    unsigned long long* __restrict__ nonzero_count) {
    unsigned int local = 0;
    const long long n_vec = (n + 3) >> 2;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n_vec;
         v += (long long)gridDim.x * blockDim.x) {
        const uint4 r = philox4x32((uint32_t)v, (uint32_t)(v >> 32), 7u, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
        const uint32_t rnd[4] = {r.x, r.y, r.z, r.w};
        const long long i0 = v << 2;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (i0 + e < n) {
                const float x = (u32_to_unit(rnd[e]) < keep_prob) ? in[i0 + e] : 0.f;
                out[i0 + e] = x;
                local += x != 0.f;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(nonzero_count, (unsigned long long)local);
}

// FedOBD block importance: sum over each block of (cur - prev)^2 in one pass over the arena
// (reference: method/fed_obd/obd_algorithm.py:129-145 cats, copies to CPU and calls
// torch.norm().item() once per block)
__global__ void __launch_bounds__(256)
block_sqdiff_kernel(const float* __restrict__ cur, const float* __restrict__ prev,
                    const long long* __restrict__ tensor_offsets, const int32_t* __restrict__ block_of_tensor,
                    int n_tensors, double* __restrict__ scores) {
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    const long long n = s_off[n_tensors];
    const int lane = threadIdx.x & 31;
    constexpr int kIters = 64;  // a warp walks 2048 contiguous elements, one atomic at the end
    const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long base = warp_id * (kIters * 32); base < n; base += n_warps * (kIters * 32)) {
        int b_cur = -1;
        long long t_end = -1;
        double acc = 0.0;
        for (int j = 0; j < kIters; ++j) {
            const long long i = base + (long long)j * 32 + lane;
            if (i >= n) break;
            if (i >= t_end) {
                const int t = find_segment(s_off, n_tensors, i);
                t_end = s_off[t + 1];
                const int b = block_of_tensor[t];
                if (b != b_cur) {
                    if (b_cur >= 0) atomicAdd(scores + b_cur, acc);
                    b_cur = b;
                    acc = 0.0;
                }
            }
            const float d = cur[i] - prev[i];
            if (b_cur >= 0) acc += (double)d * d;
        }
        const int b0 = __shfl_sync(0xffffffffu, b_cur, 0);
        if (__all_sync(0xffffffffu, b_cur == b0) && b0 >= 0) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) atomicAdd(scores + b0, acc);
        } else if (b_cur >= 0) {
            atomicAdd(scores + b_cur, acc);
        }
    }
}

}  // namespace dls

using namespace dls;

extern "C" int dls_sign_pack(const float* grad, long long n, uint32_t* packed, void* stream) {
    if (n < 1) return fail_msg("dls_sign_pack: empty input");
    sign_pack_kernel<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(grad, n, packed);
    DLS_LAUNCHED("sign_pack_kernel");
    return 0;
}

extern "C" int dls_sign_majority(const uint32_t* const* rows, int n_rows, long long n, float* vote, void* stream) {
    if (n < 1 || n_rows < 1) return fail_msg("dls_sign_majority: empty input");
    sign_majority_kernel<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(rows, n_rows, n, vote);
    DLS_LAUNCHED("sign_majority_kernel");
    return 0;
}

extern "C" int dls_sign_vote_sgd(const uint32_t* const* rows, int n_rows, long long n, float* params, float* mom,
                                 float lr, float momentum, float dampening, int nesterov, int mom_ready, void* shadow,
                                 const uint8_t* shadow_mask, const long long* tensor_offsets, int n_tensors,
                                 void* stream) {
    if (n < 1 || n_rows < 1) return fail_msg("dls_sign_vote_sgd: empty input");
    if (momentum != 0.f && mom == nullptr) return fail_msg("dls_sign_vote_sgd: momentum buffer missing");
    if (shadow != nullptr && shadow_mask != nullptr && (tensor_offsets == nullptr || n_tensors < 1))
        return fail_msg("dls_sign_vote_sgd: a shadow mask needs the tensor offsets");
    DLS_CHECK(launch_pdl(sign_vote_sgd_kernel, dim3(grid_for(n, 256, 8)), dim3(256), 0, (cudaStream_t)stream, 0, rows,
                         n_rows, n, params, mom, lr, momentum, dampening, nesterov, mom_ready,
                         reinterpret_cast<__nv_bfloat16*>(shadow), shadow_mask, tensor_offsets, n_tensors));
    DLS_LAUNCHED("sign_vote_sgd_kernel");
    return 0;
}

extern "C" int dls_dropout_mask(const float* in, long long n, float keep_prob, uint64_t seed, float* out,
                                unsigned long long* nonzero_count, void* stream) {
    if (n < 1) return fail_msg("dls_dropout_mask: empty input");
    dropout_mask_kernel<<<grid_for((n + 3) / 4, 256, 8), 256, 0, (cudaStream_t)stream>>>(in, n, keep_prob, seed, out,
                                                                                        nonzero_count);
    DLS_LAUNCHED("dropout_mask_kernel");
    return 0;
}

extern "C" int dls_block_sqdiff(const float* cur, const float* prev, const long long* tensor_offsets,
                                const int32_t* block_of_tensor, int n_tensors, int n_blocks, double* scores,
                                void* stream) {
    if (n_tensors < 1 || n_tensors > 5000 || n_blocks < 1) return fail_msg("dls_block_sqdiff: bad sizes");
    cudaStream_t s = (cudaStream_t)stream;
    DLS_CHECK(cudaMemsetAsync(scores, 0, sizeof(double) * n_blocks, s));
    const size_t smem = (size_t)(n_tensors + 1) * sizeof(long long);
    block_sqdiff_kernel<<<kNumSMs * 8, 256, smem, s>>>(cur, prev, tensor_offsets, block_of_tensor, n_tensors, scores);
    DLS_LAUNCHED("block_sqdiff_kernel");
    return 0;
}

// ---- grouped im2col for the coalition evaluator -----------------------------------------------------
// x: [B][H][W][G*C] (channels-last activations of G coalitions, C channels each), valid k x k windows.
// cols: [G][B*OH*OW][k_pad]; column order (kh, kw, c) when channels_last_k (the arena stores 4-D filters
// channels-last) else (c, kh, kw); columns past C*k*k are zero.  One thread per (g, b, oh, ow, kh, kw).
// The reference has no counterpart (it runs each coalition through an eager torch convolution,
// server/server.py:40-55); torch's unfold took 80 % of the batched evaluator before this kernel.
namespace dls {
template <typename T, bool kPadded>
__global__ void __launch_bounds__(256)
im2col_grouped_kernel(const T* __restrict__ x, int batch, int height, int width, int groups, int channels, int k,
                      int k_pad, int channels_last_k, int pad, T* __restrict__ cols) {
    // kPadded: zero padding of `pad` pixels on every border (taps outside the image contribute zeros)
    const int oh_n = height + 2 * pad - k + 1, ow_n = width + 2 * pad - k + 1;
    const long long rows = (long long)batch * oh_n * ow_n;
    const long long total = (long long)groups * rows * k * k;
    const int kk = k * k;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const int tap = (int)(t % kk);
        const long long row_g = t / kk;           // g * rows + row
        const long long row = row_g % rows;
        const int g = (int)(row_g / rows);
        const int ow = (int)(row % ow_n), oh = (int)((row / ow_n) % oh_n);
        const int b = (int)(row / ((long long)ow_n * oh_n));
        const int kh = tap / k, kw = tap - kh * k;
        const int iy = oh + kh - pad, ix = ow + kw - pad;
        T* dst_row = cols + row_g * k_pad;
        if (kPadded && (iy < 0 || iy >= height || ix < 0 || ix >= width)) {
            if (channels_last_k)
                for (int c = 0; c < channels; ++c) dst_row[tap * channels + c] = T(0);
            else
                for (int c = 0; c < channels; ++c) dst_row[c * kk + tap] = T(0);
            if (tap == kk - 1)
                for (int j = channels * kk; j < k_pad; ++j) dst_row[j] = T(0);
            continue;
        }
        const T* src = x + (((long long)b * height + iy) * width + ix) * ((long long)groups * channels) +
                       (long long)g * channels;
        if (channels_last_k) {
            T* dst = dst_row + tap * channels;
            if (sizeof(T) == 2 && (channels & 1) == 0 && (k_pad & 1) == 0) {
                const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src);
                uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
                for (int c = 0; c < channels / 2; ++c) d32[c] = s32[c];
            } else {
                for (int c = 0; c < channels; ++c) dst[c] = src[c];
            }
        } else {
            for (int c = 0; c < channels; ++c) dst_row[c * kk + tap] = src[c];
        }
        if (tap == kk - 1)
            for (int j = channels * kk; j < k_pad; ++j) dst_row[j] = T(0);
    }
}
}  // namespace dls

extern "C" int dls_im2col_grouped(const void* x, int elem_bytes, int batch, int height, int width, int groups,
                                  int channels, int k, int k_pad, int channels_last_k, void* cols, void* stream) {
    if (batch < 1 || groups < 1 || channels < 1 || k < 1 || height < k || width < k || k_pad < channels * k * k)
        return dls::fail_msg("dls_im2col_grouped: bad shape");
    if (elem_bytes != 2 && elem_bytes != 4) return dls::fail_msg("dls_im2col_grouped: 2- or 4-byte elements");
    const long long total = (long long)groups * batch * (height - k + 1) * (width - k + 1) * k * k;
    const int grid = dls::grid_for(total, 256, 16);
    if (elem_bytes == 2)
        dls::im2col_grouped_kernel<uint16_t, false><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const uint16_t*)x, batch, height, width, groups, channels, k, k_pad, channels_last_k, 0, (uint16_t*)cols);
    else
        dls::im2col_grouped_kernel<uint32_t, false><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const uint32_t*)x, batch, height, width, groups, channels, k, k_pad, channels_last_k, 0, (uint32_t*)cols);
    DLS_LAUNCHED("im2col_grouped_kernel");
    return 0;
}

extern "C" int dls_im2col_grouped_padded(const void* x, int elem_bytes, int batch, int height, int width, int groups,
                                         int channels, int k, int k_pad, int channels_last_k, int pad, void* cols,
                                         void* stream) {
    if (pad < 0 || batch < 1 || groups < 1 || channels < 1 || k < 1 || height + 2 * pad < k || width + 2 * pad < k ||
        k_pad < channels * k * k)
        return dls::fail_msg("dls_im2col_grouped_padded: bad shape");
    if (elem_bytes != 2 && elem_bytes != 4) return dls::fail_msg("dls_im2col_grouped_padded: 2- or 4-byte elements");
    const long long total =
        (long long)groups * batch * (height + 2 * pad - k + 1) * (width + 2 * pad - k + 1) * k * k;
    const int grid = dls::grid_for(total, 256, 16);
    if (elem_bytes == 2)
        dls::im2col_grouped_kernel<uint16_t, true><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const uint16_t*)x, batch, height, width, groups, channels, k, k_pad, channels_last_k, pad, (uint16_t*)cols);
    else
        dls::im2col_grouped_kernel<uint32_t, true><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const uint32_t*)x, batch, height, width, groups, channels, k, k_pad, channels_last_k, pad, (uint32_t*)cols);
    DLS_LAUNCHED("im2col_grouped_kernel");
    return 0;
}
