// Local-SGD kernels of the CUDA-graph training step.
//
// The reference trains with torch.optim.SGD inside the external Trainer (and re-implements
// the update by hand for per-step gradient exchange: simulation_lib/worker/
// gradient_worker.py:13-26,96-116): a Python loop over parameter tensors, several ATen
// launches per tensor per step.  Here parameters and momentum live in one flat fp32 arena
// per worker; one launch updates every tensor, reading each tensor's gradient through a
// pointer table (autograd allocates gradients inside the captured graph's pool, so their
// addresses are fixed but not contiguous) and the step's learning rate from a device-side
// schedule table indexed by a device-side step counter -- nothing in the captured graph
// depends on a host value, so one graph serves every step of every round.
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/dls_b200.h"

namespace dls {

constexpr int kSgdThreads = 256;

__global__ void __launch_bounds__(kSgdThreads)
sgd_step_kernel(float* __restrict__ params, float* __restrict__ mom, const float* const* __restrict__ grads,
                const long long* __restrict__ tensor_offsets, int n_tensors, const float* __restrict__ lr_table,
                int lr_len, const int32_t* __restrict__ state, float weight_decay, float momentum, float dampening,
                int nesterov, const uint8_t* __restrict__ grad_is_bf16, __nv_bfloat16* __restrict__ shadow) {
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ long long s_off[];
    for (int i = threadIdx.x; i <= n_tensors; i += blockDim.x) s_off[i] = tensor_offsets[i];
    __syncthreads();
    const long long total = s_off[n_tensors];
    int step = state[0];
    if (step >= lr_len) step = lr_len - 1;
    const float lr = lr_table[step];
    const bool mom_ready = state[1] != 0;
    const long long n_vec = (total + 3) >> 2;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n_vec;
         v += (long long)gridDim.x * blockDim.x) {
        const long long i = v << 2;
        const int t = find_segment(s_off, n_tensors, i);
        const long long j = i - s_off[t];
        const bool fast = (i + 4 <= s_off[t + 1]) && ((j & 3) == 0) && grads[t] != nullptr;
        float g[4], p[4], m[4];
        // per-tensor flags: bit 0 = bf16 gradient, bits 0|1 = the model's matrix is the bf16 shadow (rewrite
        // it), bit 2 = fp32 gradient sink that accumulates (conv wgrad partial tiles): zero it after reading
        const uint8_t flags = grad_is_bf16 != nullptr ? grad_is_bf16[t] : (uint8_t)0;
        const bool half_grad = (flags & 1) != 0;
        const bool has_shadow = (flags & 3) != 0;
        if (fast) {
            if (half_grad) {  // mixed precision: the model's matrix is the bf16 shadow, so is its gradient
                const uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(grads[t]) + j);
                const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
                const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
                g[0] = lo.x; g[1] = lo.y; g[2] = hi.x; g[3] = hi.y;
            } else {
                if (flags & 4) {
                    float4* gp4 = reinterpret_cast<float4*>(const_cast<float*>(grads[t]) + j);
                    const float4 gv = *gp4;
                    *gp4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    g[0] = gv.x; g[1] = gv.y; g[2] = gv.z; g[3] = gv.w;
                } else {
                    const float4 gv = ld_stream_f4(grads[t] + j);  // gradients are read exactly once
                    g[0] = gv.x; g[1] = gv.y; g[2] = gv.z; g[3] = gv.w;
                }
            }
            const float4 pv = *reinterpret_cast<const float4*>(params + i);
            p[0] = pv.x; p[1] = pv.y; p[2] = pv.z; p[3] = pv.w;
            if (momentum != 0.f) {
                const float4 mv = *reinterpret_cast<const float4*>(mom + i);
                m[0] = mv.x; m[1] = mv.y; m[2] = mv.z; m[3] = mv.w;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float d = g[e] + weight_decay * p[e];
                if (momentum != 0.f) {
                    m[e] = mom_ready ? momentum * m[e] + (1.f - dampening) * d : d;
                    d = nesterov ? d + momentum * m[e] : m[e];
                }
                p[e] -= lr * d;
            }
            *reinterpret_cast<float4*>(params + i) = make_float4(p[0], p[1], p[2], p[3]);
            if (momentum != 0.f) *reinterpret_cast<float4*>(mom + i) = make_float4(m[0], m[1], m[2], m[3]);
            if (has_shadow && shadow != nullptr) {  // refresh the bf16 compute copy in the same pass
                uint2 out;
                *reinterpret_cast<__nv_bfloat162*>(&out.x) = __floats2bfloat162_rn(p[0], p[1]);
                *reinterpret_cast<__nv_bfloat162*>(&out.y) = __floats2bfloat162_rn(p[2], p[3]);
                *reinterpret_cast<uint2*>(shadow + i) = out;
            }
        } else {
            for (int e = 0; e < 4; ++e) {
                const long long ie = i + e;
                if (ie >= total) break;
                const int te = find_segment(s_off, n_tensors, ie);
                const float* gp = grads[te];
                if (gp == nullptr) continue;  // parameter received no gradient this step
                const uint8_t fe = grad_is_bf16 != nullptr ? grad_is_bf16[te] : (uint8_t)0;
                const bool half_e = (fe & 1) != 0;
                const float ge = half_e ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(gp)[ie - s_off[te]])
                                        : gp[ie - s_off[te]];
                if (fe & 4) const_cast<float*>(gp)[ie - s_off[te]] = 0.f;
                float d = ge + weight_decay * params[ie];
                if (momentum != 0.f) {
                    const float mm = mom_ready ? momentum * mom[ie] + (1.f - dampening) * d : d;
                    mom[ie] = mm;
                    d = nesterov ? d + momentum * mm : mm;
                }
                params[ie] -= lr * d;
                if ((fe & 3) != 0 && shadow != nullptr) shadow[ie] = __float2bfloat16(params[ie]);
            }
        }
    }
}

__global__ void sgd_tick_kernel(int32_t* state) {
    pdl_launch_dependents();
    pdl_wait();
    state[0] += 1;  // optimizer / LR-schedule step
    state[1] = 1;   // momentum buffer is initialised from now on
    state[2] += 1;  // batch index inside the epoch (the host zeroes it per epoch)
}

// one block per output row; the row index comes from a device-side step counter
__global__ void __launch_bounds__(256)
gather_batch_kernel(const float* __restrict__ x_all, const long long* __restrict__ y_all,
                    const long long* __restrict__ order, long long n_order, const int32_t* __restrict__ state,
                    int state_index, int batch_stride, long long row_elems, int channels,
                    float* __restrict__ x_out, long long* __restrict__ y_out) {
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x;
    // batch_stride is the epoch's regular batch size: the (smaller) last batch starts at
    // index * stride too
    long long pos = (long long)state[state_index] * batch_stride + b;
    if (pos >= n_order) pos = n_order - 1;
    const long long src = order[pos];
    const float* in = x_all + src * row_elems;
    float* out = x_out + (long long)b * row_elems;
    if (channels > 1) {
        // source rows are C,H,W; the model runs channels-last, so emit H,W,C
        const long long hw = row_elems / channels;
        for (long long v = threadIdx.x; v < row_elems; v += blockDim.x) {
            const long long pix = v / channels;
            const int c = (int)(v - pix * channels);
            out[v] = in[(long long)c * hw + pix];
        }
    } else if ((row_elems & 3) == 0) {
        const long long nv = row_elems >> 2;
        for (long long v = threadIdx.x; v < nv; v += blockDim.x)
            reinterpret_cast<float4*>(out)[v] = ld_stream_f4(in + (v << 2));
    } else {
        for (long long v = threadIdx.x; v < row_elems; v += blockDim.x) out[v] = in[v];
    }
    if (threadIdx.x == 0) y_out[b] = y_all[src];
}

// loss/accuracy bookkeeping without a host round trip: one warp per sample
__global__ void __launch_bounds__(256)
metric_accumulate_kernel(const float* __restrict__ logits, const long long* __restrict__ targets,
                         const float* __restrict__ loss, int batch, int classes, double* __restrict__ acc) {
    pdl_launch_dependents();
    pdl_wait();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    int correct = 0;
    if (warp < batch) {
        const float* row = logits + (long long)warp * classes;
        float best = -INFINITY;
        int best_i = 0x7fffffff;
        for (int c = lane; c < classes; c += 32) {
            const float v = row[c];
            if (v > best) { best = v; best_i = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
        }
        correct = (lane == 0 && best_i == (int)targets[warp]) ? 1 : 0;
    }
    __shared__ int s_correct;
    if (threadIdx.x == 0) s_correct = 0;
    __syncthreads();
    if (correct) atomicAdd(&s_correct, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_correct) atomicAdd(acc + 1, (double)s_correct);
        if (blockIdx.x == 0) {
            atomicAdd(acc + 0, (double)loss[0] * batch);
            atomicAdd(acc + 2, (double)batch);
        }
    }
}

}  // namespace dls

using namespace dls;

extern "C" int dls_sgd_step(float* params, float* momentum_buf, const float* const* grads,
                            const long long* tensor_offsets, int n_tensors, const float* lr_table, int lr_len,
                            const int32_t* state, float weight_decay, float momentum, float dampening, int nesterov,
                            const uint8_t* grad_is_bf16, void* shadow_bf16, void* stream) {
    if (n_tensors < 1 || n_tensors > 5000) return fail_msg("dls_sgd_step: n_tensors must be in [1, 5000]");
    if (lr_len < 1) return fail_msg("dls_sgd_step: empty lr table");
    if (momentum != 0.f && momentum_buf == nullptr) return fail_msg("dls_sgd_step: momentum buffer missing");
    const size_t smem = (size_t)(n_tensors + 1) * sizeof(long long);
    // grid from the arena size is unknown on the host without a sync: a fixed persistent grid
    const int grid = kNumSMs * 8;
    DLS_CHECK(launch_pdl(sgd_step_kernel, dim3(grid), dim3(kSgdThreads), smem, (cudaStream_t)stream, 0, params,
                         momentum_buf, grads, tensor_offsets, n_tensors, lr_table, lr_len, state, weight_decay, momentum,
                         dampening, nesterov, grad_is_bf16, reinterpret_cast<__nv_bfloat16*>(shadow_bf16)));
    DLS_LAUNCHED("sgd_step_kernel");
    return 0;
}

extern "C" int dls_sgd_tick(int32_t* state, void* stream) {
    DLS_CHECK(launch_pdl(sgd_tick_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, 0, state));
    DLS_LAUNCHED("sgd_tick_kernel");
    return 0;
}

extern "C" int dls_gather_batch(const float* x_all, const long long* y_all, const long long* order,
                                long long n_order, const int32_t* state, int state_index, int batch,
                                int batch_stride, long long row_elems, int nhwc_channels, float* x_out,
                                long long* y_out, void* stream) {
    if (batch < 1 || n_order < 1 || batch_stride < batch) return fail_msg("dls_gather_batch: bad batch/order");
    if (nhwc_channels > 1 && row_elems % nhwc_channels) return fail_msg("dls_gather_batch: bad channel count");
    gather_batch_kernel<<<batch, 256, 0, (cudaStream_t)stream>>>(x_all, y_all, order, n_order, state, state_index,
                                                                 batch_stride, row_elems, nhwc_channels, x_out,
                                                                 y_out);
    DLS_LAUNCHED("gather_batch_kernel");
    return 0;
}

extern "C" int dls_metric_accumulate(const float* logits, const long long* targets, const float* loss, int batch,
                                     int classes, double* acc, void* stream) {
    if (batch < 1 || classes < 1) return fail_msg("dls_metric_accumulate: empty batch");
    const int warps_per_block = 8;
    const int grid = (batch + warps_per_block - 1) / warps_per_block;
    DLS_CHECK(launch_pdl(metric_accumulate_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, 0, logits, targets,
                         loss, batch, classes, acc));
    DLS_LAUNCHED("metric_accumulate_kernel");
    return 0;
}
