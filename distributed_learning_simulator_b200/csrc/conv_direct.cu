// Direct (CUDA-core) convolutions for the layers the tensor-core implicit GEMM (conv_tc.cu) does not
// take: a handful of channels (the 3-channel stems, LeNet-5's 1 / 6 / 16-channel 5x5 layers), where
// the GEMM K dimension would be mostly padding.  Every operand is addressed through element strides,
// so NCHW, channels-last and channel-sliced tensors all work without a layout copy; fp32 or bf16 per
// operand, fp32 accumulation.  Filters (<= 12 288 elements) are staged once per block in shared
// memory as fp32.
//
// Reference counterpart: cuDNN inside the external Trainer / Inferencer
// (/root/reference/simulation_lib/worker/worker.py:86-88, server/server.py:40-55).
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/dls_b200.h"

namespace dls {

constexpr int kDirectThreads = 256;
constexpr int kDirectMaxFilter = 12288;
constexpr int kDirectCo = 16;

__device__ __forceinline__ float ld_any(const void* p, long long i, int dtype) {
    return dtype ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i])
                 : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void st_any(void* p, long long i, int dtype, float v) {
    if (dtype) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
    else reinterpret_cast<float*>(p)[i] = v;
}

// one thread per output pixel, kDirectCo output channels at a time; s_w[(c, r, s)][co]
__global__ void __launch_bounds__(kDirectThreads)
conv_direct_fprop_kernel(const void* __restrict__ x, const void* __restrict__ w, const void* __restrict__ bias,
                         int bias_dtype, void* __restrict__ y, const __grid_constant__ DlsDirectConv d) {
    extern __shared__ float s_w[];
    const int q_count = d.c_in * d.r * d.s;
    for (int i = threadIdx.x; i < q_count * d.c_out; i += blockDim.x) {
        const int co = i % d.c_out, q = i / d.c_out;
        const int s = q % d.s, r = (q / d.s) % d.r, c = q / (d.s * d.r);
        s_w[i] = ld_any(w, co * d.ws[0] + c * d.ws[1] + r * d.ws[2] + s * d.ws[3], d.w_dtype);
    }
    __syncthreads();
    const long long pixels = (long long)d.n * d.oh * d.ow;
    for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < pixels;
         pix += (long long)gridDim.x * blockDim.x) {
        const int ow = (int)(pix % d.ow), oh = (int)((pix / d.ow) % d.oh), n = (int)(pix / ((long long)d.ow * d.oh));
        const int ih0 = oh * d.stride_h - d.pad_h, iw0 = ow * d.stride_w - d.pad_w;
        for (int co0 = 0; co0 < d.c_out; co0 += kDirectCo) {
            float acc[kDirectCo];
#pragma unroll
            for (int j = 0; j < kDirectCo; ++j)
                acc[j] = (bias != nullptr && co0 + j < d.c_out) ? ld_any(bias, co0 + j, bias_dtype) : 0.f;
            for (int c = 0; c < d.c_in; ++c)
                for (int r = 0; r < d.r; ++r) {
                    const int ih = ih0 + r;
                    if (ih < 0 || ih >= d.h) continue;
                    for (int s = 0; s < d.s; ++s) {
                        const int iw = iw0 + s;
                        if (iw < 0 || iw >= d.w) continue;
                        const float xv = ld_any(x, n * d.xs[0] + c * d.xs[1] + ih * d.xs[2] + iw * d.xs[3], d.x_dtype);
                        const float* wr = s_w + ((c * d.r + r) * d.s + s) * d.c_out + co0;
#pragma unroll
                        for (int j = 0; j < kDirectCo; ++j)
                            if (co0 + j < d.c_out) acc[j] = fmaf(xv, wr[j], acc[j]);
                    }
                }
#pragma unroll
            for (int j = 0; j < kDirectCo; ++j)
                if (co0 + j < d.c_out)
                    st_any(y, n * d.ys[0] + (co0 + j) * d.ys[1] + oh * d.ys[2] + ow * d.ys[3], d.y_dtype, acc[j]);
        }
    }
}

// one thread per input pixel, kDirectCo input channels at a time; s_w[(k, r, s)][c]
// here the DlsDirectConv "x" strides describe dx and the "y" strides describe dy
__global__ void __launch_bounds__(kDirectThreads)
conv_direct_dgrad_kernel(const void* __restrict__ dy, const void* __restrict__ w, void* __restrict__ dx,
                         const __grid_constant__ DlsDirectConv d) {
    extern __shared__ float s_w[];
    const int q_count = d.c_out * d.r * d.s;
    for (int i = threadIdx.x; i < q_count * d.c_in; i += blockDim.x) {
        const int c = i % d.c_in, q = i / d.c_in;
        const int s = q % d.s, r = (q / d.s) % d.r, k = q / (d.s * d.r);
        s_w[i] = ld_any(w, k * d.ws[0] + c * d.ws[1] + r * d.ws[2] + s * d.ws[3], d.w_dtype);
    }
    __syncthreads();
    const long long pixels = (long long)d.n * d.h * d.w;
    for (long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x; pix < pixels;
         pix += (long long)gridDim.x * blockDim.x) {
        const int iw = (int)(pix % d.w), ih = (int)((pix / d.w) % d.h), n = (int)(pix / ((long long)d.w * d.h));
        for (int c0 = 0; c0 < d.c_in; c0 += kDirectCo) {
            float acc[kDirectCo];
#pragma unroll
            for (int j = 0; j < kDirectCo; ++j) acc[j] = 0.f;
            for (int r = 0; r < d.r; ++r) {
                const int th = ih + d.pad_h - r;
                if (th < 0 || th % d.stride_h) continue;
                const int oh = th / d.stride_h;
                if (oh >= d.oh) continue;
                for (int s = 0; s < d.s; ++s) {
                    const int tw = iw + d.pad_w - s;
                    if (tw < 0 || tw % d.stride_w) continue;
                    const int ow = tw / d.stride_w;
                    if (ow >= d.ow) continue;
                    for (int k = 0; k < d.c_out; ++k) {
                        const float g = ld_any(dy, n * d.ys[0] + k * d.ys[1] + oh * d.ys[2] + ow * d.ys[3], d.y_dtype);
                        const float* wr = s_w + ((k * d.r + r) * d.s + s) * d.c_in + c0;
#pragma unroll
                        for (int j = 0; j < kDirectCo; ++j)
                            if (c0 + j < d.c_in) acc[j] = fmaf(g, wr[j], acc[j]);
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < kDirectCo; ++j)
                if (c0 + j < d.c_in)
                    st_any(dx, n * d.xs[0] + (c0 + j) * d.xs[1] + ih * d.xs[2] + iw * d.xs[3], d.x_dtype, acc[j]);
        }
    }
}

// Blocks stride over chunks of kWgPixels output pixels: dy rows and input patches of the chunk are staged
// in shared memory, every thread owns kJpt filter elements (j = k * Q + q) in registers across all its
// chunks and adds them to dw once at the end.  Outputs J .. J + c_out - 1 are the bias gradient.
constexpr int kWgPixels = 32;

template <int kJpt>
__global__ void __launch_bounds__(kDirectThreads)
conv_direct_wgrad_kernel(const void* __restrict__ dy, const void* __restrict__ x, float* __restrict__ dw,
                         float* __restrict__ dbias, const __grid_constant__ DlsDirectConv d) {
    extern __shared__ float s_buf[];
    const int q_count = d.c_in * d.r * d.s;
    const int j_count = d.c_out * q_count + (dbias != nullptr ? d.c_out : 0);
    float* s_dy = s_buf;                          // [kWgPixels][c_out]
    float* s_xp = s_buf + kWgPixels * d.c_out;    // [kWgPixels][q_count]
    float acc[kJpt];
    int jk[kJpt], jq[kJpt];
#pragma unroll
    for (int jj = 0; jj < kJpt; ++jj) {
        acc[jj] = 0.f;
        const int j = threadIdx.x + jj * kDirectThreads;
        if (j < d.c_out * q_count) { jk[jj] = j / q_count; jq[jj] = j % q_count; }
        else if (j < j_count) { jk[jj] = j - d.c_out * q_count; jq[jj] = -1; }
        else { jk[jj] = -1; jq[jj] = -1; }
    }
    const long long pixels = (long long)d.n * d.oh * d.ow;
    const long long chunks = (pixels + kWgPixels - 1) / kWgPixels;
    for (long long chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
        const long long p0 = chunk * kWgPixels;
        for (int i = threadIdx.x; i < kWgPixels * d.c_out; i += blockDim.x) {
            const int k = i % d.c_out;
            const long long pix = p0 + i / d.c_out;
            float v = 0.f;
            if (pix < pixels) {
                const int ow = (int)(pix % d.ow), oh = (int)((pix / d.ow) % d.oh);
                const int n = (int)(pix / ((long long)d.ow * d.oh));
                v = ld_any(dy, n * d.ys[0] + k * d.ys[1] + oh * d.ys[2] + ow * d.ys[3], d.y_dtype);
            }
            s_dy[i] = v;
        }
        for (int i = threadIdx.x; i < kWgPixels * q_count; i += blockDim.x) {
            const int q = i % q_count;
            const long long pix = p0 + i / q_count;
            float v = 0.f;
            if (pix < pixels) {
                const int ow = (int)(pix % d.ow), oh = (int)((pix / d.ow) % d.oh);
                const int n = (int)(pix / ((long long)d.ow * d.oh));
                const int s = q % d.s, r = (q / d.s) % d.r, c = q / (d.s * d.r);
                const int ih = oh * d.stride_h - d.pad_h + r, iw = ow * d.stride_w - d.pad_w + s;
                if (ih >= 0 && ih < d.h && iw >= 0 && iw < d.w)
                    v = ld_any(x, n * d.xs[0] + c * d.xs[1] + ih * d.xs[2] + iw * d.xs[3], d.x_dtype);
            }
            s_xp[i] = v;
        }
        __syncthreads();
#pragma unroll
        for (int jj = 0; jj < kJpt; ++jj) {
            if (jk[jj] < 0) continue;
            float a = acc[jj];
            if (jq[jj] >= 0) {
                for (int pp = 0; pp < kWgPixels; ++pp)
                    a = fmaf(s_dy[pp * d.c_out + jk[jj]], s_xp[pp * q_count + jq[jj]], a);
            } else {
                for (int pp = 0; pp < kWgPixels; ++pp) a += s_dy[pp * d.c_out + jk[jj]];
            }
            acc[jj] = a;
        }
        __syncthreads();
    }
#pragma unroll
    for (int jj = 0; jj < kJpt; ++jj) {
        if (jk[jj] < 0) continue;
        if (jq[jj] >= 0) {
            const int q = jq[jj];
            const int s = q % d.s, r = (q / d.s) % d.r, c = q / (d.s * d.r);
            atomicAdd(dw + jk[jj] * d.ws[0] + c * d.ws[1] + r * d.ws[2] + s * d.ws[3], acc[jj]);
        } else {
            atomicAdd(dbias + jk[jj], acc[jj]);
        }
    }
}

// Stem layers: one thread per (output pixel, 8 consecutive k) writes 16 bytes of the patch matrix
// [pixels][k_pad], k = (r * S + s) * c_in + c; the first blocks also re-pack the filter to [c_out][k_pad].
__global__ void __launch_bounds__(kDirectThreads)
conv_stem_im2col_kernel(const void* __restrict__ x, const void* __restrict__ w, __nv_bfloat16* __restrict__ patches,
                        __nv_bfloat16* __restrict__ w_packed, int k_pad, const __grid_constant__ DlsDirectConv d) {
    pdl_launch_dependents();
    pdl_wait();
    const int q_count = d.c_in * d.r * d.s;
    const int groups = k_pad / 8;
    const long long pixels = (long long)d.n * d.oh * d.ow;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < (long long)d.c_out * k_pad) {
        const int k = (int)(gid % k_pad), co = (int)(gid / k_pad);
        float v = 0.f;
        if (k < q_count) {
            const int c = k % d.c_in, s = (k / d.c_in) % d.s, r = k / (d.c_in * d.s);
            v = ld_any(w, co * d.ws[0] + c * d.ws[1] + r * d.ws[2] + s * d.ws[3], d.w_dtype);
        }
        w_packed[gid] = __float2bfloat16(v);
    }
    if (gid >= pixels * groups) return;
    const int g = (int)(gid % groups);
    const long long pix = gid / groups;
    const int ow = (int)(pix % d.ow), oh = (int)((pix / d.ow) % d.oh), n = (int)(pix / ((long long)d.ow * d.oh));
    const int ih0 = oh * d.stride_h - d.pad_h, iw0 = ow * d.stride_w - d.pad_w;
    uint4 pk;
    __nv_bfloat16* h = reinterpret_cast<__nv_bfloat16*>(&pk);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int k = g * 8 + j;
        float v = 0.f;
        if (k < q_count) {
            const int c = k % d.c_in, s = (k / d.c_in) % d.s, r = k / (d.c_in * d.s);
            const int ih = ih0 + r, iw = iw0 + s;
            if (ih >= 0 && ih < d.h && iw >= 0 && iw < d.w)
                v = ld_any(x, n * d.xs[0] + c * d.xs[1] + ih * d.xs[2] + iw * d.xs[3], d.x_dtype);
        }
        h[j] = __float2bfloat16(v);
    }
    *reinterpret_cast<uint4*>(patches + pix * k_pad + g * 8) = pk;
}

static int check_direct(const DlsDirectConv& d) {
    if (d.n < 1 || d.c_in < 1 || d.c_out < 1 || d.oh < 1 || d.ow < 1) return fail_msg("direct conv: empty problem");
    if ((long long)d.c_in * d.c_out * d.r * d.s > kDirectMaxFilter)
        return fail_msg("direct conv: filter larger than 12288 elements");
    return 0;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_conv_direct_fprop(const void* x, const void* w, const void* bias, int bias_dtype, void* y,
                                     const DlsDirectConv* d, void* stream) {
    if (int rc = check_direct(*d)) return rc;
    const size_t smem = (size_t)d->c_in * d->c_out * d->r * d->s * sizeof(float);
    if (smem > 48 * 1024)
        DLS_CHECK(cudaFuncSetAttribute(conv_direct_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const long long pixels = (long long)d->n * d->oh * d->ow;
    conv_direct_fprop_kernel<<<grid_for(pixels, kDirectThreads, 8), kDirectThreads, smem, (cudaStream_t)stream>>>(
        x, w, bias, bias_dtype, y, *d);
    DLS_LAUNCHED("conv_direct_fprop_kernel");
    return 0;
}

extern "C" int dls_conv_direct_dgrad(const void* dy, const void* w, void* dx, const DlsDirectConv* d, void* stream) {
    if (int rc = check_direct(*d)) return rc;
    const size_t smem = (size_t)d->c_in * d->c_out * d->r * d->s * sizeof(float);
    if (smem > 48 * 1024)
        DLS_CHECK(cudaFuncSetAttribute(conv_direct_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const long long pixels = (long long)d->n * d->h * d->w;
    conv_direct_dgrad_kernel<<<grid_for(pixels, kDirectThreads, 8), kDirectThreads, smem, (cudaStream_t)stream>>>(
        dy, w, dx, *d);
    DLS_LAUNCHED("conv_direct_dgrad_kernel");
    return 0;
}

extern "C" int dls_conv_direct_wgrad(const void* dy, const void* x, float* dw, float* dbias, const DlsDirectConv* d,
                                     void* stream) {
    if (int rc = check_direct(*d)) return rc;
    const int q_count = d->c_in * d->r * d->s;
    const int j_count = d->c_out * q_count + (dbias != nullptr ? d->c_out : 0);
    const size_t smem = (size_t)kWgPixels * (d->c_out + q_count) * sizeof(float);
    if (smem > 200 * 1024) return fail_msg("direct wgrad: shared-memory tile too large");
    const long long pixels = (long long)d->n * d->oh * d->ow;
    const long long chunks = (pixels + kWgPixels - 1) / kWgPixels;
    const int grid = (int)(chunks < 2 * kNumSMs ? chunks : 2 * kNumSMs);
    cudaStream_t st = (cudaStream_t)stream;
#define DLS_DIRECT_WGRAD(JPT)                                                                                    \
    do {                                                                                                         \
        if (smem > 48 * 1024)                                                                                    \
            DLS_CHECK(cudaFuncSetAttribute(conv_direct_wgrad_kernel<JPT>,                                        \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));             \
        conv_direct_wgrad_kernel<JPT><<<grid, kDirectThreads, smem, st>>>(dy, x, dw, dbias, *d);                 \
    } while (0)
    if (j_count <= 8 * kDirectThreads) DLS_DIRECT_WGRAD(8);
    else if (j_count <= 16 * kDirectThreads) DLS_DIRECT_WGRAD(16);
    else DLS_DIRECT_WGRAD(49);
#undef DLS_DIRECT_WGRAD
    DLS_LAUNCHED("conv_direct_wgrad_kernel");
    return 0;
}

extern "C" int dls_conv_stem_im2col(const void* x, const void* w, void* patches, void* w_packed, int k_pad,
                                    const DlsDirectConv* d, void* stream) {
    if (k_pad % 8 || k_pad < d->c_in * d->r * d->s) return fail_msg("stem im2col: k_pad must be a multiple of 8 >= Cin*R*S");
    if (((uintptr_t)patches | (uintptr_t)w_packed) & 15) return fail_msg("stem im2col: outputs must be 16-byte aligned");
    const long long pixels = (long long)d->n * d->oh * d->ow;
    long long work = pixels * (k_pad / 8);
    const long long w_work = (long long)d->c_out * k_pad;
    if (w_work > work) work = w_work;
    const long long blocks = (work + kDirectThreads - 1) / kDirectThreads;
    DLS_CHECK(launch_pdl(conv_stem_im2col_kernel, dim3((unsigned)blocks), dim3(kDirectThreads), 0, (cudaStream_t)stream, 0,
                         x, w, (__nv_bfloat16*)patches, (__nv_bfloat16*)w_packed, k_pad, *d));
    DLS_LAUNCHED("conv_stem_im2col_kernel");
    return 0;
}
