// Fused training-mode BatchNorm (+ residual add) (+ ReLU) for channels-last activations,
// ONE launch forward and ONE launch backward per layer.
//
// Local SGD of the CNN workloads (ResNet-18, DenseNet-40: the reference's conf/**/cifar*.yaml)
// spends two thirds of its GPU time in ATen's generic batch-norm and elementwise kernels once
// the convolutions run on tensor cores: collect_statistics + transform_input + relu (+ add)
// forward, backward_reduce + backward_elemt + relu/add backward -- six or more launches per
// layer, each on an activation tensor of at most a few MB, i.e. pure launch/latency cost.
//
// Design for Blackwell: the activation is [rows = N*H*W, C] with channels contiguous.  A CTA owns
// a few consecutive channels (W vectors of V elements) and a slice of the rows; the S CTAs that
// share a channel group form a thread-block CLUSTER and combine their partial statistics through
// distributed shared memory (cluster.sync + remote smem reads, ~0.2 us) instead of a round trip
// through global memory, atomics or a second kernel.  After the exchange every CTA holds the
// channel statistics and immediately runs the elementwise phase over its own rows (second read
// hits L2).  No scratch memory, deterministic summation order.
//   fwd:  mean/invstd -> y = relu(gamma * (x - mean) * invstd + beta [+ residual])
//   bwd:  dy' = dy * (y > 0); dbeta = sum dy'; dgamma = sum dy' * xhat;
//         dx = gamma * invstd * (dy' - mean(dy') - xhat * mean(dy' * xhat)) [, dresidual = dy']
#include <stdlib.h>

#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/dls_b200.h"

namespace cg = cooperative_groups;

namespace dls {

constexpr int kBnThreads = 512;
constexpr int kBnWarps = kBnThreads / 32;
constexpr int kBnMaxW = 8;        // channel vectors per CTA
constexpr int kBnMaxCluster = 16; // CTAs per cluster (non-portable size on B200)

template <typename T, int V> struct Vec;
template <> struct Vec<__nv_bfloat16, 8> {
    static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
        const uint4 raw = *reinterpret_cast<const uint4*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
        uint4 raw;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        *reinterpret_cast<uint4*>(p) = raw;
    }
};
template <> struct Vec<__nv_bfloat16, 4> {
    static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[4]) {
        const uint2 raw = *reinterpret_cast<const uint2*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 2; ++i) { const float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[4]) {
        uint2 raw;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 2; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        *reinterpret_cast<uint2*>(p) = raw;
    }
};
template <> struct Vec<float, 4> {
    static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
        const float4 f = *reinterpret_cast<const float4*>(p);
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
    static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};

// Raw (unconverted) vector of V elements of T, used to park a thread's rows in shared memory
// between the statistics phase and the elementwise phase (each thread re-reads only its own
// slots, so no synchronisation and no bank conflicts are involved).
template <typename T, int V> struct RawVec { using type = uint4; };
template <> struct RawVec<__nv_bfloat16, 4> { using type = uint2; };
template <typename T, int V>
__device__ __forceinline__ void raw_to_float(const typename RawVec<T, V>::type& raw, float (&v)[V]) {
    Vec<T, V>::load(reinterpret_cast<const T*>(&raw), v);
}
template <typename T, int V>
__device__ __forceinline__ typename RawVec<T, V>::type float_to_raw(const float (&v)[V]) {
    typename RawVec<T, V>::type raw;
    Vec<T, V>::store(reinterpret_cast<T*>(&raw), v);
    return raw;
}

// Work decomposition of one CTA.
struct BnTile {
    int w;            // channel vectors owned by this CTA's group (power of two <= 8)
    int my_w;         // this thread's vector inside the group
    long long col;    // element offset of this thread's first channel
    bool active;      // false for padding vectors of a ragged last group
    long long r0, r1; // row slice of this CTA
    int rows_per_iter;
    int my_r;
};

template <int V>
__device__ __forceinline__ BnTile bn_tile(long long rows, int channels, int w, int cluster_size, int cluster_rank) {
    BnTile t;
    t.w = w;
    const int group = blockIdx.x / cluster_size;
    const int cv = channels / V;
    t.my_w = threadIdx.x % w;
    const int vec = group * w + t.my_w;
    t.active = vec < cv;
    t.col = (long long)vec * V;
    const long long chunk = (rows + cluster_size - 1) / cluster_size;
    t.r0 = chunk * cluster_rank;
    t.r1 = t.r0 + chunk < rows ? t.r0 + chunk : rows;
    t.rows_per_iter = kBnThreads / w;
    t.my_r = threadIdx.x / w;
    return t;
}

// Sum a[V], b[V] over all threads of the CTA that own the same channel vector, then over the
// CTAs of the cluster (DSMEM).  On return s_tot[q][w*V + i] (q = 0 for a, 1 for b) holds the
// cluster-wide totals in EVERY CTA's shared memory.
template <int V>
__device__ __forceinline__ void bn_cluster_sum(const BnTile& t, float (&a)[V], float (&b)[V],
                                               float (*s_warp)[2 * kBnMaxW * 8], float* s_part, float* s_tot) {
    cg::cluster_group cluster = cg::this_cluster();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // lanes with equal (lane % w) own the same vector: butterfly over the other lane bits
    for (int o = 16; o >= t.w; o >>= 1) {
#pragma unroll
        for (int i = 0; i < V; ++i) {
            a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
            b[i] += __shfl_xor_sync(0xffffffffu, b[i], o);
        }
    }
    if (lane < t.w) {
#pragma unroll
        for (int i = 0; i < V; ++i) {
            s_warp[warp][lane * V + i] = a[i];
            s_warp[warp][kBnMaxW * 8 + lane * V + i] = b[i];
        }
    }
    __syncthreads();
    const int n = t.w * V;
    if (threadIdx.x < 2 * n) {
        const int q = threadIdx.x / n, j = threadIdx.x % n;
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < kBnWarps; ++k) acc += s_warp[k][q * kBnMaxW * 8 + j];
        s_part[q * kBnMaxW * 8 + j] = acc;
    }
    cluster.sync();  // every CTA's partials are visible cluster-wide
    if (threadIdx.x < 2 * n) {
        const int q = threadIdx.x / n, j = threadIdx.x % n;
        float acc = 0.f;
        const unsigned int size = cluster.num_blocks();
        for (unsigned int r = 0; r < size; ++r) acc += cluster.map_shared_rank(s_part, r)[q * kBnMaxW * 8 + j];
        s_tot[q * kBnMaxW * 8 + j] = acc;
    }
    cluster.sync();  // nobody leaves (or overwrites s_part) while a peer may still read it
}

template <typename T, int V>
__global__ void __launch_bounds__(kBnThreads)
bn_fwd_kernel(const T* __restrict__ x, const T* __restrict__ residual, long long rows, int channels,
              const float* __restrict__ gamma, const float* __restrict__ beta, float eps, int relu, int w,
              int cached, T* __restrict__ y, float* __restrict__ mean, float* __restrict__ invstd) {
    pdl_launch_dependents();
    pdl_wait();
    using Raw = typename RawVec<T, V>::type;
    extern __shared__ __align__(16) unsigned char s_dyn[];
    Raw* s_x = reinterpret_cast<Raw*>(s_dyn);
    __shared__ float s_warp[kBnWarps][2 * kBnMaxW * 8];
    __shared__ float s_part[2 * kBnMaxW * 8];
    __shared__ float s_tot[2 * kBnMaxW * 8];
    cg::cluster_group cluster = cg::this_cluster();
    const BnTile t = bn_tile<V>(rows, channels, w, (int)cluster.num_blocks(), (int)cluster.block_rank());
    float s[V], q[V];
#pragma unroll
    for (int i = 0; i < V; ++i) { s[i] = 0.f; q[i] = 0.f; }
    if (t.active) {
        const T* base = x + t.col;
        int it = 0;
#pragma unroll 4
        for (long long r = t.r0 + t.my_r; r < t.r1; r += t.rows_per_iter, ++it) {
            const Raw raw = *reinterpret_cast<const Raw*>(base + r * channels);
            if (cached) s_x[it * kBnThreads + threadIdx.x] = raw;
            float v[V];
            raw_to_float<T, V>(raw, v);
#pragma unroll
            for (int i = 0; i < V; ++i) { s[i] += v[i]; q[i] += v[i] * v[i]; }
        }
    }
    bn_cluster_sum<V>(t, s, q, s_warp, s_part, s_tot);
    if (t.active) {
        float scale[V], shift[V];
        const float inv_rows = 1.f / (float)rows;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int c = (int)t.col + i;
            const float mu = s_tot[t.my_w * V + i] * inv_rows;
            float var = s_tot[kBnMaxW * 8 + t.my_w * V + i] * inv_rows - mu * mu;
            var = var > 0.f ? var : 0.f;
            const float is = rsqrtf(var + eps);
            scale[i] = gamma[c] * is;
            shift[i] = beta[c] - mu * scale[i];
            if (cluster.block_rank() == 0 && t.my_r == 0) { mean[c] = mu; invstd[c] = is; }
        }
        int it = 0;
#pragma unroll 4
        for (long long r = t.r0 + t.my_r; r < t.r1; r += t.rows_per_iter, ++it) {
            const long long off = r * channels + t.col;
            float v[V];
            if (cached) raw_to_float<T, V>(s_x[it * kBnThreads + threadIdx.x], v);
            else Vec<T, V>::load(x + off, v);
            if (residual != nullptr) {
                float z[V];
                Vec<T, V>::load(residual + off, z);
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaf(v[i], scale[i], shift[i]) + z[i];
            } else {
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaf(v[i], scale[i], shift[i]);
            }
            if (relu) {
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaxf(v[i], 0.f);
            }
            Vec<T, V>::store(y + off, v);
        }
    }
}

template <typename T, int V>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, long long rows,
              int channels, const float* __restrict__ gamma, const float* __restrict__ mean,
              const float* __restrict__ invstd, int relu, int w, int cached, T* __restrict__ dx,
              T* __restrict__ dresidual, float* __restrict__ dgamma, float* __restrict__ dbeta) {
    pdl_launch_dependents();
    pdl_wait();
    using Raw = typename RawVec<T, V>::type;
    extern __shared__ __align__(16) unsigned char s_dyn[];
    Raw* s_g = reinterpret_cast<Raw*>(s_dyn);   // masked upstream gradient, then x, per thread slot
    __shared__ float s_warp[kBnWarps][2 * kBnMaxW * 8];
    __shared__ float s_part[2 * kBnMaxW * 8];
    __shared__ float s_tot[2 * kBnMaxW * 8];
    cg::cluster_group cluster = cg::this_cluster();
    const BnTile t = bn_tile<V>(rows, channels, w, (int)cluster.num_blocks(), (int)cluster.block_rank());
    float s1[V], s2[V], mu[V], is[V];
#pragma unroll
    for (int i = 0; i < V; ++i) { s1[i] = 0.f; s2[i] = 0.f; mu[i] = 0.f; is[i] = 0.f; }
    if (t.active) {
#pragma unroll
        for (int i = 0; i < V; ++i) { mu[i] = mean[t.col + i]; is[i] = invstd[t.col + i]; }
        int it = 0;
#pragma unroll 2
        for (long long r = t.r0 + t.my_r; r < t.r1; r += t.rows_per_iter, ++it) {
            const long long off = r * channels + t.col;
            const Raw xr = *reinterpret_cast<const Raw*>(x + off);
            float g[V], xv[V];
            Vec<T, V>::load(dy + off, g);
            raw_to_float<T, V>(xr, xv);
            if (relu) {
                float yv[V];
                Vec<T, V>::load(y + off, yv);
#pragma unroll
                for (int i = 0; i < V; ++i) g[i] = yv[i] > 0.f ? g[i] : 0.f;
            }
            if (cached) {
                s_g[(2 * it) * kBnThreads + threadIdx.x] = float_to_raw<T, V>(g);
                s_g[(2 * it + 1) * kBnThreads + threadIdx.x] = xr;
            }
#pragma unroll
            for (int i = 0; i < V; ++i) { s1[i] += g[i]; s2[i] += g[i] * (xv[i] - mu[i]) * is[i]; }
        }
    }
    bn_cluster_sum<V>(t, s1, s2, s_warp, s_part, s_tot);
    if (t.active) {
        float k[V], a[V], b[V];
        const float inv_rows = 1.f / (float)rows;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int c = (int)t.col + i;
            const float sum1 = s_tot[t.my_w * V + i], sum2 = s_tot[kBnMaxW * 8 + t.my_w * V + i];
            if (cluster.block_rank() == 0 && t.my_r == 0) { dbeta[c] = sum1; dgamma[c] = sum2; }
            k[i] = gamma[c] * is[i];
            a[i] = sum1 * inv_rows;
            b[i] = sum2 * inv_rows;
        }
        int it = 0;
#pragma unroll 2
        for (long long r = t.r0 + t.my_r; r < t.r1; r += t.rows_per_iter, ++it) {
            const long long off = r * channels + t.col;
            float g[V], xv[V];
            if (cached) {
                raw_to_float<T, V>(s_g[(2 * it) * kBnThreads + threadIdx.x], g);
                raw_to_float<T, V>(s_g[(2 * it + 1) * kBnThreads + threadIdx.x], xv);
            } else {
                Vec<T, V>::load(dy + off, g);
                Vec<T, V>::load(x + off, xv);
                if (relu) {
                    float yv[V];
                    Vec<T, V>::load(y + off, yv);
#pragma unroll
                    for (int i = 0; i < V; ++i) g[i] = yv[i] > 0.f ? g[i] : 0.f;
                }
            }
            if (dresidual != nullptr) Vec<T, V>::store(dresidual + off, g);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const float xh = (xv[i] - mu[i]) * is[i];
                g[i] = k[i] * (g[i] - a[i] - xh * b[i]);
            }
            Vec<T, V>::store(dx + off, g);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Row-major ("wide") variant for activations with many rows and few channels (ResNet 64/128
// channel stages): a CTA owns ALL channels of a slice of rows, so a warp reads whole rows
// (512 contiguous bytes per instruction) instead of a 16-byte column strip of 32 different rows.
// Partial statistics are folded inside the CTA, then across the cluster through DSMEM; one partial
// per cluster goes to a small global scratch and the elementwise kernel folds those (<= 32) while
// it builds its per-channel scale/shift table in shared memory.  Two launches each way.
constexpr int kWideThreads = 512;
constexpr int kWideCluster = 8;
constexpr int kWideMaxClusters = 32;

// block-level fold of per-thread partials over the threads that own the same channel vector
template <int V>
__device__ __forceinline__ void wide_block_fold(int cv, const float (&a)[V], const float (&b)[V], float* s_buf,
                                                float* s_part, int channels) {
    // s_buf: [kWideThreads][2V]
#pragma unroll
    for (int i = 0; i < V; ++i) {
        s_buf[threadIdx.x * 2 * V + i] = a[i];
        s_buf[threadIdx.x * 2 * V + V + i] = b[i];
    }
    __syncthreads();
    const int rows_in_block = kWideThreads / cv;  // threads past rows_in_block * cv are idle (zeros)
    for (int j = threadIdx.x; j < cv * 2 * V; j += kWideThreads) {
        const int vec = j / (2 * V), k = j % (2 * V);
        float acc = 0.f;
        for (int r = 0; r < rows_in_block; ++r) acc += s_buf[(r * cv + vec) * 2 * V + k];
        const int q = k / V, i = k % V;
        s_part[q * channels + vec * V + i] = acc;
    }
}

// kAtomic: the CTA adds its partial sums to `partials` ([2][channels], zeroed by the caller) with atomics instead of
// folding them across a cluster: no cluster barriers, no DSMEM reads, and the apply kernel reads one partial.
template <typename T, int V, bool kBackward, bool kAtomic = false>
__global__ void __launch_bounds__(kWideThreads)
bn_wide_reduce_kernel(const T* __restrict__ x, const T* __restrict__ dy, const T* __restrict__ y, long long rows,
                      int channels, const float* __restrict__ mean, const float* __restrict__ invstd, int relu,
                      float* __restrict__ partials) {
    pdl_launch_dependents();
    pdl_wait();
    extern __shared__ __align__(16) float s_wide[];
    float* s_buf = s_wide;                              // [kWideThreads][2V]
    float* s_part = s_wide + kWideThreads * 2 * V;      // [2][channels]
    cg::cluster_group cluster = cg::this_cluster();
    const int cv = channels / V;
    const int vec = threadIdx.x % cv, my_r = threadIdx.x / cv;
    const int rows_per_iter = kWideThreads / cv;
    const long long chunk = (rows + gridDim.x - 1) / gridDim.x;
    const long long r0 = chunk * blockIdx.x, r1 = r0 + chunk < rows ? r0 + chunk : rows;
    float a[V], b[V], mu[V], is[V];
#pragma unroll
    for (int i = 0; i < V; ++i) { a[i] = 0.f; b[i] = 0.f; mu[i] = 0.f; is[i] = 0.f; }
    if (kBackward) {
#pragma unroll
        for (int i = 0; i < V; ++i) { mu[i] = mean[vec * V + i]; is[i] = invstd[vec * V + i]; }
    }
    if (my_r < rows_per_iter) {
#pragma unroll 4
        for (long long r = r0 + my_r; r < r1; r += rows_per_iter) {
            const long long off = r * channels + vec * V;
            float xv[V];
            Vec<T, V>::load(x + off, xv);
            if (kBackward) {
                float g[V];
                Vec<T, V>::load(dy + off, g);
                if (relu) {
                    float yv[V];
                    Vec<T, V>::load(y + off, yv);
#pragma unroll
                    for (int i = 0; i < V; ++i) g[i] = yv[i] > 0.f ? g[i] : 0.f;
                }
#pragma unroll
                for (int i = 0; i < V; ++i) { a[i] += g[i]; b[i] += g[i] * (xv[i] - mu[i]) * is[i]; }
            } else {
#pragma unroll
                for (int i = 0; i < V; ++i) { a[i] += xv[i]; b[i] += xv[i] * xv[i]; }
            }
        }
    }
    wide_block_fold<V>(cv, a, b, s_buf, s_part, channels);
    if (kAtomic) {
        __syncthreads();
        for (int j = threadIdx.x; j < 2 * channels; j += kWideThreads) {
            const float v = s_part[j];
            if (v != 0.f) atomicAdd(partials + j, v);
        }
        return;
    }
    cluster.sync();
    if (cluster.block_rank() == 0) {
        const unsigned int size = cluster.num_blocks();
        float* out = partials + (size_t)(blockIdx.x / size) * 2 * channels;
        for (int j = threadIdx.x; j < 2 * channels; j += kWideThreads) {
            float acc = 0.f;
            for (unsigned int r = 0; r < size; ++r) acc += cluster.map_shared_rank(s_part, r)[j];
            out[j] = acc;
        }
    }
    cluster.sync();
}

template <typename T, int V, bool kBackward>
__global__ void __launch_bounds__(kWideThreads)
bn_wide_apply_kernel(const T* __restrict__ x, const T* __restrict__ aux, const T* __restrict__ y_or_res,
                     long long rows, int channels, const float* __restrict__ gamma, const float* __restrict__ beta,
                     float eps, int relu, const float* __restrict__ partials, int n_partials,
                     float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, T* __restrict__ out, T* __restrict__ dresidual) {
    pdl_launch_dependents();
    pdl_wait();
    // forward : x = input, y_or_res = residual (may be null), out = y; writes mean / invstd
    // backward: x = input, aux = dy, y_or_res = forward output (ReLU mask), out = dx; reads mean /
    //           invstd, writes dgamma / dbeta
    extern __shared__ __align__(16) float s_tab[];  // [2][channels]: fwd scale/shift, bwd a/b
    float* s_t0 = s_tab;
    float* s_t1 = s_tab + channels;
    const float inv_rows = 1.f / (float)rows;
    for (int c = threadIdx.x; c < channels; c += kWideThreads) {
        float p0 = 0.f, p1 = 0.f;
        for (int k = 0; k < n_partials; ++k) {
            p0 += partials[(size_t)k * 2 * channels + c];
            p1 += partials[(size_t)k * 2 * channels + channels + c];
        }
        if (!kBackward) {
            const float mu = p0 * inv_rows;
            float var = p1 * inv_rows - mu * mu;
            var = var > 0.f ? var : 0.f;
            const float is = rsqrtf(var + eps);
            const float sc = gamma[c] * is;
            s_t0[c] = sc;
            s_t1[c] = beta[c] - mu * sc;
            if (blockIdx.x == 0) { mean[c] = mu; invstd[c] = is; }
        } else {
            s_t0[c] = p0 * inv_rows;
            s_t1[c] = p1 * inv_rows;
            if (blockIdx.x == 0) { dbeta[c] = p0; dgamma[c] = p1; }
        }
    }
    __syncthreads();
    const int cv = channels / V;
    const int vec = threadIdx.x % cv, my_r = threadIdx.x / cv;
    const int rows_per_iter = kWideThreads / cv;
    if (my_r >= rows_per_iter) return;
    float t0[V], t1[V], mu[V], is[V], k[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
        const int c = vec * V + i;
        t0[i] = s_t0[c]; t1[i] = s_t1[c];
        mu[i] = 0.f; is[i] = 0.f; k[i] = 0.f;
        if (kBackward) { mu[i] = mean[c]; is[i] = invstd[c]; k[i] = gamma[c] * is[i]; }
    }
    const long long stride = (long long)gridDim.x * rows_per_iter;
#pragma unroll 2
    for (long long r = (long long)blockIdx.x * rows_per_iter + my_r; r < rows; r += stride) {
        const long long off = r * channels + vec * V;
        float v[V];
        Vec<T, V>::load(x + off, v);
        if (!kBackward) {
            if (y_or_res != nullptr) {
                float z[V];
                Vec<T, V>::load(y_or_res + off, z);
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaf(v[i], t0[i], t1[i]) + z[i];
            } else {
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaf(v[i], t0[i], t1[i]);
            }
            if (relu) {
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = fmaxf(v[i], 0.f);
            }
            Vec<T, V>::store(out + off, v);
        } else {
            float g[V];
            Vec<T, V>::load(aux + off, g);
            if (relu) {
                float yv[V];
                Vec<T, V>::load(y_or_res + off, yv);
#pragma unroll
                for (int i = 0; i < V; ++i) g[i] = yv[i] > 0.f ? g[i] : 0.f;
            }
            if (dresidual != nullptr) Vec<T, V>::store(dresidual + off, g);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const float xh = (v[i] - mu[i]) * is[i];
                g[i] = k[i] * (g[i] - t0[i] - xh * t1[i]);
            }
            Vec<T, V>::store(out + off, g);
        }
    }
}

// wide path applies when rows are plentiful and a row is at most 128 channel vectors wide (any count:
// the threads of a CTA that do not fit a whole number of rows stay idle)
static bool bn_use_wide(long long rows, int channels, int v) {
    static const bool disabled = [] { const char* e = getenv("DLS_BN_WIDE"); return e && e[0] == '0'; }();
    static const long long min_rows = [] {
        const char* e = getenv("DLS_BN_WIDE_MIN_ROWS");
        return e ? atoll(e) : 8192LL;
    }();
    const int cv = channels / v;
    return !disabled && cv <= 128 && rows >= min_rows;
}
static int bn_wide_clusters(long long rows, int channels, int v) {
    const int rows_per_iter = kWideThreads / (channels / v);
    long long want = rows / ((long long)rows_per_iter * kWideCluster * 4);  // >= 4 rows per thread
    if (want > kWideMaxClusters) want = kWideMaxClusters;
    if (want * kWideCluster > 2 * kNumSMs) want = 2 * kNumSMs / kWideCluster;
    return want < 1 ? 1 : (int)want;
}

template <typename K, typename... Args>
static cudaError_t wide_launch(K kernel, int grid, int cluster, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kWideThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

template <typename T, int V>
static cudaError_t bn_wide_fwd(const void* x, long long rows, int channels, const float* gamma, const float* beta,
                               const void* residual, float eps, int relu, void* y, float* mean, float* invstd,
                               float* scratch, cudaStream_t s) {
    const int clusters = bn_wide_clusters(rows, channels, V);
    const size_t smem1 = (size_t)(kWideThreads * 2 * V + 2 * channels) * sizeof(float);
    cudaError_t e = wide_launch(bn_wide_reduce_kernel<T, V, false>, clusters * kWideCluster, kWideCluster, smem1, s,
                                (const T*)x, (const T*)nullptr, (const T*)nullptr, rows, channels,
                                (const float*)nullptr, (const float*)nullptr, 0, scratch);
    if (e != cudaSuccess) return e;
    g_launch_count.fetch_add(1);
    const int grid2 = (int)((rows + (kWideThreads / (channels / V)) * 4 - 1) / ((kWideThreads / (channels / V)) * 4));
    return wide_launch(bn_wide_apply_kernel<T, V, false>, grid2 > 2 * kNumSMs ? 2 * kNumSMs : grid2, 1,
                       (size_t)2 * channels * sizeof(float), s, (const T*)x, (const T*)nullptr, (const T*)residual,
                       rows, channels, gamma, beta, eps, relu, (const float*)scratch, clusters, mean, invstd,
                       (float*)nullptr, (float*)nullptr, (T*)y, (T*)nullptr);
}

template <typename T, int V>
static cudaError_t bn_wide_bwd(const void* dy, const void* y, const void* x, long long rows, int channels,
                               const float* gamma, const float* mean, const float* invstd, int relu, void* dx,
                               void* dresidual, float* dgamma, float* dbeta, float* scratch, float* zeroed,
                               cudaStream_t s) {
    const int clusters = bn_wide_clusters(rows, channels, V);
    const size_t smem1 = (size_t)(kWideThreads * 2 * V + 2 * channels) * sizeof(float);
    cudaError_t e;
    if (zeroed != nullptr) {
        e = wide_launch(bn_wide_reduce_kernel<T, V, true, true>, clusters * kWideCluster, 1, smem1, s, (const T*)x,
                        (const T*)dy, (const T*)y, rows, channels, mean, invstd, relu, zeroed);
    } else {
        e = wide_launch(bn_wide_reduce_kernel<T, V, true>, clusters * kWideCluster, kWideCluster, smem1, s,
                        (const T*)x, (const T*)dy, (const T*)y, rows, channels, mean, invstd, relu, scratch);
    }
    if (e != cudaSuccess) return e;
    g_launch_count.fetch_add(1);
    const int grid2 = (int)((rows + (kWideThreads / (channels / V)) * 4 - 1) / ((kWideThreads / (channels / V)) * 4));
    return wide_launch(bn_wide_apply_kernel<T, V, true>, grid2 > 2 * kNumSMs ? 2 * kNumSMs : grid2, 1,
                       (size_t)2 * channels * sizeof(float), s, (const T*)x, (const T*)dy, (const T*)y, rows,
                       channels, gamma, (const float*)nullptr, 0.f, relu,
                       (const float*)(zeroed != nullptr ? zeroed : scratch), zeroed != nullptr ? 1 : clusters,
                       const_cast<float*>(mean), const_cast<float*>(invstd), dgamma, dbeta, (T*)dx, (T*)dresidual);
}

struct BnPlan { int v, w, groups, cluster; long long iters; int vec_bytes; };
constexpr size_t kBnMaxDynSmem = 200 * 1024;

static int bn_plan(long long rows, int channels, int dtype, BnPlan* p) {
    if (rows < 1 || channels < 4) return fail_msg("bn: empty input");
    if (dtype != 0 && dtype != 1) return fail_msg("bn: dtype must be 0 (fp32) or 1 (bf16)");
    if (dtype == 1 && channels % 8 == 0) p->v = 8;
    else if (channels % 4 == 0) p->v = 4;
    else return fail_msg("bn: channels must be a multiple of 4");
    const int cv = channels / p->v;
    // one channel vector per CTA unless that would mean more CTAs than two waves can hold
    p->w = 1;
    while ((cv + p->w - 1) / p->w > kNumSMs && p->w < kBnMaxW) p->w *= 2;
    p->groups = (cv + p->w - 1) / p->w;
    // split the rows over a cluster so that all CTAs fit in ONE wave (a CTA that parks its rows in
    // shared memory owns its SM) and a CTA keeps >= 64 rows
    static const int max_cluster = [] {
        const char* e = getenv("DLS_BN_MAX_CLUSTER");  // tuning knob
        const int v = e ? atoi(e) : kBnMaxCluster;
        return v < 1 ? 1 : (v > kBnMaxCluster ? kBnMaxCluster : v);
    }();
    int s = 1;
    while (s * 2 <= max_cluster && p->groups * s * 2 <= kNumSMs && rows / (s * 2) >= 64) s *= 2;
    p->cluster = s;
    const long long rows_per_cta = (rows + s - 1) / s;
    const int rows_per_iter = kBnThreads / p->w;
    p->iters = (rows_per_cta + rows_per_iter - 1) / rows_per_iter;
    p->vec_bytes = p->v * (dtype == 1 ? 2 : 4);
    return 0;
}

template <typename K, typename... Args>
static cudaError_t bn_launch(K kernel, const BnPlan& p, size_t dyn_smem, cudaStream_t stream, Args... args) {
    static thread_local const void* configured[16] = {nullptr};
    bool seen = false;
    for (const void* k : configured) seen |= (k == (const void*)kernel);
    if (!seen) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBnMaxDynSmem);
        if (e != cudaSuccess) return e;
        for (auto& k : configured) if (k == nullptr) { k = (const void*)kernel; break; }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.dynamicSmemBytes = dyn_smem;
    cfg.gridDim = dim3((unsigned)(p.groups * p.cluster));
    cfg.blockDim = dim3(kBnThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

}  // namespace dls

using namespace dls;

extern "C" int dls_bn_scratch_floats(long long rows, int channels, int dtype) {
    BnPlan p;
    if (bn_plan(rows, channels, dtype, &p)) return -1;
    return bn_use_wide(rows, channels, p.v) ? kWideMaxClusters * 2 * channels : 0;
}

extern "C" int dls_bn_fwd(const void* x, int dtype, long long rows, int channels, const float* gamma,
                          const float* beta, const void* residual, float eps, int relu, void* y, float* mean,
                          float* invstd, float* scratch, void* stream) {
    BnPlan p;
    if (int rc = bn_plan(rows, channels, dtype, &p)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    if (bn_use_wide(rows, channels, p.v)) {
        if (scratch == nullptr) return fail_msg("dls_bn_fwd: this shape needs the scratch buffer");
        if (dtype == 1 && p.v == 8) e = bn_wide_fwd<__nv_bfloat16, 8>(x, rows, channels, gamma, beta, residual, eps, relu, y, mean, invstd, scratch, s);
        else if (dtype == 1) e = bn_wide_fwd<__nv_bfloat16, 4>(x, rows, channels, gamma, beta, residual, eps, relu, y, mean, invstd, scratch, s);
        else e = bn_wide_fwd<float, 4>(x, rows, channels, gamma, beta, residual, eps, relu, y, mean, invstd, scratch, s);
        if (e != cudaSuccess) return fail("bn_wide_fwd", e);
        DLS_LAUNCHED("bn_wide_apply_kernel");
        return 0;
    }
    // park the CTA's rows in shared memory between the two phases when they fit
    size_t smem = (size_t)p.iters * kBnThreads * p.vec_bytes;
    const int cached = smem <= kBnMaxDynSmem;
    if (!cached) smem = 0;
    if (dtype == 1 && p.v == 8)
        e = bn_launch(bn_fwd_kernel<__nv_bfloat16, 8>, p, smem, s, (const __nv_bfloat16*)x,
                      (const __nv_bfloat16*)residual, rows, channels, gamma, beta, eps, relu, p.w, cached,
                      (__nv_bfloat16*)y, mean, invstd);
    else if (dtype == 1)
        e = bn_launch(bn_fwd_kernel<__nv_bfloat16, 4>, p, smem, s, (const __nv_bfloat16*)x,
                      (const __nv_bfloat16*)residual, rows, channels, gamma, beta, eps, relu, p.w, cached,
                      (__nv_bfloat16*)y, mean, invstd);
    else
        e = bn_launch(bn_fwd_kernel<float, 4>, p, smem, s, (const float*)x, (const float*)residual, rows, channels,
                      gamma, beta, eps, relu, p.w, cached, (float*)y, mean, invstd);
    if (e != cudaSuccess) return fail("bn_fwd_kernel", e);
    DLS_LAUNCHED("bn_fwd_kernel");
    return 0;
}

// Forward pass when the producing convolution already accumulated the channel statistics in its epilogue
// (stats = fp32 [2][channels]: sum, sum of squares): the reduction launch disappears, one elementwise launch
// builds scale / shift from the sums and applies normalisation (+ residual) (+ ReLU).
extern "C" int dls_bn_fwd_stats(const void* x, int dtype, long long rows, int channels, const float* gamma,
                                const float* beta, const void* residual, float eps, int relu, void* y, float* mean,
                                float* invstd, const float* stats, void* stream) {
    BnPlan p;
    if (int rc = bn_plan(rows, channels, dtype, &p)) return rc;
    if (stats == nullptr) return fail_msg("dls_bn_fwd_stats: stats is NULL");
    const int cv = channels / p.v;
    if (cv > kWideThreads) return fail_msg("dls_bn_fwd_stats: too many channels for one row per CTA pass");
    cudaStream_t s = (cudaStream_t)stream;
    const int rows_per_iter = kWideThreads / cv;
    long long grid = (rows + (long long)rows_per_iter * 4 - 1) / ((long long)rows_per_iter * 4);
    if (grid > 2 * kNumSMs) grid = 2 * kNumSMs;
    if (grid < 1) grid = 1;
    const size_t smem = (size_t)2 * channels * sizeof(float);
    cudaError_t e;
#define DLS_BN_APPLY(T, V)                                                                                         \
    e = wide_launch(bn_wide_apply_kernel<T, V, false>, (int)grid, 1, smem, s, (const T*)x, (const T*)nullptr,     \
                    (const T*)residual, rows, channels, gamma, beta, eps, relu, stats, 1, mean, invstd,            \
                    (float*)nullptr, (float*)nullptr, (T*)y, (T*)nullptr)
    if (dtype == 1 && p.v == 8) DLS_BN_APPLY(__nv_bfloat16, 8);
    else if (dtype == 1) DLS_BN_APPLY(__nv_bfloat16, 4);
    else DLS_BN_APPLY(float, 4);
#undef DLS_BN_APPLY
    if (e != cudaSuccess) return fail("bn_wide_apply_kernel", e);
    DLS_LAUNCHED("bn_wide_apply_kernel");
    return 0;
}

static int bn_bwd_impl(const void* dy, const void* y, const void* x, int dtype, long long rows, int channels,
                       const float* gamma, const float* mean, const float* invstd, int relu, void* dx,
                       void* dresidual, float* dgamma, float* dbeta, float* scratch, float* zeroed, void* stream) {
    BnPlan p;
    if (int rc = bn_plan(rows, channels, dtype, &p)) return rc;
    if (relu && y == nullptr) return fail_msg("bn bwd: the forward output is needed for the ReLU mask");
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    if (bn_use_wide(rows, channels, p.v)) {
        if (scratch == nullptr && zeroed == nullptr) return fail_msg("dls_bn_bwd: this shape needs the scratch buffer");
        if (dtype == 1 && p.v == 8) e = bn_wide_bwd<__nv_bfloat16, 8>(dy, y, x, rows, channels, gamma, mean, invstd, relu, dx, dresidual, dgamma, dbeta, scratch, zeroed, s);
        else if (dtype == 1) e = bn_wide_bwd<__nv_bfloat16, 4>(dy, y, x, rows, channels, gamma, mean, invstd, relu, dx, dresidual, dgamma, dbeta, scratch, zeroed, s);
        else e = bn_wide_bwd<float, 4>(dy, y, x, rows, channels, gamma, mean, invstd, relu, dx, dresidual, dgamma, dbeta, scratch, zeroed, s);
        if (e != cudaSuccess) return fail("bn_wide_bwd", e);
        DLS_LAUNCHED("bn_wide_apply_kernel");
        return 0;
    }
    size_t smem = (size_t)p.iters * kBnThreads * p.vec_bytes * 2;  // masked dy and x
    const int cached = smem <= kBnMaxDynSmem;
    if (!cached) smem = 0;
    if (dtype == 1 && p.v == 8)
        e = bn_launch(bn_bwd_kernel<__nv_bfloat16, 8>, p, smem, s, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,
                      (const __nv_bfloat16*)x, rows, channels, gamma, mean, invstd, relu, p.w, cached,
                      (__nv_bfloat16*)dx, (__nv_bfloat16*)dresidual, dgamma, dbeta);
    else if (dtype == 1)
        e = bn_launch(bn_bwd_kernel<__nv_bfloat16, 4>, p, smem, s, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,
                      (const __nv_bfloat16*)x, rows, channels, gamma, mean, invstd, relu, p.w, cached,
                      (__nv_bfloat16*)dx, (__nv_bfloat16*)dresidual, dgamma, dbeta);
    else
        e = bn_launch(bn_bwd_kernel<float, 4>, p, smem, s, (const float*)dy, (const float*)y, (const float*)x, rows,
                      channels, gamma, mean, invstd, relu, p.w, cached, (float*)dx, (float*)dresidual, dgamma,
                      dbeta);
    if (e != cudaSuccess) return fail("bn_bwd_kernel", e);
    DLS_LAUNCHED("bn_bwd_kernel");
    return 0;
}

extern "C" int dls_bn_bwd(const void* dy, const void* y, const void* x, int dtype, long long rows, int channels,
                          const float* gamma, const float* mean, const float* invstd, int relu, void* dx,
                          void* dresidual, float* dgamma, float* dbeta, float* scratch, void* stream) {
    return bn_bwd_impl(dy, y, x, dtype, rows, channels, gamma, mean, invstd, relu, dx, dresidual, dgamma, dbeta, scratch,
                       nullptr, stream);
}

// Same, for callers that own a zeroed fp32 [2][channels] buffer (e.g. a slice of a per-step arena cleared by one
// fill): on the row-major shapes the reduction pass adds its partial sums there with atomics and the apply pass
// reads that single partial - no cluster fold.  Shapes on the channel-major path ignore the buffer.
extern "C" int dls_bn_bwd_acc(const void* dy, const void* y, const void* x, int dtype, long long rows, int channels,
                              const float* gamma, const float* mean, const float* invstd, int relu, void* dx,
                              void* dresidual, float* dgamma, float* dbeta, float* zeroed_partials, void* stream) {
    if (zeroed_partials == nullptr) return fail_msg("dls_bn_bwd_acc: needs the zeroed partial-sum buffer");
    return bn_bwd_impl(dy, y, x, dtype, rows, channels, gamma, mean, invstd, relu, dx, dresidual, dgamma, dbeta, nullptr,
                       zeroed_partials, stream);
}
