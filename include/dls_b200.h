/* C ABI of libdls_b200.so -- the hand-written sm_100a kernels of the B200-native
 * federated-learning simulator.  Every entry point returns 0 on success or a non-zero
 * code (see dls_last_error()).  Pointers are device pointers unless named h_*; `stream`
 * is a cudaStream_t passed as void*.  Python binds these through ctypes
 * (distributed_learning_simulator_b200/ops/runtime.py). */
#ifndef DLS_B200_H
#define DLS_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* dls_last_error(void);
long long dls_launch_count(void);
void dls_reset_launch_count(void);

/* ---- aggregation (reference: simulation_lib/algorithm/fed_avg_algorithm.py:39-81,
 *      message.py:41-49 restore, server/server.py:114-120 broadcast) -------------------
 * out[i] = old_scale*old[i] + (sum_k w[k]*rows[k][i]) * inv_total  for i in [lo, hi),
 * fp64 accumulation, written to every dsts[d] (local or peer-mapped buffers: the
 * broadcast is fused into the reduction epilogue).  rows/weights are device arrays; dsts is a
 * HOST array of n_dsts device pointers (it travels in the launch arguments).
 * If flags != NULL the kernel runs a device-side barrier over n_ranks before reading
 * (peers' uploads are complete) and after writing (everyone's slice has landed):
 * peer_flags[r] points at rank r's flag block (mapped locally), epoch grows by one per
 * call.  nan_flag (uint32, device) is set to 1 if any output is NaN. */
int dls_fedavg_reduce_f32(const float* const* rows, const float* weights, int n_rows, const float* old,
                          float old_scale, double inv_total, float* const* dsts, int n_dsts, long long lo,
                          long long hi, uint32_t* const* peer_flags, int n_ranks, int rank, uint32_t epoch,
                          uint32_t* nan_flag, void* stream);
/* Same reduction, broadcast through NVLS: every output vector is ONE multimem.st to `mc_dst`, the
 * multicast address of the destination (NVSwitch replicates it into every bound GPU, the caller's
 * included), instead of one peer store per destination.  Needs the flag barriers (n_ranks >= 2). */
int dls_fedavg_reduce_f32_mc(const float* const* rows, const float* weights, int n_rows, const float* old,
                             float old_scale, double inv_total, float* mc_dst, long long lo, long long hi,
                             uint32_t* const* peer_flags, int n_ranks, int rank, uint32_t epoch, uint32_t* nan_flag,
                             void* stream);
/* Global-model broadcast on its own: n floats from src to every GPU behind mc_dst. */
int dls_multimem_bcast_f32(const float* src, float* mc_dst, long long n, void* stream);

/* Same reduction with per-row decode at load time: kind 0 fp32, 1 stochastic-255 packed
 * (u8 magnitudes + sign plane + per-tensor scale), 2 NNADQ packed (per-tensor bit width);
 * optional per-element denominators (FedDropoutAvg: weights count only non-zero
 * uploads) and per-(row,tensor) participation bitmap (FedOBD block dropout: absent
 * blocks contribute the old value).  Work is decomposed into (tensor, 8-element group)
 * items so packed groups are byte aligned.  See csrc/reduce.cu for the descriptor. */
struct DlsPackedRow {
    const void* payload;          /* fp32 row | u8 magnitudes | packed codes           */
    const uint8_t* aux;           /* sq: sign plane over the flat vector               */
    const float* scale;           /* [T]                                               */
    const float* offset;          /* [T] (nnadq)                                       */
    const int32_t* bits;          /* [T] (nnadq)                                       */
    const long long* byte_offsets;/* [T+1] (nnadq)                                     */
    const uint8_t* present;       /* [T] 1 = this row carries tensor t (NULL = all)    */
    float weight;
    int32_t kind;
    int32_t is_delta;             /* 1: row holds deltas (added to old), 0: absolute   */
    int32_t pad;
};
int dls_fedavg_reduce_packed(const struct DlsPackedRow* rows, int n_rows, const float* old,
                             const long long* tensor_offsets, const long long* group_offsets, int n_tensors,
                             int nonzero_denominator, float* out, uint32_t* nan_flag, void* stream);

/* ---- peer-mapped buffers for one-process-per-GPU runs (reference transport:
 *      simulation_lib/algorithm_factory.py:26-28, pickled CPU tensors over pipes) -------- */
int dls_ipc_alloc(long long bytes, void** out_ptr);               /* cudaMalloc + zero        */
int dls_ipc_free(void* ptr);
int dls_ipc_get_handle(void* ptr, unsigned char* handle_out64);   /* cudaIpcGetMemHandle      */
int dls_ipc_open_handle(const unsigned char* handle64, void** out_ptr);
int dls_ipc_close_handle(void* ptr);
int dls_enable_peer_access(int device, int peer);
int dls_copy_device(void* dst, const void* src, long long bytes, void* stream); /* src may be peer-mapped */

/* ---- local SGD (reference: external torch.optim.SGD inside Trainer.train(), and
 *      simulation_lib/worker/gradient_worker.py:13-26,96-116) --------------------------
 * Multi-tensor fused SGD over the flat parameter arena: gradients are read through a
 * device pointer table (one entry per tensor), weight decay, momentum/dampening/
 * nesterov, lr = lr_table[min(state[0], lr_len-1)].  state[1] == 0 means "momentum
 * buffer not initialised" (first step copies the gradient like torch.optim.SGD).
 * Mixed precision: grad_is_bf16[t] != 0 (device [n_tensors], may be NULL) marks tensors whose
 * gradient is bf16; for those the updated fp32 value is also written, rounded, to the bf16
 * shadow arena `shadow_bf16` (same offsets; may be NULL) so no cast kernel runs per step. */
int dls_sgd_step(float* params, float* momentum_buf, const float* const* grads, const long long* tensor_offsets,
                 int n_tensors, const float* lr_table, int lr_len, const int32_t* state, float weight_decay,
                 float momentum, float dampening, int nesterov, const uint8_t* grad_is_bf16, void* shadow_bf16,
                 void* stream);
/* state[0] += 1 (schedule step); state[1] = 1 (momentum ready); state[2] += 1 (batch index in epoch) */
int dls_sgd_tick(int32_t* state, void* stream);
/* x_out[b,:] = x_all[order[state[state_index]*batch_stride + b], :] for b < batch, y likewise;
 * positions past n_order repeat the last entry.  nhwc_channels > 1: each C,H,W source row is
 * written H,W,C (channels-last activations for the tensor-core convolutions) */
int dls_gather_batch(const float* x_all, const long long* y_all, const long long* order, long long n_order,
                     const int32_t* state, int state_index, int batch, int batch_stride, long long row_elems,
                     int nhwc_channels, float* x_out, long long* y_out, void* stream);
/* acc[0] += loss*batch ; acc[1] += #(argmax(logits)==y) ; acc[2] += batch   (fp64 acc) */
int dls_metric_accumulate(const float* logits, const long long* targets, const float* loss, int batch, int classes,
                          double* acc, void* stream);

/* ---- fused training-mode BatchNorm (+residual) (+ReLU), channels-last activations ---------
 *      (reference: BatchNorm/ReLU inside the external Trainer's model forward/backward;
 *      batch statistics are always used, simulation_lib/util/model.py:23) ---------------------
 * x/y/residual/dy/dx: [rows, channels] with channels contiguous (NHWC), dtype 0 = fp32,
 * 1 = bf16; channels % 4 == 0.  Few rows / many channels: one cluster launch each way, the CTAs
 * that share a channel group combine their statistics through distributed shared memory.  Many
 * rows / few channels: a row-major reduce launch (cluster DSMEM fold, one partial per cluster in
 * `scratch`) plus an elementwise launch.  scratch: dls_bn_scratch_floats() floats (0 = not needed,
 * NULL allowed).  fwd writes mean/invstd (fp32 [channels]) for bwd; bwd writes dgamma/dbeta. */
int dls_bn_scratch_floats(long long rows, int channels, int dtype);
int dls_bn_fwd(const void* x, int dtype, long long rows, int channels, const float* gamma, const float* beta,
               const void* residual, float eps, int relu, void* y, float* mean, float* invstd, float* scratch,
               void* stream);
/* Forward when the producing convolution accumulated stats = [2][channels] (sum, sum of squares of x) in its
 * epilogue (dls_conv_fprop): no reduction launch, one elementwise launch. */
int dls_bn_fwd_stats(const void* x, int dtype, long long rows, int channels, const float* gamma, const float* beta,
                     const void* residual, float eps, int relu, void* y, float* mean, float* invstd,
                     const float* stats, void* stream);
int dls_bn_bwd(const void* dy, const void* y, const void* x, int dtype, long long rows, int channels,
               const float* gamma, const float* mean, const float* invstd, int relu, void* dx, void* dresidual,
               float* dgamma, float* dbeta, float* scratch, void* stream);
/* dls_bn_bwd for callers that own a ZEROED fp32 [2][channels] buffer (one slice of an arena cleared once per step):
 * the row-major path adds its partial sums there with atomics instead of folding them across a cluster. */
int dls_bn_bwd_acc(const void* dy, const void* y, const void* x, int dtype, long long rows, int channels,
                   const float* gamma, const float* mean, const float* invstd, int relu, void* dx, void* dresidual,
                   float* dgamma, float* dbeta, float* zeroed_partials, void* stream);

/* ---- bf16 GEMM on tcgen05 tensor cores, TMEM accumulators, TMA operand loads ------------
 *      (reference: torch.nn.Linear inside the external Inferencer, server/server.py:40-55, and
 *      one eager test inference per Shapley coalition, shapley_value_algorithm.py:67-76) -------
 * c[b] = act(a[b] * B^T + bias): a [batch][m][k] bf16, b [batch | 1][n][k] bf16 (k contiguous:
 * the torch Linear weight layout), bias [batch | 1][n] fp32 or NULL, c [batch][m][n] fp32
 * (out_dtype 0) or bf16 (1).  k % 8 == 0, operands 16-byte aligned; m, n, k tails are
 * zero-filled by TMA.  b_shared = 1: every batch entry uses the same b and bias. */
int dls_gemm_bf16_tn(const void* a, const void* b, const float* bias, void* c, int out_dtype, int m, int n, int k,
                     int batch, int b_shared, int relu, void* stream);

/* ---- convolutions (reference: every nn.Conv2d of every local step and server evaluation runs
 *      in cuDNN inside the external Trainer / Inferencer, simulation_lib/worker/worker.py:86-88,
 *      server/server.py:40-55) -----------------------------------------------------------------
 * Tensor-core path (csrc/conv_tc.cu): implicit GEMM on tcgen05, one TMA box load per filter tap.
 * x [n][h][w][x_pitch] bf16 (c_in channels used), filters [c_out][r][s][c_in] bf16 (the
 * channels-last arena segment), y [n][oh][ow][y_pitch] bf16; *_pitch = 0 means dense.  c_in, c_out
 * and pitches % 8 == 0, stride 1 or 2 (stride 2: even h, w and c_in % 64 == 0), dilation 1, groups 1.
 * dls_conv_supported() tells whether a shape is eligible (else use the dls_conv_direct_* kernels). */
typedef struct DlsConvShape {
    int n, h, w, c_in, c_out, r, s, stride, pad_h, pad_w, x_pitch, y_pitch;
    int dw_pitch, dw_cols; /* wgrad only: elements per (c_out, tap) row of dw and how many of the c_in
                              columns are written (0 = dense / all): lets a channel-padded input feed an
                              unpadded gradient */
} DlsConvShape;
int dls_conv_supported(const DlsConvShape* shape);
/* y = conv(x, w) + bias; stats (fp32 [2][c_out], caller-zeroed, may be NULL) receives the per-channel
 * sum and sum of squares of the stored outputs: the statistics of the BatchNorm that follows. */
int dls_conv_fprop(const void* x, const void* w, void* y, const float* bias, float* stats,
                   const DlsConvShape* shape, void* stream);
/* dx = conv_transpose(dy, w): every element of dx is written (stride-2 parity classes without a tap
 * get zeros). */
int dls_conv_dgrad(const void* dy, const void* w, void* dx, const DlsConvShape* shape, void* stream);
/* dw [c_out][r][s][c_in].  mode 1: dw is fp32 and holds zeros (or a running sum); the kernel may split the
 * pixel range over CTAs and adds its partial tiles with vector reductions.  mode 0 / 2: every element is stored
 * exactly once (no split), as fp32 / as bf16. */
int dls_conv_wgrad(const void* dy, const void* x, void* dw, const DlsConvShape* shape, int mode, void* stream);
/* Direct (CUDA-core) convolutions for layers the tensor-core path does not take: few channels
 * (the 3-channel stem, LeNet-5's 1/6/16-channel 5x5 layers), any stride / padding.  Every tensor is
 * addressed through element strides {n, c, h, w} ({o, c, r, s} for filters), dtype 0 = fp32,
 * 1 = bf16 per operand; accumulation is fp32.  wgrad adds into fp32 dw / dbias (caller-zeroed). */
typedef struct DlsDirectConv {
    int n, c_in, h, w, c_out, r, s, oh, ow, stride_h, stride_w, pad_h, pad_w;
    long long xs[4], ws[4], ys[4];
    int x_dtype, w_dtype, y_dtype;
} DlsDirectConv;
int dls_conv_direct_fprop(const void* x, const void* w, const void* bias, int bias_dtype, void* y,
                          const DlsDirectConv* d, void* stream);
int dls_conv_direct_dgrad(const void* dy, const void* w, void* dx, const DlsDirectConv* d, void* stream);
int dls_conv_direct_wgrad(const void* dy, const void* x, float* dw, float* dbias, const DlsDirectConv* d,
                          void* stream);
/* Stem layers (c_in < 8): patches [n * oh * ow][k_pad] bf16 with k = (r * S + s) * c_in + c, zero padded to
 * k_pad (a multiple of 8), plus the filter re-packed the same way as [c_out][k_pad]; the layer then runs as a
 * 1x1 convolution over k_pad channels on the tensor-core kernels above.  x / w are addressed through element
 * strides like the direct kernels (DlsDirectConv: ys unused). */
int dls_conv_stem_im2col(const void* x, const void* w, void* patches, void* w_packed, int k_pad,
                         const DlsDirectConv* d, void* stream);

/* ---- quantizers (reference: topology/quantized_endpoint.py:76,82,89,102-104 and the
 *      external cyy_torch_algorithm.quantization; oracle in algorithm/quantization) ---- */
int dls_sq_quant(const float* flat, long long n, const long long* tensor_offsets, int n_tensors, int level,
                 uint64_t seed, uint8_t* payload, uint8_t* sign_plane, float* scale, uint32_t* scratch_absmax,
                 void* stream);
/* Same quantizer, same outputs, but for callers that cannot allocate device memory: a
 * C-ABI host whose HBM is already fully handed out (a late cudaMalloc may fail), or a
 * worker sharing its GPU with many other streams (cudaMalloc/cudaFree stall all of them).
 * The per-tensor scratch is a 16 KiB array built into the library (one copy per GPU, held
 * until exit) instead of a caller buffer, so n_tensors is limited to
 * dls_sq_quant_noalloc_max_tensors() (4096); larger models are refused with an error.
 * All calls share that array: it is cleared at the start of each call, and calls whose GPU
 * work overlaps race -- keep them on one stream.  The first call on each device resolves
 * the array's address (cudaGetSymbolAddress); issue it before any stream capture begins.
 * Nothing in this entry point allocates or frees.  No counterpart in the reference. */
int dls_sq_quant_noalloc(const float* flat, long long n, const long long* tensor_offsets, int n_tensors, int level,
                         uint64_t seed, uint8_t* payload, uint8_t* sign_plane, float* scale, void* stream);
int dls_sq_quant_noalloc_max_tensors(void);
int dls_sq_dequant(const uint8_t* payload, const uint8_t* sign_plane, const float* scale,
                   const long long* tensor_offsets, int n_tensors, long long n, float* out, void* stream);
/* stats scratch: [T][4] doubles (min, max, sum, sumsq) -- zeroed/initialised by the call */
int dls_nnadq_quant(const float* flat, const long long* tensor_offsets, const long long* group_offsets, int n_tensors,
                    float weight, double* scratch_stats, float* scale, float* offset, int32_t* bits,
                    long long* byte_offsets, uint8_t* payload, long long payload_capacity, void* stream);
int dls_nnadq_dequant(const uint8_t* payload, const float* scale, const float* offset, const int32_t* bits,
                      const long long* byte_offsets, const long long* tensor_offsets, const long long* group_offsets,
                      int n_tensors, float* out, void* stream);

/* ---- sign_SGD (reference scaffold: worker/gradient_worker.py:47-48,85-94) ----------- */
int dls_sign_pack(const float* grad, long long n, uint32_t* packed, void* stream);
/* vote[i] = sign(sum_k sign_k[i]) in {-1,0,+1}; rows are packed sign words (1 = negative) */
int dls_sign_majority(const uint32_t* const* rows, int n_rows, long long n, float* vote, void* stream);
/* signSGD step in ONE launch: majority vote over n_rows packed rows (local or peer-mapped) + momentum /
 * nesterov / p -= lr * vote on this worker's flat parameters (+ bf16 shadow: all elements, or only the tensors
 * whose shadow_mask entry is non-zero).  No counterpart kernel in the reference (its scaffold applies the
 * update per tensor in Python, worker/gradient_worker.py:96-116). */
int dls_sign_vote_sgd(const uint32_t* const* rows, int n_rows, long long n, float* params, float* mom, float lr,
                      float momentum, float dampening, int nesterov, int mom_ready, void* shadow,
                      const uint8_t* shadow_mask, const long long* tensor_offsets, int n_tensors, void* stream);

/* ---- FedDropoutAvg / FedOBD helpers (reference: method/fed_dropout_avg/worker.py:23-27,
 *      method/fed_obd/obd_algorithm.py:129-145) ---------------------------------------- */
int dls_dropout_mask(const float* in, long long n, float keep_prob, uint64_t seed, float* out,
                     unsigned long long* nonzero_count, void* stream);
/* scores[b] = sum over block b of (cur-prev)^2 ; block_of_tensor maps tensor -> block */
int dls_block_sqdiff(const float* cur, const float* prev, const long long* tensor_offsets,
                     const int32_t* block_of_tensor, int n_tensors, int n_blocks, double* scores, void* stream);

/* ---- coalition GEMM for the Shapley engines (reference: method/shapley_value/
 *      shapley_value_algorithm.py:67-76 builds each coalition model on the CPU) -----------------
 * C[s] = act(A[s] * W_s^T + b_s) with W_s = sum_i coefs[s][i] * W_i built inside the kernel as the
 * B-operand producer of the tcgen05 GEMM (the coalition model is never written to memory).
 * worker_ptrs: DEVICE array of n_workers (<= 64) fp32 arena base pointers; the layer's weight
 * [n_layer][k_w] (row-major) starts at element w_offset of every arena, its bias at bias_offset (-1:
 * none).  coefs: device fp32 [n_sets][n_workers].  a: bf16 [n_sets or 1][m][k_a], k_a % 8 == 0,
 * k_a >= k_w (extra columns are ignored).  stacked == 0: c is [n_sets][m][n_layer] (batch = coalition;
 * a_shared selects one A for all).  stacked != 0: ONE GEMM, c is [m][n_sets * n_layer] (column
 * block s holds coalition s; A is shared by construction).  ptrs_aligned: every arena base is
 * 16-byte aligned (enables 128-bit weight loads where w_offset / k_w allow). */
int dls_subset_gemm_bf16(const void* a, int a_shared, const float* const* worker_ptrs, int n_workers,
                         long long w_offset, long long bias_offset, const float* coefs, int n_sets, int n_layer,
                         int stacked, void* c, int out_dtype, int m, int k_a, int k_w, int relu, int ptrs_aligned,
                         void* stream);

/* Grouped im2col for the coalition evaluator: x [B][H][W][groups*channels] (channels-last), valid k x k
 * windows -> cols [groups][B*OH*OW][k_pad], column order (kh, kw, c) when channels_last_k else (c, kh, kw),
 * zero beyond channels*k*k.  elem_bytes 2 (bf16/fp16) or 4. */
int dls_im2col_grouped(const void* x, int elem_bytes, int batch, int height, int width, int groups, int channels,
                       int k, int k_pad, int channels_last_k, void* cols, void* stream);
/* Same with `pad` pixels of zero padding on every border (rows = B * (H + 2 pad - k + 1) * (W + 2 pad - k + 1)). */
int dls_im2col_grouped_padded(const void* x, int elem_bytes, int batch, int height, int width, int groups,
                              int channels, int k, int k_pad, int channels_last_k, int pad, void* cols, void* stream);

/* ---- graph family: fed_gnn / fed_gcn / fed_aas (reference: worker/graph_worker.py:205-216,
 *      252-269, 271-321) ------------------------------------------------------ */
/* Ordered (stable) compaction of index pairs.  Pair i is kept iff mask[m(a[i])] && (b == NULL ||
 * mask[m(b[i])]) && Bernoulli(keep_prob; Philox(seed, i)), with m(v) = n_id ? n_id[v] : v.
 * out_a / out_b / out_pos (each may be NULL except out_a; out_b iff b) receive the kept values and
 * their original positions, *count (device) the number kept.  scratch: device ints,
 * dls_index_compact_scratch_ints(n) of them.  Two launches. */
long long dls_index_compact_scratch_ints(long long n);
int dls_index_compact(const long long* a, const long long* b, long long n, const long long* n_id,
                      const uint8_t* mask, float keep_prob, unsigned long long seed, long long* out_a,
                      long long* out_b, long long* out_pos, long long* count, int* scratch, void* stream);
/* out[k, :] = x[rows[k], :]; rows are row_bytes wide (a multiple of 4).  `out` may be a peer-mapped
 * buffer: the boundary embeddings are then written straight into the consumer GPU's memory. */
int dls_row_gather(const void* x, const long long* rows, long long n_rows, long long row_bytes, void* out,
                   void* stream);
/* Halo merge.  Forward (foreign_in == NULL): row r with g = n_id[r] is copied from x when
 * train_mask[g], else from emb[perm[j]] where sorted_ids[j] == g, else zero-filled; foreign_out[r]
 * (may be NULL) records which rows were foreign.  Backward (foreign_in != NULL): out[r] = foreign_in[r]
 * ? 0 : x[r]. */
int dls_halo_merge(const void* x, const long long* n_id, long long n_rows, long long row_bytes,
                   const uint8_t* train_mask, const long long* sorted_ids, const long long* perm, long long n_ids,
                   const void* emb, const uint8_t* foreign_in, uint8_t* foreign_out, void* out, void* stream);
/* GCN normalisation and aggregation: dinv[i] = (1 + in-degree(i))^-1/2 (deg_scratch: n_nodes ints),
 * out = D^-1/2 (A + I) D^-1/2 h (+ bias) with A given as COO (from[e] -> to[e]); fp32 vector
 * reductions, so the sum order over a node's neighbours is not fixed. */
int dls_gcn_dinv(const long long* dst, long long n_edges, long long n_nodes, int* deg_scratch, float* dinv,
                 void* stream);
int dls_gcn_aggregate(const float* h, const long long* from, const long long* to, long long n_edges,
                      long long n_nodes, int channels, const float* dinv, const float* bias, float* out,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif
