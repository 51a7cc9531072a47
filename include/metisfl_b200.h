/*
 * metisfl_b200 — C ABI of the sm_100a kernel library (libmetisfl_b200_kernels.so).
 *
 * Every entry point takes raw device pointers and a CUDA stream (as void*; pass
 * torch.cuda.current_stream().cuda_stream) and only ENQUEUES work: nothing here
 * synchronises the stream, so the calls can be overlapped and captured by the caller.
 * Return value: 0 on success, a cudaError_t (>0) or a negative argument-error code
 * otherwise; mfl_last_error() returns the matching message for the calling thread.
 *
 * Tensor element types use the wire-format numbering of metisfl.DType.Type
 * (reference metisfl/proto/model.proto:16-27, plus FLOAT16/BFLOAT16 extensions).
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum mfl_dtype {
  MFL_INT8 = 0, MFL_INT16 = 1, MFL_INT32 = 2, MFL_INT64 = 3,
  MFL_UINT8 = 4, MFL_UINT16 = 5, MFL_UINT32 = 6, MFL_UINT64 = 7,
  MFL_FLOAT32 = 8, MFL_FLOAT64 = 9, MFL_FLOAT16 = 10, MFL_BFLOAT16 = 11
};

#define MFL_MAX_SRC 64      /* sources folded per launch; callers chunk beyond this */
#define MFL_MAX_PEERS 16    /* GPUs in one NVSwitch domain handled by the fused kernels */
#define MFL_MAX_SEGMENTS 8  /* dtype-homogeneous segments of one model arena */

const char* mfl_last_error(void);
int mfl_sm_count(int device);

/* ---- K1: weighted community aggregation (replaces federated_average.cc:15-145) ------------
 * dst[i] = cast_dst( (accumulate ? acc : 0) + sum_l w[l] * cast_f32(src[l][i]) ) * post_scale
 * exact != 0 selects the reference's per-term truncation semantics:
 *   dst_T[i] = sum_l T(double(src_l[i]) * w[l])   accumulated in T  (federated_average.cc:25-35)
 * acc_f32 (optional) is an fp32 running sum used for streaming aggregation (FedStride-style):
 *   acc += sum_l w_l x_l ; dst (optional) = cast(acc * post_scale).
 */
int mfl_weighted_sum(void* dst, int dst_dtype, const void* const* srcs, const double* weights,
                     int n_src, int src_dtype, size_t numel, float* acc_f32, int accumulate,
                     double post_scale, int exact, void* stream);

/* ---- K3: rolling update (federated_rolling_average_base.cc:175-293, FedRec/FedStride) -----
 * S += w_new * x_new - w_old * x_old (x_old may be NULL); C = cast(S / z).  S is fp32 (or fp64
 * for FLOAT64 models).  FedAsync mixing: C = (1-a) C + a x is mfl_weighted_sum with 2 sources.
 */
int mfl_rolling_update(void* community, int dtype, float* scaled_sum, const void* x_new, double w_new,
                       const void* x_old, double w_old, double z, size_t numel, void* stream);

/* One dtype-homogeneous segment of a model arena (shared by the fused kernels below). */
typedef struct {
  uint64_t offset_bytes;   /* start of the segment inside every slot / output buffer */
  uint64_t numel;
  int32_t dtype;           /* element type in the slots and in the outputs */
  int32_t exact;           /* reference truncation semantics for this segment */
} mfl_segment;

/* ---- K3 fused over NVLink: one asynchronous community update in one launch on the executor GPU ----------
 * state = decay * state + w_new * x_new - w_old * x_old   (fp32, local to the caller)
 * C     = cast(state * post_scale), stored into EVERY rank's community buffer: through mc_out (a multicast
 *         address, one multimem.st per 16 bytes) when non-NULL, else with peer stores to out_bufs[0..world).
 * FedRec (federated_recency.cc:27-94): decay 1, post_scale 1/z, x_old = the learner's previous contribution.
 * FedAsync (extension): decay 1 - a_t, w_new = a_t, post_scale 1, x_old NULL.
 * x_new / x_old may be peer-mapped slots of another GPU.  Integer ("exact") segments copy x_new.
 * Ordering: if wait_flag is non-NULL the kernel first waits until *wait_flag >= wait_value (the slot's owner
 * released it with mfl_flag_signal); when every CTA has fenced its stores the last one publishes new_version into
 * version_flags[0..world) (release, system scope).  done_counter is a zero-initialised device word owned by the
 * caller (it resets itself).  Pass version_flags = done_counter = NULL for a purely local update.
 * state_elem_offsets[s] is the element offset of segment s inside `state` (multiples of 4). */
int mfl_async_fold(const void* x_new, double w_new, const void* x_old, double w_old, double decay,
                   double post_scale, float* state, const uint64_t* state_elem_offsets,
                   const mfl_segment* segments, int n_segments, void* const* out_bufs, void* mc_out, int world,
                   unsigned long long* const* version_flags, unsigned long long new_version,
                   const unsigned long long* wait_flag, unsigned long long wait_value,
                   unsigned int* done_counter, int n_blocks, void* stream);
/* Monotonic 64-bit counters in (peer-mapped) device memory: release-store on the stream / bounded wait on the
 * stream.  Never enqueue a wait and the signal it waits for on the SAME GPU in different streams - order those
 * with CUDA events instead (two kernels of one GPU are not guaranteed to run concurrently). */
int mfl_flag_signal(unsigned long long* flag, unsigned long long value, void* stream);
int mfl_flag_wait(const unsigned long long* flag, unsigned long long value, void* stream);
/* Copy the newest community model (version v = *version_flag, buffer v % n_comm) into a learner's working arena
 * and, optionally, segment 0 as fp32 master weights.  status[0] receives v; status[1] counts CTAs that observed
 * more than n_comm - 2 newer versions by the time they finished (a possibly torn copy - retry). */
int mfl_snapshot_install(const void* const* community_bufs, int n_comm, const unsigned long long* version_flag,
                         unsigned long long* status, void* dst, size_t nbytes, float* master_f32,
                         const mfl_segment* segment0, int n_blocks, void* stream);
/* In-kernel waits that ran out of their 20 s budget on the current device since the library was loaded
 * (synchronises with the device; for tests and error paths). */
int mfl_spin_timeouts(void);

/* ---- K1 fused over NVLink: reduce-scatter + all-gather (+ fp32 install) in ONE launch -----
 * Every rank launches the same call.  Rank r owns shard r of each segment, reads it from every
 * source slot (peer-mapped pointers), applies weights, and stores the result into EVERY rank's
 * output buffer; in-kernel flag barriers over the signal pads order the phases.
 */
int mfl_fused_aggregate(const void* const* src_slots, const double* weights, int n_src,
                        void* const* out_bufs, uint32_t* const* signal_pads, int rank, int world,
                        const mfl_segment* segments, int n_segments,
                        float* install_f32, /* optional local fp32 copy of segment 0 */
                        int n_blocks, void* stream);

/* NVLS variant: the same launch contract and arguments as mfl_fused_aggregate, plus MULTICAST addresses of the
 * symmetric slot / community buffers (cuMulticast* objects bound on every GPU; torch.distributed._symmetric_memory
 * provides them).  Floating-point segments are reduced INSIDE the NVSwitch (multimem.ld_reduce on mc_src_slots[j],
 * j = local slot index, n_mc = learners per GPU <= MFL_MAX_MC_SRC) and multicast-stored to every GPU
 * (multimem.st on mc_out): per GPU 1/world of the model crosses NVLink each way instead of (world-1)/world.
 * The switch only adds, so every learner of the federation must participate and the slots must already hold
 * w_l * theta_l (mfl_optimizer_step's publish_scale); `weights` is used only by integer ("exact") segments and
 * the < 8-element tails, which keep the peer-pointer path.  No reference counterpart (controller.cc:428-518
 * sums on the CPU). */
#define MFL_MAX_MC_SRC 8
int mfl_fused_aggregate_nvls(const void* const* src_slots, const double* weights, int n_src, void* const* out_bufs,
                             uint32_t* const* signal_pads, int rank, int world, const mfl_segment* segments,
                             int n_segments, float* install_f32, int n_blocks, void* stream,
                             const void* const* mc_src_slots, int n_mc, void* mc_out);
size_t mfl_signal_pad_bytes(int n_blocks);

/* ---- K2: fused optimizer step + publish (keras_model_ops.py:245-283, fed_prox.py:50-60) ---
 * One launch over the flat parameter arena.  master (fp32) is updated in place; model (working
 * copy in model_dtype) and, if non-NULL, publish (peer-visible slot) receive the new parameters
 * scaled by publish_scale (1 for plain FedAvg; n_l for pre-scaled in-switch reduction).
 */
typedef struct {
  int32_t kind;            /* 0 vanilla SGD (+L1/L2), 1 momentum SGD, 2 FedProx, 3 Adam, 4 AdamW */
  float lr, l1, l2, momentum, mu, beta1, beta2, eps, weight_decay;
  int32_t step;            /* 1-based Adam step for bias correction */
  float grad_scale;        /* multiply grads (loss-scale / accumulation) */
  float publish_scale;
} mfl_optim_cfg;

int mfl_optimizer_step(float* master, void* model, int model_dtype, const void* grad, int grad_dtype,
                       float* state1, float* state2, const float* anchor /* FedProx theta* */,
                       void* publish, int publish_dtype, size_t numel, const mfl_optim_cfg* cfg,
                       int zero_grad, void* stream);

/* ---- K4: tensor quantifier (proto_tensor_serde.h:35-50) ------------------------------------
 * counts[2*v] = non-zeros, counts[2*v+1] = zeros of variable v = elements [offsets[v], offsets[v+1]).
 * offsets is a DEVICE array of n_vars+1 uint64 element offsets; partials is caller scratch of
 * mfl_quantify_scratch_bytes(n_vars) bytes (per-block partial counts, reduced deterministically).
 */
size_t mfl_quantify_scratch_bytes(int n_vars);
int mfl_quantify(const void* data, int dtype, const uint64_t* offsets, int n_vars,
                 unsigned long long* counts, void* scratch, void* stream);

/* Quantifier variant that allocates nothing (this entry point has no counterpart in the reference).
 *
 * For hosts that cannot obtain device scratch: plain-C controller plug-ins that only hold the buffers
 * their application gave them, processes whose GPU memory is already owned by a framework's caching
 * allocator (a raw cudaMalloc there can fail or fragment the pool), and code on the round's critical
 * path that must avoid the device-wide synchronisation inside cudaMalloc/cudaFree.  Same inputs and
 * outputs as mfl_quantify() minus the scratch argument; the per-CTA partial records go to a
 * file-scope `static __device__` array of MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES bytes that every GPU
 * loading this library holds until the process exits.
 *
 *  - Limit: a call whose scratch need (mfl_quantify_scratch_bytes) exceeds that array is refused
 *    with an error; use mfl_quantify() for it.
 *  - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *    that device current) and caches it; make that first call outside of any CUDA stream capture.
 *    No call allocates or frees device memory, and none calls into an allocating library.
 *  - All calls on a device share the one array, which keeps whatever the previous call left in it.
 *    Nothing needs clearing (every record is rewritten before it is read), but two calls whose
 *    kernels overlap on the GPU - different streams, or different host threads on different
 *    streams - race on it: issue them on one stream or order them with events. */
#define MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES 16384
int mfl_quantify_noalloc(const void* data, int dtype, const uint64_t* offsets, int n_vars,
                         unsigned long long* counts, void* stream);

/* ---- K12: device pack/unpack between wire bytes and typed arenas ---------------------------
 * Batched cast/copy: for each tensor t, dst+dst_off[t] <- cast(src+src_off[t]) over numel[t]
 * elements (tables are DEVICE arrays).  byte_swap != 0 reverses element bytes (BIG_ENDIAN wire).
 */
int mfl_pack_cast(void* dst, int dst_dtype, const void* src, int src_dtype,
                  const uint64_t* dst_off_bytes, const uint64_t* src_off_bytes,
                  const uint64_t* numel, int n_tensors, int byte_swap, void* stream);

/* ---- Peer-memory fabric (cudaIpc): symmetric arena shared by the GPU processes of one box --- */
int mfl_fabric_alloc(void** ptr, size_t bytes, void* ipc_handle_out /* 64 bytes */);
int mfl_fabric_open(void** ptr, const void* ipc_handle /* 64 bytes */);
int mfl_fabric_close(void* ptr);
int mfl_fabric_free(void* ptr);
int mfl_fabric_enable_peers(int device, int world);
/* cross-GPU barrier as its own launch (K13); same signal pads as the fused kernels */
int mfl_fabric_barrier(uint32_t* const* signal_pads, int rank, int world, int channel, void* stream);

/* ---- K10: bf16 GEMM on tcgen05 tensor cores, TMEM accumulator, TMA-fed (csrc/kernels/gemm_tcgen05.cu)
 * C[m,n] = act(A[m,k] . B[n,k]^T + bias[n]); A and B are K-major (row-major with k contiguous) bf16,
 * lda/ldb/ldc in elements (lda, ldb multiples of 8); C is bf16 or fp32; act: 0 none, 1 ReLU, 2 GELU(tanh).
 * Replaces the framework matmul under the learners' Linear layers (keras_model_ops.py:156). */
int mfl_gemm_bf16(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, const void* b, int64_t ldb,
                  const void* bias, int bias_dtype, int m, int n, int k, int act, void* stream);
/* Same kernel with per-operand majors: a_mn = 0 -> A is [m, k] row-major (reduction index contiguous); a_mn = 1 -> A is
 * passed as its transpose, [k, m] row-major (lda = its row pitch).  Likewise b_mn for B ([n, k] or [k, n]).  The backward
 * GEMMs of a Linear layer (dX = dY . W, dW = dY^T . X) therefore need no transposed copies: the TMA loads 64 x 64 boxes
 * and the instruction descriptor transposes in the tensor core. */
int mfl_gemm_bf16_ex(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb,
                     int b_mn, const void* bias, int bias_dtype, int m, int n, int k, int act, void* stream);
/* Large problems (m >= 256, enough 256 x BN tiles to occupy the SM pairs) run on a CTA-PAIR kernel
 * (csrc/kernels/gemm_2sm.cuh): tcgen05.mma.cta_group::2, M = 256 per instruction, each CTA stages its half of B.
 * mode: -1 automatic (default; env MFL_GEMM_2SM), 0 never, 1 wherever legal.  Returns the previous mode. */
int mfl_gemm_set_pair_mode(int mode);
/* The same product with an optional split of the reduction over more SMs: ws = ZERO-initialised fp32 scratch of at
 * least m * n floats (handed back zeroed).  Used when the output has few tiles and k is long (weight gradients of
 * square Linear layers); bf16 output only.  Partial tiles are added into ws with TMA reduce-add stores. */
int mfl_gemm_bf16_ws(void* c, int c_dtype, int64_t ldc, const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb,
                     int b_mn, int m, int n, int k, float* ws, size_t ws_bytes, void* stream);

/* ---- K11: implicit-GEMM convolution forward on the same tcgen05 pipeline; the A operand is gathered by
 * im2col-mode TMA (cuTensorMapEncodeIm2col), never materialised.  NHWC bf16: x[n,h,w,c], w[k,r,s,c],
 * y[n,oh,ow,k]; c % 64 == 0, k % 8 == 0.  Stride-1 data gradients reuse it with flipped, transposed
 * filters (metisfl_b200/ops/conv.py).  Replaces the framework convolutions under keras_model_ops.py:156. */
int mfl_conv2d_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k, int r, int s,
                          int stride, int pad, void* stream);

/* ---- K15: weight-stationary 3x3 / stride-1 / pad-1 convolution on tcgen05 (csrc/kernels/conv3x3_halo.cu): all
 * filter taps of a CTA's 64 output channels stay in shared memory, one zero-padded halo tile per output tile is
 * loaded by tiled TMA and read nine times through shifted shared-memory descriptors.  NHWC bf16: x[n,h,w,c],
 * w[k,3,3,c], y[n,h,w,k]; c in {64, 128}, k % 64 == 0, even w <= 126 (mfl_conv3x3_supported).  Data gradients
 * use it with the filters transposed and flipped.  stats: NULL, or [2][k] fp32 to which the per-channel sum and
 * sum of squares of the outputs are added.  Replaces the framework convolutions (keras_model_ops.py:156). */
int mfl_conv3x3_supported(int n, int h, int wd, int c, int k);
int mfl_conv3x3_bf16(void* y, const void* x, const void* w, float* stats, int n, int h, int wd, int c, int k,
                     void* stream);
/* K15 as a data gradient: dx[n,h,w,c] from dy[n,h,w,k] and the FORWARD filters w[k,3,3,c] (the kernel reads them
 * transposed and flipped in place; no repacked copy). */
int mfl_conv3x3_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k, void* stream);

/* ---- K11 gradients (csrc/kernels/gemm_tcgen05.cu).  Same NHWC bf16 tensors as the forward call.
 * _ws variants take an fp32 scratch (any size; used for split-K when the output has few tiles: partial tiles are
 * added into it with TMA reduce-add and mfl_splitk_finalize casts to bf16).  Without scratch no split-K is used.
 *  - dgrad: dx[n,h,w,c] = sum over taps of dy . w; stride 1 and 2; ONE launch walks the stride^2 parity classes of
 *    input pixels, each with its own tap subset, and stores through a 5-D strided TMA map.
 *  - wgrad: dw[k,r,s,c] (+)= dy^T . im2col(x), split over output pixels across the whole grid; ws must hold
 *    mfl_conv2d_wgrad_workspace_bytes(c,k,r,s) bytes and be ZERO on entry (the call leaves it zero again). */
int mfl_conv2d_fprop_bf16_ws(void* y, const void* x, const void* w, float* ws, size_t ws_bytes, int n, int h, int wd,
                             int c, int k, int r, int s, int stride, int pad, void* stream);
int mfl_conv2d_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k, int r, int s,
                          int stride, int pad, void* stream);
int mfl_conv2d_dgrad_bf16_ws(void* dx, const void* dy, const void* w, float* ws, size_t ws_bytes, int n, int h, int wd,
                             int c, int k, int r, int s, int stride, int pad, void* stream);
size_t mfl_conv2d_wgrad_workspace_bytes(int c, int k, int r, int s);
int mfl_conv2d_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c, int k, int r,
                          int s, int stride, int pad, int accumulate, void* stream);
int mfl_splitk_finalize(float* ws, void* out_bf16, int rows, int cols, int64_t ld, int accumulate, void* stream);

/* ---- K15w: 3x3 / stride-1 / pad-1 weight gradient from halo tiles (csrc/kernels/conv3x3_wgrad.cu): one x halo tile
 * and one dy tile per chunk of output pixels; the three taps of a filter row are ONE N=192 tcgen05 instruction whose
 * B descriptor walks overlapping 64-column blocks.  Same workspace contract as mfl_conv2d_wgrad_bf16. */
int mfl_conv3x3_wgrad_supported(int n, int h, int wd, int c, int k);
int mfl_conv3x3_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c, int k,
                           int accumulate, void* stream);

/* ---- K15t: 3x3 / stride-1 / pad-1 convolution with PIXELS as the tensor-core N dimension
 * (csrc/kernels/conv3x3_t.cu): filters stream as the M operand, the halo tile is the N operand, the epilogue
 * transposes back to NHWC.  Measured best of the hand-written kernels for 128-channel layers on 16x16 images (the
 * Python dispatch routes >= 128 channels, width >= 16 here).  fprop and dgrad (forward filters, read transposed +
 * flipped in place). */
int mfl_conv3x3_t_supported(int n, int h, int wd, int c_reduce, int c_out);
int mfl_conv3x3_t_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k, void* stream);
int mfl_conv3x3_t_dgrad_bf16(void* dx, const void* dy, const void* w, int n, int h, int wd, int c, int k, void* stream);

/* ---- K11s: the network stem (csrc/kernels/conv_stem.cu): 3x3 / stride 1 / pad 1 with c <= 4 input channels, where
 * an implicit GEMM would pad the reduction 16x.  Direct convolution, filters in registers / shared memory. */
int mfl_conv_stem_supported(int c, int k);
int mfl_conv_stem_fprop_bf16(void* y, const void* x, const void* w, int n, int h, int wd, int c, int k, void* stream);
int mfl_conv_stem_wgrad_bf16(void* dw, const void* dy, const void* x, float* ws, int n, int h, int wd, int c, int k,
                             int accumulate, void* stream);
/* The stem on the tensor cores: patches[n*h*w, 32] (bf16; column (r*3+s)*c + ch, zero padded; c <= 3).  Forward is
 * then mfl_gemm_bf16(y, patches, w32) with w32 = the [k, 9c] filters padded to 32 columns, the weight gradient
 * mfl_gemm_bf16_ws(dw32, dy as MN-major A, patches as MN-major B) - see metisfl_b200/ops/conv.py: _StemGemmFn. */
int mfl_conv_stem_im2col_bf16(void* patches, const void* x, int n, int h, int wd, int c, void* stream);

/* ---- K16: fused softmax cross-entropy for a language-model head (csrc/kernels/cross_entropy.cu): per-row loss and
 * d(loss)/d(logits) in one pass pair over bf16 logits, no fp32 copy.  logits [rows, vocab] bf16 with row pitch ld,
 * targets [rows] int64 (rows with target == ignore_index contribute nothing), dlogits [rows, vocab] bf16 with pitch
 * ldd receives (softmax - onehot) * scale, row_loss [rows] fp32 receives logsumexp - logit[target].
 * Replaces the framework loss under keras_model_ops.py:156. */
int mfl_lm_cross_entropy(void* dlogits, int64_t ldd, float* row_loss, const void* logits, int64_t ld,
                         const long long* targets, int rows, int vocab, float scale, long long ignore_index,
                         void* stream);

/* ---- K17: fp32 column sums of a bf16 [rows, cols] matrix (bias gradient of a Linear layer), csrc/kernels/colsum.cu.
 * partial is [mfl_colsum_slices(rows), cols] fp32 and is overwritten; the caller adds the slices. */
int mfl_colsum_slices(int rows);
int mfl_colsum_bf16(float* partial, const void* x, int64_t ld, int rows, int cols, void* stream);

/* ---- K18: fused residual-add + LayerNorm (csrc/kernels/layernorm.cu), bf16 activations [rows, d], d in
 * {256, 512, 768, 1024}; gamma / beta are bf16 (param_bf16 != 0) or fp32, may be NULL.
 * forward : r = a + b (b may be NULL: then r_out may be NULL and r = a), h = LN(r); mean / rstd [rows] fp32 are saved.
 * backward: dr_out = (dr_in ? dr_in : 0) + LN'(dh); partial [mfl_ln_partial_rows(rows)][2][d] fp32 receives per-CTA
 *           dgamma / dbeta rows (the caller sums over the first axis).
 * Replaces the framework normalisation layers under keras_model_ops.py:156. */
int mfl_ln_supported(int d);
int mfl_ln_partial_rows(int rows);
int mfl_add_ln_fwd(void* r_out, void* h_out, float* mean, float* rstd, const void* a, const void* b, const void* gamma,
                   const void* beta, int param_bf16, int rows, int d, float eps, void* stream);
int mfl_add_ln_bwd(void* dr_out, float* partial, const void* dh, const void* dr_in, const void* r, const void* gamma,
                   int param_bf16, const float* mean, const float* rstd, int rows, int d, void* stream);

/* ---- K14: fused BatchNorm2d (+ residual add) (+ ReLU), NHWC bf16 activations viewed as [m, c] rows
 * (m = N*H*W, c a power of two in [8, 2048]); fp32 affine / statistics (csrc/kernels/bn_relu.cu).
 * forward : training != 0: batch statistics (saved to save_mean / save_invstd, running statistics and
 *           num_batches_tracked updated in place), else running statistics;
 *           y = relu?(bn(x) + residual?) - residual may be NULL.
 * backward: g = relu ? dy * (y > 0) : dy; dx = BN backward of g; dres (may be NULL) = g;
 *           dgamma / dbeta (may be NULL) are overwritten, or accumulated into when accumulate != 0.
 * workspace: mfl_bn_workspace_bytes(m, c) bytes of device memory, contents irrelevant on entry.
 * barrier  : NULL -> three launches per direction (partials, finalize, apply).  Otherwise MFL_BN_BARRIER_WORDS device
 *            uint32 words (two-level arrival counters + generation) that were zero when first used (the kernels keep them consistent): the three phases then run as ONE
 *            cooperative launch separated by grid barriers - a dependent launch costs more than a whole phase of the
 *            small layers.  One barrier must not be shared by calls that can run concurrently.
 * Replaces the framework normalisation / activation kernels under keras_model_ops.py:156. */
#define MFL_BN_BARRIER_WORDS (64 + 32 * 32)
size_t mfl_bn_workspace_bytes(int m, int c);
int mfl_bn_fwd(void* y, const void* x, const void* residual, const float* gamma, const float* beta,
               float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
               float* save_invstd, float* workspace, size_t workspace_bytes, unsigned int* barrier, int m, int c,
               float eps, float momentum, int training, int relu, void* stream);
int mfl_bn_bwd(void* dx, void* dres, float* dgamma, float* dbeta, int accumulate, const void* dy, const void* y,
               const void* x, const float* gamma, const float* save_mean, const float* save_invstd,
               float* workspace, size_t workspace_bytes, unsigned int* barrier, int m, int c, int relu, void* stream);

/* ---- K5-K9: RNS-CKKS on the GPU (csrc/ckks/ckks_kernels.cu) ----------------------------------
 * Behavioural spec: encryption/palisade/ckks_scheme.cc:110-252 (Encrypt / ComputeWeightedAverage /
 * Decrypt) with the ciphertext layout of csrc/ckks/ckks_common.h: [ct][poly 0..1][limb][N] uint64 in
 * evaluation (NTT) form.  All table pointers are DEVICE pointers prepared once per crypto context
 * (metisfl_b200/ops/ckks.py); polynomials of one limb (N = 16384 x 8 B = 128 KB) are transformed
 * entirely in shared memory, one CTA per row. */
typedef struct {
  uint64_t q[4], n_inv[4], n_inv_shoup[4];
  const uint64_t* psi[4];         /* bit-reversed powers of the 2N-th root, per limb */
  const uint64_t* psi_shoup[4];
  const uint64_t* ipsi[4];
  const uint64_t* ipsi_shoup[4];
  int32_t log_n, n_limbs;
} mfl_ckks_ctx;

/* K5: in-place negacyclic NTT / inverse NTT of n_rows rows; row r uses limb (r % n_limbs). */
int mfl_ckks_ntt(const mfl_ckks_ctx* ctx, uint64_t* rows, size_t n_rows, int inverse, void* stream);
/* K6: canonical-embedding encode: values -> rounded scaled coefficients, coeffs[ct][2*slots]
 * (re parts then im parts; coefficient i*gap and i*gap + N/2 of the plaintext polynomial). */
int mfl_ckks_encode(const mfl_ckks_ctx* ctx, const double* values, size_t n_values, uint32_t slots, double scale,
                    const double* ksi_pows /* (2N+1) x {cos,sin} */, const uint32_t* rot_group, double* coeffs,
                    void* stream);
/* K7: RLWE public-key encryption of encoded coefficients; pk = [2][limb][N] with Shoup companions.  key256: eight
 * HOST words = a fresh 256-bit key from the OS entropy source for every call; the kernel expands it with the ChaCha20
 * block function (one stream each for the ternary mask u and the two error polynomials). */
int mfl_ckks_encrypt(const mfl_ckks_ctx* ctx, const double* coeffs, uint32_t n_ct, uint32_t slots,
                     const uint64_t* pk, const uint64_t* pk_shoup, const uint32_t* key256, uint64_t* ct_out,
                     void* stream);
/* K7: c0 + c1*s per limb, inverse NTT -> residues[ct][limb][N] (coefficient form). */
int mfl_ckks_decrypt(const mfl_ckks_ctx* ctx, const uint64_t* ct, uint32_t n_ct, const uint64_t* sk,
                     const uint64_t* sk_shoup, uint64_t* residues, void* stream);
/* K6: CRT (Garner) combine + canonical-embedding decode -> values. garner_inv[j*4+i] = q_j^-1 mod q_i. */
int mfl_ckks_decode(const mfl_ckks_ctx* ctx, const uint64_t* residues, uint32_t n_ct, uint32_t slots,
                    double inv_scale, const uint64_t* garner_inv /* host, 16 */, const double* ksi_pows,
                    const uint32_t* rot_group, double* values, size_t n_values, void* stream);
/* K8: out = sum_l ct_l (.) w_l mod q (limb-wise scalar multiply-add of whole ciphertext blobs).
 * w[l*4+limb], w_shoup likewise (host arrays).  words = n_ct * 2 * n_limbs * N. */
int mfl_ckks_weighted_sum(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                          const uint64_t* w_shoup, int n_src, size_t words, uint64_t* out, void* stream);
/* K8 fused over NVLink: rank r reduces shard r of the ciphertext words from every source (peer loads)
 * and stores it into every rank's output (peer stores); same barrier protocol as mfl_fused_aggregate. */
int mfl_ckks_fused_weighted_sum(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                                const uint64_t* w_shoup, int n_src, size_t words, uint64_t* const* outs,
                                uint32_t* const* signal_pads, int rank, int world, int n_blocks, void* stream);
/* The same with the result stored through mc_out, the MULTICAST address of the symmetric output buffers (one
 * multimem.st per 16 bytes lands in every GPU instead of `world` peer stores); mc_out == NULL is the call above. */
int mfl_ckks_fused_weighted_sum_mc(const mfl_ckks_ctx* ctx, const uint64_t* const* cts, const uint64_t* w,
                                   const uint64_t* w_shoup, int n_src, size_t words, uint64_t* const* outs,
                                   uint64_t* mc_out, uint32_t* const* signal_pads, int rank, int world, int n_blocks,
                                   void* stream);
/* K9: rescale - drop the last limb: out_l = (c_l - [c_last]_{q_l}) * q_last^-1 mod q_l.
 * qlast_inv / qlast_inv_shoup: host arrays per remaining limb. */
int mfl_ckks_rescale(const mfl_ckks_ctx* ctx, const uint64_t* ct_in, uint32_t n_ct, const uint64_t* qlast_inv,
                     const uint64_t* qlast_inv_shoup, uint64_t* ct_out, void* stream);

#ifdef __cplusplus
}
#endif
