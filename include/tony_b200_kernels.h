/* tony_b200 kernel library — public C ABI (libtony_b200_kernels.so, sm_100a only).
 *
 * Every function returns 0 on success or a TONY_ERR_* code; tony_last_error() gives the
 * text of the most recent failure on the calling thread's process.  All device pointers
 * are raw CUDA device addresses, `stream` is a cudaStream_t passed as void*, and every
 * launch is asynchronous on that stream.  A C ABI (rather than a torch extension) is
 * deliberate: TonY launches TensorFlow, PyTorch, MXNet and Horovod jobs alike
 * (reference: tony-core/src/main/java/com/linkedin/tony/runtime/*.java), so the data
 * plane must be loadable from any of them (ctypes / cffi / dlopen).
 *
 * The reference ships no device code; each entry point below cites
 * the stock op sequence it replaces in the jobs the reference launches.
 */
#ifndef TONY_B200_KERNELS_H_
#define TONY_B200_KERNELS_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes -------------------------------------------------------------------- */
#define TONY_OK 0
#define TONY_ERR_INVALID 1
#define TONY_ERR_CUDA 2
#define TONY_ERR_UNSUPPORTED 3
#define TONY_ERR_TOO_LARGE 4
#define TONY_ERR_TIMEOUT 5

#define TONY_MAX_RANKS 8   /* one NVSwitch domain */

const char* tony_last_error(void);
int tony_version(void);
/* Number of kernels this library has launched in this process (for bench accounting). */
unsigned long long tony_launch_count(void);
/* Credit n launches that did not go through this library's host entry points: the replay of a
 * CUDA graph that captured them (the counter is bench accounting only). */
void tony_launch_count_add(unsigned long long n);

/* ---- GEMM epilogue flags ------------------------------------------------------------ */
#define TONY_EPI_BIAS 0x001u        /* + bias[n]                                   */
#define TONY_EPI_GELU_ERF 0x002u    /* exact GELU (BERT)                           */
#define TONY_EPI_GELU_TANH 0x004u   /* tanh GELU (GPT-2 "gelu_new")                */
#define TONY_EPI_RESIDUAL 0x008u    /* + aux[m,n]                                  */
#define TONY_EPI_DGELU_ERF 0x010u   /* * gelu_erf'(aux[m,n])   (backward)          */
#define TONY_EPI_DGELU_TANH 0x020u  /* * gelu_tanh'(aux[m,n])  (backward)          */
#define TONY_EPI_ACCUMULATE 0x040u  /* C += result (gradient accumulation)         */
#define TONY_EPI_SAVE_PREACT 0x080u /* aux_out[m,n] = value before the activation  */
#define TONY_EPI_RELU 0x100u        /* max(x, 0)                                   */
#define TONY_EPI_DRELU 0x200u       /* * (aux[m,n] > 0)        (backward)          */

/* C[m,n] = epi(alpha * sum_k A(m,k) B(n,k)), bf16 in/out, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A stored [M,K] row-major (ld = lda);  1: stored [K,M] row-major.
 *   b_mn_major = 0: B stored [N,K] row-major (ld = ldb);  1: stored [K,N] row-major.
 * Linear layer: forward (0,0), dgrad (0,1), wgrad (1,1) — no transposes are materialised.
 * `workspace` (optional, fp32 scratch) enables split-K for shapes that cannot fill the GPU.
 * Replaces cuBLASLt GEMM + separate bias/GELU/residual elementwise kernels
 * (the nn.Linear / F.gelu calls in the jobs TonY launches, e.g.
 *  tony-examples/mnist-pytorch/mnist_distributed.py:129-136). */
int tony_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                   int b_mn_major, long long lda, long long ldb, long long ldc, unsigned int flags,
                   float alpha, const void* bias, const void* aux, long long ld_aux, void* aux_out,
                   void* workspace, size_t workspace_bytes, void* stream);

/* tony_gemm_bf16 that ALSO produces rowsum_out[m] = sum_k A(m,k) (bf16 [M], optionally added to
 * its previous contents): in a weight-gradient GEMM dW = dY^T X the A operand is dY^T, so this is
 * the bias gradient, computed by one extra N=16 MMA per k-step against a shared-memory tile of ones
 * instead of a separate pass over dY.  Only the CTA-pair kernel carries it (M >= 256 and N >= 256);
 * other shapes return TONY_ERR_UNSUPPORTED and the caller runs tony_colsum_bf16. */
int tony_gemm_bf16_rowsum(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                          int b_mn_major, long long lda, long long ldb, long long ldc, unsigned int flags,
                          float alpha, const void* bias, const void* aux, long long ld_aux, void* aux_out,
                          void* workspace, size_t workspace_bytes, void* rowsum_out, int rowsum_accumulate,
                          void* stream);

/* Extended form: tile/split overrides (0 = auto; force_splits = 1 disables split-K),
 * B-readiness flags and reserved copy CTAs used by the all-gather-fused GEMM,
 * and an SM budget (max_ctas, 0 = all SMs) so the GEMM can share the GPU with a
 * concurrently running communication kernel. */
int tony_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K, int a_mn_major,
                      int b_mn_major, long long lda, long long ldb, long long ldc,
                      unsigned int flags, float alpha, const void* bias, const void* aux,
                      long long ld_aux, void* aux_out, void* workspace, size_t workspace_bytes,
                      int force_block_n, int force_splits, const unsigned int* b_ready,
                      unsigned int b_ready_epoch, int copy_ctas, int max_ctas, void* stream);

/* ---- LayerNorm (replaces ATen layer_norm fwd/bwd) ----------------------------------- */
/* y = (x - mean) * rstd * gamma + beta over the last dim H (H % 256 == 0, H <= 4096 fwd,
 * H <= 2048 bwd); mean/rstd (fp32, [rows]) are saved for backward. */
int tony_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int rows, int H, float eps, void* stream);
/* dx, dgamma, dbeta.  `scratch` holds per-CTA fp32 partials for dgamma/dbeta:
 * tony_layernorm_bwd_scratch_bytes(H) bytes.  accumulate != 0 adds into dgamma/dbeta.
 * Scratch contract shared by tony_layernorm_bwd*, tony_colsum_bf16, tony_act_bwd_colsum_bf16, tony_bn_fwd and
 * tony_bn_bwd: the first 4096 bytes are arrival counters that must be ZERO before the first call with a buffer
 * (cudaMemset once after allocating it); every call leaves them zero, so nothing is cleared per call.  Calls
 * sharing a buffer must be ordered on one stream.  TONY_B200_CLEAR_COUNTERS=1 clears them on every call. */
size_t tony_layernorm_bwd_scratch_bytes(int H);
int tony_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                       const float* rstd, void* dx, void* dgamma, void* dbeta, int rows, int H,
                       float* scratch, size_t scratch_bytes, int accumulate, void* stream);
/* Same, with dx += add: the gradient of a residual branch that bypassed the LayerNorm (pre-LN
 * transformer blocks), so autograd needs no separate elementwise add to merge the two. */
int tony_layernorm_bwd_add(const void* dy, const void* x, const void* gamma, const float* mean,
                           const float* rstd, const void* add, void* dx, void* dgamma, void* dbeta, int rows,
                           int H, float* scratch, size_t scratch_bytes, int accumulate, void* stream);

/* ---- softmax cross-entropy, forward + backward in one pass --------------------------- */
/* logits [rows, V] bf16 (row stride ld) are overwritten with dlogits =
 * (softmax - onehot) * grad_scale * (*grad_scale_dev if non-null); loss[row] (fp32) =
 * -log softmax[target]; rows whose target == ignore_index get loss 0 and zero gradient.
 * Replaces log_softmax + nll_loss + their backward
 * (F.cross_entropy, tony-examples/mnist-pytorch/mnist_distributed.py:81). */
int tony_softmax_xent_fwd_bwd(void* logits, const long long* targets, float* loss, int rows, int V,
                              long long ld, long long ignore_index, float grad_scale,
                              const float* grad_scale_dev, void* stream);

/* ---- column sum (bias gradient) ------------------------------------------------------ */
/* out[c] (bf16) (+)= sum_r x[r,c];  scratch: tony_colsum_scratch_bytes(cols) bytes fp32. */
size_t tony_colsum_scratch_bytes(int cols);
int tony_colsum_bf16(const void* x, int rows, int cols, long long ld, void* out, float* scratch,
                     size_t scratch_bytes, int accumulate, void* stream);

/* Backward of a fused Linear + activation whose incoming gradient comes from outside the library, in one
 * pass: dpre[r,c] = dy[r,c] * act'(pre[r,c]) and dbias[c] (+)= sum_r dpre[r,c] (dbias may be NULL).
 * act: 1 ReLU (pre may also be the ReLU output), 2 exact GELU, 3 tanh GELU.  Same scratch as tony_colsum_bf16.
 * Replaces the eager fp32 passes of torch's gelu_backward / threshold_backward plus sum(dim=0) in the towers of
 * the ps/worker jobs (tony-examples/mnist-tensorflow/mnist_distributed.py:127-129). */
int tony_act_bwd_colsum_bf16(const void* dy, const void* pre, void* dpre, int act, int rows, int cols, long long ld,
                             void* dbias, float* scratch, size_t scratch_bytes, int accumulate, void* stream);

/* packed[r, p*H + c] = {a, b, c}[p][r, c] (three [rows, H] matrices with their own row strides -> [rows, 3H]) and
 * dbias[3H] (+)= column sums of packed, in one pass; H a multiple of 256; scratch as tony_colsum_bf16 for 3H columns.
 * Serves the fused QKV projection's backward when the attention core is a library that returns dq / dk / dv apart. */
int tony_pack3_rows_colsum(const void* a, const void* b, const void* c, long long lda, long long ldb, long long ldc,
                           void* packed, int rows, int H, void* dbias, float* scratch, size_t scratch_bytes,
                           int accumulate, void* stream);

/* Same result as tony_colsum_bf16, but the partial-sum scratch is a fixed 2 MiB device
 * array built into the library, so the call performs no allocation and needs no scratch
 * pointer.  For host programs that cannot allocate device memory themselves: a framework
 * that has already reserved the whole GPU for its own pool (a cudaMalloc from a dlopen'ed
 * helper then fails), or a hot path where the implicit device synchronisation of
 * cudaMalloc is not acceptable.  cols <= TONY_COLSUM_STATIC_MAX_COLS, otherwise
 * TONY_ERR_TOO_LARGE.  All calls on a GPU share that one array: issue them in order on a
 * single stream; calls that overlap on the GPU race.  The first call on each device also
 * resolves the array's device address — make that call before starting a stream capture.
 * This entry point has no counterpart in the reference (which has no device code). */
#define TONY_COLSUM_STATIC_MAX_COLS 8192
int tony_colsum_bf16_static(const void* x, int rows, int cols, long long ld, void* out,
                            int accumulate, void* stream);

/* ---- implicit-GEMM 2-D convolution on tcgen05 (TMA im2col mode), NHWC activations, KRSC filters ---- */
/* x: [N, H, W, C] bf16 (a channels-last activation), w: [K, R, S, C] bf16, y: [N, P, Q, K] with
 * P = (H + 2*pad - R) / stride + 1.  C % 8 == 0 and K % 8 == 0, R*S <= 16, stride 1 or 2 (same in both
 * dimensions), symmetric padding.  No im2col matrix is materialised: the TMA unit gathers each filter
 * tap's pixels (zero-filling the padding) into the tensor cores' shared-memory layout.
 * Replaces cuDNN's convolution forward / backward-data / backward-filter in the CNN jobs the
 * reference launches (tony-examples/horovod-on-tony/tensorflow2_mnist.py:48-49,
 * tony-examples/mnist-tensorflow/mnist_distributed.py:87-97). */
typedef struct {
    int N, H, W, C;   /* input activation */
    int K, R, S;      /* filter: K output channels, R x S taps */
    int stride, pad;
    /* padding after the last row / column; -1 = same as `pad`.  Even-sized filters need pad_hi = pad - 1
     * to keep the output size. */
    int pad_hi;
    /* 0 = pad both dimensions; 1 = pad rows only (pad_w = 0).  With x_stride_w below this lets a layer
     * fold its horizontal taps into the channel dimension (see models/resnet.py: the 7x7 stem). */
    int pad_h_only;
    /* element strides of x between neighbouring pixels / rows / images; 0 = dense NHWC.  x_stride_w may
     * be SMALLER than C (overlapping pixel windows): the "channels" of pixel w are then the C elements
     * starting at w * x_stride_w.  Multiples of 8 elements. */
    long long x_stride_w, x_stride_h, x_stride_n;
} tony_conv_desc;
int tony_conv2d_out_size(int in, int filter, int stride, int pad);
/* stat_sum / stat_sqsum (both or neither; fp32 [K], caller-zeroed): the epilogue adds the per-channel
 * sum and sum of squares of the bf16 output (BatchNorm batch statistics) with one atomic per column
 * per warp, so the BatchNorm that follows needs no statistics pass over y. */
int tony_conv2d_fprop_bf16(const tony_conv_desc* d, const void* x, const void* w, void* y, float* stat_sum,
                           float* stat_sqsum, void* stream);
/* dx [N, H, W, C] from dy [N, P, Q, K]; flags: 0 or TONY_EPI_ACCUMULATE (dx += ...).  A stride-2 layer
 * runs one launch per output parity class, each a stride-1 convolution over dy. */
int tony_conv2d_dgrad_bf16(const tony_conv_desc* d, const void* dy, const void* w, void* dx, unsigned int flags,
                           void* stream);
/* dw [K, R, S, C] from dy and x (needs C % 64 == 0); split-K over the output pixels with fp32
 * partials in `workspace` (tony_conv2d_wgrad_workspace_bytes); flags: 0 or TONY_EPI_ACCUMULATE. */
size_t tony_conv2d_wgrad_workspace_bytes(const tony_conv_desc* d);
int tony_conv2d_wgrad_bf16(const tony_conv_desc* d, const void* dy, const void* x, void* dw, unsigned int flags,
                           void* workspace, size_t workspace_bytes, void* stream);

/* 2x2 space-to-depth of an image batch for a 7x7 / stride-2 / 3-channel stem: x [N, 3, H, W] bf16 (NCHW,
 * H and W even) -> xs [N, H/2, W/2 + 3, 16] with xs[n, i, 2 + j, (dy*2 + dx)*4 + c] = x[n, c, 2i+dy, 2j+dx],
 * zero elsewhere.  The stem then is a 4 x 4 / stride-1 convolution that tony_conv2d_fprop_bf16 runs with
 * R = 4, S = 1, C = 64 and x_stride_w = 16 (the 4 horizontal taps are 64 contiguous elements). */
int tony_stem_space_to_depth(const void* x, void* xs, int N, int H, int W, void* stream);
/* Same transform with the 4-pixel horizontal window written out: xw [N, H/2, W/2, 64] (dense NHWC input
 * for tony_conv2d_fprop_bf16 with R = 4, S = 1, C = 64; 4x the bytes of xs). */
int tony_stem_space_to_depth_window(const void* x, void* xw, int N, int H, int W, void* stream);

/* ---- fused attention on a packed QKV activation (tcgen05 + TMA, head_dim 64) ------------ */
/* qkv: [batch*seq, 3, heads, 64] bf16 (the fused QKV projection's output, read in place);
 * out: [batch*seq, heads*64] bf16; lse: [batch*heads, seq] fp32 log2-sum-exp saved for backward.
 * seq must be a multiple of 128.  Replaces the library attention core (cuDNN / flash kernels
 * behind scaled_dot_product_attention) of the stock PyTorch jobs the reference launches; no
 * reference counterpart. */
int tony_attention_fwd_bf16(const void* qkv, void* out, float* lse, int batch, int seq, int heads,
                            int head_dim, int causal, float scale, void* stream);
/* dqkv: packed gradient [batch*seq, 3, heads, 64], every element written (no atomics, no
 * zero-fill); delta: [batch*heads, seq] fp32 scratch.  Three launches: delta, dq, dk+dv. */
int tony_attention_bwd_bf16(const void* qkv, const void* out, const void* dout, const float* lse,
                            float* delta, void* dqkv, int batch, int seq, int heads, int head_dim,
                            int causal, float scale, void* stream);

/* ---- packing ------------------------------------------------------------------------- */
/* out[r, 0..2, :] = (a[r,:], b[r,:], c[r,:]): interleave the dq/dk/dv of a library attention
 * backward into the packed gradient of a fused QKV projection in one launch (strides in
 * elements).  Replaces three zero-filled select_backward tensors + adds of stock autograd. */
int tony_pack3_rows(const void* a, const void* b, const void* c, long long lda, long long ldb,
                    long long ldc, void* out, int rows, int row_elems, void* stream);

/* ---- fused input embedding (word + token-type + position) ------------------------------ */
/* out[t,:] = word[ids[t],:] (+ type[tt[t],:]) (+ pos[t mod seq,:]); type / pos may be NULL.
 * ids / tt are int64.  Replaces three gathers + two adds of a stock embedding stack. */
int tony_embedding_sum_fwd(const void* word, const long long* ids, const void* type, const long long* tt,
                           const void* pos, void* out, long long tokens, int seq, int H, void* stream);
/* One pass over dy -> all three gradients.  dword (may be NULL): bf16x2 reductions into
 * [vocab_rows, H], zero-filled first unless word_accumulate; dtype (may be NULL, n_types <= 4):
 * masked column sums through `scratch`; dpos (may be NULL): [seq, H] sums over the batch. */
size_t tony_embedding_bwd_scratch_bytes(int seq, int n_types, int H);
int tony_embedding_sum_bwd(const void* dy, const long long* ids, const long long* tt, void* dword,
                           long long vocab_rows, int word_accumulate, void* dtype, int n_types,
                           int type_accumulate, void* dpos, int pos_accumulate, float* scratch,
                           size_t scratch_bytes, int batch, int seq, int H, void* stream);

/* ---- row-sharded embedding bag over peer memory (recommender, ps ownership) ------------- */
/* Row r of every field lives on rank r % world at local row r / world; `table[r]` / `grad[r]`
 * are rank r's shard ([fields, rows_per_shard, row_elems] bf16) mapped into this process
 * (tony_symm_alloc + tony_ipc_open, or the VMM heap).  A row is [deep_elems | rest]. */
typedef struct {
    void* table[TONY_MAX_RANKS];
    void* grad[TONY_MAX_RANKS];
    int rank, world;
} tony_embbag_comm;
/* deep: [batch, fields * deep_elems]; rest: [batch, fields, row_elems - deep_elems].  Rows come
 * straight from the owners' memory (lookup + dispatch in one kernel).  The caller orders this
 * after every owner's last table update (tony_symm_barrier). */
int tony_embbag_gather(const tony_embbag_comm* comm, const long long* ids, long long batch, int fields,
                       long long rows_per_shard, int row_elems, int deep_elems, void* deep, void* rest,
                       void* stream);
/* Adds every gradient row into the owner's gradient shard with system-scope bf16x2 reductions
 * (combine + sparse accumulation in one kernel).  Shards must be zero before the first scatter of
 * a step and the owner may read them only after a barrier that follows every rank's scatter. */
int tony_embbag_scatter(const tony_embbag_comm* comm, const long long* ids, long long batch, int fields,
                        long long rows_per_shard, int row_elems, int deep_elems, const void* ddeep,
                        const void* drest, void* stream);

/* ---- fused BatchNorm (+ residual) (+ ReLU), channels-last bf16, training mode ------------- */
/* x / y / residual: [M = N*H*W, C] row-major (a channels-last activation); C a power of two in
 * [8, 2048].  Forward: batch statistics (mean, rstd out, fp32; running_mean / running_var updated
 * when non-NULL), then y = relu?(gamma*(x-mean)*rstd + beta (+ residual)).  Two launches. */
size_t tony_bn_scratch_bytes(int C);
int tony_bn_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y, void* mask,
                float* mean, float* rstd, float* running_mean, float* running_var, long long M, int C,
                float eps, float momentum, int relu, float* scratch, size_t scratch_bytes, void* stream);
/* Same forward when the per-channel sum and sum of squares of x already exist (fp32 [C], produced by
 * tony_conv2d_fprop_bf16's epilogue): ONE launch, no statistics pass over x and no scratch. */
int tony_bn_fwd_from_sums(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                          void* mask, float* mean, float* rstd, float* running_mean, float* running_var, long long M,
                          int C, float eps, float momentum, int relu, const float* stat_sum, const float* stat_sqsum,
                          void* stream);
/* Backward of the same: dx, dgamma, dbeta (bf16) and, when dres != NULL, the residual branch's
 * gradient.  `mask` is the M*C/8-byte ReLU bit mask the forward wrote (one bit per element;
 * required when relu != 0, pass it to tony_bn_fwd as well).  Two launches. */
int tony_bn_bwd(const void* dy, const void* x, const void* mask, const void* gamma, const float* mean,
                const float* rstd, void* dx, void* dres, void* dgamma, void* dbeta, long long M, int C,
                int relu, float* scratch, size_t scratch_bytes, void* stream);

/* 3x3 / stride 2 / pad 1 max pooling, channels-last bf16 [N, H, W, C] (C % 8 == 0).  idx: one byte
 * per output element (position 0..8 of the first maximum), consumed by the gather-style backward. */
int tony_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream);
int tony_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, void* stream);

/* ---- optimizers on flat buffers ------------------------------------------------------ */
/* AdamW / Adam over n elements: fp32 master weights + moments, bf16 gradients in, bf16
 * weights out.  wd_mask (optional) has one byte per 256 elements: 0 = no weight decay.
 * bc1 = 1 - beta1^t, bc2 = 1 - beta2^t.  Replaces torch.optim.AdamW(fused=True)
 * plus the bf16 re-cast of the parameters. */
int tony_adamw_flat(float* master, float* exp_avg, float* exp_avg_sq, const void* grad_bf16,
                    void* param_bf16, size_t n, float lr, float beta1, float beta2, float eps,
                    float weight_decay, float bc1, float bc2, float grad_scale,
                    const unsigned char* wd_mask, size_t mask_offset_elems, void* stream);
/* SGD with momentum (+ optional Nesterov) on the same layout. */
int tony_sgd_flat(float* master, float* momentum_buf, const void* grad_bf16, void* param_bf16,
                  size_t n, float lr, float momentum, float weight_decay, int nesterov,
                  float grad_scale, const unsigned char* wd_mask, size_t mask_offset_elems,
                  void* stream);

/* ---- symmetric (peer-mapped) memory -------------------------------------------------- */
/* Device allocations that other ranks on the same NVSwitch domain can map.  The 64-byte
 * handle is exchanged out of band (the launcher's rendezvous or torch.distributed). */
#define TONY_IPC_HANDLE_BYTES 64
int tony_symm_alloc(size_t bytes, void** ptr);
int tony_symm_free(void* ptr);
int tony_ipc_get_handle(void* ptr, unsigned char* handle_out);
int tony_ipc_open(const unsigned char* handle, void** peer_ptr);
int tony_ipc_close(void* peer_ptr);

/* ---- fused gradient collective + optimizer (K1 / K2 in SURVEY.md section 8) ---------- */
#define TONY_OPT_ADAMW 0
#define TONY_OPT_SGD 1
typedef struct tony_fused_comm {
    int rank, world;
    const void* grad[TONY_MAX_RANKS]; /* bf16 gradient buffers of every rank (peer-mapped)   */
    void* param[TONY_MAX_RANKS];      /* bf16 parameter buffers of every rank (peer-mapped)  */
    unsigned int* flags[TONY_MAX_RANKS]; /* each rank's flag block: 2*TONY_MAX_RANKS words   */
    unsigned int* counter;            /* local: one word, zero-initialised                   */
    int* status;                      /* local: set to TONY_ERR_TIMEOUT if a peer never came */
} tony_fused_comm;

/* One launch per gradient bucket [begin, begin+count) (elements; count % (world*8) == 0):
 * wait for every rank's bucket -> each rank reduces its 1/world shard straight out of
 * the peers' gradient buffers over NVLink (fp32 accumulation) -> scales -> applies the
 * optimizer to its shard of the fp32 master state (ZeRO-1 sharding, state index
 * state_offset + i) -> writes the new bf16 weights into every rank's parameter buffer
 * (broadcast != 0; the DDP / Horovod path) or only its own (broadcast == 0; the
 * parameter-server path, where the all-gather is fused into the next forward GEMM).
 * Replaces ncclAllReduce + grad unscale + optimizer.step() (+ DDP bucket copies):
 * tony-examples/mnist-pytorch/mnist_distributed.py:113-126,
 * tony-examples/horovod-on-tony/tensorflow2_mnist.py:73-76. */
int tony_fused_reduce_optim(const tony_fused_comm* comm, size_t begin, size_t count,
                            float* master, float* state1, float* state2, size_t state_offset,
                            int optimizer, float lr, float beta1_or_momentum, float beta2,
                            float eps, float weight_decay, float bc1, float bc2, float grad_scale,
                            const unsigned char* wd_mask, int broadcast, unsigned int epoch,
                            int num_ctas, void* stream);

/* Dedicated parameter servers (the reference's literal ps/worker split, tony.ps.instances = P,
 * tony.worker.instances = W on P + W GPUs: tony-core/src/main/resources/tony-default.xml:160-170,204-223):
 * the bucket is cut into num_owners = P shards; the rank with owner_index in [0, P) reduces every rank's
 * gradients for its shard (workers contribute, servers keep zero gradients), applies the optimizer to its
 * shard of the fp32 state and stores the new bf16 weights into its OWN parameter buffer; ranks with
 * owner_index = -1 (workers) only take part in the two flag phases.  Workers pull the weights in their
 * next forward with tony_allgather_gemm_bf16 over the P shards. */
int tony_fused_reduce_optim_owners(const tony_fused_comm* comm, size_t begin, size_t count, float* master,
                                   float* state1, float* state2, size_t state_offset, int optimizer, float lr,
                                   float beta1_or_momentum, float beta2, float eps, float weight_decay,
                                   float bc1, float bc2, float grad_scale, const unsigned char* wd_mask,
                                   int num_owners, int owner_index, unsigned int epoch, int num_ctas,
                                   void* stream);

/* Plain collectives on symmetric buffers (validation + hvd.allreduce/broadcast shims).
 * out-of-place all-reduce(sum) * scale of `count` bf16 elements at element offset `begin`
 * of every rank's `grad[]` buffer, result written to every rank's `param[]` buffer. */
int tony_symm_allreduce_bf16(const tony_fused_comm* comm, size_t begin, size_t count, float scale,
                             unsigned int epoch, int num_ctas, void* stream);
/* Stand-alone collectives over peer memory (K0c): input = every rank's `grad[]` buffer, output = its
 * `param[]` buffer, over the same byte range [begin_bytes, begin_bytes + bytes) on every rank
 * (bytes % (world*16) == 0).  fp32 != 0 sums 32-bit floats, else bf16 (fp32 accumulation either way);
 * all-gather and broadcast move raw bytes (any dtype).  Each rank pulls over NVLink with 16-byte loads
 * and writes only its own output (all-reduce also stores its reduced shard to every peer).  Replaces
 * ncclReduceScatter / ncclAllGather / ncclBroadcast in Horovod-style jobs
 * (tony-examples/horovod-on-tony/tensorflow2_mnist.py:85-86 broadcast_variables,
 * tensorflow2_keras_mnist.py:75 metric averaging). */
#define TONY_COLL_ALLREDUCE 0
#define TONY_COLL_REDUCE_SCATTER 1
#define TONY_COLL_ALLGATHER 2
#define TONY_COLL_BROADCAST 3
int tony_symm_collective(const tony_fused_comm* comm, int op, int fp32, size_t begin_bytes, size_t bytes,
                         float scale, int root, unsigned int epoch, int num_ctas, void* stream);
/* All-rank barrier on the GPU (arrive + wait on flag words). */
int tony_symm_barrier(const tony_fused_comm* comm, unsigned int epoch, void* stream);

/* ---- NVLS (in-switch) form of the fused kernel ----------------------------------------- */
/* Symmetric heap on CUDA VMM + NVSwitch multicast (csrc/symm_vmm.cu).  Allocation protocol:
 * tony_vmm_granularity -> tony_vmm_alloc (returns an exportable fd) -> exchange fds ->
 * tony_vmm_import / tony_vmm_map for peers -> rank 0 tony_mc_create, others tony_vmm_import
 * of its fd -> every rank tony_mc_add_device -> barrier -> tony_mc_bind -> barrier ->
 * tony_vmm_map(mc handle) gives the multicast address. */
int tony_nvls_supported(void);
int tony_vmm_granularity(int world, size_t bytes, size_t* padded_out);
int tony_vmm_alloc(size_t padded_bytes, unsigned long long* handle_out, int* fd_out);
int tony_vmm_import(int fd, unsigned long long* handle_out);
int tony_vmm_map(unsigned long long handle, size_t padded_bytes, void** ptr_out);
int tony_vmm_unmap(void* ptr, size_t padded_bytes);
int tony_vmm_release(unsigned long long handle);
int tony_mc_create(size_t padded_bytes, int world, unsigned long long* handle_out, int* fd_out);
int tony_mc_add_device(unsigned long long mc_handle);
int tony_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, size_t padded_bytes);

typedef struct tony_nvls_comm {
    int rank, world;
    const void* grad_mc;        /* multicast address of the bf16 gradient buffers          */
    void* param_mc;             /* multicast address of the bf16 parameter buffers         */
    void* param_local;          /* this rank's own parameter buffer (unicast)              */
    unsigned int* flags_mc;     /* multicast address of the flag block (>= 64 words)       */
    unsigned int* flags_local;  /* this rank's copy of the flag block                      */
    unsigned int* counter;      /* local: one word, zero-initialised                       */
    int* status;                /* local: set to TONY_ERR_TIMEOUT if a peer never came     */
} tony_nvls_comm;

/* Same contract as tony_fused_reduce_optim, but the reduction happens INSIDE the NVSwitch:
 * each rank issues multimem.ld_reduce (fp32 accumulation in the switch) over its shard of the
 * multicast gradient address, updates its optimizer shard, and publishes the new bf16 weights
 * to every rank with one multimem.st; arrivals are counted with multimem.red on a multicast
 * flag word.  `call_index` = 1, 2, 3, ... counts NVLS collectives on this comm (all ranks must
 * agree).  optimizer = -1 gives a plain all-reduce (sum * grad_scale written to param). */
int tony_fused_reduce_optim_nvls(const tony_nvls_comm* comm, size_t begin, size_t count, float* master,
                                 float* state1, float* state2, size_t state_offset, int optimizer,
                                 float lr, float beta1_or_momentum, float beta2, float eps,
                                 float weight_decay, float bc1, float bc2, float grad_scale,
                                 const unsigned char* wd_mask, int broadcast, unsigned int call_index,
                                 int num_ctas, void* stream);

/* Broadcast through the switch: the root publishes `count` bf16-sized elements at element offset `begin`
 * of its own parameter buffer to every rank's parameter buffer with multimem.st (16 bytes per store leave
 * the root once; the NVSwitch replicates them).  All ranks call it with the same call_index. */
int tony_nvls_broadcast(const tony_nvls_comm* comm, size_t begin, size_t count, int root,
                        unsigned int call_index, int num_ctas, void* stream);

/* ---- all-gather fused into a GEMM (K3) ---------------------------------------------- */
/* Y = epi(X . W^T) where W [N,K] is row-sharded over `world` ranks (rows
 * [r*N/world, (r+1)*N/world) live on rank r at w_shard[r]).  Copy CTAs pull the remote
 * shards over NVLink into w_full (local, [N,K]) slab by slab and publish per-n-block
 * flags; GEMM CTAs in the same launch consume each slab as soon as it lands, local
 * shard first.  Replaces ncclAllGather + cuBLASLt GEMM (the ps "pull variables" step:
 * tony-examples/mnist-tensorflow/mnist_distributed.py:213-241). */
int tony_allgather_gemm_bf16(const void* X, const void* const* w_shard, void* w_full, void* Y,
                             int M, int N, int K, int rank, int world, unsigned int flags,
                             const void* bias, void* aux_out, unsigned int* ready_flags,
                             unsigned int epoch, int copy_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TONY_B200_KERNELS_H_ */
