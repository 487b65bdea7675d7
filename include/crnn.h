/*
 * crnn.h — C-ABI of the B200-native char-RNN training path (libcrnn.so).
 *
 * The reference (Abhishek8394/distributed_char_rnn) has no FFI seam: its hot path is the
 * single TensorFlow call `mon_sess.run([summaries, cost, final_state, train_op], feed)` at
 * train.py:199 over the graph built in model.py:8-98.  Each entry point below names the
 * slice of that graph it replaces.  Python binds these with ctypes
 * (distributed_char_rnn_b200/_lib.py); INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *   - plain C types only; every pointer named d_* is a DEVICE pointer owned by the caller
 *     (a torch tensor's data_ptr()).  The library allocates no device memory: the caller
 *     sizes one workspace with crnn_workspace_bytes() and passes it to every call.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and the
 *     calls return without synchronising.
 *   - return value 0 = ok; otherwise a negative code and crnn_last_error() describes it.
 *   - a handle is bound to one GPU and one thread at a time.
 *
 * Data layout (all fp32 unless stated)
 *   - parameters, gradients, Adam m/v: ONE flat buffer each, tensors in
 *     tf.trainable_variables() order (model.py:48-54,72: softmax_w [H,V], softmax_b [V],
 *     embedding [V,H], then per layer the cell kernel(s) [2H,G*H] / bias(es)), row-major,
 *     each tensor's offset rounded up to 64 floats; crnn_param_layout() reports offsets.
 *     The gradient buffer has CRNN_GRAD_TAIL extra floats after crnn_flat_size():
 *       [P+0] sum over (b,t) of |dX[b,t]|^2   (the un-deduplicated embedding slices that
 *             tf.global_norm sees, model.py:91)        [P+1] cost (model.py:85)
 *     so that ONE all-reduce of P+CRNN_GRAD_TAIL floats carries everything ranks exchange.
 *   - ids: int32 [B,S] with a row stride (ld) so a batch can be a column block of the
 *     loader's [B, num_batches*S] matrix (utils.py:79-82) without a copy.
 *   - RNN state: [L][n][B][H], n = 2 (c then h) for lstm, 1 otherwise (train.py:194-197).
 */
#ifndef CRNN_H_
#define CRNN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRNN_MODEL_RNN  0   /* BasicRNNCell  h' = tanh([x,h]W + b)            (model.py:16-17) */
#define CRNN_MODEL_GRU  1   /* GRUCell       r,u = sigmoid(..); reset-before-matmul (model.py:18-19) */
#define CRNN_MODEL_LSTM 2   /* LSTMCell      gate order i,j,f,o; forget_bias 1.0   (model.py:20-21) */

#define CRNN_PREC_FP32  0   /* fp32 FFMA contractions: the 1e-4 parity mode                    */
#define CRNN_PREC_BF16  1   /* bf16 tcgen05 contractions, fp32 accumulate / state / master weights */
#define CRNN_PREC_BF16X3 2  /* fp32-grade tensor-core mode: every operand is a (hi, lo) bf16 pair and each
                             * contraction issues hi*hi + lo*hi + hi*lo into the fp32 accumulator (error
                             * ~2^-17 per product instead of bf16's 2^-9); accurate activations.  Meets the
                             * same 1e-4 parity bar as CRNN_PREC_FP32 against the fp32 oracle.            */

#define CRNN_GRAD_TAIL  2

/* return codes */
#define CRNN_OK            0
#define CRNN_ERR_ARG      -1   /* bad argument                                                   */
#define CRNN_ERR_CUDA     -2   /* a CUDA call failed, or there is no sm_100 device               */
#define CRNN_ERR_UNSUP    -3   /* shape / precision combination this build has no kernel plan for */

/* phases of one step, for crnn_profile_* (bench.py's per-kernel roofline figures) */
#define CRNN_PHASE_EMBED_GATHER  0
#define CRNN_PHASE_XPROJ_GEMM    1
#define CRNN_PHASE_REC_FWD       2
#define CRNN_PHASE_LOGITS_GEMM   3
#define CRNN_PHASE_SOFTMAX_XENT  4
#define CRNN_PHASE_PROJ_GRADS    5
#define CRNN_PHASE_REC_BWD       6
#define CRNN_PHASE_WEIGHT_GRADS  7
#define CRNN_PHASE_INPUT_GRADS   8
#define CRNN_PHASE_EMBED_SCATTER 9
#define CRNN_PHASE_CLIP_ADAM     10
#define CRNN_PHASE_MISC          11
#define CRNN_NUM_PHASES          12

typedef struct crnn_config {
    int32_t model;        /* CRNN_MODEL_*                     --model      train.py:37 */
    int32_t rnn_size;     /* H                                --rnn_size   train.py:39 */
    int32_t num_layers;   /* L                                --num_layers train.py:41 */
    int32_t seq_length;   /* S                                --seq_length train.py:44 */
    int32_t batch_size;   /* B                                --batch_size train.py:46 */
    int32_t vocab_size;   /* V  (loader.vocab_size)                        train.py:85 */
    int32_t precision;    /* CRNN_PREC_*                                              */
    int32_t device;       /* CUDA device ordinal                                      */
    /* DropoutWrapper / tf.nn.dropout of the training graph (model.py:31-34,58-59).  0 means 1.0
     * (no dropout), so a config zero-filled past `device` behaves like the reference's defaults. */
    float   input_keep_prob;   /* --input_keep_prob   train.py:60                      */
    float   output_keep_prob;  /* --output_keep_prob  train.py:58                      */
    int32_t dropout_seed;      /* masks = f(seed, training-call index, stage, element)  */
    int32_t reserved;
} crnn_config;

typedef struct crnn_handle crnn_handle;

/* Replaces Model.__init__ (model.py:9-103) as far as shapes go: validates the config
 * (unknown model -> error like model.py:25), fixes the flat layout and the kernel tilings. */
int  crnn_create(const crnn_config* cfg, crnn_handle** out);
void crnn_destroy(crnn_handle* h);
const char* crnn_last_error(void);
const char* crnn_version(void);

/* Flat-buffer layout = tf.trainable_variables() (model.py:88). */
int64_t crnn_flat_size(const crnn_handle* h);                 /* P (padded), in floats  */
int32_t crnn_num_tensors(const crnn_handle* h);
int     crnn_param_layout(const crnn_handle* h, int32_t index, char* name, int32_t name_cap,
                          int64_t* offset, int64_t* rows, int64_t* cols);
int64_t crnn_state_size(const crnn_handle* h);                /* floats: L*n*B*H        */
int64_t crnn_workspace_bytes(const crnn_handle* h);

/* tf.gradients(cost, tvars) (model.py:54-91): embedding lookup, unrolled cells, logits,
 * sequence loss, BPTT.  Writes d_grads[0..P+CRNN_GRAD_TAIL) (zeroed inside) and
 * d_state_out (= final_state, train.py:199).  d_state_in may equal d_state_out.       */
int crnn_forward_backward(crnn_handle* h, const int32_t* d_x, const int32_t* d_y, int64_t ld_xy,
                          const float* d_state_in, float* d_state_out,
                          const float* d_params, float* d_grads, void* d_workspace, void* stream);

/* Overlap of the replicas' gradient exchange with BPTT's tail (replaces the per-variable gRPC pushes of the reference's
 * workers, train.py:126-133: there each gradient travels as soon as its op has run).  `event` is a caller-owned
 * cudaEvent_t (NULL: off).  Every later crnn_forward_backward records it, on the library's own stream, at the point
 * where the gradients of cells first_layer .. num_layers-1 (a contiguous tail of the flat buffer, see
 * crnn_param_layout) are final - for the top layers that is well before the call's work ends - so the caller can
 * start all-reducing that bucket on a side stream while the lower layers are still in BPTT.                         */
int crnn_set_bucket_event(crnn_handle* h, int32_t first_layer, void* event);

/* clip_by_global_norm + AdamOptimizer.apply_gradients (model.py:91-98).  `step` is the
 * 1-based Adam step t (beta powers beta^t); grads are read as grad_mult * d_grads
 * (grad_mult = 1/world after a sum all-reduce, 1 for a single process).
 * d_gnorm_out (device, 1 float, may be NULL) receives the global norm.               */
int crnn_clip_adam(crnn_handle* h, float* d_params, const float* d_grads, float* d_m, float* d_v,
                   float lr, float grad_clip, int64_t step, float grad_mult,
                   float* d_gnorm_out, void* d_workspace, void* stream);

/* crnn_clip_adam for callers that have no device buffer to spare and must not allocate one.
 * (No counterpart in the reference: TensorFlow owns all memory there.)  Who needs it: code that
 * runs after the process has handed every free byte of HBM to some other pool (a framework's
 * caching allocator, a memory-capped serving sandbox), and code on a path where a cudaMalloc is
 * not acceptable because it serialises the whole device or may simply fail - an optimiser hook
 * called from inside somebody else's training loop, a fault-recovery path that has to apply one
 * last update while memory is exhausted.  Same arguments and arithmetic as crnn_clip_adam minus
 * the workspace: the global-norm partial sums go to a fixed 64 KB array that is part of the
 * library image on every GPU that loads it (until the process exits).
 *   - limit: crnn_flat_size(h) <= 8192 * 16384 = 134,217,728 floats; larger models get an error
 *     that says so, and should use crnn_clip_adam.
 *   - the call makes no cudaMalloc / cudaFree and uses no allocating library routine.
 *   - the first call on each GPU looks up the array's device address (cudaGetSymbolAddress) and
 *     caches it; make that first call outside any stream capture.
 *   - all calls share the one array.  Nothing in it has to be zero on entry (each call rewrites
 *     every partial it later reads), but two calls whose kernels overlap on the GPU - different
 *     streams, different handles, different threads - race on it: serialise them.            */
int crnn_clip_adam_noalloc(crnn_handle* h, float* d_params, const float* d_grads, float* d_m, float* d_v,
                           float lr, float grad_clip, int64_t step, float grad_mult,
                           float* d_gnorm_out, void* stream);

/* The whole train_op of train.py:199 for a single process: forward_backward + clip_adam. */
int crnn_train_step(crnn_handle* h, const int32_t* d_x, const int32_t* d_y, int64_t ld_xy,
                    float* d_state, float* d_params, float* d_grads, float* d_m, float* d_v,
                    float lr, float grad_clip, int64_t step, void* d_workspace, void* stream);

/* Forward only (model.py:54-77): logits [B*S, V] with rows ordered b*S+t like the
 * reference's reshape at model.py:73, optional probs = softmax(logits) (model.py:77),
 * and final_state.  This is what Model.sample (model.py:105-140) drives at B=S=1.     */
int crnn_forward(crnn_handle* h, const int32_t* d_x, int64_t ld_x,
                 const float* d_state_in, float* d_state_out, const float* d_params,
                 float* d_logits, float* d_probs, void* d_workspace, void* stream);

/* Measurement hooks (no reference counterpart: the reference only prints wall-clock
 * time/batch, train.py:204-208).  crnn_launch_count: kernels this library has launched in
 * this process.  Profiling: when enabled, CUDA events are recorded on the launching stream
 * at every phase boundary of crnn_forward_backward / crnn_clip_adam; crnn_profile_read
 * synchronises on the last one and adds the elapsed milliseconds per phase into
 * ms_out[CRNN_NUM_PHASES] (call it after every step while profiling).                  */
long long   crnn_launch_count(void);
int         crnn_profile_enable(crnn_handle* h, int32_t on);
int         crnn_profile_read(crnn_handle* h, double* ms_out, int32_t reset);
const char* crnn_profile_phase_name(int32_t phase);
/* While profiling is enabled every launch of the persistent recurrence kernels is bracketed by
 * events on the stream it is launched on (layers may run on concurrent internal streams).
 * ms_out[2] / launches_out[2]: summed kernel durations and launch counts, [0] forward, [1] BPTT. */
int         crnn_profile_read_kernels(crnn_handle* h, double* ms_out, int64_t* launches_out, int32_t reset);

/* The dropout multipliers (0 or 1/keep products) that training call number `call_index`
 * (0-based count of crnn_forward_backward calls on this handle) applies at `site`:
 * site l in [0,L) = input of cell l (cell 0: embedding mask x input mask; cell l: output mask of
 * cell l-1 x input mask of cell l), site L = output of the top cell as seen by the projection.
 * d_out: [S,B,H] floats, time-major.  For tests and for reproducing a run on the CPU.          */
int crnn_dropout_mask(const crnn_handle* h, int64_t call_index, int32_t site, float* d_out, void* stream);

/* ---- single operators (each is one slice of the graph; used by the parity tests) ---- */

/* tf.nn.embedding_lookup + split/squeeze (model.py:55,62-63): out[t,b,:] = E[x[b,t],:]. */
int crnn_embedding_gather(const float* d_table, int32_t V, int32_t H, const int32_t* d_x, int64_t ld_x,
                          int32_t B, int32_t S, float* d_out_sbh, void* stream);
/* its gradient: IndexedSlices scatter-add into the dense [V,H] table (must be zeroed by the
 * caller) and the slices' sum of squares (d_sqnorm: device double, accumulated).       */
int crnn_embedding_scatter_add(float* d_dtable, int32_t V, int32_t H, const int32_t* d_x, int64_t ld_x,
                               int32_t B, int32_t S, const float* d_dout_sbh, double* d_sqnorm, void* stream);
/* sequence_loss_by_example + its gradient on materialised logits [N,V] (model.py:80-85):
 * in place logits -> dlogits = (softmax - onehot)/N; d_loss_sum (device double) += sum xent.
 * labels[i] for row i = t*B+b is y[b*ld_y + t].                                         */
int crnn_softmax_xent(float* d_logits_tb, int32_t B, int32_t S, int32_t V, const int32_t* d_y, int64_t ld_y,
                      double* d_loss_sum, void* stream);
/* C[M,N] = alpha-less GEMM used everywhere on the path: C (+)= op(A) op(B) (+ bias[N]).
 * ta/tb: 0 = as stored row-major, 1 = transposed.  accumulate: 0 overwrite, 1 add.    */
int crnn_gemm_f32(int32_t ta, int32_t tb, int32_t M, int32_t N, int32_t K,
                  const float* d_A, int64_t lda, const float* d_B, int64_t ldb,
                  float* d_C, int64_t ldc, const float* d_bias, int32_t accumulate, void* stream);

/* The tensor-core form of the same contraction (CRNN_PREC_BF16): bf16 operands, fp32
 * accumulate (tcgen05.mma kind::f16), C fp32.  op(A) is [M,K]: a_mn = 0 -> A stored [M,K]
 * row-major, a_mn = 1 -> stored [K,M] row-major.  op(B) is [K,N]: b_mn = 0 -> stored [N,K]
 * row-major, b_mn = 1 -> stored [K,N] row-major.  Leading dimensions are in elements and
 * must be multiples of 8 (16-byte rows for TMA).                                          */
int crnn_gemm_bf16(int32_t a_mn, int32_t b_mn, int32_t M, int32_t N, int32_t K,
                   const void* d_A_bf16, int64_t lda, const void* d_B_bf16, int64_t ldb,
                   float* d_C, int64_t ldc, const float* d_bias, int32_t accumulate, void* stream);

/* The same with split operands (CRNN_PREC_BF16X3): op(A) = A_hi + A_lo, op(B) = B_hi + B_lo, each
 * twin laid out exactly like its hi array; computes A_hi*B_hi + A_lo*B_hi + A_hi*B_lo.           */
int crnn_gemm_bf16x3(int32_t a_mn, int32_t b_mn, int32_t M, int32_t N, int32_t K,
                     const void* d_A_hi, const void* d_A_lo, int64_t lda,
                     const void* d_B_hi, const void* d_B_lo, int64_t ldb,
                     float* d_C, int64_t ldc, const float* d_bias, int32_t accumulate, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CRNN_H_ */
