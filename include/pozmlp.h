/*
 * pozmlp.h -- C ABI of the B200-native NeuralNetworkModel hot path.
 *
 * The reference (derinworks/penr-oz-neural-network-torch) is pure Python and has no FFI of its
 * own; its hot path sits behind the methods of class NeuralNetworkModel.  Each entry point below
 * names the reference code it replaces (file:line in the reference repository).  The Python class
 * in penr_oz_neural_network_torch_b200/neural_net_model.py binds these with ctypes and is the
 * only caller; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; poz_last_error() then returns a
 *     thread-local message.  Nothing aborts the process.
 *   - pointers are borrowed for the duration of the call; the context owns device parameters,
 *     optimizer state and workspace.
 *   - "stack" = n_linear Linear layers, layer l mapping sizes[l] -> sizes[l+1], each optionally
 *     followed by one activation (POZ_ACT_*).  Weights use the reference layout W[in][out]
 *     row-major (neural_net_model.py:104), bias[out] (:106-108).
 *   - host parameter/optimizer images are always float64 (the reference's dtype); the context
 *     converts to its compute dtype on upload and back on download.
 *   - device batch buffers use the compute dtype of the context: float (POZ_DTYPE_F32), double
 *     (POZ_DTYPE_F64) or __nv_bfloat16 (POZ_DTYPE_BF16); labels are int32.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - all entry points are re-entrant across contexts; one context must not be used from two
 *     threads at once (the reference holds a per-model lock around train, main.py:264-283).
 */
#ifndef POZMLP_H
#define POZMLP_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct poz_ctx poz_ctx;

enum { POZ_ACT_NONE = 0, POZ_ACT_RELU = 1, POZ_ACT_SIGMOID = 2, POZ_ACT_TANH = 3, POZ_ACT_SOFTMAX = 4 };
enum { POZ_LOSS_MSE = 0, POZ_LOSS_CE = 1 };
enum { POZ_DTYPE_F32 = 0, POZ_DTYPE_F64 = 1, POZ_DTYPE_BF16 = 2 };
enum { POZ_OPT_SGD = 0, POZ_OPT_ADAM = 1 };

/* Hyper-parameters of one train() call: the keyword arguments of
 * NeuralNetworkModel.train (neural_net_model.py:410-411) plus the optimizer choice (:279-283). */
typedef struct poz_train_config {
    int       optimizer;        /* POZ_OPT_* */
    double    learning_rate;    /* lr_t = learning_rate * decay_rate**epoch, neural_net_model.py:464 */
    double    decay_rate;
    double    dropout_rate;     /* F.dropout p on every hidden entry, neural_net_model.py:393-395 */
    double    l2_lambda;        /* cost += l2 * sum(w**2) over weights, neural_net_model.py:482-483 */
    double    beta1, beta2, epsilon;   /* Adam param_groups, neural_net_model.py:445-448 */
    long long epochs;           /* number of steps this call will run (sizes the cost ring) */
    long long progress_stride;  /* weight-update ratio recorded when epoch % stride == 0 (:504); 0 = never */
    unsigned long long dropout_seed;   /* Philox key for in-kernel masks (perf mode) */
    long long global_batch;     /* data-parallel: loss mean is over this many rows; 0 = local batch */
} poz_train_config;

/* ---- library ---------------------------------------------------------------------------- */
const char* poz_last_error(void);
int         poz_version(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
long long   poz_launch_count(void);

/* Per-kernel device timing (new; used by bench.py for the roofline of the dominant kernel).
 * Between begin and end every kernel launch of the context is bracketed by CUDA events on the
 * launching stream; end synchronises and writes {"<kernel tag>": {"launches": n, "ms": total}, ...}. */
int poz_profile_begin(poz_ctx* ctx);
int poz_profile_end(poz_ctx* ctx, char* json_out, long long cap);

/* ---- context = device image of one NeuralNetworkModel -------------------------------------
 * Replaces the parameter containers Layer.params / MultiLayerPerceptron.params
 * (neural_net_model.py:24-31,262-267) and torch.optim.Adam's state (:279-283).
 * drop_z[l] / drop_a[l]: the reference's `hidden` flag (neural_net_model.py:255) of the Linear
 * entry / of the activation entry that follows it -- the places where dropout is applied. */
int poz_ctx_create(int device, int n_linear, const int* sizes, const int* act_codes,
                   const unsigned char* drop_z, const unsigned char* drop_a,
                   int has_bias, int loss_code, int dtype_code, poz_ctx** out);
/* General stacks: an optional EmbeddingLayer head
 * (neural_net_model.py:55-68), FlattenLayer entries (:70-91, a pure reinterpretation of the rows),
 * and an optional BatchNormLayer (:122-170) between a Linear and its activation.
 * row_mult: rows of the layer's input and output = batch * row_mult (flatten entries merge `ratio`
 * rows into one); the last layer must have row_mult 1.  drop_*: how many dropout applications
 * (hidden entries, including the flatten entries folded into the tensor) hit the recorded tensor.
 * With an embedding head x_dev of every entry point is int32 indices [batch][tokens]. */
typedef struct poz_layer_desc {
    int    in, out;              /* Linear in -> out */
    int    act;                  /* POZ_ACT_* */
    int    has_bn;
    double bn_eps, bn_momentum;
    int    drop_z, drop_bn, drop_a;
    int    row_mult;
} poz_layer_desc;
typedef struct poz_stack_desc {
    int n_linear;
    const poz_layer_desc* layers;
    int has_bias, loss_code, dtype_code;
    int emb_vocab, emb_dim;      /* 0, 0 = no embedding head */
    int emb_drop;                /* dropout applications on the (flattened) embedding output */
    int tokens;                  /* indices per sample; rows of the embedding output = batch * tokens */
} poz_stack_desc;
int poz_ctx_create_ex(int device, const poz_stack_desc* desc, poz_ctx** out);
/* Upload (download = 0) or download (1) one tensor as a float64 host image.
 * kind: 0 Linear weights, 1 Linear bias, 2 embedding table, 3 batchnorm gain, 4 batchnorm bias,
 *       5 batchnorm running mean, 6 batchnorm running variance (kinds 5-6: buffer 0 only).
 * buffer: 0 value, 1 Adam exp_avg, 2 Adam exp_avg_sq, 3 gradient of the last diagnostics step (download). */
int poz_tensor_io(poz_ctx* ctx, int kind, int layer, int buffer, int download, double* host);
int poz_ctx_destroy(poz_ctx* ctx);
/* Grow the workspace for batches of up to max_batch rows (otherwise grown on demand). */
int poz_ctx_reserve(poz_ctx* ctx, long long max_batch, int for_training);

/* Layer.state_dict setter / getter (neural_net_model.py:33-45): float64 host images. */
int poz_params_upload(poz_ctx* ctx, int layer, const double* w, const double* b);
int poz_params_download(poz_ctx* ctx, int layer, double* w, double* b);
/* torch.optim.Adam state_dict()/load_state_dict() (neural_net_model.py:338-341,351-354):
 * exp_avg / exp_avg_sq per parameter tensor and the shared step count. */
int poz_adam_upload(poz_ctx* ctx, int layer, const double* m_w, const double* v_w,
                    const double* m_b, const double* v_b);
int poz_adam_download(poz_ctx* ctx, int layer, double* m_w, double* v_w, double* m_b, double* v_b);
int poz_adam_set_step(poz_ctx* ctx, long long step);
int poz_adam_get_step(poz_ctx* ctx, long long* step);

/* ---- inference: NeuralNetworkModel._forward without target + compute_output
 * (neural_net_model.py:371-396; LinearLayer.forward :110-120; activations :172-188).
 * x_dev [batch][sizes[0]] row-major, out_dev [batch][sizes[n_linear]]; both device memory. */
int poz_forward(poz_ctx* ctx, const void* x_dev, long long batch, void* out_dev, void* stream);
/* Same with the loss (neural_net_model.py:398-406).  target_dev: int32 labels [batch] for
 * POZ_LOSS_CE, compute-dtype [batch][out] for POZ_LOSS_MSE.  cost_host receives the scalar
 * (the call synchronises the stream to deliver it). */
int poz_forward_loss(poz_ctx* ctx, const void* x_dev, const void* target_dev, long long batch,
                     void* out_dev, double* cost_host, void* stream);

/* ---- training: the loop body of NeuralNetworkModel.train (neural_net_model.py:459-500) ------
 * poz_train_begin  = the per-call set-up (:444-448, epoch counter reset).
 * poz_train_step   = one epoch: forward (:480), L2 (:482-483), backward (:495), optimizer
 *                    (:496-500).  row_idx_dev (int32 [batch], may be NULL) gathers the batch rows
 *                    from x_dev/target_dev in place of the reference's list indexing (:460-461).
 *                    keep_masks (may be NULL): host array of 2*n_linear device pointers
 *                    (uint8 [batch][out] keep-masks for the z and a dropout points of each layer,
 *                    NULL entries = draw in-kernel) -- used for bit-parity with a recorded
 *                    torch mask stream.  diagnostics != 0 keeps every entry's output, its
 *                    gradient and the weight gradients of this step (:491-493) for download.
 * poz_train_costs  = cost.item() of epochs [first, first+count) (:503); synchronises.
 * poz_train_update_ratios = sum statistics behind weight_upd_ratio (:506-507). */
int poz_train_begin(poz_ctx* ctx, const poz_train_config* cfg);
int poz_train_step(poz_ctx* ctx, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                   long long batch, const void* const* keep_masks, int diagnostics, void* stream);
int poz_train_costs(poz_ctx* ctx, long long first, long long count, double* costs_host, void* stream);
/* The same step from HOST buffers -- the tensor copy of neural_net_model.py:472 plus the step plus the
 * cost.item() of :503 in one call.  x_host [batch][in] in the context's compute dtype (int32 token indices
 * for an embedding head), target_host int32 labels [batch] or [batch][out]; both should be pinned so the
 * copies are asynchronous.  The batch goes to one of POZ_HOST_LAG + 1 device slots on an internal copy
 * stream, so the copy of this call overlaps the kernels of the steps before it; the step itself runs on
 * `stream` (replayed from a CUDA graph once a slot has been seen twice; use a non-default stream for that).
 * *done_cost receives the cost of the host step issued POZ_HOST_LAG calls earlier (NaN on the first
 * POZ_HOST_LAG calls), read back on its own stream: the call returns when THAT step has finished, so the
 * host runs POZ_HOST_LAG steps ahead of the GPU and neither a slow host link nor host-side launch
 * latency leaves the GPU idle.  The host buffers of a call may be reused once POZ_HOST_LAG further calls
 * (or poz_train_host_flush) have returned.  poz_train_host_flush waits for the outstanding steps and
 * returns their costs, oldest first, right-aligned in last_costs[POZ_HOST_LAG] (NaN where fewer steps
 * were outstanding). */
#define POZ_HOST_LAG 2
int poz_train_step_host(poz_ctx* ctx, const void* x_host, const void* target_host, long long batch,
                        double* done_cost, void* stream);
int poz_train_host_flush(poz_ctx* ctx, double* last_costs);
/* ratios_host [count][n_linear]: std(w_new - w_old) / (std(w_new) + 1e-8), unbiased std. */
int poz_train_update_ratios(poz_ctx* ctx, long long first_record, long long count,
                            double* ratios_host, void* stream);

/* Diagnostics of the last step run with diagnostics != 0 (feeds the unchanged
 * _record_training_overall_progress, neural_net_model.py:532-583).  `entry` counts the
 * reference's layer list: for linear l, which = 0 is the Linear entry, which = 1 its activation,
 * which = 2 its batchnorm; layer = -1 is the embedding head (and the flatten entries after it). */
int poz_diag_activation(poz_ctx* ctx, int layer, int which, double* host_out);
int poz_diag_activation_grad(poz_ctx* ctx, int layer, int which, double* host_out, int* has_grad);
int poz_diag_weight_grad(poz_ctx* ctx, int layer, double* w_grad, double* b_grad);

/* Device-side statistics of the same tensors, replacing the mean / std / saturation / 100-bin density
 * histograms that _record_training_overall_progress computes on the host (neural_net_model.py:542-583),
 * so that only a few hundred bytes per tensor leave the device.
 * source: 0 = entry output, 1 = its gradient, 2 = the layer's weight matrix, 3 = its gradient.
 * poz_diag_moments -> out6 = {count (0 if the tensor does not exist), sum, sum of squares, min, max,
 *   saturated fraction}; rule 1: |x| > threshold (:558,:561), rule 2: x <= 0 (:559), rule 3: row maximum >
 *   threshold (:560), rule 4: row L2 norm > threshold (:556), 0: none.
 * poz_diag_histogram -> counts per bin for the caller's bin edges (torch.histogram's: bin i is
 *   [edges[i], edges[i+1]), the last one closed). */
int poz_diag_moments(poz_ctx* ctx, int layer, int which, int source, int rule, double threshold, double* out6);
int poz_diag_histogram(poz_ctx* ctx, int layer, int which, int source, const double* edges, int bins,
                       double* counts);

/* ---- data parallel (new: the reference is single-process) --------------------------------
 * The flat gradient image (all layers, compute dtype... always float32/float64) that the host
 * all-reduces with NCCL between the backward and the update of a step. */
int poz_grad_buffer(poz_ctx* ctx, void** dev_ptr, long long* n_elems, int* dtype_code, void* stream);
/* Per-layer bucket: makes `stream` wait until layer `layer`'s weight gradient of the running
 * backward is complete, sums its batch-split slabs on `stream`, and returns the layer's [W;b] slice
 * of the flat buffer -- so the host can all-reduce layer l on a communication stream while the
 * backward of layers l-1..0 is still running on the compute stream. */
int poz_grad_layer(poz_ctx* ctx, int layer, void** dev_ptr, long long* n_elems, void* stream);
/* Split step for data parallelism: backward leaves SUM-ready local gradients (already scaled by
 * 1/global_batch) in the flat buffer; the caller all-reduces; poz_train_apply runs the update. */
int poz_train_backward(poz_ctx* ctx, const void* x_dev, const void* target_dev,
                       const int* row_idx_dev, long long batch, const void* const* keep_masks,
                       int diagnostics, void* stream);
int poz_train_apply(poz_ctx* ctx, void* stream);

/* ---- unit operators: ONE entry of the reference's layer list per call -------------------------------
 * What the drop-in's Layer objects run when their forward() is called directly, as the reference's tests do
 * (test_neural_net_model.py:13-66), and what NeuralNetworkModel._forward chains when it is asked for every
 * entry's output without a target (neural_net_model.py:389-396).  Stateless: device pointers in the given
 * dtype (POZ_DTYPE_F32 / POZ_DTYPE_F64) in and out, row-major, nothing allocated, nothing kept.
 *   poz_op_linear      LinearLayer.forward (:110-120): y[rows][out] = x[rows][in] @ w[in][out], then += b[out]
 *                      (b_dev may be NULL: bias_algo "" gives a layer without bias, :105).
 *   poz_op_activation  ReluLayer / SigmoidLayer / TanhLayer / SoftmaxLayer.forward (:172-188), POZ_ACT_*;
 *                      softmax is over the last dimension (cols).  POZ_ACT_NONE copies (Layer.forward, :48-53).
 *   poz_op_embedding   EmbeddingLayer.forward (:67-68): y[n][dim] = table[idx[n]][:], idx int32 in [0, vocab).
 *   poz_op_batchnorm   BatchNormLayer.forward (:158-170).  training != 0: mean and UNBIASED variance of the batch
 *                      (over the rows, per column) normalise it and move run_mean / run_var by `momentum`
 *                      (:161-166); otherwise the running statistics are used (:168).  scalar_stats != 0: one mean /
 *                      variance over every element, broadcast to all columns -- what the reference computes for a
 *                      1-D input in training mode, where `dim = ()` reduces over everything (:160-162). */
int poz_op_linear(int device, int dtype_code, const void* x_dev, const void* w_dev, const void* b_dev, long long rows,
                  int in, int out, void* y_dev, void* stream);
int poz_op_activation(int device, int dtype_code, int act, const void* x_dev, long long rows, int cols, void* y_dev,
                      void* stream);
int poz_op_embedding(int device, int dtype_code, const void* table_dev, int vocab, int dim, const int* idx_dev,
                     long long n, void* y_dev, void* stream);
int poz_op_batchnorm(int device, int dtype_code, int training, int scalar_stats, const void* x_dev, long long rows, int cols,
                     const void* gain_dev, const void* bias_dev, void* run_mean_dev, void* run_var_dev, double eps,
                     double momentum, void* y_dev, void* stream);

/* ---- tcgen05 GEMM building block (new; the dense contractions of LinearLayer.forward,
 * neural_net_model.py:117, and of its autograd backward, :495, for wide layers) -----------------
 * bf16 row-major device matrices, fp32 accumulation in tensor memory.  Exposed so that the three
 * operand-layout variants can be tested and benchmarked in isolation; bf16 contexts call the same
 * launchers internally.
 *   variant 0 (forward)   out bf16 [m][n] = act(a[m][k] * b[k][n] + bias[n])         (bias may be NULL)
 *   variant 1 (backward-x) out bf16 [m][n] = (a[m][k] * b[n][k]^T) .* act'(aux[m][n])  (aux = activation outputs)
 *   variant 2 (backward-w) out fp32 [splits][m][n] = partial sums over k of a[k][m]^T * b[k][n]
 * n and k (and m for variant 2) must be multiples of 8. */
int poz_tc_gemm(int device, int variant, const void* a, const void* b, long long m, int n, long long k,
                void* out, const float* bias, const void* aux, int act, int splits, void* stream);

/* Split-bf16 variant of the same building block: float32 device matrices in and out; each operand
 * is carried as three bf16 planes (hi + mid + lo) and the six significant plane products are
 * accumulated in tensor memory, giving float32-grade results at tensor-core rate (used by float32
 * contexts for layers wide enough to tile).  Same variants as poz_tc_gemm; out is float32
 * ([splits][m][n] for variant 2); aux (variant 1, may be NULL) holds activation outputs as float32.
 * Allocates temporary plane buffers: a test / benchmark entry, not a hot-path one. */
int poz_tc3_gemm(int device, int variant, const float* a, const float* b, long long m, int n, long long k,
                 float* out, const float* bias, const float* aux, int act, int splits, void* stream);

/* ---- inference without any allocation ---------------------------------------------------------
 * No counterpart in the reference (it has no notion of device memory at all).
 *
 * For callers that must not allocate GPU memory when they need a prediction -- for example a server
 * whose own memory pool already owns the device, or a loop that cannot tolerate the device-wide
 * pause an allocation may cause -- and that therefore cannot create a poz_ctx.  The caller passes
 * every input and output as device pointers it already holds: float32 weights W[in][out] row-major
 * and biases b[out] per layer (biases_dev, or single entries of it, may be NULL), x_dev
 * [batch][sizes[0]], out_dev [batch][sizes[n_linear]].  Result = poz_forward on a float32 context
 * holding the same parameters.
 *
 * The hidden activations go through a fixed POZ_NOALLOC_SCRATCH_BYTES array that is part of the
 * library image: every GPU that loads the library keeps those 8 MiB until the process ends.  Limit:
 * no layer output wider than POZ_NOALLOC_MAX_WIDTH (larger requests fail with a message; use a
 * context for those); any batch size (rows are processed in slices that fit the array).
 * The array is shared by all calls and keeps the previous call's contents; nothing in it needs to be
 * zero, so nothing is cleared.  Calls that run at the same time on one GPU overwrite each other's
 * activations: serialise them.  The first call on a device looks up the array's device address; make
 * that first call before starting any stream capture on that device.  The function calls no
 * cudaMalloc / cudaFree and no allocating library routine.  `stream` must be a stream of `device`
 * (or NULL); the caller's current device is the same after the call as before it. */
#define POZ_NOALLOC_SCRATCH_BYTES (8ll << 20)
#define POZ_NOALLOC_MAX_WIDTH 4096
int poz_forward_noalloc(int device, int n_linear, const int* sizes, const int* act_codes,
                        const float* const* weights_dev, const float* const* biases_dev,
                        const float* x_dev, long long batch, float* out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* POZMLP_H */
