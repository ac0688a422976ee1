/*
 * tdq_b200.h -- C-ABI of the B200-native PINN hot path (libtdq_b200.so).
 *
 * This is the drop-in boundary for ONE path of tensordiffeq: the per-step
 * collocation residual evaluation, its gradient, and the optimizer updates
 * that consume it.  Every entry point names the reference interface it
 * replaces (paths relative to the tensordiffeq repository).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; the message is
 *     available from tdq_last_error(h) (h may be NULL for tdq_create failures);
 *   - pointers are HOST pointers unless a name ends in _dev; the caller owns
 *     them, the library copies on call and keeps nothing after return;
 *   - all device memory lives behind the handle; one handle = one GPU, one
 *     CUDA stream; a handle is not thread-safe; multi-GPU = one process (or
 *     thread) and one handle per rank;
 *   - arithmetic is fp32 throughout (reference: models.py:46-49, keras default);
 *   - flat parameter vectors use the reference layout of utils.py:7-29
 *     (per layer: kernel [in,out] row-major, then bias).
 */
#ifndef TDQ_B200_H
#define TDQ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tdq_handle tdq_handle;

/* ---- residual / loss-integrand program --------------------------------
 * The reference hands the solver a Python callback (f_model, models.py:40,157;
 * deriv_model, boundaries.py:238-239) that builds the residual with nested
 * tf.gradients.  Here the host traces that callback once into a straight-line
 * program over the Taylor-jet channels the kernel carries per point; the kernel
 * evaluates it (and its reverse sweep) per collocation point.
 * Slot i of the program is the result of ops[i].  */
enum {
    TDQ_OP_CONST = 0,   /* imm                                               */
    TDQ_OP_JET = 1,     /* a = jet channel (0 = u, then per input dimension
                           its orders 1..K in tdq_term_spec.orders order)    */
    TDQ_OP_COORD = 2,   /* a = input dimension                               */
    TDQ_OP_TARGET = 3,  /* per-point target value (BC/IC value, data)        */
    TDQ_OP_LAMBDA = 4,  /* per-point self-adaptive weight                    */
    TDQ_OP_VAR = 5,     /* a = index of a trainable PDE coefficient          */
    TDQ_OP_ADD = 6, TDQ_OP_SUB = 7, TDQ_OP_MUL = 8, TDQ_OP_DIV = 9, /* a, b  */
    TDQ_OP_NEG = 10, TDQ_OP_SIN = 11, TDQ_OP_COS = 12, TDQ_OP_EXP = 13,
    TDQ_OP_TANH = 14, TDQ_OP_SQRT = 15, TDQ_OP_LOG = 16,             /* a    */
    TDQ_OP_POWI = 17,   /* a ** (int)imm                                     */
    TDQ_OP_POWF = 18    /* a ** imm                                          */
};
#define TDQ_MAX_OPS 160
#define TDQ_MAX_VARS 8
#define TDQ_MAX_DIMS 3
#define TDQ_MAX_ORDER 4

typedef struct { int32_t op, a, b; float imm; } tdq_op;

/* One loss term = one point set + one program (replaces one entry of `bcs`
 * or the residual in CollocationSolverND.update_loss, models.py:80-192;
 * MSE / g_MSE of utils.py:38-45 are expressed inside the program).
 *   pair == 0: loss += sum_i slot[out_slot](x_i) / n_global
 *   pair == 1: points 2i / 2i+1 are an (upper, lower) couple and
 *              loss += sum_i (E(x_2i) - E(x_2i+1))^2 / n_global, E = slot[out_slot]
 *              (periodicBC, models.py:100-111).                              */
typedef struct {
    int32_t n_dims;                   /* input dimension d of the network     */
    int32_t orders[TDQ_MAX_DIMS];     /* highest pure derivative per dimension */
    const tdq_op* ops;
    int32_t n_ops;
    int32_t out_slot;
    int32_t f_slot;                   /* slot reported as "f" by tdq_predict / tdq_term_residual, or -1 */
    int32_t pair;
    int32_t mixed;                    /* 1: the program also reads the mixed partial u_xy (n_dims == 2, orders <= 2) as
                                         the jet channel after the pure ones, index 1 + orders[0] + orders[1]      */
} tdq_term_spec;

/* keras.optimizers.Adam as the reference builds it (models.py:35-36): TF's
 * ResourceApplyAdam.  sign = +1 descends, -1 ascends (negated lambda
 * gradients, fit.py:101,104). */
typedef struct { float lr, beta1, beta2, eps; int32_t sign; } tdq_adam_cfg;

/* ---- life cycle: replaces neural_net(layer_sizes) (networks.py:10-20) and the
 * state CollocationSolverND.compile builds (models.py:17-69).
 * stream: a cudaStream_t to run on, or NULL for a private stream. */
int tdq_create(const int32_t* layer_sizes, int32_t n_layers, int32_t device, void* stream, tdq_handle** out);
int tdq_destroy(tdq_handle* h);
const char* tdq_last_error(const tdq_handle* h);
/* number of SMs / kernels launched so far / bytes of device scratch held */
int tdq_info(const tdq_handle* h, int32_t* n_sm, int64_t* n_launches, int64_t* scratch_bytes);

/* utils.py:7-29 set_weights / get_weights */
int tdq_set_params(tdq_handle* h, const float* flat, int64_t n);
int tdq_get_params(tdq_handle* h, float* flat, int64_t n);
/* 1 = trainable, 0 = frozen by the weight optimizer (DiscoveryModel never
 * updates the output bias, models.py:339,346) */
int tdq_set_param_mask(tdq_handle* h, const float* mask, int64_t n);
/* trainable PDE coefficients (DiscoveryModel.vars, models.py:301) */
int tdq_set_vars(tdq_handle* h, const float* v, int32_t n);
int tdq_get_vars(tdq_handle* h, float* v, int32_t n);

/* ---- terms */
int tdq_add_term(tdq_handle* h, const tdq_term_spec* spec, int32_t* term_id);
/* X: [n][d] row-major fp32 (models.py:49 casts the LHS columns to float32).
 * n_global: denominator of the mean (n, or the global point count when this
 * rank holds a shard; for a paired term, the number of couples).  May be called again
 * (same or different n). */
int tdq_set_term_points(tdq_handle* h, int32_t term, const float* X, int64_t n, int64_t n_global);
/* Streaming point sets (collocation points resampled between steps; the reference would rebuild its X_f
 * tensors, models.py:49-51): tdq_prefetch_term_points starts an asynchronous upload of n = the term's current
 * count points ([n][d] row-major, pinned host memory for a truly asynchronous copy) into the term's shadow
 * buffer, on a stream of its own, and returns; steps issued meanwhile keep using the current set.
 * tdq_swap_term_points waits for that upload and makes the shadow buffer the current set (the previous one
 * becomes the shadow).  X must stay valid and unchanged until tdq_swap_term_points returns. */
int tdq_prefetch_term_points(tdq_handle* h, int32_t term, const float* X, int64_t n);
int tdq_swap_term_points(tdq_handle* h, int32_t term);
/* minibatching (models.py:148-155, fit.py:86-89): evaluate only points [start, start+count) of the
 * stored set, with mean denominator n_global; count < 0 restores the full set. */
int tdq_set_term_range(tdq_handle* h, int32_t term, int64_t start, int64_t count, int64_t n_global);
int tdq_set_term_target(tdq_handle* h, int32_t term, const float* target, int64_t n);
/* utils.py:99-112 initialize_weights_loss: one lambda vector per adaptive term;
 * resets that lambda's Adam moments */
int tdq_set_term_lambda(tdq_handle* h, int32_t term, const float* lam, int64_t n);
int tdq_get_term_lambda(tdq_handle* h, int32_t term, float* lam, int64_t n);
int tdq_get_term_lambda_grad(tdq_handle* h, int32_t term, float* grad, int64_t n);
/* slot[f_slot] at every point of the term, from the last evaluation */
int tdq_get_term_residual(tdq_handle* h, int32_t term, float* f, int64_t n);

/* ---- the hot path.
 * models.py:195-199 grad() and :260-272 loss_and_flat_grad(): loss, d loss/d
 * weights (flat layout), d loss / d vars; lambda gradients stay on the device
 * (tdq_get_term_lambda_grad).  Any output pointer may be NULL. */
int tdq_loss_grad(tdq_handle* h, float* loss, float* grad_flat, float* var_grads);
/* fit.py:83-113 apply_grads, n_steps times: evaluate, Adam on the weights,
 * Adam (sign) on every lambda, Adam on vars.  Any cfg may be NULL (that group
 * is then left alone).  loss_hist (NULL or n_steps floats) receives the loss
 * of every step, evaluated before that step's update like the reference. */
int tdq_adam_steps(tdq_handle* h, const tdq_adam_cfg* w, const tdq_adam_cfg* lam, const tdq_adam_cfg* vars,
                   int32_t n_steps, float* loss_hist);
/* fit.py:86-111 with batch_sz set: n_epochs passes over n_batches consecutive slices of batch_sz points of the
 * listed terms (the points were stored in the frozen permuted order, models.py:148-155), one Adam step per batch,
 * BC / IC terms whole in every batch, the slice's mean taken over batch_sz.  The whole schedule is enqueued without
 * a host round trip per batch.  loss_hist (NULL or n_epochs floats): the loss of each epoch's last batch
 * (fit.py:109).  The terms' ranges are reset to the full sets on return. */
int tdq_adam_epochs(tdq_handle* h, const tdq_adam_cfg* w, const tdq_adam_cfg* lam, const tdq_adam_cfg* vars,
                    const int32_t* terms, int32_t n_terms, int64_t batch_sz, int32_t n_batches, int32_t n_epochs,
                    float* loss_hist);
int tdq_adam_reset(tdq_handle* h);   /* models.py:244-247: fresh optimizers */
/* optimizers.py:214-402 eager_lbfgs(opfunc, x, state, maxIter, learningRate) as fit.py:49-51
 * calls it: history 50, no line search.  f_hist: max_iter+1 floats.  The iteration (two-loop recursion, step,
 * stop tests of :357-380) runs on the device; the host looks at its state every 16 iterations. */
int tdq_lbfgs(tdq_handle* h, int32_t max_iter, float learning_rate, float* f_hist, int32_t* n_hist, int32_t* n_eval);
/* fit.py:65-80 lbfgs_train -> tfp.optimizer.lbfgs_minimize(func, init_params, max_iterations=newton_iter,
 * tolerance=1e-20) (the newton_eager=False branch of fit.py:52-58): L-BFGS with 10 correction pairs and the
 * Hager-Zhang line search, tfp's documented defaults (tfp itself is third-party and absent: the algorithm is restated
 * from its published sources, see oracle/hz_lbfgs.py).  f_hist: max_iter+1 floats, the loss at the start and after
 * every accepted step.  status: 0 max_iter reached, 1 max|g| <= tolerance, 2 the line search found no step. */
int tdq_lbfgs_hz(tdq_handle* h, int32_t max_iter, float tolerance, float* f_hist, int32_t* n_hist, int32_t* n_eval,
                 int32_t* status);
/* models.py:274-284 predict: u = network output, f = slot[f_slot] of `spec` at the M points X[M][d] */
int tdq_predict(tdq_handle* h, const tdq_term_spec* spec, const float* X, int64_t m, float* u, float* f);

/* ---- for callers that must not allocate device memory.
 * No counterpart in the reference (TensorFlow owns every allocation there).
 * Loss and gradients of ONE term, handle-free: every input and output is a DEVICE buffer the
 * caller already owns, and all working memory (padded weights, per-CTA gradient slabs, per-CTA
 * activation stash) comes from a 32 MiB array built into the library's device image, so no
 * cudaMalloc / cudaFree happens here or behind it.  Meant for plug-ins living inside another
 * program's memory pool, and for loops that may never touch the driver's allocator.
 *   layer_sizes/n_layers, spec : host, as for tdq_create / tdq_add_term
 *   params_dev [P]             : flat parameters, utils.py:7-29 layout
 *   vars_dev [n_vars] or NULL  : trainable PDE coefficients
 *   x_cols_dev [d][n]          : one contiguous column per input variable (models.py:49)
 *   target_dev, lam_dev [n]    : or NULL when the program does not read them
 *   gvec_dev [P + 16]   (out)  : [0,P) d loss/d params, [P,P+8) d loss/d vars, [P+8] loss, rest 0
 *   lam_grad_dev [n]    (out)  : d loss/d lambda, or NULL
 *   err/err_len                : receives the message when the return value is non-zero
 * Work is enqueued on `stream` of the CURRENT device; results are valid once the stream has
 * drained.  Networks / terms whose scratch exceeds the array are refused.  The first call on
 * each device resolves the array's address -- make it at start-up, before any stream capture.
 * All calls share the array: calls that overlap on the GPU (other streams / threads) race. */
int tdq_loss_grad_static(const int32_t* layer_sizes, int32_t n_layers, const tdq_term_spec* spec,
                         const float* params_dev, const float* vars_dev, int32_t n_vars,
                         const float* x_cols_dev, const float* target_dev, const float* lam_dev,
                         int64_t n, int64_t n_global, float* gvec_dev, float* lam_grad_dev,
                         void* stream, char* err, int32_t err_len);

/* ---- multi-GPU: replaces the implicit NcclAllReduce of tf.distribute.MirroredStrategy
 * (models.py:219, fit.py:149-153).  nccl_unique_id: the 128 bytes of an ncclUniqueId
 * made on rank 0 with tdq_comm_unique_id and shared by the host layer.  After this
 * call tdq_loss_grad / tdq_adam_steps SUM-all-reduce the packed (grad, var grads,
 * loss) vector across ranks. */
int tdq_comm_unique_id(void* out128);
int tdq_comm_init(tdq_handle* h, const void* nccl_unique_id, int32_t rank, int32_t world);
/* how the ranks exchange gradients, NUL-terminated into buf[cap]: "single", "nccl" (reduce kernel -> ncclAllReduce
 * -> Adam kernel) or "nvlink-peer ..." (on one node with peer access: ONE kernel sums the local slabs, exchanges the
 * partial sums through the peers' memory over NVLink, adds them in rank order and applies the Adam step).
 * TDQ_NO_PEER=1 in the environment forces "nccl".  Reporting only. */
int tdq_comm_mode(tdq_handle* h, char* buf, int32_t cap);

/* ---- timing helper for callers without a CUDA binding: elapsed device time of
 * everything enqueued on the handle's stream between the two marks (ms). */
int tdq_mark_begin(tdq_handle* h);
int tdq_mark_end(tdq_handle* h, float* ms);
int tdq_sync(tdq_handle* h);
/* device time (ms) accumulated in the fused jet kernel of `term` since the last call with reset!=0,
 * measured with CUDA events around every launch, and the number of launches timed */
int tdq_kernel_time(tdq_handle* h, int32_t term, int32_t reset, float* ms, int64_t* launches);
int tdq_set_kernel_timing(tdq_handle* h, int32_t enable);
/* name of the kernel tdq_loss_grad / tdq_adam_steps launch for `term` (the tensor-core kernel
 * "fused_jet_tc_kernel<k0,k1,k2,warps>" for 128-wide nets with at most 4 jet channels, else the FP32 FMA kernel
 * "fused_jet_kernel<k0,k1,k2,width,tile,bwd>"), NUL-terminated into buf[cap].  Reporting only. */
int tdq_term_kernel(tdq_handle* h, int32_t term, char* buf, int32_t cap);
#ifdef __cplusplus
}
#endif
#endif /* TDQ_B200_H */
