/*
 * nlbac.h — C ABI of the B200 (sm_100a) NLBAC neural-ODE hot path.
 *
 * The reference (LiqunZhao/...-NLBAC) is pure Python and has no FFI; the surfaces these entry
 * points replace are Python call sites (paths relative to the reference root; U =
 * NLBAC_Unicycle_RL_training/Unicycle_RL_training, C = NLBAC_SimulatedCarsFollowing_RL_training/
 * Simulated_Car_Following_RL_training, P = NLBAC_pvtol_RL_training/Pvtol_RL_training):
 *
 *   nlbac_node_*      torchdiffeq.odeint(func, y0, t, method=...) with func = NeuralODEModel
 *                     (U sac_cbf_clf/model.py:177-217,252; U sac_cbf_clf/sac_cbf_clf.py:453,577;
 *                      C model.py:179-206,245; P model.py:179-222,259) and autograd through it.
 *   nlbac_mlp_*       QNetwork / LyaNetwork / BarrierNetwork / GaussianPolicy forward+backward
 *                     (U model.py:37-133; BU model.py:67-86).
 *   nlbac_adam_step, nlbac_soft_update
 *                     torch.optim.Adam.step as used by train_step (U model.py:257) and
 *                     utils.soft_update (U sac_cbf_clf/utils.py:75-79).
 *   nlbac_constraints_*   the CBF/CLF constraint block of get_policy_loss_2
 *                     (U sac_cbf_clf.py:429-504; C :462-523; BU :424-460).
 *   nlbac_get_state / nlbac_get_obs   DynamicsModel.get_state / get_obs (U dynamics.py:27-69;
 *                     C dynamics.py:25-96; P dynamics.py:25-153).
 *
 * Conventions: every function returns 0 on success, <0 for an invalid argument, >0 for a CUDA
 * error (cudaError_t); nlbac_last_error() returns the message of the last failure on the calling
 * thread.  All tensor pointers are DEVICE pointers owned by the caller, float32, row-major
 * contiguous [batch, feature]; nothing is retained after the call returns.  Work is enqueued
 * asynchronously on the given stream (a cudaStream_t passed as void*).  No hidden allocation:
 * workspaces are sized by the *_workspace_bytes queries and supplied by the caller.  Time grids
 * are HOST pointers (they are tiny and drive the launch schedule).
 * Threading: entry points may be called concurrently from several host threads on DISTINCT streams; the library keeps
 * no per-call state.  Process-wide state is limited to: per-device caches of kernel attributes / SM count / the
 * no-alloc scratch address (set once, under a lock or idempotently), the dlopen'ed NCCL table, and -- DEBUG ONLY --
 * the event list of nlbac_kernel_timing, which is NOT thread-safe: enable it from one thread with no other thread
 * inside the library (bench.py does), never in production.  nlbac_node_solve_bwd_noalloc shares one scratch array per
 * device: calls that overlap on one GPU race.  Shapes: the NODE entry points have compiled instances for the four
 * shapes of node_cfg.cuh only (anything else returns -2 "no compiled kernel instance"); the MLP entry points accept
 * n_in <= 32, n_out <= 8 and any hidden width that is a multiple of 32 in [32, 1024] (--hidden_size), else -1.
 */
#ifndef NLBAC_H_
#define NLBAC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NLBAC_ABI_VERSION 2   /* 2: nlbac_node_tc_bwd_workspace_bytes takes the precision tier; nlbac_comm_*, *_noalloc, umma_peak_probe */

typedef struct nlbac_comm nlbac_comm;   /* NCCL communicator handle, see nlbac_comm_init below */

/* model kinds */
#define NLBAC_KIND_AFFINE   0   /* x' = f_net(x) + g_net(x) u      (U, P, BU, BP) */
#define NLBAC_KIND_BLACKBOX 1   /* x' = net([x, u, const])          (C, quadrotor)  */

/* solver methods (torchdiffeq names) */
#define NLBAC_EULER    0
#define NLBAC_MIDPOINT 1
#define NLBAC_RK4      2   /* 3/8 rule, like torchdiffeq's 'rk4' */
#define NLBAC_DOPRI5   3

/* arithmetic of the hidden-layer contractions */
#define NLBAC_PREC_FP32 0  /* FP32 FMA pipe; parity bar rel 1e-5 */
#define NLBAC_PREC_BF16 1  /* reserved: bf16 operands measured at 2.7e-2 on d-theta -> not built */
#define NLBAC_PREC_TF32 2  /* tcgen05 kind::tf32 hidden layers (forward and backward; dW on kind::f16 bf16), fp32 accumulate; bar rel 1e-2 */
#define NLBAC_PREC_3XTF32 3 /* 3xTF32 split (hi*hi + lo*hi + hi*lo) on tcgen05 (forward); fp32-class, bar rel 1e-5 */

typedef struct nlbac_node_desc {
  int kind;        /* NLBAC_KIND_* */
  int n_s;         /* integrated state columns */
  int n_u;         /* action columns (constants of the flow) */
  int n_c;         /* extra constant columns after the action (cars: 1, the time feature) */
  int hidden;      /* hidden width (100 affine, 64 black-box) */
  int n_f_layers;  /* Linear layers in f_net / net */
  int n_g_layers;  /* Linear layers in g_net (0 for black-box) */
} nlbac_node_desc;

const char* nlbac_last_error(void);
int nlbac_abi_version(void);

/* 0 if this (desc, precision) pair has a compiled kernel instance, <0 otherwise. */
int nlbac_node_supported(const nlbac_node_desc* desc, int precision);

/* Number of float parameters of the model, in state-dict order (W then b per Linear, f_net then g_net). */
int64_t nlbac_node_param_count(const nlbac_node_desc* desc);

/* Floats in the packed (kernel-side) parameter image produced by nlbac_node_pack. */
int64_t nlbac_node_packed_floats(const nlbac_node_desc* desc);

/* Pack the 2*(n_f_layers+n_g_layers) parameter tensors (W0,b0,W1,b1,... in state-dict order; W is
 * [out,in] like nn.Linear.weight) into the padded/transposed image the kernels stream through
 * shared memory.  `params` is a HOST array of DEVICE pointers. */
int nlbac_node_pack(const nlbac_node_desc* desc, const float* const* params, float* packed, void* stream);

/* Packed image for the tensor-core tiers (NLBAC_PREC_TF32 / NLBAC_PREC_3XTF32): thin layers + one UMMA K-major
 * core-matrix image (hi and lo halves) per hidden matrix.  Pass THIS image as `packed` to nlbac_node_solve_fwd
 * when precision is a tensor-core tier, and to nlbac_node_solve_bwd with the same tier (its reverse sweep then runs on
 * tcgen05 too: TF32 = 1e-2 tier, 3XTF32 = fp32-class); a backward with NLBAC_PREC_FP32 takes the FP32 image of
 * nlbac_node_pack whatever the forward's tier was. */
int64_t nlbac_node_tc_packed_floats(const nlbac_node_desc* desc);
int nlbac_node_tc_pack(const nlbac_node_desc* desc, const float* const* params, float* packed, void* stream);

/* Workspace of the tensor-core backward (precision NLBAC_PREC_TF32 or NLBAC_PREC_3XTF32 in nlbac_node_solve_bwd,
 * `packed` = the tensor-core image): schedule tables + gradient replicas + (when parameter gradients are wanted) the
 * bf16 activation / cotangent images of one chunk of trajectories (<= 24 GB; the call loops over chunks; the 3xTF32
 * tier writes every image as a hi / lo pair, twice the bytes per trajectory). */
int nlbac_node_tc_bwd_workspace_bytes(const nlbac_node_desc* desc, int64_t B, int n_grid, int n_t, int method,
                                      int precision, int want_param_grads, size_t* bwd_bytes);

/* Workspace sizes in bytes for a solve over `n_grid` integration grid points (n_grid-1 steps). */
int nlbac_node_workspace_bytes(const nlbac_node_desc* desc, int64_t B, int n_grid, int n_t, int method,
                               size_t* fwd_bytes, size_t* ckpt_bytes, size_t* bwd_bytes);

/* Fixed-grid solve (euler / midpoint / rk4).
 *   y0      [B, D]   D = n_s + n_u + n_c
 *   t       host, n_t output times; grid host, n_grid integration points (grid==t when no step_size)
 *   y_out   [n_t, B, D]; y_out[0] = y0
 *   ckpt    NULL, or ckpt_bytes of device memory receiving the stage inputs the backward re-uses
 *   ws      fwd_bytes of device memory (schedule tables) */
int nlbac_node_solve_fwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                         const float* t, int n_t, const float* grid, int n_grid, int method, int precision,
                         float* y_out, float* ckpt, void* ws, void* stream);

/* Reverse sweep through the discrete steps of the solve above.
 *   grad_out     [n_t, B, D] cotangent of y_out
 *   grad_y0      [B, D] or NULL (overwritten)
 *   grad_params  [param_count] flat, state-dict order, or NULL (overwritten; sums over the batch) */
int nlbac_node_solve_bwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                         const float* t, int n_t, const float* grid, int n_grid, int method, int precision,
                         const float* ckpt, const float* grad_out, float* grad_y0, float* grad_params,
                         void* ws, void* stream);

/* nlbac_node_solve_bwd for callers that may not allocate (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as nlbac_node_solve_bwd with precision NLBAC_PREC_FP32, minus `ws`: the scratch (step
 * schedule + the 8 gradient replicas the CTAs reduce into, ~1.7 MB for the shipped shapes) lives in a file-scope
 * `static __device__` array of NLBAC_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that loads the library
 * holds until exit.  The entry point calls no cudaMalloc / cudaFree / allocating routine; the caller still passes every
 * input and output.  A solve whose scratch (nlbac_node_workspace_bytes' bwd_bytes) exceeds the array, or any precision
 * other than NLBAC_PREC_FP32 (the tensor-core tier needs GBs of operand images), is rejected with -1 and a message.
 * The array is shared by every call on that device and keeps the last call's contents: the entry point clears what must
 * start at zero on the caller's stream before use, and calls that overlap on one GPU (two streams, two threads) RACE --
 * serialise them. */
#define NLBAC_NOALLOC_SCRATCH_BYTES (2u << 20)
int nlbac_node_solve_bwd_noalloc(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                                 const float* t, int n_t, const float* grid, int n_grid, int method,
                                 const float* ckpt, const float* grad_out, float* grad_y0, float* grad_params,
                                 void* stream);

/* Adaptive dopri5 with torchdiffeq semantics (one step size for the whole batch from an RMS norm over all
 * of [B, D]; FSAL; 4th-order dense output at the requested times).  The forward SYNCHRONISES the stream once
 * per attempted step (accept/reject is a host decision, as in torchdiffeq).  `t` host, strictly increasing.
 *   ckpt      NULL or ckpt_bytes: stage inputs of every accepted step (needed by the backward)
 *   step_dt   host out [max_steps]; out_step / out_x host out [n_t]: which accepted step and which fraction of
 *             it produced each output (inputs of the backward). */
int nlbac_node_dopri5_workspace_bytes(const nlbac_node_desc* desc, int64_t B, int n_t, int max_steps,
                                      size_t* fwd_bytes, size_t* ckpt_bytes, size_t* bwd_bytes);
int nlbac_node_dopri5_fwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                          const float* t, int n_t, float rtol, float atol, int max_steps, int precision,
                          float* y_out, float* ckpt, void* ws, int* n_accepted, int* n_rejected, float* step_dt,
                          int* out_step, float* out_x, void* stream);
int nlbac_node_dopri5_bwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B, int n_t,
                          int n_accepted, const float* step_dt, const int* out_step, const float* out_x,
                          int precision, const float* ckpt, const float* grad_out, float* grad_y0,
                          float* grad_params, void* ws, void* stream);

/* The same solve with the step-size controller ON THE DEVICE: accept / reject, the next step size, the output
 * bookkeeping and the error flags live in `state` (device memory, nlbac_node_dopri5_state_bytes, 8-byte aligned) and
 * are advanced by a one-thread kernel after each attempt in the same double arithmetic as the host loop above, so
 * the step sequence is identical.  Nothing is read back and nothing is allocated: forward and backward can be
 * captured in a CUDA graph (torchdiffeq cannot: it reads error_ratio on the host, rk_common.py
 * _adaptive_step).  `attempts` attempts (3 launches each) are enqueued; once the last output time has been passed
 * the remaining ones return at once.  resume = 1 enqueues `attempts` more on the same `ws` / `state` / `ckpt`.
 * nlbac_node_dopri5_state_read synchronises and reports; err: 0 ok, 1 non-finite error ratio, 2 max_steps accepted
 * steps exhausted, 3 stalled (100 * max_steps attempts), 4 step-size underflow, 5 not finished within the attempts
 * enqueued so far (enqueue more with resume = 1).  The backward takes the step table from `state`; after err != 0
 * its results are meaningless (but it stays in bounds).  ws / ckpt sizes: nlbac_node_dopri5_workspace_bytes.
 * Sharded batch: with `comm` non-NULL the three squared-norm sums are all-reduced over the ranks
 * (3 doubles, on `stream`) before every controller step and divided by B_global * D, so every rank takes the step
 * sequence of the unsharded batch; every rank must enqueue the same number of attempts.  comm = NULL: B_global unused. */
int nlbac_node_dopri5_state_bytes(int n_t, int max_steps, size_t* bytes);
int nlbac_node_dopri5_fwd_dev(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                              const float* t, int n_t, float rtol, float atol, int max_steps, int attempts,
                              int resume, int precision, float* y_out, float* ckpt, void* ws, void* state,
                              nlbac_comm* comm, int64_t B_global, void* stream);
int nlbac_node_dopri5_state_read(const void* state, int n_t, int max_steps, int* n_accepted, int* n_rejected,
                                 int* done, int* err, float* step_dt, void* stream);
int nlbac_node_dopri5_bwd_dev(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B, int n_t,
                              int max_steps, const void* state, int precision, const float* ckpt,
                              const float* grad_out, float* grad_y0, float* grad_params, void* ws, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Actor / critic / Lyapunov / barrier MLPs: in -> hidden -> hidden -> out, ReLU (QNetwork is two of
 * them, U sac_cbf_clf/model.py:37-64; LyaNetwork :67-83; BarrierNetwork BU model.py:67-86;
 * GaussianPolicy :86-115 with out = [mean | log_std]).
 *   params   HOST array of 6 DEVICE pointers W1,b1,W2,b2,W3,b3 (nn.Linear layout, W is [out,in])
 *   act      [2,B,hidden] post-ReLU activations written by the forward for the backward (or NULL)
 *   ws       bwd_bytes of scratch;  grad_params HOST array of 6 DEVICE pointers (overwritten) or NULL
 * ------------------------------------------------------------------------------------------------ */
typedef struct nlbac_mlp_desc {
  int n_in;    /* <= 32 */
  int hidden;  /* multiple of 32, <= 1024 (reference: 256) */
  int n_out;   /* <= 8 */
} nlbac_mlp_desc;

int nlbac_mlp_workspace_bytes(const nlbac_mlp_desc* desc, int64_t B, size_t* act_bytes, size_t* bwd_bytes);
int nlbac_mlp_fwd(const nlbac_mlp_desc* desc, const float* const* params, const float* x, int64_t B, float* out,
                  float* act, void* stream);
int nlbac_mlp_bwd(const nlbac_mlp_desc* desc, const float* const* params, const float* x, int64_t B, const float* act,
                  const float* grad_out, float* grad_x, float* const* grad_params, void* ws, void* stream);

/* tanh-squashed Gaussian head of GaussianPolicy.sample (U model.py:117-128) with the N(0,1) draw `eps`
 * injected: mean_logstd [B, 2*n_u] -> action [B,n_u], log_prob [B,1], mean_out [B,n_u]. */
int nlbac_policy_head_fwd(const float* mean_logstd, const float* eps, const float* action_scale,
                          const float* action_bias, int64_t B, int n_u, float* action, float* log_prob,
                          float* mean_out, void* stream);
int nlbac_policy_head_bwd(const float* mean_logstd, const float* eps, const float* action_scale, int64_t B, int n_u,
                          const float* grad_action, const float* grad_log_prob, const float* grad_mean_out,
                          float* grad_mean_logstd, void* stream);

/* ------------------------------------------------------------------------------------------------
 * CBF / CLF constraint block of get_policy_loss_2: rows -> where(>0) -> batch sum / batch_size -> req[m]
 * (U sac_cbf_clf.py:472-504; C :474-523; BU :424-460).  One fused elementwise + reduction kernel.
 *   UNICYCLE  a = p(x_t) [B,2], b = p(x_{t+1}) [B,2] (grad);  rows: n_hazards CBF rows, then the CLF row
 *   CARS      a = x_t [B,10], b = x_{t+1} [B,10] (grad), c = x_{t+2} [B,10] (grad); rows: cbf23, cbf34, CLF
 *   BARRIER   a = B(x_t,u_t) detached [B], b = B(x_{t+1},u_{t+1}) [B] (grad); rows: barrier, CLF
 *   PVTOL     a = x_t [B,7], b, c, e = x_{t+1}, x_{t+2}, x_{t+3} [B,7] (grad; columns 0 = x, 1 = y, 6 = safety operator's x);
 *             rows: n_hazards third-order CBF rows, operator (two sides), y_max, y_min, CLF  (P sac_cbf_clf.py:556-697);
 *             four states -> nlbac_constraints4_fwd/bwd
 * CLF row = (Vnext - V) * lya_inv_dt + gamma_l * V;  V [B] detached, Vnext [B] (grad).
 * ------------------------------------------------------------------------------------------------ */
#define NLBAC_CONSTRAINT_UNICYCLE 0
#define NLBAC_CONSTRAINT_CARS     1
#define NLBAC_CONSTRAINT_BARRIER  2
#define NLBAC_CONSTRAINT_PVTOL    3
#define NLBAC_MAX_HAZARDS 15
typedef struct nlbac_constraint_desc {
  int kind;
  int m;                  /* rows: n_hazards+1 (UNICYCLE), 3 (CARS), 2 (BARRIER), n_hazards+5 (PVTOL) */
  int n_hazards;
  float dt;               /* env.dt (0.02) */
  float gamma_b;
  float gamma_l;
  float lya_inv_dt;       /* 1/dt for U/BU (U :474), 1 for C (:474) and P/BP (P :478) */
  float batch_size;       /* args.batch_size, the divisor the reference uses (U :503-504) */
  float collision_radius; /* U: 1.05 * hazards_radius; C: 4.5 */
  float hazards[2 * NLBAC_MAX_HAZARDS];
  /* PVTOL only (P sac_cbf_clf.py:591, :633-636): 0.9 * env.operator_dist, env.y_max, env.y_min, delta_y (10) */
  float operator_dist, y_max, y_min, delta_y;
} nlbac_constraint_desc;

int nlbac_constraints_fwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                          const float* V, const float* Vnext, int64_t B, float* req, void* stream);
int nlbac_constraints_bwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                          const float* V, const float* Vnext, int64_t B, const float* grad_req, float* grad_b,
                          float* grad_c, float* grad_Vnext, void* stream);

/* Same with a fourth state tensor e = x_{t+3} (and its gradient); e / grad_e may be NULL for the kinds that do not use
 * them.  nlbac_constraints_fwd/bwd are these with e = NULL. */
int nlbac_constraints4_fwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                           const float* e, const float* V, const float* Vnext, int64_t B, float* req, void* stream);
int nlbac_constraints4_bwd(const nlbac_constraint_desc* desc, const float* a, const float* b, const float* c,
                           const float* e, const float* V, const float* Vnext, int64_t B, const float* grad_req,
                           float* grad_b, float* grad_c, float* grad_e, float* grad_Vnext, void* stream);

/* lookahead point p(x) = [x + l_p cos(theta), y + l_p sin(theta)] of a [B, n_s] state (theta = column 2)
 * (U sac_cbf_clf.py:429-437, 455-469). */
int nlbac_lookahead_fwd(const float* state, int n_s, float l_p, int64_t B, float* p, void* stream);
int nlbac_lookahead_bwd(const float* state, int n_s, float l_p, int64_t B, const float* grad_p, float* grad_state,
                        void* stream);

/* DynamicsModel.get_state / get_obs (U dynamics.py:27-69; C :25-96; P :25-153; BU :73-198).
 *   UNICYCLE obs[B,7]  <-> state[B,3];  CARS obs[B,10] <-> state[B,10];  PVTOL obs[B,11] <-> state[B,7]
 * get_state does its arithmetic in float64 like the reference's numpy round trip; state_dyn (PVTOL, [B,6]) may be NULL. */
#define NLBAC_ENV_UNICYCLE 0
#define NLBAC_ENV_CARS     1
#define NLBAC_ENV_PVTOL    2
int nlbac_get_state(int env, const float* obs, int64_t B, float* state, float* state_dyn, void* stream);
int nlbac_get_obs_fwd(int env, const float* state, int64_t B, float* obs, void* stream);
int nlbac_get_obs_bwd(int env, const float* state, int64_t B, const float* grad_obs, float* grad_state, void* stream);

/* torch.optim.Adam.step (defaults, as used at U model.py:257 / sac_cbf_clf.py:249-290) on one flat tensor;
 * `step` is the 1-based step count.  utils.soft_update (U utils.py:75-79): target = target*(1-tau) + source*tau. */
int nlbac_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, int step, float lr,
                    float beta1, float beta2, float eps, void* stream);
int nlbac_soft_update(float* target, const float* source, int64_t n, float tau, void* stream);

/* Multi-tensor forms: ONE launch for all parameter tensors of an optimiser (same step count and hyper-parameters,
 * as torch.optim.Adam keeps them per group) or of a target network (U sac_cbf_clf.py:249-290, 315-317).  The
 * pointer arrays are HOST arrays of device pointers, n_tensors <= NLBAC_MAX_TENSORS. */
#define NLBAC_MAX_TENSORS 32
int nlbac_adam_step_multi(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                          float* const* exp_avg_sq, const int64_t* numels, int step, float lr, float beta1,
                          float beta2, float eps, void* stream);
/* Same update with the step counter kept ON THE DEVICE: step_dev = DEVICE int32[2], zero-initialised by the caller --
 * [0] the step count, [1] a ticket word the kernel uses and leaves at zero.  With `bump` != 0 the launch works with
 * step_dev[0] + 1 and writes it back from its last block (no separate increment launch); with bump = 0 (further chunks
 * of the same optimiser step) it reads step_dev[0].  The kernel forms the bias corrections itself: nothing depends on
 * host state, so the call can be captured in a CUDA graph and replayed. */
int nlbac_adam_step_multi_dev(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avg,
                              float* const* exp_avg_sq, const int64_t* numels, int* step_dev, int bump, float lr,
                              float beta1, float beta2, float eps, void* stream);
int nlbac_soft_update_multi(int n_tensors, float* const* targets, const float* const* sources, const int64_t* numels,
                            float tau, void* stream);

/* Scalar glue of update_parameters as single launches, float32, same operation order as the reference:
 *   nlbac_td_target        out = reward + mask * gamma * (min(q1, q2) - alpha * logp)   (U sac_cbf_clf.py:234-236; q2 / logp
 *                          NULL: the Lyapunov / barrier target reward + mask * gamma * q1, :237-238); alpha: DEVICE [1]
 *   nlbac_mse_mean_*       F.mse_loss(pred, target) (mean over n elements) and its gradient w.r.t. pred (:244-246)
 *   nlbac_sac_actor_term_* mean(alpha * logp - min(q1, q2)) (:270) and its gradients (torch.min semantics: the smaller
 *                          input takes the gradient, a tie splits it).  The mean kernels are one 256-thread CTA: meant
 *                          for the reference's B = 128 / 256 minibatches (any n is correct). */
int nlbac_td_target(const float* q1, const float* q2, const float* logp, const float* alpha, const float* reward,
                    const float* mask, float gamma, int64_t B, float* out, void* stream);
int nlbac_mse_mean_fwd(const float* pred, const float* target, int64_t n, float* loss, void* stream);
int nlbac_mse_mean_bwd(const float* pred, const float* target, int64_t n, const float* grad_loss, float* grad_pred,
                       void* stream);
int nlbac_sac_actor_term_fwd(const float* logp, const float* q1, const float* q2, const float* alpha, int64_t B,
                             float* loss, void* stream);
int nlbac_sac_actor_term_bwd(const float* q1, const float* q2, const float* alpha, int64_t B, const float* grad_loss,
                             float* grad_logp, float* grad_q1, float* grad_q2, void* stream);

/* The tail of get_policy_loss_2 / backup_get_policy_loss_2 (U sac_cbf_clf.py:506-530, :611-640) on DEVICE-resident
 * bookkeeping: multiplier update when due (updates != NULL: the device-resident update counter,
 * due = *updates % interval == 0; else *due > 0) (lambda_i <- clamp(lambda_i + rho req_i, 0.01,
 * lambda_max)), rho <- min(rho * aug_ratio, 200), the CBF/CLF ratio |mean(c[:-1])| / |c[-1]| (clf_last: the last row is
 * the CLF row; use_ratio / has_floor + ratio_floor as the variants set them) and
 * loss = sum_i lambda_i w_i c_i + rho / 2 w_i^2 c_i^2, c = req - cost_limit, in ONE single-thread launch; float64
 * scalars like the reference's Python floats.  req [m] float (m <= 32), lambda [m] / rho / due double, loss [1], coef
 * [2m] (saved for the backward: d loss / d req_i = coef_i + 2 coef_{m+i} c_i; the ratio is a constant, as in the
 * reference's float(ratio)). */
int nlbac_lagrangian_fwd(const float* req, int m, int clf_last, double* lambda, double* rho, const double* due,
                         const int64_t* updates, int interval, double cost_limit, double lambda_max, double aug_ratio, int use_ratio, int has_floor,
                         double ratio_floor, float* loss, float* coef, void* stream);
int nlbac_lagrangian_bwd(const float* req, const float* coef, int m, double cost_limit, const float* grad_loss,
                         float* grad_req, void* stream);

/* Device-resident replay memory (U replay_memory.py:4-28, U sac_cbf_clf.py:201-229): transitions live
 * as float32 rows [capacity, width] on the device; a minibatch is ONE gather launch, out[i] = ring[idx[i]].
 * idx: DEVICE array of B row indices (int64), chosen by the host with the reference's own `random.sample`. */
int nlbac_replay_gather(const float* ring, int width, const int64_t* idx, int64_t B, float* out, void* stream);
/* The same gather with every field (columns [offsets[f], offsets[f] + widths[f]) of a row, ascending) written to its own
 * contiguous [B, widths[f]] buffer outs[f] (NULL: field skipped) -- e.g. straight into the static input buffers of a
 * captured update, instead of one strided copy per field.  offsets / widths / outs: HOST arrays, n_fields <=
 * NLBAC_REPLAY_MAX_FIELDS. */
#define NLBAC_REPLAY_MAX_FIELDS 16
int nlbac_replay_gather_split(const float* ring, int width, const int64_t* idx, int64_t B, int n_fields,
                              const int* offsets, const int* widths, float* const* outs, void* stream);

/* Vectorised environment stepping: E independent copies of the reference's UnicycleEnv advanced in
 * one launch, float64 like the reference's numpy state.  Replaces, per environment, UnicycleEnv.reset
 * (U envs/unicycle_env.py:166-184) and UnicycleEnv.step / _step / get_obs / obs_compass / goal_met (U :52-165,
 * :263-279, :302-321).  Persistent per-environment state, owned by the caller: state [E,3], last_goal_dist [E],
 * episode_step [E] (int32).  mask ([E] bytes, NULL = all) selects the environments to reset.  step outputs: obs [E,7],
 * reward / constraint [E], center / next_center [E,2] (the Lyapunov inputs before / after the step), done / goal_met
 * [E] bytes, and the reference's info dict as arrays: num_safety_violation [E] (int32, 0 = key absent), safety_cost [E]. */
int nlbac_env_unicycle_reset(double* state, double* last_goal_dist, int32_t* episode_step, const uint8_t* mask,
                             int64_t E, double* obs, void* stream);
int nlbac_env_unicycle_step(double* state, double* last_goal_dist, int32_t* episode_step, const double* action,
                            int64_t E, double* obs, double* reward, double* constraint, double* center,
                            double* next_center, uint8_t* done, uint8_t* goal_met, int32_t* num_safety_violation,
                            double* safety_cost, void* stream);

/* The same for SimulatedCarsEnv (C envs/simulated_cars_env.py: reset :158-177, step :42-146, _get_reward :148-155,
 * _get_obs :197-213).  Per-environment state: state [E,10] = [x_1 v_1 ... x_5 v_5], t [E], episode_step [E].  reset draws
 * ONE initial-velocity offset per environment (the reference: 3.0 + np.random.normal(0, 0.5), shared by the cars);
 * the caller supplies it as vel_noise [E] so that runs are reproducible.  action [E] (car 4's acceleration).  step
 * outputs: obs [E,10], reward / constraint [E], previous / next positions [E,4] (cars 3 and 4: x, v), done [E],
 * info: reached [E] bytes, num_safety_violation [E], safety_cost [E] (goal_met is always False in the reference). */
int nlbac_env_cars_reset(double* state, double* t, int32_t* episode_step, const double* vel_noise, const uint8_t* mask,
                         int64_t E, double* obs, void* stream);
int nlbac_env_cars_step(double* state, double* t, int32_t* episode_step, const double* action, int64_t E, double* obs,
                        double* reward, double* constraint, double* previous_positions, double* next_positions,
                        uint8_t* done, uint8_t* reached, int32_t* num_safety_violation, double* safety_cost,
                        void* stream);

/* The same for PvtolEnv (P envs/pvtol_env.py: reset :226-246, step / _step :66-216, get_obs :361-377, obs_compass
 * :379-405).  state [E,7] = [x y theta vx vy thrust operator_x], episode_step [E], action [E,2] = (thrust rate, theta
 * rate).  step outputs: obs [E,11] (the reference returns it twice), reward / constraint [E], lya_pre_term [E,11] (the
 * observation BEFORE the step), done / goal_met [E] bytes, and the info dict as [E,4] arrays in the order (obstacles,
 * safety operator, y_max, y_min): num_safety_violation (int32) and safety_cost. */
int nlbac_env_pvtol_reset(double* state, int32_t* episode_step, const uint8_t* mask, int64_t E, double* obs,
                          void* stream);
int nlbac_env_pvtol_step(double* state, int32_t* episode_step, const double* action, int64_t E, double* obs,
                         double* reward, double* constraint, double* lya_pre_term, uint8_t* done, uint8_t* goal_met,
                         int32_t* num_safety_violation, double* safety_cost, void* stream);

/* Data-parallel plumbing (the reference's only analogue is the unused mpi4py scaffold,
 * utils/mpi_pytorch.py:20-34).  One NCCL communicator per process (one process per GPU); nlbac_allreduce_grads is ONE
 * in-place fp32 SUM all-reduce of a flat device buffer (the per-optimiser gradient buffer the backward kernels write, or
 * the [m] constraint sums), enqueued on `stream` with no host synchronisation.  nlbac_comm_unique_id fills 128 bytes on
 * rank 0 that the caller broadcasts to the other ranks (any transport) before nlbac_comm_init, with the rank's GPU
 * current.  Returns 0, < 0 for bad arguments, 1000 + ncclResult_t for NCCL errors (text via nlbac_last_error).  NCCL is
 * resolved at run time (dlopen of libnccl.so.2); the handle is the only state the library keeps. */
int nlbac_comm_unique_id(void* id128);
int nlbac_comm_init(int world, int rank, const void* id128, nlbac_comm** comm);
int nlbac_allreduce_grads(nlbac_comm* comm, float* flat, size_t n, void* stream);
int nlbac_allreduce_sum_f64(nlbac_comm* comm, double* buf, size_t n, void* stream);
int nlbac_comm_destroy(nlbac_comm* comm);

/* Self-test of the tcgen05 operand layouts the tensor-core NODE kernels use (csrc/umma.cuh): one 128x112 tile.
 * mode 0 (kind::tf32, both operands K-major):   D[t][j] = sum_k X[t][k] Y[j][k],  X [128,104], Y [104,104];
 * mode 1 (kind::f16 on bf16 images, MN-major):  D[j][k] = sum_t X[t][j] Y[t][k],  X, Y [128,104].
 * mode 2 (kind::tf32, A operand in tensor memory written by tcgen05.st): same product as mode 0.
 * D [128,112] fp32 (rows/columns >= 104 are padding). */
int nlbac_umma_selftest(int mode, const float* X, const float* Y, float* D, void* stream);

/* Diagnostic descriptor sweep (tools/umma_sweep.py): params10 = {a_mn, b_mn, lboA, sboA, stepA, lboB, sboB, stepB,
 * nsteps, operand_kind (0 fp32 row-owner images, 1 fp32 weight image for B, 2 bf16 row-owner images)}. */
int nlbac_umma_probe(const int* params10, const float* X, const float* Y, float* D, void* stream);

/* Opt-in per-kernel timing of the NODE path for bench.py's roofline: while enabled, every launch of the solve
 * kernels is bracketed by CUDA events on the launching stream.  nlbac_kernel_timing_read waits for them and returns
 * summed milliseconds and launch counts per slot {0 forward solve, 1 reverse sweep, 2 dW GEMM, 3 gradient fold},
 * then clears the records.  Not thread-safe; intended for measurement only. */
int nlbac_kernel_timing(int enable);
int nlbac_kernel_timing_read(double* ms4, int* launches4);

/* Measured-peak helper for the roofline: register-resident FP32 FMA loop.  Returns the number of
 * FLOPs one launch executes in *flops; the caller times it with CUDA events. */
int nlbac_fp32_peak_probe(int iters, float* sink, double* flops, void* stream);

/* Measured tensor-pipe peak: one CTA per SM, ONE thread issues tcgen05.mma back to back over operands
 * resident in shared / tensor memory (no registers, no HBM), 13 k-steps per accumulator like a hidden layer.
 * mode 0 kind::tf32 SS M=128 N=256; 1 kind::tf32 TS (A in tensor memory) N=256; 2 kind::tf32 SS N=112 (the NODE kernels'
 * own tile); 3 kind::tf32 TS N=112; 4 kind::f16 on bf16 SS N=256 (cross-check against cuBLAS).  *flops = FLOPs of one
 * launch (2 M N K per MMA); the caller times it with CUDA events (burst: one short launch; sustained: seconds). */
int nlbac_umma_peak_probe(int mode, int iters, double* flops, void* stream);

/* Same, for register-fed outer-product loops shaped like the GEMM micro-kernel: variant 0 scalar FFMA with
 * three distinct register operands, variant 1 packed fma.rn.f32x2 (sm_100 FFMA2). */
int nlbac_fma_shape_probe(int variant, int iters, float* sink, double* flops, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NLBAC_H_ */
