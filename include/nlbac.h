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
 */
#ifndef NLBAC_H_
#define NLBAC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NLBAC_ABI_VERSION 1

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
#define NLBAC_PREC_BF16 1  /* bf16 tcgen05 tensor cores, fp32 accumulate; parity bar rel 1e-2 */

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

/* Measured-peak helper for the roofline: register-resident FP32 FMA loop.  Returns the number of
 * FLOPs one launch executes in *flops; the caller times it with CUDA events. */
int nlbac_fp32_peak_probe(int iters, float* sink, double* flops, void* stream);

/* Same, for register-fed outer-product loops shaped like the GEMM micro-kernel: variant 0 scalar FFMA with
 * three distinct register operands, variant 1 packed fma.rn.f32x2 (sm_100 FFMA2). */
int nlbac_fma_shape_probe(int variant, int iters, float* sink, double* flops, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NLBAC_H_ */
