// C ABI of the fused actor / critic / Lyapunov / barrier MLPs (include/nlbac.h: nlbac_mlp_*, nlbac_policy_head_*).
#include "../../include/nlbac.h"
#include "mlp_kernels.cuh"

namespace nlbac {

// tanh-squashed Gaussian head of GaussianPolicy.sample (U sac_cbf_clf/model.py:110-128), eps injected.
struct HeadArgs {
  const float* ml;      // [B, 2*n_u]  (mean | raw log_std) = last-layer output of the policy MLP
  const float* eps;     // [B, n_u]
  const float* scale;   // [n_u] action_scale
  const float* bias;    // [n_u] action_bias
  float* action;        // [B, n_u]
  float* log_prob;      // [B, 1]
  float* mean_out;      // [B, n_u]  tanh(mean) * scale + bias
  long long B;
  int n_u;
};
constexpr float LOG_SIG_MAX_F = 2.f, LOG_SIG_MIN_F = -20.f, HEAD_EPS_F = 1e-6f;
constexpr float HALF_LOG_2PI = 0.91893853320467274178f;

__global__ void policy_head_fwd_kernel(HeadArgs a) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  const int nu = a.n_u;
  float lp = 0.f;
  for (int i = 0; i < nu; ++i) {
    const float m = a.ml[b * 2 * nu + i];
    const float ls = fminf(fmaxf(a.ml[b * 2 * nu + nu + i], LOG_SIG_MIN_F), LOG_SIG_MAX_F);
    const float s = expf(ls);
    const float x = fmaf(s, a.eps[b * nu + i], m);
    const float y = tanhf(x);
    const float sc = a.scale[i], bi = a.bias[i];
    a.action[b * nu + i] = fmaf(y, sc, bi);
    a.mean_out[b * nu + i] = fmaf(tanhf(m), sc, bi);
    const float d = x - m;
    lp += -(d * d) / (2.f * s * s) - ls - HALF_LOG_2PI - logf(sc * (1.f - y * y) + HEAD_EPS_F);
  }
  a.log_prob[b] = lp;
}

struct HeadBwdArgs {
  const float *ml, *eps, *scale;
  const float* g_action;   // [B, n_u] or null
  const float* g_logp;     // [B, 1] or null
  const float* g_mean;     // [B, n_u] or null
  float* g_ml;             // [B, 2*n_u]
  long long B;
  int n_u;
};
__global__ void policy_head_bwd_kernel(HeadBwdArgs a) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  const int nu = a.n_u;
  const float gl = a.g_logp ? a.g_logp[b] : 0.f;
  for (int i = 0; i < nu; ++i) {
    const float m = a.ml[b * 2 * nu + i];
    const float lr = a.ml[b * 2 * nu + nu + i];
    const float ls = fminf(fmaxf(lr, LOG_SIG_MIN_F), LOG_SIG_MAX_F);
    const float s = expf(ls), e = a.eps[b * nu + i];
    const float x = fmaf(s, e, m);
    const float y = tanhf(x);
    const float sc = a.scale[i];
    const float one_m = 1.f - y * y;
    const float q = sc * one_m + HEAD_EPS_F;
    const float ga = a.g_action ? a.g_action[b * nu + i] : 0.f;
    const float gx = ga * sc * one_m + gl * (2.f * y * sc * one_m / q);
    float gm = gx;
    if (a.g_mean) {
      const float tm = tanhf(m);
      gm += a.g_mean[b * nu + i] * sc * (1.f - tm * tm);
    }
    const bool pass = lr >= LOG_SIG_MIN_F && lr <= LOG_SIG_MAX_F;
    a.g_ml[b * 2 * nu + i] = gm;
    a.g_ml[b * 2 * nu + nu + i] = pass ? (gx * e * s - gl) : 0.f;
  }
}

static int check_mlp(const nlbac_mlp_desc* d) {
  NLBAC_REQUIRE(d != nullptr, "null MLP descriptor");
  NLBAC_REQUIRE(d->n_in >= 1 && d->n_in <= MLP_MAX_IN, "MLP n_in=%d outside [1,%d]", d->n_in, MLP_MAX_IN);
  NLBAC_REQUIRE(d->n_out >= 1 && d->n_out <= MLP_MAX_OUT, "MLP n_out=%d outside [1,%d]", d->n_out, MLP_MAX_OUT);
  NLBAC_REQUIRE(d->hidden >= 32 && d->hidden % 32 == 0 && d->hidden <= 1024, "MLP hidden=%d must be a multiple of 32 in [32,1024]",
                d->hidden);
  return 0;
}

}  // namespace nlbac

using namespace nlbac;

extern "C" {

int nlbac_mlp_workspace_bytes(const nlbac_mlp_desc* desc, int64_t B, size_t* act_bytes, size_t* bwd_bytes) {
  if (int rc = check_mlp(desc)) return rc;
  NLBAC_REQUIRE(B >= 0, "negative batch");
  if (act_bytes) *act_bytes = (size_t)2 * B * desc->hidden * 4;
  if (bwd_bytes) *bwd_bytes = (size_t)2 * B * desc->hidden * 4;
  return 0;
}

#ifndef NLBAC_MLP_STAGE
#define NLBAC_MLP_STAGE 1         // layer 2 of the forward through the transposing shared-memory stage
#endif
#ifndef NLBAC_MLP_STAGE_LARGE
#define NLBAC_MLP_STAGE_LARGE 1   // the staged layer 2 for the 8-row (large-batch) tiles too (B = 65536: 748 -> 513 us)
#endif

int nlbac_mlp_fwd(const nlbac_mlp_desc* desc, const float* const* params, const float* x, int64_t B, float* out,
                  float* act, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:mlp_fwd");
  if (int rc = check_mlp(desc)) return rc;
  NLBAC_REQUIRE(B >= 0, "negative batch");
  if (B == 0) return 0;
  NLBAC_REQUIRE(params && x && out, "null pointer");
  MlpArgs a;
  a.W1 = params[0]; a.b1 = params[1]; a.W2 = params[2]; a.b2 = params[3]; a.W3 = params[4]; a.b3 = params[5];
  a.x = x; a.out = out; a.act = act; a.B = B; a.n_in = desc->n_in; a.H = desc->hidden; a.n_out = desc->n_out;
  // row tile: 8 rows per CTA once that fills the SMs, 2 rows per CTA below (B = 128 -> 64 CTAs instead of 16)
  auto launch = [&]<int TM>() -> int {
    size_t smem = (size_t)(TM * MLP_MAX_IN + 2 * TM * a.H) * 4;
    a.stage = (NLBAC_MLP_STAGE && a.H <= 2 * MLP_NT && a.H % MLP_KC == 0 && reinterpret_cast<uintptr_t>(a.W2) % 16 == 0 && (TM == MLP_TM_SMALL || NLBAC_MLP_STAGE_LARGE) &&
               smem + mlp_stage_floats(a.H) * 4 <= 200 * 1024) ? 1 : 0;
    if (a.stage) smem += mlp_stage_floats(a.H) * 4;
    NLBAC_CUDA(cudaFuncSetAttribute(mlp_fwd_kernel<TM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mlp_fwd_kernel<TM><<<(unsigned)((B + TM - 1) / TM), MLP_NT, smem, (cudaStream_t)stream>>>(a);
    NLBAC_CUDA(cudaGetLastError());
    return 0;
  };
  if (int rc = (B < MLP_SMALL_B) ? launch.template operator()<MLP_TM_SMALL>() : launch.template operator()<MLP_TM_LARGE>())
    return rc;
  return 0;
}

int nlbac_mlp_bwd(const nlbac_mlp_desc* desc, const float* const* params, const float* x, int64_t B, const float* act,
                  const float* grad_out, float* grad_x, float* const* grad_params, void* ws, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:mlp_bwd");
  if (int rc = check_mlp(desc)) return rc;
  NLBAC_REQUIRE(B >= 0, "negative batch");
  NLBAC_REQUIRE(B == 0 || (params && x && act && grad_out && ws), "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const int H = desc->hidden, NI = desc->n_in, NO = desc->n_out;
  if (B == 0) {
    if (grad_params) {
      const size_t n[6] = {(size_t)H * NI, (size_t)H, (size_t)H * H, (size_t)H, (size_t)NO * H, (size_t)NO};
      for (int i = 0; i < 6; ++i) NLBAC_CUDA(cudaMemsetAsync(grad_params[i], 0, n[i] * 4, st));
    }
    return 0;
  }
  MlpBwdArgs a;
  a.W1 = params[0]; a.W2 = params[2]; a.W3 = params[4];
  a.act = act; a.grad_out = grad_out; a.grad_x = grad_x; a.B = B; a.n_in = NI; a.H = H; a.n_out = NO;
  a.d1 = static_cast<float*>(ws);
  a.d2 = a.d1 + (size_t)B * H;
  auto launch = [&]<int TM>() -> int {
    size_t smem = (size_t)(TM * MLP_MAX_OUT + 2 * TM * H) * 4;
    a.stage = (NLBAC_MLP_STAGE && H <= 2 * MLP_NT && H % MLP_KC == 0 && reinterpret_cast<uintptr_t>(a.W2) % 16 == 0 &&
               (TM == MLP_TM_SMALL || NLBAC_MLP_STAGE_LARGE) && smem + mlp_bwd_stage_floats(H) * 4 <= 200 * 1024) ? 1 : 0;
    if (a.stage) smem += mlp_bwd_stage_floats(H) * 4;
    NLBAC_CUDA(cudaFuncSetAttribute(mlp_bwd_rows_kernel<TM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mlp_bwd_rows_kernel<TM><<<(unsigned)((B + TM - 1) / TM), MLP_NT, smem, st>>>(a);
    NLBAC_CUDA(cudaGetLastError());
    return 0;
  };
  if (int rc = (B < MLP_SMALL_B) ? launch.template operator()<MLP_TM_SMALL>() : launch.template operator()<MLP_TM_LARGE>())
    return rc;
  if (grad_params) {
    const float* h1 = act;
    const float* h2 = act + (size_t)B * H;
    AtbArgs g;
    g.B = B;
    g.p[0] = {a.d1, x, grad_params[0], H, NI, grad_params[1]};        // dW1 = d1^T x,   db1 = colsum d1
    g.p[1] = {a.d2, h1, grad_params[2], H, H, grad_params[3]};        // dW2 = d2^T h1,  db2 = colsum d2
    g.p[2] = {grad_out, h2, grad_params[4], NO, H, grad_params[5]};   // dW3 = g^T h2,   db3 = colsum g
    ColSumArgs c;
    c.B = B;
    c.p[0] = {a.d1, grad_params[1], H};
    c.p[1] = {a.d2, grad_params[3], H};
    c.p[2] = {grad_out, grad_params[5], NO};
    if (B >= MLP_SPLIT_MIN_B) {   // split over the batch: chunks meet in pre-zeroed outputs (mlp_kernels.cuh)
      const size_t sizes[6] = {(size_t)H * NI, (size_t)H, (size_t)H * H, (size_t)H, (size_t)NO * H, (size_t)NO};
      for (int i = 0; i < 6; ++i) NLBAC_CUDA(cudaMemsetAsync(grad_params[i], 0, sizes[i] * 4, st));
      int nsplit = (int)(B / 2048);
      if (nsplit > 64) nsplit = 64;
      if (nsplit < 1) nsplit = 1;
      AtbSplitArgs gs;
      gs.B = B; gs.nsplit = nsplit;
      for (int i = 0; i < 3; ++i) gs.p[i] = g.p[i];
      mlp_atb_split_kernel<<<dim3((H + 63) / 64, (H + 63) / 64, 3 * nsplit), 256, 0, st>>>(gs);
      NLBAC_CUDA(cudaGetLastError());
      ColSumSplitArgs cs;
      cs.B = B; cs.nsplit = nsplit * 2;
      for (int i = 0; i < 3; ++i) cs.p[i] = c.p[i];
      mlp_colsum_split_kernel<<<dim3((H + 31) / 32, 3, cs.nsplit), 256, 0, st>>>(cs);
      NLBAC_CUDA(cudaGetLastError());
      return 0;
    }
    mlp_atb_kernel<<<dim3((H + 31) / 32, (H + 31) / 32, 3), 256, 0, st>>>(g);   // weight AND bias gradients
    NLBAC_CUDA(cudaGetLastError());
  }
  return 0;
}

int nlbac_policy_head_fwd(const float* mean_logstd, const float* eps, const float* action_scale,
                          const float* action_bias, int64_t B, int n_u, float* action, float* log_prob,
                          float* mean_out, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:policy_head_fwd");
  NLBAC_REQUIRE(mean_logstd && eps && action_scale && action_bias && action && log_prob && mean_out, "null pointer");
  NLBAC_REQUIRE(B >= 0 && n_u >= 1 && 2 * n_u <= MLP_MAX_OUT, "bad sizes (B=%lld n_u=%d)", (long long)B, n_u);
  if (B == 0) return 0;
  HeadArgs a{mean_logstd, eps, action_scale, action_bias, action, log_prob, mean_out, B, n_u};
  policy_head_fwd_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(a);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_policy_head_bwd(const float* mean_logstd, const float* eps, const float* action_scale, int64_t B, int n_u,
                          const float* grad_action, const float* grad_log_prob, const float* grad_mean_out,
                          float* grad_mean_logstd, void* stream) {
  ::nlbac::NvtxRange nvtx_range_("nlbac:policy_head_bwd");
  NLBAC_REQUIRE(mean_logstd && eps && action_scale && grad_mean_logstd, "null pointer");
  NLBAC_REQUIRE(B >= 0 && n_u >= 1 && 2 * n_u <= MLP_MAX_OUT, "bad sizes (B=%lld n_u=%d)", (long long)B, n_u);
  if (B == 0) return 0;
  HeadBwdArgs a{mean_logstd, eps, action_scale, grad_action, grad_log_prob, grad_mean_out, grad_mean_logstd, B, n_u};
  policy_head_bwd_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(a);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
