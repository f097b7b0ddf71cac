// C ABI of the NODE solve (include/nlbac.h): descriptor dispatch, schedule tables, launches.
#include <cstdarg>
#include <cstring>
#include <vector>

#include "../../include/nlbac.h"
#include "node_bwd.cuh"
#include "node_pack.cuh"

namespace nlbac {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return (int)e;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

template <class C>
static bool matches(const nlbac_node_desc* d) {
  return d->kind == C::KIND && d->n_s == C::NS && d->n_u == C::NU && d->n_c == C::NC && d->hidden == C::H &&
         d->n_f_layers == C::NF && d->n_g_layers == C::NG;
}

// invoke f.template operator()<Cfg>() for the compiled instance matching the descriptor
template <class F>
static int dispatch(const nlbac_node_desc* d, F&& f) {
  if (d == nullptr) { set_error("null descriptor"); return -1; }
  if (matches<CfgU>(d)) return f.template operator()<CfgU>();
  if (matches<CfgP>(d)) return f.template operator()<CfgP>();
  if (matches<CfgC>(d)) return f.template operator()<CfgC>();
  if (matches<CfgQ>(d)) return f.template operator()<CfgQ>();
  set_error("no compiled kernel instance for NODE shape kind=%d n_s=%d n_u=%d n_c=%d hidden=%d layers=%d/%d",
            d->kind, d->n_s, d->n_u, d->n_c, d->hidden, d->n_f_layers, d->n_g_layers);
  return -2;
}

// ---- schedule: what FixedGridODESolver.integrate does with (t, grid), as tables -----------------
struct Schedule {
  std::vector<float> dts;
  std::vector<int> out_step;
  std::vector<float> out_w;
};
static int build_schedule(const float* t, int n_t, const float* grid, int n_grid, Schedule* s) {
  NLBAC_REQUIRE(n_t >= 1 && n_grid >= 2, "need at least one output time and one step (n_t=%d n_grid=%d)", n_t, n_grid);
  NLBAC_REQUIRE(grid[0] == t[0], "grid must start at t[0]");
  const int n_steps = n_grid - 1;
  s->dts.resize(n_steps);
  s->out_step.assign(n_t, -1);
  s->out_w.assign(n_t, 0.f);
  int j = 1;
  for (int i = 0; i < n_steps; ++i) {
    const float t0 = grid[i], t1 = grid[i + 1];
    NLBAC_REQUIRE(t1 > t0, "integration grid must be strictly increasing");
    s->dts[i] = t1 - t0;
    while (j < n_t && t1 >= t[j]) {
      s->out_step[j] = i;
      s->out_w[j] = (t[j] == t0) ? 0.f : (t[j] == t1 ? 1.f : (t[j] - t0) / (t1 - t0));
      ++j;
    }
  }
  NLBAC_REQUIRE(j == n_t, "output times extend past the integration grid");
  return 0;
}
static size_t sched_bytes(int n_steps, int n_t) {
  return (size_t)round_up(n_steps * 4 + n_t * 8, 16);
}
static int upload_schedule(const Schedule& s, void* ws, cudaStream_t st, const float** dts, const int** ostep,
                           const float** ow) {
  const int n_steps = (int)s.dts.size(), n_t = (int)s.out_step.size();
  char* p = static_cast<char*>(ws);
  NLBAC_CUDA(cudaMemcpyAsync(p, s.dts.data(), n_steps * 4, cudaMemcpyHostToDevice, st));
  NLBAC_CUDA(cudaMemcpyAsync(p + n_steps * 4, s.out_step.data(), n_t * 4, cudaMemcpyHostToDevice, st));
  NLBAC_CUDA(cudaMemcpyAsync(p + n_steps * 4 + n_t * 4, s.out_w.data(), n_t * 4, cudaMemcpyHostToDevice, st));
  *dts = reinterpret_cast<const float*>(p);
  *ostep = reinterpret_cast<const int*>(p + n_steps * 4);
  *ow = reinterpret_cast<const float*>(p + n_steps * 4 + n_t * 4);
  return 0;
}

template <class C, int T>
static int launch_fwd(const FwdArgs& a, cudaStream_t st) {
  using G = FwdGeom<C, T>;
  static bool attr_done = false;
  if (!attr_done) {
    NLBAC_CUDA(cudaFuncSetAttribute(node_fwd_kernel<C, T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)G::SMEM_BYTES));
    attr_done = true;
  }
  const long long n_tiles = (a.B + T - 1) / T;
  const int grid = (int)(n_tiles < sm_count() ? n_tiles : sm_count());
  node_fwd_kernel<C, T><<<grid, G::NT, G::SMEM_BYTES, st>>>(a);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

template <class C, int TB>
static int launch_bwd(const BwdArgs& a, cudaStream_t st) {
  using G = BwdGeom<C, TB>;
  static bool attr_done = false;
  if (!attr_done) {
    NLBAC_CUDA(cudaFuncSetAttribute(node_bwd_kernel<C, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)G::SMEM_BYTES));
    attr_done = true;
  }
  const long long n_tiles = (a.B + TB - 1) / TB;
  const int grid = (int)(n_tiles < sm_count() ? n_tiles : sm_count());
  node_bwd_kernel<C, TB><<<grid, G::NT, G::SMEM_BYTES, st>>>(a);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

// register-resident FMA loop: 8 independent chains per thread
__global__ void fp32_peak_kernel(int iters, float* sink) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f,
        a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float m = 0.999f, c = 1e-3f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
      a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
  }
  const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == 123.456f) sink[0] = s;
}

// Register-only outer-product loops shaped like the GEMM micro-kernel (acc[i][j] += a[i] * w[j], all three
// operands distinct registers): variant 0 = scalar FFMA, variant 1 = packed fma.rn.f32x2 (FFMA2) with the
// a operand duplicated into a 64-bit pair.  They measure what the FMA pipe sustains when it is fed from the
// register file rather than from an immediate (the plain peak probe).
template <int VARIANT>
__global__ void __launch_bounds__(256, 1) fma_shape_kernel(int iters, float* sink) {
  constexpr int TT = 8, NN = 12;
  float a[TT], w[NN], acc[TT][NN];
#pragma unroll
  for (int i = 0; i < TT; ++i) a[i] = 1e-3f * (threadIdx.x + i);
#pragma unroll
  for (int j = 0; j < NN; ++j) w[j] = 1e-3f * (blockIdx.x + j);
#pragma unroll
  for (int i = 0; i < TT; ++i)
#pragma unroll
    for (int j = 0; j < NN; ++j) acc[i][j] = 0.f;
  for (int it = 0; it < iters; ++it) {
    if (VARIANT == 0) {
#pragma unroll
      for (int i = 0; i < TT; ++i)
#pragma unroll
        for (int j = 0; j < NN; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    } else {
#pragma unroll
      for (int i = 0; i < TT; ++i) {
        unsigned long long aa;
        asm volatile("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a[i]));
#pragma unroll
        for (int j = 0; j < NN; j += 2) {
          unsigned long long ww, cc;
          asm volatile("mov.b64 %0, {%1, %2};" : "=l"(ww) : "f"(w[j]), "f"(w[j + 1]));
          asm volatile("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(acc[i][j]), "f"(acc[i][j + 1]));
          asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(cc) : "l"(aa), "l"(ww));
          asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(acc[i][j]), "=f"(acc[i][j + 1]) : "l"(cc));
        }
      }
    }
    // keep a and w live and changing without adding FMA-pipe work worth mentioning
    a[it & 7] += 1e-7f;
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < TT; ++i)
#pragma unroll
    for (int j = 0; j < NN; ++j) s += acc[i][j];
  if (s == 123.456f) sink[0] = s;
}

}  // namespace nlbac

using namespace nlbac;

extern "C" {

const char* nlbac_last_error(void) { return g_err; }
int nlbac_abi_version(void) { return NLBAC_ABI_VERSION; }

int nlbac_node_supported(const nlbac_node_desc* desc, int precision) {
  if (precision != NLBAC_PREC_FP32) { set_error("precision %d not compiled", precision); return -3; }
  return dispatch(desc, [&]<class C>() { return 0; });
}

int64_t nlbac_node_param_count(const nlbac_node_desc* desc) {
  int64_t n = -1;
  dispatch(desc, [&]<class C>() { n = C::NPARAM; return 0; });
  return n;
}

int64_t nlbac_node_packed_floats(const nlbac_node_desc* desc) {
  int64_t n = -1;
  dispatch(desc, [&]<class C>() { n = C::PACKED; return 0; });
  return n;
}

int nlbac_node_pack(const nlbac_node_desc* desc, const float* const* params, float* packed, void* stream) {
  NLBAC_REQUIRE(params && packed, "null pointer");
  return dispatch(desc, [&]<class C>() { return launch_pack<C>(params, packed, (cudaStream_t)stream); });
}

int nlbac_node_workspace_bytes(const nlbac_node_desc* desc, int64_t B, int n_grid, int n_t, int method,
                               size_t* fwd_bytes, size_t* ckpt_bytes, size_t* bwd_bytes) {
  NLBAC_REQUIRE(B >= 0 && n_grid >= 2 && n_t >= 1, "bad sizes");
  NLBAC_REQUIRE(method >= NLBAC_EULER && method <= NLBAC_RK4, "fixed-grid methods only (got %d)", method);
  return dispatch(desc, [&]<class C>() {
    const int n_steps = n_grid - 1;
    const size_t sb = sched_bytes(n_steps, n_t);
    if (fwd_bytes) *fwd_bytes = sb;
    if (ckpt_bytes) *ckpt_bytes = (size_t)n_steps * method_stages(method) * (size_t)B * C::NS * 4;
    if (bwd_bytes) *bwd_bytes = sb + (size_t)NREP * C::NGRAD * 4;
    return 0;
  });
}

int nlbac_node_solve_fwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                         const float* t, int n_t, const float* grid, int n_grid, int method, int precision,
                         float* y_out, float* ckpt, void* ws, void* stream) {
  NLBAC_REQUIRE(packed && y0 && t && grid && y_out && ws, "null pointer");
  NLBAC_REQUIRE(precision == NLBAC_PREC_FP32, "precision %d not compiled", precision);
  NLBAC_REQUIRE(method >= NLBAC_EULER && method <= NLBAC_RK4, "fixed-grid methods only (got %d)", method);
  NLBAC_REQUIRE(B >= 0, "negative batch");
  Schedule s;
  if (int rc = build_schedule(t, n_t, grid, n_grid, &s)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  return dispatch(desc, [&]<class C>() {
    if (B == 0) return 0;
    FwdArgs a;
    a.packed = packed; a.y0 = y0; a.y_out = y_out; a.ckpt = ckpt; a.B = B;
    a.n_steps = n_grid - 1; a.n_t = n_t; a.method = method;
    if (int rc = upload_schedule(s, ws, st, &a.dts, &a.out_step, &a.out_w)) return rc;
    // largest tile that still gives every SM at least one tile
    const long long sms = sm_count();
    if ((B + 255) / 256 >= sms) return launch_fwd<C, 256>(a, st);
    if ((B + 127) / 128 >= sms) return launch_fwd<C, 128>(a, st);
    if ((B + 63) / 64 >= sms) return launch_fwd<C, 64>(a, st);
    return launch_fwd<C, 32>(a, st);
  });
}

int nlbac_node_solve_bwd(const nlbac_node_desc* desc, const float* packed, const float* y0, int64_t B,
                         const float* t, int n_t, const float* grid, int n_grid, int method, int precision,
                         const float* ckpt, const float* grad_out, float* grad_y0, float* grad_params,
                         void* ws, void* stream) {
  NLBAC_REQUIRE(packed && y0 && t && grid && ckpt && grad_out && ws, "null pointer");
  NLBAC_REQUIRE(precision == NLBAC_PREC_FP32, "precision %d not compiled", precision);
  NLBAC_REQUIRE(method >= NLBAC_EULER && method <= NLBAC_RK4, "fixed-grid methods only (got %d)", method);
  NLBAC_REQUIRE(B >= 0, "negative batch");
  Schedule s;
  if (int rc = build_schedule(t, n_t, grid, n_grid, &s)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  return dispatch(desc, [&]<class C>() {
    if (B == 0) {
      if (grad_params) NLBAC_CUDA(cudaMemsetAsync(grad_params, 0, (size_t)C::NPARAM * 4, st));
      return 0;
    }
    BwdArgs a;
    a.packed = packed; a.y0 = y0; a.ckpt = ckpt; a.grad_out = grad_out; a.grad_y0 = grad_y0; a.B = B;
    a.n_steps = n_grid - 1; a.n_t = n_t; a.method = method;
    if (int rc = upload_schedule(s, ws, st, &a.dts, &a.out_step, &a.out_w)) return rc;
    float* rep = nullptr;
    if (grad_params) {
      rep = reinterpret_cast<float*>(static_cast<char*>(ws) + sched_bytes(a.n_steps, n_t));
      NLBAC_CUDA(cudaMemsetAsync(rep, 0, (size_t)NREP * C::NGRAD * 4, st));
    }
    a.grad_rep = rep;
    int rc = ((B + 63) / 64 >= sm_count()) ? launch_bwd<C, 64>(a, st) : launch_bwd<C, 32>(a, st);
    if (rc) return rc;
    if (grad_params) {
      GradTable gt = build_grad_table<C>();
      fold_grad_kernel<<<dim3(8, gt.n), 256, 0, st>>>(gt, rep, C::NGRAD, grad_params);
      NLBAC_CUDA(cudaGetLastError());
    }
    return 0;
  });
}

int nlbac_fma_shape_probe(int variant, int iters, float* sink, double* flops, void* stream) {
  NLBAC_REQUIRE(iters > 0 && sink && (variant == 0 || variant == 1), "bad arguments");
  const int blocks = sm_count() * 2, threads = 256;
  if (variant == 0) fma_shape_kernel<0><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
  else fma_shape_kernel<1><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
  NLBAC_CUDA(cudaGetLastError());
  if (flops) *flops = 2.0 * 8 * 12 * (double)iters * blocks * threads;
  return 0;
}

int nlbac_fp32_peak_probe(int iters, float* sink, double* flops, void* stream) {
  NLBAC_REQUIRE(iters > 0 && sink, "bad arguments");
  const int blocks = sm_count() * 8, threads = 256;
  fp32_peak_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
  NLBAC_CUDA(cudaGetLastError());
  if (flops) *flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
  return 0;
}

}  // extern "C"
