// dopri5 (adaptive Dormand–Prince 5(4)) with torchdiffeq 0.2.3 semantics:
// one step size for the WHOLE batch, chosen from an RMS norm over every element of [B, D]; FSAL; the
// last step is not clipped — outputs are read from the 4th-order dense-output polynomial.
//
// Forward: the host drives accept/reject (it needs one scalar per attempted step, exactly like
// torchdiffeq's `error_ratio <= 1`), each attempt is ONE launch that evaluates the 6 new stages of
// every trajectory with the same fused right-hand side as the fixed-grid kernel.
// Backward: reverse sweep over the accepted steps' checkpointed stage inputs with the FSAL carry.
#pragma once
#include "node_bwd.cuh"

namespace nlbac {

// Dormand–Prince tableau (rk_common / dopri5.py)
__device__ constexpr float DP_BETA[6][6] = {
    {1.f / 5, 0, 0, 0, 0, 0},
    {3.f / 40, 9.f / 40, 0, 0, 0, 0},
    {44.f / 45, -56.f / 15, 32.f / 9, 0, 0, 0},
    {19372.f / 6561, -25360.f / 2187, 64448.f / 6561, -212.f / 729, 0, 0},
    {9017.f / 3168, -355.f / 33, 46732.f / 5247, 49.f / 176, -5103.f / 18656, 0},
    {35.f / 384, 0, 500.f / 1113, 125.f / 192, -2187.f / 6784, 11.f / 84}};
__device__ constexpr float DP_CERR[7] = {(float)(35.0 / 384 - 1951.0 / 21600), 0.f,
                                         (float)(500.0 / 1113 - 22642.0 / 50085), (float)(125.0 / 192 - 451.0 / 720),
                                         (float)(-2187.0 / 6784 + 12231.0 / 42400), (float)(11.0 / 84 - 649.0 / 6300),
                                         (float)(-1.0 / 60)};
__device__ constexpr float DP_CMID[7] = {(float)(6025192743.0 / 30085553152.0 / 2), 0.f,
                                         (float)(51252292925.0 / 65400821598.0 / 2),
                                         (float)(-2691868925.0 / 45128329728.0 / 2),
                                         (float)(187940372067.0 / 1594534317056.0 / 2),
                                         (float)(-1776094331.0 / 19743644256.0 / 2),
                                         (float)(11237099.0 / 235043384.0 / 2)};

// Controller state of the device-driven solve (nlbac_node_dopri5_fwd_dev): one instance in device memory, followed
// by the arrays t_out[n_t] (double), step_dt[max_steps], out_step[n_t], out_x[n_t], pend_j[n_t], pend_x[n_t].
struct Dp5Dev {
  double dt, t_cur, t_prev, h0, d1;
  float dt_last;
  int cur, n_acc, n_rej, j, done, err, n_pend;
};
struct Dp5DevArrays {
  double* t_out;
  float* step_dt;
  int* out_step;
  float* out_x;
  int* pend_j;
  float* pend_x;
};
__host__ __device__ inline size_t dp5_dev_bytes(int n_t, int max_steps) {
  return 128 + (size_t)n_t * 8 + ((size_t)max_steps + 4 * (size_t)n_t) * 4 + 64;
}
__host__ __device__ inline Dp5DevArrays dp5_dev_arrays(void* state, int n_t, int max_steps) {
  char* p = static_cast<char*>(state) + 128;
  Dp5DevArrays r;
  r.t_out = reinterpret_cast<double*>(p); p += (size_t)n_t * 8;
  r.step_dt = reinterpret_cast<float*>(p); p += (size_t)max_steps * 4;
  r.out_step = reinterpret_cast<int*>(p); p += (size_t)n_t * 4;
  r.out_x = reinterpret_cast<float*>(p); p += (size_t)n_t * 4;
  r.pend_j = reinterpret_cast<int*>(p); p += (size_t)n_t * 4;
  r.pend_x = reinterpret_cast<float*>(p);
  return r;
}
static_assert(sizeof(Dp5Dev) <= 128, "Dp5Dev header");

struct Dp5Args {
  const float* packed;
  const float* y0;     // [B, D] (constants of the flow)
  const float* y;      // [B, NS] state at the start of the step
  const float* k1;     // [B, NS] f(y) (FSAL)
  float* y1;           // [B, NS] candidate end state
  float* k7;           // [B, NS] f(y1)
  float* krec;         // [7, B, NS] stages of this attempt (dense output)
  float* xrec;         // [6, B, NS] stage inputs x2..x7 of this attempt, or null
  double* acc;         // [3] sums of squares (mode 0: d0,d1; mode 1: d2; mode 2: error ratio in acc[0])
  long long B;
  float dt, rtol, atol;
  int mode;            // 0: f0 = f(y0) + norms; 1: initial-step probe; 2: one attempted step
  // device-driven solve: dt, the ping-pong parity and the checkpoint slot come from *dev (y = ybuf[0] = ws base,
  // xrec = the checkpoint BASE); a finished solve makes the launch a no-op
  const Dp5Dev* dev;
};

template <class C, int T>
__global__ void __launch_bounds__(FwdGeom<C, T>::NT, 1) dopri5_kernel(Dp5Args a) {
  using G = FwdGeom<C, T>;
  extern __shared__ __align__(16) float smem[];
  float* sA = smem + G::OFF_A;
  float* sX = smem + G::OFF_X;
  float* sRes = smem + G::OFF_RES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
  __shared__ double sred[3][8];
  const int tid = threadIdx.x;
  const float *yp = a.y, *k1p = a.k1;
  float *y1p = a.y1, *k7p = a.k7, *xrecp = a.xrec;
  float dtv = a.dt;
  if (a.dev != nullptr) {
    if (a.mode == 1) dtv = (float)a.dev->h0;
    if (a.mode == 2) {
      if (a.dev->done) return;
      const size_t n = (size_t)a.B * C::NS;
      const int cur = a.dev->cur;
      float* base = const_cast<float*>(a.y);
      yp = base + (size_t)cur * n; y1p = base + (size_t)(cur ^ 1) * n;
      k1p = base + (size_t)(2 + cur) * n; k7p = base + (size_t)(2 + (cur ^ 1)) * n;
      if (xrecp != nullptr) xrecp += (size_t)a.dev->n_acc * 6 * n;
      dtv = (float)a.dev->dt;
    }
  }
  const long long n_tiles = (a.B + T - 1) / T;
  const int nev = a.mode == 2 ? 6 : 1;
  long long my_tiles = 0;
  if ((long long)blockIdx.x < n_tiles) my_tiles = (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  WeightRing<G::NSLOT, C::NHID, C::HID_FLOATS> ring;
  ring.slots = smem + G::OFF_RING;
  ring.full = bars;
  ring.seq = 0;
  ring.total = (uint32_t)(my_tiles * nev * C::NHID);
  if (tid == 0) {
    for (int i = 0; i <= G::NSLOT; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (my_tiles == 0) return;
  if (tid == 0) {
    mbar_expect_tx(&bars[G::NSLOT], C::RES_FWD * 4);
    tma_load_1d(sRes, a.packed, C::RES_FWD * 4, &bars[G::NSLOT]);
  }
  ring.start(a.packed, [](uint32_t s) { return C::off_hid_t(s % C::NHID); });
  mbar_wait(&bars[G::NSLOT], 0);

  double s0 = 0.0, s1 = 0.0, s2 = 0.0;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long b = tile * T + tid;
    const bool live = tid < T && b < a.B;
    float y[C::NS], cst[C::NCONST], x[C::NS], kn[C::NS];
    float k[7][C::NS];
#pragma unroll
    for (int d = 0; d < C::NS; ++d) { y[d] = 0.f; k[0][d] = 0.f; }
#pragma unroll
    for (int d = 0; d < C::NCONST; ++d) cst[d] = 0.f;
    if (live) {
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) cst[d] = a.y0[b * C::D + C::NS + d];
#pragma unroll
      for (int d = 0; d < C::NS; ++d) y[d] = yp[b * C::NS + d];
      if (a.mode != 0) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) k[0][d] = k1p[b * C::NS + d];
      }
    }
    // one call site of the right-hand side for all three modes
#pragma unroll 1
    for (int i = 1; i <= nev; ++i) {
      if (a.mode == 0) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) x[d] = y[d];
      } else if (a.mode == 1) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) x[d] = y[d] + dtv * k[0][d];
      } else {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) {
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 6; ++j)
            if (j < i) s = fmaf(k[j][d], DP_BETA[i - 1][j] * dtv, s);
          x[d] = y[d] + s;
        }
        if (xrecp != nullptr && live) {
#pragma unroll
          for (int d = 0; d < C::NS; ++d) xrecp[((long long)(i - 1) * a.B + b) * C::NS + d] = x[d];
        }
      }
      eval_rhs_tile<C, T>(a.packed, sA, sX, sRes, ring, x, cst, kn);
#pragma unroll
      for (int d = 0; d < C::NS; ++d) k[i][d] = kn[d];
    }
    if (live) {
      if (a.mode == 0) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) {
          k7p[b * C::NS + d] = kn[d];
          const float sc = a.atol + fabsf(y[d]) * a.rtol;
          const float q0 = y[d] / sc, q1 = kn[d] / sc;
          s0 += (double)q0 * q0;
          s1 += (double)q1 * q1;
        }
#pragma unroll
        for (int d = 0; d < C::NCONST; ++d) {
          const float q0 = cst[d] / (a.atol + fabsf(cst[d]) * a.rtol);
          s0 += (double)q0 * q0;
        }
      } else if (a.mode == 1) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) {
          const float q = (kn[d] - k[0][d]) / (a.atol + fabsf(y[d]) * a.rtol);
          s2 += (double)q * q;
        }
      } else {  // x == y1 (Dormand–Prince: c_sol equals the last beta row), kn == f(y1)
#pragma unroll
        for (int d = 0; d < C::NS; ++d) {
          y1p[b * C::NS + d] = x[d];
          k7p[b * C::NS + d] = kn[d];
          float e = 0.f;
#pragma unroll
          for (int j = 0; j < 7; ++j) e = fmaf(k[j][d], DP_CERR[j] * dtv, e);
          const float tol = a.atol + a.rtol * fmaxf(fabsf(y[d]), fabsf(x[d]));
          const float q = e / tol;
          s0 += (double)q * q;
#pragma unroll
          for (int j = 0; j < 7; ++j) a.krec[((long long)j * a.B + b) * C::NS + d] = k[j][d];
        }
      }
    }
  }
  // block reduction -> one double atomic per sum per CTA
  double v[3] = {s0, s1, s2};
#pragma unroll
  for (int q = 0; q < 3; ++q) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    if ((tid & 31) == 0) sred[q][tid >> 5] = v[q];
  }
  __syncthreads();
  if (tid < 3) {
    double s = 0.0;
    for (int w = 0; w < G::NT / 32; ++w) s += sred[tid][w];
    if (s != 0.0) atomicAdd(a.acc + tid, s);
  }
}

// dense output at fraction xf of the last accepted step (interp._interp_fit/_interp_evaluate)
struct Dp5InterpArgs {
  const float* y0;    // [B, D]
  const float* ya;    // [B, NS] step start
  const float* yb;    // [B, NS] step end
  const float* krec;  // [7, B, NS]
  float* out;         // [B, D]
  long long B;
  float dt, xf;
  int ns, d;
};
__device__ __forceinline__ float dp5_dense(const float* ya, const float* yb, const float* krec, long long o,
                                           long long st, float dt, float x) {
  const float y0 = ya[o], y1 = yb[o], f0 = krec[o], f1 = krec[6 * st + o];
  float ymid = 0.f;
#pragma unroll
  for (int j = 0; j < 7; ++j) ymid = fmaf(krec[j * st + o], DP_CMID[j] * dt, ymid);
  ymid += y0;
  const float ca = 2 * dt * (f1 - f0) - 8 * (y1 + y0) + 16 * ymid;
  const float cb = dt * (5 * f0 - 3 * f1) + 18 * y0 + 14 * y1 - 32 * ymid;
  const float cc = dt * (f1 - 4 * f0) - 11 * y0 - 5 * y1 + 16 * ymid;
  const float cd = dt * f0;
  float tot = y0 + x * cd;
  float xp = x;
  xp *= x; tot += xp * cc;
  xp *= x; tot += xp * cb;
  xp *= x; tot += xp * ca;
  return tot;
}
__global__ void dopri5_interp_kernel(Dp5InterpArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.B * a.d) return;
  const long long b = i / a.d;
  const int c = (int)(i - b * a.d);
  if (c >= a.ns) { a.out[i] = a.y0[i]; return; }
  a.out[i] = dp5_dense(a.ya, a.yb, a.krec, b * a.ns + c, a.B * a.ns, a.dt, a.xf);
}

// Device-driven variant: writes every output time the LAST attempt passed (none after a rejected step).
struct Dp5InterpDevArgs {
  const float* y0;     // [B, D]
  const float* base;   // workspace base: ybuf[0], ybuf[1], kbuf[0], kbuf[1], krec
  float* y_out;        // [n_t, B, D]
  const Dp5Dev* dev;
  const int* pend_j;
  const float* pend_x;
  long long B;
  int ns, d;
};
__global__ void dopri5_interp_dev_kernel(Dp5InterpDevArgs a) {
  const int n_pend = a.dev->n_pend;
  if (n_pend == 0) return;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.B * a.d) return;
  const long long b = i / a.d;
  const int c = (int)(i - b * a.d);
  const size_t n = (size_t)a.B * a.ns;
  const int cur = a.dev->cur;
  const float* ya = a.base + (size_t)(cur ^ 1) * n;
  const float* yb = a.base + (size_t)cur * n;
  const float* krec = a.base + 4 * n;
  const float dt = a.dev->dt_last;
  for (int p = 0; p < n_pend; ++p) {
    float* out = a.y_out + (size_t)a.pend_j[p] * a.B * a.d;
    out[i] = c >= a.ns ? a.y0[i] : dp5_dense(ya, yb, krec, b * a.ns + c, (long long)n, dt, a.pend_x[p]);
  }
}

// The step-size controller as ONE device thread, in the same double arithmetic as the host-driven loop of
// nlbac_node_dopri5_fwd (torchdiffeq rk_common._adaptive_step / misc._select_initial_step / _optimal_step_size).
// phase -1: reset (t_chunk holds up to 16 output times starting at t_off); 0 / 1: the two norms of the initial step;
// 2: accept / reject of one attempt; 3: close (flags an exhausted attempt budget).
struct Dp5CtrlArgs {
  Dp5Dev* s;
  double* acc;
  double cnt;
  double t_chunk[16];
  int phase, n_t, max_steps, t_off, t_cnt;
};
__global__ void dopri5_ctrl_kernel(Dp5CtrlArgs a) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  Dp5Dev* s = a.s;
  Dp5DevArrays A = dp5_dev_arrays(s, a.n_t, a.max_steps);
  if (a.phase == -1) {
    for (int i = 0; i < a.t_cnt; ++i) A.t_out[a.t_off + i] = a.t_chunk[i];
    if (a.t_off == 0) {
      s->dt = 0.0; s->t_cur = s->t_prev = a.t_chunk[0]; s->h0 = 0.0; s->d1 = 0.0; s->dt_last = 0.f;
      s->cur = 0; s->n_acc = 0; s->n_rej = 0; s->j = 1; s->done = 0; s->err = 0; s->n_pend = 0;
      a.acc[0] = a.acc[1] = a.acc[2] = 0.0;
    }
    for (int i = 0; i < a.t_cnt; ++i) { A.out_step[a.t_off + i] = -1; A.out_x[a.t_off + i] = 0.f; }
    return;
  }
  if (a.phase == 0) {
    const double d0 = sqrt(a.acc[0] / a.cnt), d1 = sqrt(a.acc[1] / a.cnt);
    s->h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    s->d1 = d1;
    a.acc[0] = a.acc[1] = a.acc[2] = 0.0;
    return;
  }
  if (a.phase == 1) {
    const double h0 = s->h0, d1 = s->d1, d2 = sqrt(a.acc[2] / a.cnt) / h0;
    const double h1 = (d1 <= 1e-15 && d2 <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / fmax(d1, d2), 1.0 / 5.0);
    s->dt = fmin(100.0 * h0, h1);
    a.acc[0] = a.acc[1] = a.acc[2] = 0.0;
    return;
  }
  if (a.phase == 3) {
    s->n_pend = 0;
    if (!s->done) { s->err = 5; }
    return;
  }
  s->n_pend = 0;
  if (s->done) return;
  if (s->err == 5) s->err = 0;   // resumed after an exhausted attempt budget
  const double ratio = sqrt(a.acc[0] / a.cnt);
  a.acc[0] = 0.0;
  if (!isfinite(ratio)) { s->err = 1; s->done = 1; return; }
  double dt = s->dt;
  if (ratio <= 1.0) {
    A.step_dt[s->n_acc] = (float)dt;
    s->dt_last = (float)dt;
    s->n_acc += 1;
    s->t_prev = s->t_cur;
    s->t_cur = s->t_cur + dt;
    s->cur ^= 1;
    int j = s->j, np = 0;
    while (j < a.n_t && !(A.t_out[j] > s->t_cur)) {
      const float xf = (float)((A.t_out[j] - s->t_prev) / (s->t_cur - s->t_prev));
      A.pend_j[np] = j; A.pend_x[np] = xf; A.out_step[j] = s->n_acc - 1; A.out_x[j] = xf;
      ++np; ++j;
    }
    s->j = j; s->n_pend = np;
    if (j == a.n_t) s->done = 1;
  } else {
    s->n_rej += 1;
  }
  if (ratio == 0.0) dt *= 10.0;
  else {
    const double dfactor = ratio < 1.0 ? 1.0 : 0.2;
    dt *= fmin(10.0, fmax(0.9 / pow(ratio, 0.2), dfactor));
  }
  s->dt = dt;
  if (!s->done) {
    if (s->n_acc >= a.max_steps) { s->err = 2; s->done = 1; }
    else if (s->n_acc + s->n_rej >= 100 * a.max_steps) { s->err = 3; s->done = 1; }
    else if (!(s->t_cur + dt > s->t_cur)) { s->err = 4; s->done = 1; }
  }
}

struct Dp5BwdArgs {
  const float* packed;
  const float* y0;        // [B, D]
  const float* xrec;      // [n_acc, 6, B, NS] stage inputs x2..x7 of every accepted step
  const float* grad_out;  // [n_t, B, D]
  float* grad_y0;         // [B, D] or null
  float* grad_rep;        // [NREP, NGRAD] zeroed, or null
  const float* dts;       // [n_acc]
  const int* out_step;    // [n_t] accepted-step index whose dense output produced y_out[j]
  const float* out_x;     // [n_t] fraction inside that step
  long long B;
  int n_acc, n_t;
  const int* n_acc_dev;   // device-driven solve: the accepted-step count lives in Dp5Dev
};

// Reverse sweep.  Per accepted step: k1 = f(y) is shared with the previous step's k7 (FSAL), so its
// cotangent is carried backwards and the VJP at y0 happens once, at the very end (pseudo-step s = -1).
template <class C, int TB>
__global__ void __launch_bounds__(BwdGeom<C, TB>::NT, 1) dopri5_bwd_kernel(Dp5BwdArgs a) {
  using G = BwdGeom<C, TB>;
  extern __shared__ __align__(16) float smem[];
  float* sRes = smem + G::OFF_RES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
  const int tid = threadIdx.x;
  const long long n_tiles = (a.B + TB - 1) / TB;
  long long my_tiles = 0;
  if ((long long)blockIdx.x < n_tiles) my_tiles = (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
  WeightRing<G::NSLOT, G::CYC, C::HID_FLOATS> ring;
  ring.slots = smem + G::OFF_RING;
  ring.full = bars;
  ring.seq = 0;
  const int n_acc = a.n_acc_dev != nullptr ? *a.n_acc_dev : a.n_acc;
  ring.total = (uint32_t)(my_tiles * (6 * n_acc + 1) * G::CYC);
  if (tid == 0) {
    for (int i = 0; i <= G::NSLOT; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (my_tiles == 0) return;
  if (tid == 0) {
    mbar_expect_tx(&bars[G::NSLOT], C::RES_ALL * 4);
    tma_load_1d(sRes, a.packed, C::RES_ALL * 4, &bars[G::NSLOT]);
  }
  ring.start(a.packed, [](uint32_t s) { return G::pkt_off(s); });
  mbar_wait(&bars[G::NSLOT], 0);
  float* grep = a.grad_rep ? a.grad_rep + (size_t)(blockIdx.x % NREP) * C::NGRAD : nullptr;

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long b = tile * TB + tid;
    const bool live = tid < TB && b < a.B;
    float adj[C::NS], carry[C::NS], cst[C::NCONST], dcst[C::NCONST], ay[C::NS], a1[C::NS];
    float dks[7][C::NS];
#pragma unroll
    for (int d = 0; d < C::NS; ++d) { adj[d] = 0.f; carry[d] = 0.f; ay[d] = 0.f; a1[d] = 0.f; }
#pragma unroll
    for (int d = 0; d < C::NCONST; ++d) { cst[d] = 0.f; dcst[d] = 0.f; }
    if (live) {
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) cst[d] = a.y0[b * C::D + C::NS + d];
    }
    int jo = a.n_t - 1;
#pragma unroll 1
    for (int s = n_acc - 1; s >= -1; --s) {
      float dt = 0.f;
      int i_hi = 0, i_lo = 0;
      if (s >= 0) {
        dt = a.dts[s];
        i_hi = 6; i_lo = 1;
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
          for (int d = 0; d < C::NS; ++d) dks[i][d] = 0.f;
#pragma unroll
        for (int d = 0; d < C::NS; ++d) { ay[d] = 0.f; a1[d] = adj[d]; }
        while (jo >= 1 && a.out_step[jo] == s) {
          if (live) {
            const float x = a.out_x[jo], x2 = x * x, x3 = x2 * x, x4 = x3 * x;
            const float A0 = 1.f - 11.f * x2 + 18.f * x3 - 8.f * x4, A1 = -5.f * x2 + 14.f * x3 - 8.f * x4;
            const float Am = 16.f * x2 - 32.f * x3 + 16.f * x4;
            const float F0 = x - 4.f * x2 + 5.f * x3 - 2.f * x4, F1 = x2 - 3.f * x3 + 2.f * x4;
            const float* g = a.grad_out + ((long long)jo * a.B + b) * C::D;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) {
              const float gv = g[d];
              ay[d] = fmaf(A0 + Am, gv, ay[d]);
              a1[d] = fmaf(A1, gv, a1[d]);
#pragma unroll
              for (int i = 0; i < 7; ++i) dks[i][d] = fmaf(dt * DP_CMID[i] * Am, gv, dks[i][d]);
              dks[0][d] = fmaf(dt * F0, gv, dks[0][d]);
              dks[6][d] = fmaf(dt * F1, gv, dks[6][d]);
            }
#pragma unroll
            for (int d = 0; d < C::NCONST; ++d) dcst[d] += g[C::NS + d];
          }
          --jo;
        }
#pragma unroll
        for (int d = 0; d < C::NS; ++d) dks[6][d] += carry[d];
      }
#pragma unroll 1
      for (int i = i_hi; i >= i_lo; --i) {
        float x[C::NS], dk[C::NS], dx[C::NS];
#pragma unroll
        for (int d = 0; d < C::NS; ++d) { x[d] = 0.f; dk[d] = 0.f; }
        if (live) {
          const float* p = (s >= 0) ? a.xrec + (((long long)s * 6 + (i - 1)) * a.B + b) * C::NS : a.y0 + b * C::D;
#pragma unroll
          for (int d = 0; d < C::NS; ++d) { x[d] = p[d]; dk[d] = (s >= 0) ? dks[i][d] : carry[d]; }
        }
        vjp_rhs_tile<C, TB>(a.packed, smem, ring, grep, x, cst, dk, dx, dcst);
        if (s >= 0) {
          if (i == 6) {
#pragma unroll
            for (int d = 0; d < C::NS; ++d) dx[d] += a1[d];
          }
#pragma unroll
          for (int d = 0; d < C::NS; ++d) ay[d] += dx[d];
#pragma unroll
          for (int j = 0; j < 6; ++j)
            if (j < i) {
              const float bij = dt * DP_BETA[i - 1][j];
#pragma unroll
              for (int d = 0; d < C::NS; ++d) dks[j][d] = fmaf(bij, dx[d], dks[j][d]);
            }
        } else {
#pragma unroll
          for (int d = 0; d < C::NS; ++d) adj[d] += dx[d];
        }
      }
      if (s >= 0) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) { carry[d] = dks[0][d]; adj[d] = ay[d]; }
      }
    }
    if (live && a.grad_y0 != nullptr) {
      const float* g = a.grad_out + b * C::D;
      float* o = a.grad_y0 + b * C::D;
#pragma unroll
      for (int d = 0; d < C::NS; ++d) o[d] = adj[d] + g[d];
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) o[C::NS + d] = dcst[d] + g[C::NS + d];
    }
  }
}

}  // namespace nlbac
