// K1 — fused fixed-grid NODE solve, forward, FP32-exact (FP32 FMA pipe).
//
// Replaces torchdiffeq.odeint(NeuralODEModel, y0, t, method='euler'|'midpoint'|'rk4') as called at
// U sac_cbf_clf/model.py:252, U sac_cbf_clf/sac_cbf_clf.py:453 (and the 17 sibling call sites): every
// solver step and stage of a tile of T trajectories runs inside one persistent CTA; the RK state and
// stages of a trajectory live in the registers of its owner thread; the hidden HxH matrices stream
// through a shared-memory ring via TMA bulk copies; the thin first/last layers stay resident.
#pragma once
#include "node_blocks.cuh"

namespace nlbac {

constexpr int METHOD_EULER = 0, METHOD_MIDPOINT = 1, METHOD_RK4 = 2, METHOD_DOPRI5 = 3;
__host__ __device__ constexpr int method_stages(int m) { return m == METHOD_EULER ? 1 : (m == METHOD_MIDPOINT ? 2 : 4); }

struct FwdArgs {
  const float* packed;
  const float* y0;      // [B, D]
  float* y_out;         // [n_t, B, D]
  float* ckpt;          // [n_steps, S, B, NS] or null
  const float* dts;     // [n_steps]
  const int* out_step;  // [n_t]   (entry 0 unused)
  const float* out_w;   // [n_t]
  long long B;
  int n_steps, n_t, method;
};

template <class C, int T>
struct FwdGeom {
  static constexpr bool SMALL = (T % 32 != 0);               // narrow geometry: 8 row-threads x HP/4 column groups
  static constexpr int TG = SMALL ? NODE_SMALL_TILE : 32;   // row-threads per column group (regular: one warp)
  static constexpr int TT = T / TG;   // trajectories per thread
  using CM = std::conditional_t<SMALL, ColMapN<C::HP>, ColMap<C::HP>>;
  static constexpr int NGRP = SMALL ? C::HP / 4 : 8;        // column groups
  static constexpr int NT = (TG * NGRP + 31) / 32 * 32;     // small: 8 x 26 = 208 -> 224 (threads past NGRP idle in the GEMMs)
  static_assert(T % TG == 0 && TT >= 1, "tile size");
  static constexpr int NSLOT = C::NHID <= 2 ? C::NHID : 2;
  static constexpr int OFF_A = 0;
  static constexpr int OFF_X = OFF_A + T * C::RS;
  static constexpr int OFF_RES = OFF_X + T * C::INP0;
  static constexpr int OFF_RING = OFF_RES + C::RES_FWD;
  static constexpr int OFF_BAR = OFF_RING + NSLOT * C::HID_FLOATS;
  static constexpr size_t SMEM_BYTES = (size_t)OFF_BAR * 4 + (NSLOT + 1) * 8;
  static_assert(NT >= T && NT <= 1024, "one owner thread per trajectory");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// Evaluate the NODE right-hand side for the tile: owner threads pass x (and the constants) in
// registers and receive k = dx/dt in registers.
template <class C, int T, class Ring>
__device__ __forceinline__ void eval_rhs_tile(const float* __restrict__ packed, float* sA, float* sX,
                                              const float* __restrict__ sRes, Ring& ring, const float (&x)[C::NS],
                                              const float (&cst)[C::NCONST], float (&k)[C::NS]) {
  using G = FwdGeom<C, T>;
  const int tid = threadIdx.x;
  const bool owner = tid < T;
  const int tg = tid % G::TG, ng = tid / G::TG;
  const bool gact = ng < G::NGRP;
  using CM = typename G::CM;
  auto off_t = [](uint32_t s) { return C::off_hid_t(s % C::NHID); };

  if (owner) {
    float* xr = sX + tid * C::INP0;
#pragma unroll
    for (int d = 0; d < C::INP0; ++d) {
      float v = 0.f;
      if (d < C::NS) v = x[d];
      else if (C::KIND == KIND_BLACKBOX && d < C::D) v = cst[d - C::NS];
      xr[d] = v;
    }
  }
  __syncthreads();

  float fo[C::outp(0)];
#pragma unroll(NODE_ROLL_CHAINS_FWD ? 1 : 2)
  for (int c = 0; c < C::NCH; ++c) {
    float acc[G::TT][CM::NN];
    // first layer: sX -> sA
    if (gact) {
      gemm_tile<G::TT, G::TG, C::HP, (G::TT <= NODE_FWD_SMALL_TT ? NODE_GEMM_UNROLL_FWD_SMALL : NODE_GEMM_UNROLL_FWD), CM>(sX, C::INP0, sRes + C::off_wt0(c), sRes + C::off_b0(c), C::INP0, tg, ng, acc);
      store_relu<G::TT, G::TG, C::HP, CM>(sA, C::RS, tg, ng, acc);
    }
    __syncthreads();
    // hidden layers, in place
#pragma unroll 1
    for (int l = 1; l <= C::nhc(c); ++l) {
      const float* W = ring.acquire();
      if (gact) gemm_tile<G::TT, G::TG, C::HP, (G::TT <= NODE_FWD_SMALL_TT ? NODE_GEMM_UNROLL_FWD_SMALL : NODE_GEMM_UNROLL_FWD), CM>(sA, C::RS, W, sRes + C::off_bh(c, l), C::H, tg, ng, acc);
      __syncthreads();
      ring.release(packed, off_t);
      if (gact) store_relu<G::TT, G::TG, C::HP, CM>(sA, C::RS, tg, ng, acc);
      __syncthreads();
    }
    // thin last layer, one thread per trajectory
    if (owner) {
      if (c == 0) {
        thin_out<C::H, C::outp(0)>(sA + tid * C::RS, sRes + C::off_wtl(0), sRes + C::off_bl(0), fo);
      } else {
        float go[C::outp(C::NCH - 1)];
        thin_out<C::H, C::outp(C::NCH - 1)>(sA + tid * C::RS, sRes + C::off_wtl(C::NCH - 1),
                                            sRes + C::off_bl(C::NCH - 1), go);
#pragma unroll
        for (int i = 0; i < C::NS; ++i) {
          float v = fo[i];
#pragma unroll
          for (int j = 0; j < C::NU; ++j) v = fmaf(go[i * C::NU + j], cst[j], v);
          k[i] = v;
        }
      }
    }
    __syncthreads();
  }
  if (C::NCH == 1 && owner) {
#pragma unroll
    for (int i = 0; i < C::NS; ++i) k[i] = fo[i];
  }
}

template <class C, int T>
__global__ void __launch_bounds__(FwdGeom<C, T>::NT, 1) node_fwd_kernel(FwdArgs a) {
  using G = FwdGeom<C, T>;
  extern __shared__ __align__(16) float smem[];
  float* sA = smem + G::OFF_A;
  float* sX = smem + G::OFF_X;
  float* sRes = smem + G::OFF_RES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
  const int tid = threadIdx.x;

  const long long n_tiles = (a.B + T - 1) / T;
  const int S = method_stages(a.method);
  long long my_tiles = 0;
  if ((long long)blockIdx.x < n_tiles) my_tiles = (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;

  WeightRing<G::NSLOT, C::NHID, C::HID_FLOATS> ring;
  ring.slots = smem + G::OFF_RING;
  ring.full = bars;
  ring.seq = 0;
  ring.total = (uint32_t)(my_tiles * a.n_steps * S * C::NHID);

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

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long b = tile * T + tid;
    const bool owner = tid < T;
    const bool live = owner && b < a.B;
    float y[C::NS], cst[C::NCONST], k1[C::NS], k2[C::NS], k3[C::NS], k4[C::NS], x[C::NS];
#pragma unroll
    for (int d = 0; d < C::NS; ++d) k1[d] = k2[d] = k3[d] = k4[d] = 0.f;
#pragma unroll
    for (int d = 0; d < C::NS; ++d) y[d] = 0.f;
#pragma unroll
    for (int d = 0; d < C::NCONST; ++d) cst[d] = 0.f;
    if (live) {
      const float* yr = a.y0 + b * C::D;
#pragma unroll
      for (int d = 0; d < C::NS; ++d) y[d] = yr[d];
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) cst[d] = yr[C::NS + d];
      float* o = a.y_out + b * C::D;  // y_out[0] = y0
#pragma unroll
      for (int d = 0; d < C::NS; ++d) o[d] = y[d];
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) o[C::NS + d] = cst[d];
    }
    int jout = 1;
    for (int step = 0; step < a.n_steps; ++step) {
      const float dt = a.dts[step];
      float y1[C::NS];
      auto ck = [&](int stage, const float(&xs)[C::NS]) {
        if (a.ckpt != nullptr && live) {
          float* p = a.ckpt + (((long long)step * S + stage) * a.B + b) * C::NS;
#pragma unroll
          for (int d = 0; d < C::NS; ++d) p[d] = xs[d];
        }
      };
      // one call site for the right-hand side keeps the instruction footprint small
#pragma unroll 1
      for (int stage = 0; stage < S; ++stage) {
        if (stage == 0) {
#pragma unroll
          for (int d = 0; d < C::NS; ++d) x[d] = y[d];
        } else if (a.method == METHOD_MIDPOINT) {
          const float half = 0.5f * dt;
#pragma unroll
          for (int d = 0; d < C::NS; ++d) x[d] = y[d] + k1[d] * half;
        } else {  // rk4 = 3/8 rule (torchdiffeq rk4_alt_step_func)
          const float third = 1.0f / 3.0f;
#pragma unroll
          for (int d = 0; d < C::NS; ++d) {
            if (stage == 1) x[d] = y[d] + dt * k1[d] * third;
            else if (stage == 2) x[d] = y[d] + dt * (k2[d] - k1[d] * third);
            else x[d] = y[d] + dt * (k1[d] - k2[d] + k3[d]);
          }
        }
        ck(stage, x);
        float kn[C::NS];
        eval_rhs_tile<C, T>(a.packed, sA, sX, sRes, ring, x, cst, kn);
#pragma unroll
        for (int d = 0; d < C::NS; ++d) {
          if (stage == 0) k1[d] = kn[d];
          else if (stage == 1) k2[d] = kn[d];
          else if (stage == 2) k3[d] = kn[d];
          else k4[d] = kn[d];
        }
      }
      if (a.method == METHOD_EULER) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) y1[d] = y[d] + dt * k1[d];
      } else if (a.method == METHOD_MIDPOINT) {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) y1[d] = y[d] + dt * k2[d];
      } else {
#pragma unroll
        for (int d = 0; d < C::NS; ++d) y1[d] = y[d] + (k1[d] + 3.f * (k2[d] + k3[d]) + k4[d]) * dt * 0.125f;
      }
      // outputs that fall inside this step (linear interpolation, like FixedGridODESolver.integrate)
      while (jout < a.n_t && a.out_step[jout] == step) {
        if (live) {
          const float w = a.out_w[jout];
          float* o = a.y_out + ((long long)jout * a.B + b) * C::D;
#pragma unroll
          for (int d = 0; d < C::NS; ++d) o[d] = (w == 1.f) ? y1[d] : ((w == 0.f) ? y[d] : y[d] + w * (y1[d] - y[d]));
#pragma unroll
          for (int d = 0; d < C::NCONST; ++d) o[C::NS + d] = cst[d];
        }
        ++jout;
      }
#pragma unroll
      for (int d = 0; d < C::NS; ++d) y[d] = y1[d];
    }
  }
}

}  // namespace nlbac
