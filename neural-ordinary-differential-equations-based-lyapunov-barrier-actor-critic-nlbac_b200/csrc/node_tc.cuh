// K1-TC — fused fixed-grid NODE solve, forward, hidden layers on the 5th-generation tensor cores
// (tcgen05.mma kind::tf32, accumulators in TMEM).  Same call sites as node_fwd.cuh (U model.py:252,
// U sac_cbf_clf.py:453, ...); precision tiers:
//   SPLIT = false  one TF32 pass per hidden layer            -> declared 1e-2 tier (measured ~3e-3 on d-theta)
//   SPLIT = true   3xTF32 (hi*hi + lo*hi + hi*lo)            -> fp32-class
//
// One persistent CTA runs G independent "groups" of 128 threads; a group owns a tile of 128 trajectories:
// thread = trajectory = TMEM lane.  Per right-hand-side evaluation the owner thread computes the thin first
// layer on the FP32 pipe and writes its activation row straight into the K-major core-matrix image
// (umma.cuh); ONE thread issues the 13 (39) MMAs of a hidden layer [128 x 112 x 104]; the owner reads its
// accumulator row back with tcgen05.ld, applies bias + ReLU and either rewrites the activation image in
// place or folds the thin last layer in registers.  The G groups share one TMA weight ring (one packet =
// one hidden matrix image, 46.6 KB), so while one group is in its epilogue the other groups' MMAs run.
// Shared memory holds two activation images; a THIRD group keeps its A operand in tensor memory instead (row = lane,
// feature = column, written with tcgen05.st by the owner threads; self-test mode 2 in tc_abi.cu), which costs no
// shared memory: 3 x 128 accumulator columns + 128 operand columns = all 512 TMEM columns of the SM.
#pragma once
#include "node_fwd.cuh"
#include "umma.cuh"

namespace nlbac {

#ifndef NLBAC_TC_L0_UNROLL
#define NLBAC_TC_L0_UNROLL 1   // column-group loop of the thin first layer in the tensor-core kernels
#endif
constexpr int TC_L0_UNROLL = NLBAC_TC_L0_UNROLL;


template <class C>
struct TcCfg {
  static constexpr int ROWS = 128;                 // trajectories per tile = MMA M
  static constexpr int NP = round_up(C::H, 16);    // MMA N (100 -> 112, 64 -> 64)
  static constexpr int KP = round_up(C::H, 8);     // MMA K extent (100 -> 104, 64 -> 64)
  static constexpr int KSTEPS = KP / 8;
  static constexpr int NCHUNK = KP / 4;            // 16-byte chunks along K
  static constexpr int ACH = ROWS * 16;            // activation image: bytes per chunk
  static constexpr int WCH = NP * 16;              // weight image: bytes per chunk
  static constexpr int A_BYTES = NCHUNK * ACH;
  static constexpr int W_FLOATS = NCHUNK * NP * 4;
  static constexpr int W_BYTES = W_FLOATS * 4;
  static constexpr int TCOLS = 128;                // TMEM columns reserved per group
  // Hidden biases ride inside the GEMM when the K extent has a spare column (H = 100 -> KP = 104): activation
  // column H is a constant 1 (first-layer bias image b0[H] = 1, weight image W[H][H] = 1 keeps it alive from layer
  // to layer) and weight column H holds the bias, so the epilogue is ReLU only.  H = 64 has no spare column.
  static constexpr bool FOLD = KP > C::H;
  static_assert(NP <= TCOLS && KP == C::HP, "tile shape");
  // packed image: [RES_ALL (thin layers fwd + bwd, node_cfg.cuh) | U_hi[NHID] | U_lo[NHID] | U_T[NHID]]
  //   U_hi / U_lo: B operand of the forward GEMM  (rows j = out, K = in);  U_T / U_Tlo: B operand of dX (rows k = in, K = out)
  __host__ __device__ static constexpr int off_hi(int h) { return C::RES_ALL + h * W_FLOATS; }
  __host__ __device__ static constexpr int off_lo(int h) { return C::RES_ALL + (C::NHID + h) * W_FLOATS; }
  __host__ __device__ static constexpr int off_tr(int h) { return C::RES_ALL + (2 * C::NHID + h) * W_FLOATS; }
  // U_Tlo: residual of the transposed image (3xTF32 reverse sweep: dX = d_hi W_hi + d_lo W_hi + d_hi W_lo)
  __host__ __device__ static constexpr int off_trlo(int h) { return C::RES_ALL + (3 * C::NHID + h) * W_FLOATS; }
  static constexpr int PACKED = C::RES_ALL + 4 * C::NHID * W_FLOATS;
};

// round-to-nearest TF32 (the MMA itself truncates the low 13 mantissa bits: rounding first halves the error and
// removes its bias); the result is exactly representable, so hi + (x - hi) == x for the 3xTF32 split
__host__ __device__ __forceinline__ float tf32_hi(float x) {
#ifdef __CUDA_ARCH__
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
#else
  return x;
#endif
}

// Activation-side rounding.  cvt.rna.tf32.f32 costs four instructions per element (add, |x| >= inf test, select,
// mask); on the hot path two cheaper forms give the same rounding for finite values:
//   tf32_bias: bits + 0x1000  -- the MMA then drops the low 13 bits itself, which completes round-to-nearest
//              (ties away from zero).  ONE instruction.  Inf / NaN inputs come out as NaN (a solve that has
//              already overflowed), which the exact form would have preserved as Inf.
//   tf32_rn:   (bits + 0x1000) & ~0x1FFF  -- exactly representable, needed where hi is subtracted again (3xTF32).
__device__ __forceinline__ float tf32_bias(float x) { return __uint_as_float(__float_as_uint(x) + 0x1000u); }
__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

// ReLU and the TF32 rounding bias in ONE integer instruction (VIADDMNMX): max(bits + 0x1000, 0) as signed integers
// -- negative floats are negative integers and stay so after the add
__device__ __forceinline__ float relu_tf32_bias(float x) {
  return __int_as_float(__viaddmax_s32(__float_as_int(x), 0x1000, 0));
}

// Eight fp32 accumulators as four packed pairs: the thin layers run on fma.rn.f32x2 (SASS FFMA2, one issue slot for
// two IEEE FMAs -- common.cuh), which halves their share of the epilogue's instruction stream.
struct Acc8 { unsigned long long p[4]; };
__device__ __forceinline__ void acc8_init(Acc8& a, const float4& lo, const float4& hi) {
  a.p[0] = pack2(lo.x, lo.y); a.p[1] = pack2(lo.z, lo.w); a.p[2] = pack2(hi.x, hi.y); a.p[3] = pack2(hi.z, hi.w);
}
__device__ __forceinline__ void acc8_zero(Acc8& a) { a.p[0] = a.p[1] = a.p[2] = a.p[3] = 0ull; }
// a[0..8) += s * w[0..8)
__device__ __forceinline__ void acc8_fma(Acc8& a, float s, const float4& wa, const float4& wb) {
  const unsigned long long ss = pack2(s, s);
  a.p[0] = fma2(ss, pack2(wa.x, wa.y), a.p[0]); a.p[1] = fma2(ss, pack2(wa.z, wa.w), a.p[1]);
  a.p[2] = fma2(ss, pack2(wb.x, wb.y), a.p[2]); a.p[3] = fma2(ss, pack2(wb.z, wb.w), a.p[3]);
}
__device__ __forceinline__ void acc8_get(const Acc8& a, float (&v)[8]) {
  unpack2(a.p[0], v[0], v[1]); unpack2(a.p[1], v[2], v[3]); unpack2(a.p[2], v[4], v[5]); unpack2(a.p[3], v[6], v[7]);
}
__device__ __forceinline__ float pair_sum(unsigned long long a) { float lo, hi; unpack2(a, lo, hi); return lo + hi; }

// weight image of one hidden matrix: W[j][k] (nn.Linear [out][in]) -> float index (k/4)*NP*4 + j*4 + k%4
template <class C>
__global__ void pack_tc_kernel(const float* __restrict__ W, const float* __restrict__ b, float* __restrict__ hi,
                               float* __restrict__ lo, float* __restrict__ tr, float* __restrict__ trlo,
                               float* __restrict__ b0_one) {
  using T = TcCfg<C>;
  if (T::FOLD && b0_one != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *b0_one = 1.f;   // a_1[H] = relu(1) = 1
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T::W_FLOATS; i += gridDim.x * blockDim.x) {
    const int c = i / (T::NP * 4), j = (i / 4) % T::NP, k = 4 * c + (i & 3);
    const bool in = j < C::H && k < C::H;
    float w = in ? W[j * C::H + k] : 0.f;
    if (T::FOLD && k == C::H) w = j < C::H ? b[j] : (j == C::H ? 1.f : 0.f);
    const float h = tf32_hi(w);
    hi[i] = h;
    lo[i] = w - h;
    const float wt = in ? W[k * C::H + j] : 0.f;                         // row j of the image = column j of W
    const float ht = tf32_hi(wt);
    tr[i] = ht;
    trlo[i] = wt - ht;
  }
}

// TMEM columns of a CTA: one 128-column accumulator per group + one 128-column operand block per TMEM-operand group
// (allocations are powers of two)
__host__ __device__ constexpr int tmem_cols(int groups, int tmem_groups, int nimg = 1) {
  const int need = (groups + tmem_groups * nimg) * 128;
  return need <= 128 ? 128 : (need <= 256 ? 256 : 512);
}

template <class C, bool SPLIT>
struct TcFwdGeom {
  using T = TcCfg<C>;
  static constexpr int GS = SPLIT ? 1 : 2;         // groups whose activation image(s) live in shared memory
  static constexpr int GT = 1;                     // groups whose activation operand(s) live in tensor memory (3xTF32: the
                                                   // hi and lo blocks, 2 x 128 columns -> 2 accumulators + 2 blocks = 512)
  static constexpr int G = GS + GT;                // groups (tiles in flight) per CTA
  static constexpr int NT = G * 128;               // no producer warp: see ring_prefetch
  static constexpr int NIMG = SPLIT ? 2 : 1;       // activation images per group (hi, lo)
  static constexpr int TMEM_COLS = tmem_cols(G, GT, NIMG);
  static constexpr int NSLOT = 2;
  static constexpr int OFF_A = 0;                                   // bytes
  static constexpr int OFF_W = OFF_A + GS * NIMG * T::A_BYTES;
  static constexpr int OFF_RES = OFF_W + NSLOT * T::W_BYTES;
  static constexpr int OFF_BAR = OFF_RES + C::RES_FWD * 4;
  static constexpr int NBAR = 2 * NSLOT + G + 1;   // full[NSLOT], empty[NSLOT], done[G], res
  static constexpr size_t SMEM_BYTES = OFF_BAR + NBAR * 8 + 16;
  static constexpr int PKT_PER_EVAL = C::NHID * (SPLIT ? 2 : 1);
  static_assert(OFF_W % 128 == 0 && OFF_RES % 16 == 0 && OFF_BAR % 8 == 0, "alignment");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
  // global offset (floats) of ring packet p
  __device__ static int pkt_off(uint32_t p) {
    const int q = p % PKT_PER_EVAL;
    if (SPLIT) return (q & 1) ? T::off_lo(q >> 1) : T::off_hi(q >> 1);
    return T::off_hi(q);
  }
};

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }


// store 4 consecutive features [c, c+4) of activation row r into the image(s)
template <bool SPLIT, int ACH, int A_BYTES>
__device__ __forceinline__ void store_act4(unsigned char* img, int r, int c, float v0, float v1, float v2, float v3) {
  unsigned char* p = img + (c >> 2) * ACH + (r >> 3) * 128 + (r & 7) * 16;
  if (SPLIT) {
    const float h0 = tf32_rn(v0), h1 = tf32_rn(v1), h2 = tf32_rn(v2), h3 = tf32_rn(v3);
    *reinterpret_cast<float4*>(p) = make_float4(h0, h1, h2, h3);
    *reinterpret_cast<float4*>(p + A_BYTES) = make_float4(v0 - h0, v1 - h1, v2 - h2, v3 - h3);
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(tf32_bias(v0), tf32_bias(v1), tf32_bias(v2), tf32_bias(v3));
  }
}

// Visit the accumulator row in batches of 4 x 8 columns: up to four tcgen05.ld in flight, ONE wait per batch.
// body(q, u, g8, v): q = batch (rolled loop over the full batches; the remainder is a separate, guard-free tail),
// u = 0..3 (unrolled), g8 = 4q + u = column group, v = 8 columns.
struct NoPost { __device__ __forceinline__ void operator()(int) const {} };
// post(q) runs once after the (up to four) column groups of batch q
#ifndef NLBAC_TC_PIPE
#define NLBAC_TC_PIPE 1     // 1: accumulator loads software-pipelined in half-batches of 16 columns (below)
#endif
#ifndef NLBAC_TC_BARWAIT
#define NLBAC_TC_BARWAIT 0  // 1: one thread per group waits on the MMA-done mbarrier, the other 127 block in bar.sync
                            // (measured on one box, round 2: sweep 68.2 -> 69.0 ms; the extra barrier hop costs more
                            // than the polling instructions it removes, so the default stays 0)
#endif

template <int NG8, bool LOAD, class F, class P = NoPost>
__device__ __forceinline__ void tc_batches(uint32_t tacc, F&& body, P&& post = P()) {
  constexpr int NFULL = NG8 / 4, REM = NG8 % 4;
  if constexpr (LOAD && NLBAC_TC_PIPE) {
    // Half-batches of 2 x 8 columns, double-buffered in registers: while the owner thread works on one half the
    // tcgen05.ld of the next is in flight (same 32 live registers as four loads + one wait, but the TMEM read latency
    // is hidden behind ~50 instructions of epilogue instead of being exposed four times per layer).
    float va[2][8], vb[2][8];
    if (NFULL > 0) { umma::ld8f(tacc, va[0]); umma::ld8f(tacc + 8, va[1]); }
#pragma unroll 1
    for (int q = 0; q < NFULL; ++q) {
      umma::wait_ld16(va[0], va[1]);
      umma::ld8f(tacc + 8 * (4 * q + 2), vb[0]);
      umma::ld8f(tacc + 8 * (4 * q + 3), vb[1]);
      body(q, 0, 4 * q, va[0]);
      body(q, 1, 4 * q + 1, va[1]);
      umma::wait_ld16(vb[0], vb[1]);
      if (q + 1 < NFULL) {
        umma::ld8f(tacc + 8 * (4 * q + 4), va[0]);
        umma::ld8f(tacc + 8 * (4 * q + 5), va[1]);
      } else if (REM > 0) {
        umma::ld8f(tacc + 8 * (4 * NFULL), va[0]);
        if (REM > 1) umma::ld8f(tacc + 8 * (4 * NFULL + 1), va[1]);
      }
      body(q, 2, 4 * q + 2, vb[0]);
      body(q, 3, 4 * q + 3, vb[1]);
      post(q);
    }
    if (REM > 0) {
      if (NFULL == 0) {
        umma::ld8f(tacc, va[0]);
        if (REM > 1) umma::ld8f(tacc + 8, va[1]);
      }
      if (REM > 2) umma::ld8f(tacc + 8 * (4 * NFULL + 2), vb[0]);
      umma::wait_ld16(va[0], vb[0]);
      if (REM > 1) umma::wait_ld16(va[1], vb[0]);
      body(NFULL, 0, 4 * NFULL, va[0]);
      if (REM > 1) body(NFULL, 1, 4 * NFULL + 1, va[1]);
      if (REM > 2) body(NFULL, 2, 4 * NFULL + 2, vb[0]);
      post(NFULL);
    }
  } else {
#pragma unroll 1
  for (int q = 0; q < NFULL; ++q) {
    float v[4][8];
    if (LOAD) {
#pragma unroll
      for (int u = 0; u < 4; ++u) umma::ld8(tacc + 8 * (4 * q + u), v[u]);
      umma::wait_ld();
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) body(q, u, 4 * q + u, v[u]);
    post(q);
  }
  if (REM > 0) {
    float v[REM > 0 ? REM : 1][8];
    if (LOAD) {
#pragma unroll
      for (int u = 0; u < REM; ++u) umma::ld8(tacc + 8 * (4 * NFULL + u), v[u]);
      umma::wait_ld();
    }
#pragma unroll
    for (int u = 0; u < REM; ++u) body(NFULL, u, 4 * NFULL + u, v[u]);
    post(NFULL);
  }
  }
}

// Compact form of tc_batches for kernels bound by INSTRUCTION FETCH (the 3xTF32 reverse sweep: 193 KB of SASS, ncu
// stalled_no_instruction 5.7 per issued instruction): a rolled loop over PAIRS of column groups, so every epilogue body
// is inlined twice (+ once for an odd tail) instead of five times.  Same body / post contract: q = g8 / 4 (mask word),
// u = g8 % 4 (byte inside it), post(q) once after the last group of word q.
template <int NG8, bool LOAD, class F, class P = NoPost>
__device__ __forceinline__ void tc_pairs(uint32_t tacc, F&& body, P&& post = P()) {
  constexpr int NPAIR = NG8 / 2;
#pragma unroll 1
  for (int p = 0; p < NPAIR; ++p) {
    float v0[8], v1[8];
    if (LOAD) {
      umma::ld8f(tacc + 16 * p, v0);
      umma::ld8f(tacc + 16 * p + 8, v1);
      umma::wait_ld16(v0, v1);
    }
    const int g = 2 * p;
    body(g >> 2, g & 3, g, v0);
    body(g >> 2, (g & 3) + 1, g + 1, v1);
    if ((g & 3) == 2) post(g >> 2);
  }
  if (NG8 & 1) {
    float v0[8];
    if (LOAD) {
      umma::ld8f(tacc + 8 * (NG8 - 1), v0);
      umma::wait_ld8(v0);
    }
    body((NG8 - 1) >> 2, (NG8 - 1) & 3, NG8 - 1, v0);
  }
  if ((NG8 & 3) != 0) post((NG8 - 1) >> 2);
}

// ... and ONE column group per iteration (every body inlined exactly once; one exposed TMEM-load wait per 8 columns)
template <int NG8, bool LOAD, class F, class P = NoPost>
__device__ __forceinline__ void tc_singles(uint32_t tacc, F&& body, P&& post = P()) {
#pragma unroll 1
  for (int g = 0; g < NG8; ++g) {
    float v0[8];
    if (LOAD) {
      umma::ld8f(tacc + 8 * g, v0);
      umma::wait_ld8(v0);
    }
    body(g >> 2, g & 3, g, v0);
    if ((g & 3) == 3 || g == NG8 - 1) post(g >> 2);
  }
}

// Shared-memory / barrier handles of one group
struct TcGroupCtx {
  unsigned char* sA;       // activation image(s) of this group
  unsigned char* sW;       // weight ring base
  const float* sRes;       // thin layers (RES_FWD layout)
  uint64_t *full, *empty, *done;
  uint32_t tacc;           // TMEM address of this thread's accumulator row (lane quadrant + group column base)
  uint32_t tmma;           // TMEM address of the group's accumulator tile (lane 0)
  uint32_t seq;            // ring packets consumed so far (uniform over the CTA's groups)
  uint32_t nd;             // completed hidden layers of this group (phase of `done`)
  uint64_t adesc, wdesc;   // UMMA descriptors of the activation image and of ring slot 0 (k-step 0)
  uint32_t total;          // ring packets this CTA will consume in all
  const float* packed;     // global weight image (ring source)
  bool keeper;             // this thread re-fills the weight ring
  uint32_t ta_row;         // TMEM address of this thread's activation row   (groups with the A operand in TMEM)
  uint32_t ta_mma;         // TMEM address of the group's activation operand (lane 0)
  bool tmemA;
  int lt, g;
#ifdef NLBAC_TC_PROF   // -DNLBAC_TC_PROF: cycle breakdown of the MMA round trip (tools/tc_roundtrip.py), debug builds only
  long long pf_t, pf_epi, pf_full, pf_issue, pf_done, pf_n;
#endif
};

// 8 consecutive features [8 g8, 8 g8 + 8) of the owner's activation row -> A operand of the group's next MMA
template <class C, bool SPLIT>
__device__ __forceinline__ void act_store8(const TcGroupCtx& cx, int g8, const float (&v)[8]) {
  using T = TcCfg<C>;
  if (cx.tmemA) {
    if (SPLIT) {   // hi block at ta_row, lo block 128 columns further (TcBwdGeom<C, true>)
      float h[8], l[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) { h[i] = tf32_rn(v[i]); l[i] = v[i] - h[i]; }
      umma::st8(cx.ta_row + 8 * g8, h);
      umma::st8(cx.ta_row + T::TCOLS + 8 * g8, l);
    } else {
      float r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = tf32_bias(v[i]);
      umma::st8(cx.ta_row + 8 * g8, r);
    }
  } else {
    store_act4<SPLIT, T::ACH, T::A_BYTES>(cx.sA, cx.lt, 8 * g8, v[0], v[1], v[2], v[3]);
    store_act4<SPLIT, T::ACH, T::A_BYTES>(cx.sA, cx.lt, 8 * g8 + 4, v[4], v[5], v[6], v[7]);
  }
}
// same for values that already carry their TF32 rounding (relu_tf32_bias)
template <class C>
__device__ __forceinline__ void act_store8_raw(const TcGroupCtx& cx, int g8, const float (&r)[8]) {
  using T = TcCfg<C>;
  if (cx.tmemA) {
    umma::st8(cx.ta_row + 8 * g8, r);
  } else {
    unsigned char* p = cx.sA + (g8 * 2) * T::ACH + (cx.lt >> 3) * 128 + (cx.lt & 7) * 16;
    *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
    *reinterpret_cast<float4*>(p + T::ACH) = make_float4(r[4], r[5], r[6], r[7]);
  }
}
// operand written -> visible to the MMA the group's elected thread issues next
template <bool SPLIT>
__device__ __forceinline__ void act_publish(const TcGroupCtx& cx) {
  if (cx.tmemA) umma::wait_st();
  else fence_proxy_async();
  umma::fence_before_sync();
  group_sync(cx.g);
}

// The KSTEPS MMAs of one hidden layer.  Descriptors differ from k-step to k-step only in the 14-bit start address,
// so they are formed by adding to the low word; one out-of-line copy keeps the (fully unrolled) issue sequence
// short on the critical path without multiplying it over the call sites.
template <class C>
__device__ __noinline__ void tc_issue(uint32_t tmma, uint64_t ad, uint64_t wd, uint32_t acc0) {
  using T = TcCfg<C>;
  constexpr uint32_t idesc = umma::make_idesc_tf32(128, T::NP, false, false);
#pragma unroll
  for (int ks = 0; ks < T::KSTEPS; ++ks)
    umma::mma_tf32(tmma, ad + (uint64_t)(ks * ((2 * T::ACH) >> 4)), wd + (uint64_t)(ks * ((2 * T::WCH) >> 4)), idesc,
                   ks > 0 ? 1u : acc0);
}

// The weight ring has no warp of its own: 12 warps leave 168 registers per thread, a 13th would put four warps on
// one scheduler and cap every thread at 128.  Instead, at the top of each hidden layer one thread of group 0 (the
// "ring keeper", lane 0 of the group's SECOND warp, so that it never delays the MMA issuer in the first) requests
// packet p + NSLOT - 1, i.e. it re-fills the slot whose previous packet every group has finished multiplying
// (empty barrier, armed by tcgen05.commit).
template <class GEO>
__device__ __forceinline__ void ring_prefetch(const TcGroupCtx& cx, uint32_t p) {
  using T = typename GEO::T;
  if (p < cx.total) {
    const uint32_t slot = p % GEO::NSLOT;
    if (p >= (uint32_t)GEO::NSLOT) mbar_wait(&cx.empty[slot], ((p / GEO::NSLOT) - 1) & 1);
    mbar_expect_tx(&cx.full[slot], T::W_BYTES);
    tma_load_1d(const_cast<unsigned char*>(cx.sW) + slot * T::W_BYTES, cx.packed + GEO::pkt_off(p), T::W_BYTES,
                &cx.full[slot]);
  }
}

// Wait for the group's MMAs: every thread on the mbarrier (default), or -- NLBAC_TC_BARWAIT -- the issuing thread alone
// with the rest of the group blocked in the hardware barrier (no polling instructions: SYNCS + BRA + NANOSLEEP are 16 %
// of the executed instructions of the reverse sweep, but they issue in otherwise idle slots; measured slower).
__device__ __forceinline__ void tc_wait_done(TcGroupCtx& cx) {
#if NLBAC_TC_BARWAIT
  if (cx.lt == 0) {
    mbar_wait(cx.done, cx.nd & 1);
    umma::fence_after_sync();
    umma::fence_before_sync();
  }
  group_sync(cx.g);
#else
  mbar_wait(cx.done, cx.nd & 1);
#endif
  ++cx.nd;
  umma::fence_after_sync();
}

template <class C>
__device__ __noinline__ void tc_issue_ts(uint32_t tmma, uint32_t ta, uint64_t wd, uint32_t acc0 = 0u) {   // A operand in tensor memory
  using T = TcCfg<C>;
  constexpr uint32_t idesc = umma::make_idesc_tf32(128, T::NP, false, false);
#pragma unroll
  for (int ks = 0; ks < T::KSTEPS; ++ks)
    umma::mma_tf32_ts(tmma, ta + 8 * ks, wd + (uint64_t)(ks * ((2 * T::WCH) >> 4)), idesc, ks > 0 ? 1u : acc0);
}

// Right-hand side for the group's tile: owner thread passes x / constants, receives k = dx/dt.
template <class C, bool SPLIT>
__device__ __forceinline__ void eval_rhs_tc(TcGroupCtx& cx, const float (&x)[C::NS], const float (&cst)[C::NCONST],
                                            float (&k)[C::NS]) {
  using T = TcCfg<C>;
  using G = TcFwdGeom<C, SPLIT>;
  const int lt = cx.lt;
  float fo[C::outp(0)];
#pragma unroll
  for (int c = 0; c < C::NCH; ++c) {
    // ---- thin first layer on the FP32 pipe -> activation image ----
    {
      const float* Wt0 = cx.sRes + C::off_wt0(c);
      const float* b0 = cx.sRes + C::off_b0(c);
#pragma unroll(TC_L0_UNROLL)
      for (int g8 = 0; g8 < T::KP / 8; ++g8) {
        Acc8 acc;
        acc8_init(acc, *reinterpret_cast<const float4*>(b0 + 8 * g8), *reinterpret_cast<const float4*>(b0 + 8 * g8 + 4));
#pragma unroll
        for (int d = 0; d < C::IN0; ++d) {
          const float xv = d < C::NS ? x[d < C::NS ? d : 0] : cst[d >= C::NS ? d - C::NS : 0];
          acc8_fma(acc, xv, *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8),
                   *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8 + 4));
        }
        float v[8];
        acc8_get(acc, v);
        if (SPLIT) {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
          act_store8<C, SPLIT>(cx, g8, v);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = relu_tf32_bias(v[i]);
          act_store8_raw<C>(cx, g8, v);
        }
      }
    }
    act_publish<SPLIT>(cx);
    const int n = C::nhc(c);
    float o[C::OUTP_MAX];
#pragma unroll 1
    for (int l = 1; l <= n; ++l) {
      if (cx.keeper) {
        ring_prefetch<G>(cx, cx.seq + G::NSLOT - 1);
        if (SPLIT) ring_prefetch<G>(cx, cx.seq + G::NSLOT);
      }
      if (lt == 0) {  // ---- MMA issue: one thread per group ----
        umma::fence_after_sync();
        {
          const uint32_t slot = cx.seq % G::NSLOT;
          mbar_wait(&cx.full[slot], (cx.seq / G::NSLOT) & 1);
          umma::fence_after_sync();
          const uint64_t wd = cx.wdesc + (uint64_t)(slot * (T::W_BYTES >> 4));
          if (cx.tmemA) {
            tc_issue_ts<C>(cx.tmma, cx.ta_mma, wd, 0u);                                   // A_hi W_hi
            if (SPLIT) tc_issue_ts<C>(cx.tmma, cx.ta_mma + T::TCOLS, wd, 1u);             // A_lo W_hi
          } else {
            tc_issue<C>(cx.tmma, cx.adesc, wd, 0u);
            if (SPLIT) tc_issue<C>(cx.tmma, cx.adesc + (uint64_t)(T::A_BYTES >> 4), wd, 1u);
          }
          umma::commit(&cx.empty[slot]);
        }
        if (SPLIT) {
          const uint32_t s2 = cx.seq + 1;
          const uint32_t slot = s2 % G::NSLOT;
          mbar_wait(&cx.full[slot], (s2 / G::NSLOT) & 1);
          umma::fence_after_sync();
          const uint64_t wd2 = cx.wdesc + (uint64_t)(slot * (T::W_BYTES >> 4));                                  // A_hi W_lo
          if (cx.tmemA) tc_issue_ts<C>(cx.tmma, cx.ta_mma, wd2, 1u);
          else tc_issue<C>(cx.tmma, cx.adesc, wd2, 1u);
          umma::commit(&cx.empty[slot]);
        }
        umma::commit(cx.done);
      }
      cx.seq += SPLIT ? 2 : 1;
      tc_wait_done(cx);
      const float* bias = cx.sRes + C::off_bh(c, l);
      constexpr int NG8 = T::KP / 8;
      if (l < n) {  // ---- epilogue: bias + ReLU -> activation image (in place) ----
        tc_batches<NG8, true>(cx.tacc, [&](int, int, int g8, float (&v)[8]) {
          if (!T::FOLD) {
            const float4 b0 = *reinterpret_cast<const float4*>(bias + 8 * g8);
            const float4 b1 = *reinterpret_cast<const float4*>(bias + 8 * g8 + 4);
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] += bb[i];
          }
          if (SPLIT) {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.f);
            act_store8<C, SPLIT>(cx, g8, v);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = relu_tf32_bias(v[i]);
            act_store8_raw<C>(cx, g8, v);
          }
        });
        act_publish<SPLIT>(cx);
      } else {  // ---- last hidden layer: bias + ReLU, thin output layer folded in registers ----
        const int OUTP = C::outp(c);
        const float* WtL = cx.sRes + C::off_wtl(c);
        const float* bL = cx.sRes + C::off_bl(c);
        unsigned long long o2[C::OUTP_MAX / 2];
#pragma unroll
        for (int j = 0; j < C::OUTP_MAX; j += 2) o2[j / 2] = j < OUTP ? pack2(bL[j], bL[j + 1]) : 0ull;
        tc_batches<NG8, true>(cx.tacc, [&](int, int, int g8, float (&v)[8]) {
          if (!T::FOLD) {
            const float4 b0 = *reinterpret_cast<const float4*>(bias + 8 * g8);
            const float4 b1 = *reinterpret_cast<const float4*>(bias + 8 * g8 + 4);
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] += bb[i];
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int kk = 8 * g8 + i;
            if (kk < C::H) {
              const float h = fmaxf(v[i], 0.f);
              const unsigned long long hh = pack2(h, h);
#pragma unroll
              for (int j = 0; j < C::OUTP_MAX; j += 4)
                if (j < OUTP) {
                  const float4 w = *reinterpret_cast<const float4*>(WtL + kk * OUTP + j);
                  o2[j / 2] = fma2(hh, pack2(w.x, w.y), o2[j / 2]);
                  o2[j / 2 + 1] = fma2(hh, pack2(w.z, w.w), o2[j / 2 + 1]);
                }
            }
          }
        });
#pragma unroll
        for (int j = 0; j < C::OUTP_MAX; j += 2) unpack2(o2[j / 2], o[j], o[j + 1]);
      }
    }
    if (c == 0) {
#pragma unroll
      for (int j = 0; j < C::outp(0); ++j) fo[j] = o[j];
    } else {
#pragma unroll
      for (int i = 0; i < C::NS; ++i) {
        float v = fo[i];
#pragma unroll
        for (int j = 0; j < C::NU; ++j) v = fmaf(o[i * C::NU + j], cst[j], v);
        k[i] = v;
      }
    }
  }
  if (C::NCH == 1) {
#pragma unroll
    for (int i = 0; i < C::NS; ++i) k[i] = fo[i];
  }
}

template <class C, bool SPLIT>
__global__ void __launch_bounds__(TcFwdGeom<C, SPLIT>::NT, 1) node_tc_fwd_kernel(FwdArgs a) {
  using T = TcCfg<C>;
  using G = TcFwdGeom<C, SPLIT>;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  float* sRes = reinterpret_cast<float*>(smem_raw + G::OFF_RES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + G::OFF_BAR);
  uint64_t* full = bars;
  uint64_t* empty = bars + G::NSLOT;
  uint64_t* done = bars + 2 * G::NSLOT;
  uint64_t* resbar = bars + 2 * G::NSLOT + G::G;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + G::NBAR);
  const int tid = threadIdx.x, warp = tid / 32;

  const int S = method_stages(a.method);
  const long long n_tiles = (a.B + T::ROWS - 1) / T::ROWS;
  const long long n_pairs = (n_tiles + G::G - 1) / G::G;           // CTA-level work items (G tiles each)
  long long my_iters = 0;
  if ((long long)blockIdx.x < n_pairs) my_iters = (n_pairs - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const uint32_t total = (uint32_t)(my_iters * a.n_steps * S * G::PKT_PER_EVAL);

  if (tid == 0) {
    for (int i = 0; i < G::NSLOT; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], G::G); }
    for (int i = 0; i < G::G; ++i) mbar_init(&done[i], 1);
    mbar_init(resbar, 1);
    fence_barrier_init();
  }
  if (warp == 0) umma::tmem_alloc(tslot, G::TMEM_COLS);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;

  if (my_iters > 0) {
    TcGroupCtx cx;
    cx.g = tid / 128;
    cx.lt = tid % 128;
    cx.tmemA = cx.g >= G::GS;
    cx.sA = smem_raw + G::OFF_A + (cx.tmemA ? 0 : cx.g) * G::NIMG * T::A_BYTES;
    cx.sW = smem_raw + G::OFF_W;
    cx.sRes = sRes;
    cx.full = full; cx.empty = empty; cx.done = &done[cx.g];
    cx.tmma = tbase + cx.g * T::TCOLS;
    cx.tacc = cx.tmma + ((uint32_t)((warp % 4) * 32) << 16);
    cx.ta_mma = tbase + (G::G + (cx.tmemA ? cx.g - G::GS : 0) * G::NIMG) * T::TCOLS;
    cx.ta_row = cx.ta_mma + ((uint32_t)((warp % 4) * 32) << 16);
    cx.seq = 0; cx.nd = 0;
    cx.adesc = umma::make_desc(smem_u32(cx.sA), T::ACH, 128);
    cx.wdesc = umma::make_desc(smem_u32(cx.sW), T::WCH, 128);
    cx.total = total; cx.packed = a.packed;
    cx.keeper = tid == 32;
    if (cx.keeper) {   // thin layers once, then the first packets of the weight ring
      mbar_expect_tx(resbar, C::RES_FWD * 4);
      tma_load_1d(sRes, a.packed, C::RES_FWD * 4, resbar);
      for (uint32_t p = 0; p + 1 < (uint32_t)G::NSLOT; ++p) ring_prefetch<G>(cx, p);
    }
    mbar_wait(resbar, 0);

    for (long long it = 0; it < my_iters; ++it) {
      const long long tile = (it * gridDim.x + blockIdx.x) * G::G + cx.g;
      const long long b = tile * T::ROWS + cx.lt;
      const bool live = b < a.B;     // dead rows / dead tiles still run (the ring is shared by the groups)
      float y[C::NS], cst[C::NCONST], k1[C::NS], k2[C::NS], k3[C::NS], k4[C::NS], x[C::NS];
#pragma unroll
      for (int d = 0; d < C::NS; ++d) k1[d] = k2[d] = k3[d] = k4[d] = y[d] = 0.f;
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) cst[d] = 0.f;
      if (live) {
        const float* yr = a.y0 + b * C::D;
#pragma unroll
        for (int d = 0; d < C::NS; ++d) y[d] = yr[d];
#pragma unroll
        for (int d = 0; d < C::NCONST; ++d) cst[d] = yr[C::NS + d];
        float* o = a.y_out + b * C::D;
#pragma unroll
        for (int d = 0; d < C::NS; ++d) o[d] = y[d];
#pragma unroll
        for (int d = 0; d < C::NCONST; ++d) o[C::NS + d] = cst[d];
      }
      int jout = 1;
      for (int step = 0; step < a.n_steps; ++step) {
        const float dt = a.dts[step];
        float y1[C::NS];
#pragma unroll 1
        for (int stage = 0; stage < S; ++stage) {
          if (stage == 0) {
#pragma unroll
            for (int d = 0; d < C::NS; ++d) x[d] = y[d];
          } else if (a.method == METHOD_MIDPOINT) {
            const float half = 0.5f * dt;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) x[d] = y[d] + k1[d] * half;
          } else {
            const float third = 1.0f / 3.0f;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) {
              if (stage == 1) x[d] = y[d] + dt * k1[d] * third;
              else if (stage == 2) x[d] = y[d] + dt * (k2[d] - k1[d] * third);
              else x[d] = y[d] + dt * (k1[d] - k2[d] + k3[d]);
            }
          }
          if (a.ckpt != nullptr && live) {
            float* p = a.ckpt + (((long long)step * S + stage) * a.B + b) * C::NS;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) p[d] = x[d];
          }
          float kn[C::NS];
          eval_rhs_tc<C, SPLIT>(cx, x, cst, kn);
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
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, G::TMEM_COLS);
}

}  // namespace nlbac
