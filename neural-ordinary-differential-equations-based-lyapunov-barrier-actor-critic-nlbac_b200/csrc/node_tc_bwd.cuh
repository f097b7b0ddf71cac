// K2-TC — backward of the fixed-grid NODE solve on the 5th-generation tensor cores, TF32 tier (1e-2).
//
// Two passes:
//   pass A  node_tc_bwd_kernel: the reverse sweep of node_bwd.cuh with the hidden-layer GEMMs on tcgen05
//           (kind::tf32): per (step, stage) the tile's activations are recomputed from the checkpointed stage
//           input, the cotangent is pulled back through the layers (dX GEMMs against the transposed weight
//           images), the owner thread advances the Runge-Kutta adjoint recurrences in registers.  Every
//           activation a_i and every pre-activation cotangent d_i is ALSO written, rounded to bf16, to a
//           global "image" that is already in the UMMA MN-major core-matrix layout (fully coalesced stores).
//   pass B  node_tc_dw_kernel: parameter gradients as streaming GEMMs over those images,
//           dW = sum over (trajectory, step, stage) of d^T a, on tcgen05 kind::f16 (bf16 operands, fp32
//           accumulation in TMEM; tools/prec_emul.py: bf16 here adds 1e-4 to a 3.4e-3 tier).  A ones column in
//           every activation image yields the bias gradients from the same GEMM.  HBM-bound by construction.
#pragma once
#include "node_bwd.cuh"
#include "node_tc.cuh"
#include <cuda_bf16.h>

namespace nlbac {

// Layout of the bf16 images of one (tile of 128 trajectories, stage evaluation), in bytes.
//   big activation image  a_i   [128 x IMGW]  IMGW = round_up(H + 1, 16): features, a ones column at H, zeros
//   big cotangent image   d_i   [128 x KP]
//   thin images           xin [128 x 16] (stage input incl. constants for black-box nets, ones column at IN0),
//                         dOut_c [128 x 16] (cotangent of the chain output)
// element (t, c) of an image lives at (c / 8) * 2048 + (t / 8) * 128 + (t % 8) * 16 + (c % 8) * 2.
// SPLIT (3xTF32 tier): every image exists twice, bf16 hi and bf16 lo = bf16(v - hi) (16 mantissa bits together), the lo
// block LO bytes after the hi block; pass B then forms d_hi a_hi + d_hi a_lo + d_lo a_hi.
template <class C, bool SPLIT = false>
struct ImgMap {
  using T = TcCfg<C>;
  static constexpr int IMGW = round_up(C::H + 1, 16);              // MMA N of an activation image (multiple of 16)
  // chunks (8 columns = 2048 B) actually written by pass A and streamed by pass B; the MMA addresses IMGW / 8 (or 2)
  // chunks, the unloaded ones only feed result columns / rows that the epilogue drops
  static constexpr int ACT_CH = round_up(C::H + 1, 8) / 8, DEL_CH = T::KP / 8;
  static constexpr int XIN_CH = round_up(C::IN0 + 1, 8) / 8, DOUT_CH = round_up(C::OUTP_MAX, 8) / 8;
  static constexpr int ACT_B = ACT_CH * 2048, DEL_B = DEL_CH * 2048, XIN_B = XIN_CH * 2048, DOUT_B = DOUT_CH * 2048;
  static_assert(C::IN0 < 16 && C::OUTP_MAX <= 16, "thin images hold 16 columns");
  __host__ __device__ static constexpr int nl1(int c) { return C::nhc(c) + 1; }   // activations / cotangents per chain
  __host__ __device__ static constexpr int chain_bytes(int c) { return nl1(c) * (ACT_B + DEL_B) + DOUT_B; }
  __host__ __device__ static constexpr int off_xin() { return 0; }
  __host__ __device__ static constexpr int chain_base(int c) { return XIN_B + (c == 0 ? 0 : chain_bytes(0)); }
  __host__ __device__ static constexpr int off_act(int c, int i /*1..nl1*/) { return chain_base(c) + (i - 1) * ACT_B; }
  __host__ __device__ static constexpr int off_del(int c, int i) { return chain_base(c) + nl1(c) * ACT_B + (i - 1) * DEL_B; }
  __host__ __device__ static constexpr int off_dout(int c) { return chain_base(c) + nl1(c) * (ACT_B + DEL_B); }
  static constexpr int STRIDE1 = XIN_B + chain_bytes(0) + (C::NCH == 2 ? chain_bytes(1) : 0);
  static constexpr int LO = STRIDE1;                         // byte offset of the lo block
  static constexpr int STRIDE = (SPLIT ? 2 : 1) * STRIDE1;
  // GEMM pairs of pass B: per chain  (d_1 | xin), (d_{l+1} | a_l) l = 1..nhc, (dOut | a_{nl1})
  __host__ __device__ static constexpr int npairs(int c) { return C::nhc(c) + 2; }
  static constexpr int NPAIR = npairs(0) + (C::NCH == 2 ? npairs(1) : 0);
};

struct __align__(16) Pack8 { __nv_bfloat162 a, b, c, d; };

// 8 consecutive columns [c0, c0+8) of row t of a bf16 image (one 16-byte store; a warp writes 512 contiguous bytes)
__device__ __forceinline__ void store_img8(unsigned char* img, int t, int c0, const float (&v)[8]) {
  Pack8 p;
  p.a = __floats2bfloat162_rn(v[0], v[1]); p.b = __floats2bfloat162_rn(v[2], v[3]);
  p.c = __floats2bfloat162_rn(v[4], v[5]); p.d = __floats2bfloat162_rn(v[6], v[7]);
  *reinterpret_cast<Pack8*>(img + (c0 >> 3) * 2048 + (t >> 3) * 128 + (t & 7) * 16) = p;
}

// same with ReLU applied by the conversion itself (cvt.rn.relu.bf16x2.f32: negative inputs and NaN give +0)
__device__ __forceinline__ uint32_t pack_relu_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void store_img8_relu(unsigned char* img, int t, int c0, const float (&v)[8]) {
  uint4 p;
  p.x = pack_relu_bf16x2(v[0], v[1]); p.y = pack_relu_bf16x2(v[2], v[3]);
  p.z = pack_relu_bf16x2(v[4], v[5]); p.w = pack_relu_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(img + (c0 >> 3) * 2048 + (t >> 3) * 128 + (t & 7) * 16) = p;
}

// hi / lo bf16 pair of eight values: hi = bf16(v), lo = bf16(v - hi)
__device__ __forceinline__ void split_bf16x8(const float (&v)[8], uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const __nv_bfloat162 hb = __floats2bfloat162_rn(v[2 * p], v[2 * p + 1]);
    const float2 hf = __bfloat1622float2(hb);
    const __nv_bfloat162 lb = __floats2bfloat162_rn(v[2 * p] - hf.x, v[2 * p + 1] - hf.y);
    h[p] = *reinterpret_cast<const uint32_t*>(&hb);
    l[p] = *reinterpret_cast<const uint32_t*>(&lb);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}
// image store of either tier: SPLIT writes the hi block and, LO bytes further, the lo block; RELU applies max(v, 0) first
template <bool SPLIT, bool RELU, int LO>
__device__ __forceinline__ void img_store8(unsigned char* img, int t, int c0, const float (&v)[8]) {
  if (!SPLIT) {
    if (RELU) store_img8_relu(img, t, c0, v);
    else store_img8(img, t, c0, v);
  } else {
    float r[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = RELU ? fmaxf(v[i], 0.f) : v[i];
    uint4 hi, lo;
    split_bf16x8(r, hi, lo);
    unsigned char* p = img + (c0 >> 3) * 2048 + (t >> 3) * 128 + (t & 7) * 16;
    *reinterpret_cast<uint4*>(p) = hi;
    *reinterpret_cast<uint4*>(p + LO) = lo;
  }
}

// SPLIT = 3xTF32 reverse sweep (fp32-class): every operand image exists twice (hi = TF32-rounded, lo = residual) and a
// hidden-layer product is hi*hi + lo*hi + hi*lo, i.e. three MMA sets over TWO ring packets (W_hi then W_lo).  One group
// keeps its hi / lo operand images in shared memory (2 x 53 KB), one keeps them in tensor memory (2 x 128 columns):
// 2 accumulators + 2 operand blocks = all 512 TMEM columns.
template <class C, bool SPLIT = false>
struct TcBwdGeom {
  using T = TcCfg<C>;
  static constexpr int NSLOT = 2;
  static constexpr int NIMG = SPLIT ? 2 : 1;
  static constexpr int fixed_bytes = NSLOT * T::W_BYTES + C::RES_ALL * 4 + 256;
  static constexpr int GS = (!SPLIT && fixed_bytes + 2 * T::A_BYTES <= 227 * 1024) ? 2 : 1;   // operand image(s) in shared memory
  static constexpr int GT = 1;                                                                  // operand in tensor memory
  static constexpr int G = GS + GT;
  static constexpr int NT = G * 128;
  static constexpr int TMEM_COLS = tmem_cols(G, GT, NIMG);
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W = OFF_A + GS * NIMG * T::A_BYTES;
  static constexpr int OFF_RES = OFF_W + NSLOT * T::W_BYTES;
  static constexpr int OFF_BAR = OFF_RES + C::RES_ALL * 4;
  static constexpr int NBAR = 2 * NSLOT + G + 1;
  static constexpr size_t SMEM_BYTES = OFF_BAR + NBAR * 8 + 16;
  static constexpr int PKT_PER_LAYER = SPLIT ? 2 : 1;
  static constexpr int PKT_PER_EVAL = 2 * C::NHID * PKT_PER_LAYER;   // every hidden matrix once forward, once transposed
  static constexpr int NLAY = C::NL_MAX - 1;         // activations per chain (masks kept in registers)
  static_assert(OFF_W % 128 == 0 && OFF_RES % 16 == 0 && OFF_BAR % 8 == 0, "alignment");
  static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
  static_assert(TMEM_COLS <= 512, "tensor memory budget");
  __device__ static int pkt_off(uint32_t p) {
    int q = p % PKT_PER_EVAL, c = 0;
    constexpr int per0 = 2 * C::nhc(0) * PKT_PER_LAYER;
    if (C::NCH == 2 && q >= per0) { c = 1; q -= per0; }
    const int n = C::nhc(c);
    const int lo = SPLIT ? (q & 1) : 0;      // second packet of a layer = the residual image
    q /= PKT_PER_LAYER;
    const int hb = C::hid_base(c);
    if (q < n) return lo ? T::off_lo(hb + q) : T::off_hi(hb + q);
    const int h = hb + (n - 1 - (q - n));
    return lo ? T::off_trlo(h) : T::off_tr(h);
  }
};

// the ones column of an activation image (bias gradients) sits at column H
template <class C, bool SPLIT = false>
__device__ __forceinline__ void finish_act_img(unsigned char* img, int t) {
  if (C::H == TcCfg<C>::KP) {   // no padding column inside the K extent: the ones column opens a new chunk
    float v[8] = {1.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    img_store8<SPLIT, false, ImgMap<C, SPLIT>::LO>(img, t, C::H, v);
  }
}

// one hidden-layer GEMM of the group's tile: A = the group's activation image, B = next ring packet
template <class C, class GEO, bool SPLIT = false>
__device__ __forceinline__ void tc_layer_mma(TcGroupCtx& cx) {
  using T = TcCfg<C>;
  constexpr int NSLOT = GEO::NSLOT;
  if (cx.keeper) {
    ring_prefetch<GEO>(cx, cx.seq + NSLOT - 1);
    if (SPLIT) ring_prefetch<GEO>(cx, cx.seq + NSLOT);
  }
#ifdef NLBAC_TC_PROF
  long long p0 = clock64(), p1 = p0, p2 = p0;
  cx.pf_epi += p0 - cx.pf_t;
#endif
  if (cx.lt == 0) {
    umma::fence_after_sync();
    const uint32_t slot = cx.seq % NSLOT;
    mbar_wait(&cx.full[slot], (cx.seq / NSLOT) & 1);
    umma::fence_after_sync();
#ifdef NLBAC_TC_PROF
    p1 = clock64();
#endif
    const uint64_t wd = cx.wdesc + (uint64_t)(slot * (T::W_BYTES >> 4));
    if (cx.tmemA) {
      tc_issue_ts<C>(cx.tmma, cx.ta_mma, wd, 0u);                                       // A_hi W_hi
      if (SPLIT) tc_issue_ts<C>(cx.tmma, cx.ta_mma + T::TCOLS, wd, 1u);                 // A_lo W_hi
    } else {
      tc_issue<C>(cx.tmma, cx.adesc, wd, 0u);
      if (SPLIT) tc_issue<C>(cx.tmma, cx.adesc + (uint64_t)(T::A_BYTES >> 4), wd, 1u);
    }
    umma::commit(&cx.empty[slot]);
    if (SPLIT) {                                                                         // A_hi W_lo
      const uint32_t s2 = cx.seq + 1, slot2 = s2 % NSLOT;
      mbar_wait(&cx.full[slot2], (s2 / NSLOT) & 1);
      umma::fence_after_sync();
      const uint64_t wd2 = cx.wdesc + (uint64_t)(slot2 * (T::W_BYTES >> 4));
      if (cx.tmemA) tc_issue_ts<C>(cx.tmma, cx.ta_mma, wd2, 1u);
      else tc_issue<C>(cx.tmma, cx.adesc, wd2, 1u);
      umma::commit(&cx.empty[slot2]);
    }
    umma::commit(cx.done);
#ifdef NLBAC_TC_PROF
    p2 = clock64();
#endif
  }
  cx.seq += SPLIT ? 2 : 1;
  tc_wait_done(cx);
#ifdef NLBAC_TC_PROF
  cx.pf_t = clock64();
  cx.pf_full += p1 - p0; cx.pf_issue += p2 - p1; cx.pf_done += cx.pf_t - p2; ++cx.pf_n;
#endif
}

__device__ __forceinline__ void tc_publish(TcGroupCtx& cx) { act_publish<false>(cx); }

// ReLU masks: 8 bits per group of 8 columns, 4 groups per word, 4 words per layer -- kept in registers and
// addressed with selects so that the column loops can stay rolled (instruction-cache footprint)
__device__ __forceinline__ uint32_t sel4(const uint32_t (&m)[4], int q) {
  return q == 0 ? m[0] : (q == 1 ? m[1] : (q == 2 ? m[2] : m[3]));
}
__device__ __forceinline__ void or4(uint32_t (&m)[4], int q, uint32_t w) {
  m[0] |= q == 0 ? w : 0u; m[1] |= q == 1 ? w : 0u; m[2] |= q == 2 ? w : 0u; m[3] |= q == 3 ? w : 0u;
}

// ReLU mask of eight PRE-activations: bit i = !(sign bit of v[i]), gathered by funnel shifts (one instruction per
// element).  An exactly-zero pre-activation therefore passes its cotangent (relu'(0) = 1 here, 0 in PyTorch): a
// measure-zero set inside a tier whose operand rounding already moves pre-activations across zero; the padded
// columns and dead rows where exact zeros do occur carry zero cotangents anyway.
__device__ __forceinline__ uint32_t pos_bits8(const float (&v)[8]) {
  uint32_t lo = 0u, hi = 0u;
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    lo = __funnelshift_l(__float_as_uint(v[i]), lo, 1);
    hi = __funnelshift_l(__float_as_uint(v[i + 4]), hi, 1);
  }
  return (lo | (hi << 4)) ^ 0xFFu;
}

// Epilogue of a forward layer of the sweep for 8 pre-activations v (bias included): ReLU mask bits -> mk, the
// activation -> the group's MMA operand (ReLU + TF32 rounding in one integer instruction) and, rounded to bf16 with
// the ReLU folded into the conversion, -> the HBM image.  The ones column of the image (bias gradients, column H)
// needs no special case where the bias is folded (TcCfg::FOLD): the pre-activation of column H is the constant 1.
template <class C, bool SPLIT = false>
__device__ __forceinline__ void relu_publish8(const TcGroupCtx& cx, uint32_t& mword, int u, int g8,
                                              float (&v)[8], unsigned char* ia) {
  using T = TcCfg<C>;
  mword |= pos_bits8(v) << (8 * u);     // the batch's mask word; the caller files it once per batch (or4)
  if (ia != nullptr) img_store8<SPLIT, true, ImgMap<C, SPLIT>::LO>(ia, cx.lt, 8 * g8, v);
  float r[8];
  if (SPLIT) {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = fmaxf(v[i], 0.f);
    act_store8<C, true>(cx, g8, r);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = relu_tf32_bias(v[i]);
    act_store8_raw<C>(cx, g8, r);
  }
}

// One stage of the reverse sweep for the group's tile (mirrors vjp_rhs_tile of node_bwd.cuh).
// img: global image block of this (tile, stage evaluation) or nullptr (no parameter gradients wanted).
#ifndef NLBAC_TC_ROLL_CHAINS
#define NLBAC_TC_ROLL_CHAINS 1
#endif
constexpr bool TC_ROLL_CHAINS = NLBAC_TC_ROLL_CHAINS != 0;   // (macros are not expanded inside #pragma unroll)
#ifndef NLBAC_TC_SPLIT_COMPACT
#define NLBAC_TC_SPLIT_COMPACT 1   // the 3xTF32 sweep visits the accumulator with tc_pairs (instruction-fetch bound)
#endif
template <int NG8, bool LOAD, bool COMPACT>
struct TcVisit {
  template <class F, class P = NoPost>
  __device__ __forceinline__ void operator()(uint32_t tacc, F&& body, P&& post = P()) const {
    if constexpr (COMPACT && NLBAC_TC_SPLIT_COMPACT == 2) tc_singles<NG8, LOAD>(tacc, body, post);
    else if constexpr (COMPACT) tc_pairs<NG8, LOAD>(tacc, body, post);
    else tc_batches<NG8, LOAD>(tacc, body, post);
  }
};
#ifndef NLBAC_TC_COMPACT_ALL
#define NLBAC_TC_COMPACT_ALL 0
#endif
#define TC_VISIT(LOAD) TcVisit<NG8, LOAD, ((SPLIT && NLBAC_TC_SPLIT_COMPACT) || NLBAC_TC_COMPACT_ALL)>()

template <class C, bool SPLIT = false>
__device__ __forceinline__ void vjp_rhs_tc(TcGroupCtx& cx, unsigned char* img, const float (&x)[C::NS],
                                           const float (&cst)[C::NCONST], const float (&dk)[C::NS],
                                           float (&dx)[C::NS], float (&dcst)[C::NCONST]) {
  using T = TcCfg<C>;
  using G = TcBwdGeom<C, SPLIT>;
  using IM = ImgMap<C, SPLIT>;
  constexpr int NG8 = T::KP / 8;
  static_assert(NG8 <= 16, "mask words");
  const int lt = cx.lt;
  const bool wimg = img != nullptr;
  unsigned long long dxa2[C::IN0];   // dL/d(stage input), even / odd hidden units in the two halves
#pragma unroll
  for (int d = 0; d < C::IN0; ++d) dxa2[d] = 0ull;
  if (wimg) {   // stage input image (shared by the chains): [x | const | 1 | 0...]
    float v0[8], v1[8];
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      float val = 0.f;
      if (d < C::NS) val = x[d < C::NS ? d : 0];
      else if (d < C::IN0) val = cst[(d >= C::NS && d < C::IN0) ? d - C::NS : 0];
      else if (d == C::IN0) val = 1.f;
      if (d < 8) v0[d] = val; else v1[d - 8] = val;
    }
    img_store8<SPLIT, false, IM::LO>(img + IM::off_xin(), lt, 0, v0);
    if (IM::XIN_CH > 1) img_store8<SPLIT, false, IM::LO>(img + IM::off_xin(), lt, 8, v1);
  }
  // 3xTF32 sweep (NLBAC_TC_ROLL_CHAINS): the f / g chains as a ROLLED loop (per-chain sizes and offsets become run-time
  // values): 141 -> 102 KB of SASS and 237 -> 224 ms at the bench size, because that sweep is instruction-fetch bound
  // (DESIGN.md 5).  The TF32 sweep (104 -> 79 KB) measured neutral (103.6 vs 102.3 ms) and stays unrolled.
#pragma unroll((SPLIT && TC_ROLL_CHAINS) ? 1 : 2)
  for (int c = 0; c < C::NCH; ++c) {
    const int n = C::nhc(c);
    const int OUT = C::out(c), OUTP = C::outp(c);
    uint32_t mk[G::NLAY][4];
#pragma unroll
    for (int i = 0; i < G::NLAY; ++i)
#pragma unroll
      for (int w = 0; w < 4; ++w) mk[i][w] = 0u;
    uint32_t mword = 0u;
    // ---------------- thin first layer -> a_1 ----------------
    {
      const float* Wt0 = cx.sRes + C::off_wt0(c);
      const float* b0 = cx.sRes + C::off_b0(c);
      unsigned char* ia = img + IM::off_act(c, 1);
      TC_VISIT(false)(cx.tacc, [&](int q, int u, int g8, float (&v)[8]) {
        Acc8 acc;
        acc8_init(acc, *reinterpret_cast<const float4*>(b0 + 8 * g8), *reinterpret_cast<const float4*>(b0 + 8 * g8 + 4));
#pragma unroll
        for (int d = 0; d < C::IN0; ++d) {
          const float xv = d < C::NS ? x[d < C::NS ? d : 0] : cst[d >= C::NS ? d - C::NS : 0];
          acc8_fma(acc, xv, *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8),
                   *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8 + 4));
        }
        acc8_get(acc, v);
        relu_publish8<C, SPLIT>(cx, mword, u, g8, v, wimg ? ia : nullptr);
      }, [&](int q) { or4(mk[0], q, mword); mword = 0u; });
      if (wimg) finish_act_img<C, SPLIT>(ia, lt);
    }
    tc_publish(cx);
    // cotangent of the chain output
    float dout[C::OUTP_MAX];
    unsigned long long dc2[C::NU > 0 ? C::NU : 1];   // chain 1: dL/du_j, even / odd hidden units in the two halves
#pragma unroll
    for (int o = 0; o < C::OUTP_MAX; ++o) dout[o] = 0.f;
#pragma unroll
    for (int j = 0; j < (C::NU > 0 ? C::NU : 1); ++j) dc2[j] = 0ull;
    if (c == 0) {
#pragma unroll
      for (int o = 0; o < C::NS; ++o) dout[o] = dk[o];
    } else {
#pragma unroll
      for (int i = 0; i < C::NS; ++i)
#pragma unroll
        for (int j = 0; j < C::NU; ++j) dout[i * C::NU + j] = dk[i] * cst[j];
    }
    if (wimg) {
      float v0[8], v1[8];
#pragma unroll
      for (int o = 0; o < 16; ++o) {
        const float val = o < C::OUTP_MAX ? dout[o < C::OUTP_MAX ? o : 0] : 0.f;
        if (o < 8) v0[o] = val; else v1[o - 8] = val;
      }
      img_store8<SPLIT, false, IM::LO>(img + IM::off_dout(c), lt, 0, v0);
      if (IM::DOUT_CH > 1) img_store8<SPLIT, false, IM::LO>(img + IM::off_dout(c), lt, 8, v1);
    }
    // ---------------- hidden layers forward: a_{l+1} = relu(H_l a_l + b_l) ----------------
#pragma unroll
    for (int l = 1; l <= G::NLAY - 1; ++l) {
      if (l <= n) {
        tc_layer_mma<C, G, SPLIT>(cx);
        const float* bias = cx.sRes + C::off_bh(c, l);
        unsigned char* ia = img + IM::off_act(c, l + 1);
        if (l < n) {
          TC_VISIT(true)(cx.tacc, [&](int q, int u, int g8, float (&v)[8]) {
            if (!T::FOLD) {
              const float4 ba = *reinterpret_cast<const float4*>(bias + 8 * g8);
              const float4 bb = *reinterpret_cast<const float4*>(bias + 8 * g8 + 4);
              v[0] += ba.x; v[1] += ba.y; v[2] += ba.z; v[3] += ba.w; v[4] += bb.x; v[5] += bb.y; v[6] += bb.z; v[7] += bb.w;
            }
            relu_publish8<C, SPLIT>(cx, mword, u, g8, v, wimg ? ia : nullptr);
          }, [&](int q) { or4(mk[l], q, mword); mword = 0u; });
        } else {
          // last hidden activation: forward value of g(x) (action gradient) and d_{n+1} = relu' * WL^T dOut
          const float* WnL = cx.sRes + C::off_wnl(c);
          unsigned char* id = img + IM::off_del(c, n + 1);
          TC_VISIT(true)(cx.tacc, [&](int q, int u, int g8, float (&v)[8]) {
            if (!T::FOLD) {
              const float4 ba = *reinterpret_cast<const float4*>(bias + 8 * g8);
              const float4 bb = *reinterpret_cast<const float4*>(bias + 8 * g8 + 4);
              v[0] += ba.x; v[1] += ba.y; v[2] += ba.z; v[3] += ba.w; v[4] += bb.x; v[5] += bb.y; v[6] += bb.z; v[7] += bb.w;
            }
            float dl[8];
            Acc8 da;
            acc8_zero(da);
            if (c == 0) {
              // d_{n+1}[k] = sum_o dOut[o] WL[o][k]: two 16-byte weight loads per output row and column group
#pragma unroll
              for (int o = 0; o < C::OUTP_MAX; ++o)
                if (o < OUT)
                  acc8_fma(da, dout[o], *reinterpret_cast<const float4*>(WnL + o * C::HP + 8 * g8),
                           *reinterpret_cast<const float4*>(WnL + o * C::HP + 8 * g8 + 4));
            } else {
              // chain g, output (i, j) -> dOut = dk[i] u_j.  With q_j[k] = sum_i dk[i] WL[(i,j)][k]:
              //   d_{n+1}[k] = sum_j u_j q_j[k]        and        dL/du_j += sum_k a_{n+1}[k] q_j[k]
              // (the second is the forward value g(x) contracted with dk without ever forming g(x))
              unsigned long long a2[4];
#pragma unroll
              for (int p = 0; p < 4; ++p) a2[p] = pack2(fmaxf(v[2 * p], 0.f), fmaxf(v[2 * p + 1], 0.f));
#pragma unroll
              for (int j = 0; j < C::NU; ++j) {
                Acc8 qa;
                acc8_zero(qa);
#pragma unroll
                for (int i = 0; i < C::NS; ++i)
                  acc8_fma(qa, dk[i], *reinterpret_cast<const float4*>(WnL + (i * C::NU + j) * C::HP + 8 * g8),
                           *reinterpret_cast<const float4*>(WnL + (i * C::NU + j) * C::HP + 8 * g8 + 4));
                const unsigned long long uu = pack2(cst[j], cst[j]);
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                  da.p[p] = fma2(uu, qa.p[p], da.p[p]);
                  dc2[j] = fma2(a2[p], qa.p[p], dc2[j]);
                }
              }
            }
            acc8_get(da, dl);
#pragma unroll
            for (int i = 0; i < 8; ++i) dl[i] = (v[i] > 0.f && 8 * g8 + i < C::H) ? dl[i] : 0.f;
            act_store8<C, SPLIT>(cx, g8, dl);
            if (wimg) {
              img_store8<SPLIT, false, IM::LO>(id, lt, 8 * g8, dl);
              img_store8<SPLIT, true, IM::LO>(ia, lt, 8 * g8, v);
            }
          });
        }
        if (wimg) finish_act_img<C, SPLIT>(ia, lt);
        tc_publish(cx);
      }
    }
    if (c == 1) {   // dL/du_j += sum_i g(x)[i][j] dk[i]   (bias of the last layer included)
      const float* bL = cx.sRes + C::off_bl(c);
#pragma unroll
      for (int j = 0; j < C::NU; ++j) {
        float acc = pair_sum(dc2[j]);
#pragma unroll
        for (int i = 0; i < C::NS; ++i) acc = fmaf(bL[i * C::NU + j], dk[i], acc);
        dcst[j] += acc;
      }
    }
    // ---------------- hidden layers backward: d_l = relu'(a_l) * H_l^T d_{l+1} ----------------
#pragma unroll
    for (int l = G::NLAY - 1; l >= 1; --l) {
      if (l <= n) {
        tc_layer_mma<C, G, SPLIT>(cx);
        const float* Wt0 = cx.sRes + C::off_wt0(c);
        unsigned char* id = img + IM::off_del(c, l);
        TC_VISIT(true)(cx.tacc, [&](int q, int u, int g8, float (&v)[8]) {
          const uint32_t bits = (sel4(mk[l - 1], q) >> (8 * u)) & 0xFFu;
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = ((bits >> i) & 1u) ? v[i] : 0.f;
          if (wimg) img_store8<SPLIT, false, IM::LO>(id, lt, 8 * g8, v);
          if (l > 1) {
            act_store8<C, SPLIT>(cx, g8, v);
          } else {   // d_1: thin first layer, dx = W0^T d_1 (rows of the transposed image; padded columns are zero)
            unsigned long long d2[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) d2[p] = pack2(v[2 * p], v[2 * p + 1]);
#pragma unroll
            for (int d = 0; d < C::IN0; ++d) {
              const float4 wa = *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8);
              const float4 wb = *reinterpret_cast<const float4*>(Wt0 + d * C::HP + 8 * g8 + 4);
              dxa2[d] = fma2(d2[0], pack2(wa.x, wa.y), dxa2[d]); dxa2[d] = fma2(d2[1], pack2(wa.z, wa.w), dxa2[d]);
              dxa2[d] = fma2(d2[2], pack2(wb.x, wb.y), dxa2[d]); dxa2[d] = fma2(d2[3], pack2(wb.z, wb.w), dxa2[d]);
            }
          }
        });
        if (l > 1) tc_publish(cx);
      }
    }
  }
#pragma unroll
  for (int d = 0; d < C::NS; ++d) dx[d] = pair_sum(dxa2[d]);
  if constexpr (C::KIND == KIND_BLACKBOX) {
#pragma unroll
    for (int d = 0; d < C::NCONST; ++d) dcst[d] += pair_sum(dxa2[C::NS + d]);
  }
}

// ------------------------------------------------------------------------------------------------
// pass B: parameter gradients as streaming bf16 GEMMs over the images (one GEMM pair per blockIdx.y)
// ------------------------------------------------------------------------------------------------
struct DwPair {
  int offD, bytesD;   // cotangent image (A operand, M = its feature columns)
  int offA, bytesA;   // activation image (B operand, N = its columns)
  int N;              // MMA N
  int rows, cols;     // valid extent of the result: rows of W, input columns of W
  int goff_w, goff_b, ldw;
};
template <class C>
__host__ __device__ inline DwPair dw_pair(int pr) {
  using IM = ImgMap<C>;
  int c = 0;
  if (C::NCH == 2 && pr >= IM::npairs(0)) { c = 1; pr -= IM::npairs(0); }
  const int n = C::nhc(c);
  DwPair p;
  if (pr == 0) {
    p = {IM::off_del(c, 1), IM::DEL_B, IM::off_xin(), IM::XIN_B, 16, C::H, C::IN0, C::goff_w(c, 0), C::goff_b(c, 0), C::IN0};
  } else if (pr <= n) {
    p = {IM::off_del(c, pr + 1), IM::DEL_B, IM::off_act(c, pr), IM::ACT_B, IM::IMGW, C::H, C::H, C::goff_w(c, pr),
         C::goff_b(c, pr), C::H};
  } else {
    p = {IM::off_dout(c), IM::DOUT_B, IM::off_act(c, n + 1), IM::ACT_B, IM::IMGW, C::out(c), C::H, C::goff_w(c, n + 1),
         C::goff_b(c, n + 1), C::H};
  }
  return p;
}

template <class C>
struct DwGeom {
  using IM = ImgMap<C>;
  static constexpr int NST = 3;
  static constexpr int DBUF = 16 * 2048;            // M = 128 feature rows are addressed (13 or 2 chunks are loaded)
  static constexpr int ABUF = (IM::IMGW / 8) * 2048;   // the MMA addresses IMGW columns; ACT_B of them are loaded
  static constexpr int STAGE = DBUF + ABUF;
  static constexpr int OFF_BAR = NST * STAGE;
  static constexpr size_t SMEM_BYTES = OFF_BAR + (2 * NST + 1) * 8 + 16;
  static constexpr int NT = 160;
  static_assert(STAGE % 128 == 0 && SMEM_BYTES <= 227 * 1024, "pass-B staging");
};

// The body of pass B for ONE GEMM pair: CTA `slab` of `nslab` streams blocks slab, slab + nslab, ... of `images`.
// Called by every thread of the CTA (threads beyond DwGeom::NT only take part in the two block barriers), either from
// node_tc_dw_kernel or from the dW-role CTAs of node_tc_bwd_kernel.
template <class C>
__device__ __forceinline__ void dw_role(unsigned char* smem_raw, const unsigned char* __restrict__ images, long long n_items,
                                        float* __restrict__ grad_rep, int pair_idx, int slab, int nslab) {
  using G = DwGeom<C>;
  using IM = ImgMap<C>;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + G::OFF_BAR);
  uint64_t* full = bars;
  uint64_t* empty = bars + G::NST;
  uint64_t* done = bars + 2 * G::NST;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 2 * G::NST + 1);
  const int tid = threadIdx.x, warp = tid / 32;
  const DwPair pr = dw_pair<C>(pair_idx);
  long long my_items = 0;
  if ((long long)slab < n_items && slab < nslab) my_items = (n_items - slab + nslab - 1) / nslab;

  if (tid == 0) {
    for (int i = 0; i < G::NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) umma::tmem_alloc(tslot, 128);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;

  if (my_items > 0 && tid < G::NT) {
    if (warp == 4) {
      if (tid % 32 == 0) {   // TMA producer
        for (long long i = 0; i < my_items; ++i) {
          const int s = (int)(i % G::NST);
          if (i >= G::NST) mbar_wait(&empty[s], (uint32_t)((i / G::NST) - 1) & 1);
          const unsigned char* blk = images + (size_t)(slab + i * nslab) * IM::STRIDE;
          mbar_expect_tx(&full[s], pr.bytesD + pr.bytesA);
          tma_load_1d(smem_raw + s * G::STAGE, blk + pr.offD, pr.bytesD, &full[s]);
          tma_load_1d(smem_raw + s * G::STAGE + G::DBUF, blk + pr.offA, pr.bytesA, &full[s]);
        }
      }
      __syncwarp();
    } else {
      if (tid == 0) {        // MMA issuer: D[j][n] += sum_t d[t][j] a[t][n], K = 128 trajectories = 8 x 16
        const uint32_t idesc = umma::make_idesc_bf16(128, pr.N, true, true);
        for (long long i = 0; i < my_items; ++i) {
          const int s = (int)(i % G::NST);
          mbar_wait(&full[s], (uint32_t)(i / G::NST) & 1);
          umma::fence_after_sync();
          const uint32_t a0 = smem_u32(smem_raw + s * G::STAGE), b0 = a0 + G::DBUF;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma::mma_bf16(tbase, umma::make_desc(a0 + ks * 256, 128, 2048), umma::make_desc(b0 + ks * 256, 128, 2048),
                           idesc, (i > 0 || ks > 0) ? 1u : 0u);
          umma::commit(&empty[s]);
        }
        umma::commit(done);
      }
      mbar_wait(done, 0);
      umma::fence_after_sync();
      const int j = tid;   // result row = feature of the cotangent image
      float* grep = grad_rep + (size_t)(slab % NREP) * C::NGRAD;
      for (int c0 = 0; c0 < pr.N; c0 += 8) {
        float v[8];
        umma::ld8(tbase + ((uint32_t)(warp * 32) << 16) + c0, v);
        umma::wait_ld();
        if (j < pr.rows) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int col = c0 + i;
            if (col < pr.cols) atomicAdd(grep + pr.goff_w + j * pr.ldw + col, v[i]);
            else if (col == pr.cols) atomicAdd(grep + pr.goff_b + j, v[i]);
          }
        }
      }
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 128);
}

template <class C>
__global__ void __launch_bounds__(DwGeom<C>::NT, 1) node_tc_dw_kernel(const unsigned char* __restrict__ images,
                                                                      long long n_items, float* __restrict__ grad_rep) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  dw_role<C>(smem_raw, images, n_items, grad_rep, (int)blockIdx.y, (int)blockIdx.x, (int)gridDim.x);
}

// ------------------------------------------------------------------------------------------------
// pass B of the 3xTF32 tier: dW = sum d^T a from hi / lo bf16 image pairs (d_hi a_hi + d_hi a_lo + d_lo a_hi, 16 mantissa
// bits per operand), fp32-class.  The tensor core accumulates into TMEM with truncation (measured, DESIGN.md 5: the
// d-theta error of the TF32 tier's pass B grows linearly with the length of the accumulation chain), so here a chain is
// ONE image block long (24 MMAs): two TMEM accumulators alternate, and four drain warps add each finished block into
// fp32 registers (round-to-nearest) while the MMAs of the next block run.  Warp 4 = TMA producer, warp 5 = MMA issuer.
// ------------------------------------------------------------------------------------------------
template <class C>
struct DwSplitGeom {
  using IM = ImgMap<C, true>;
  static constexpr int NST = 2;
  static constexpr int STAGE = 2 * IM::DEL_B + 2 * IM::ACT_B + 4096;   // d_hi | d_lo | a_hi | a_lo, packed; the MMA's
  // address ranges (16 chunks of d, IMGW / 8 chunks of a) overrun into the neighbour / the pad: rows and columns that
  // the epilogue drops (every output row / column depends only on its own operand row)
  static constexpr int OFF_BAR = NST * STAGE;
  static constexpr int NBAR = 2 * NST + 4;
  static constexpr size_t SMEM_BYTES = OFF_BAR + NBAR * 8 + 16;
  static constexpr int NT = 192;
  static constexpr int NG8 = IM::IMGW / 8;
  static_assert(STAGE % 128 == 0 && SMEM_BYTES <= 227 * 1024, "pass-B staging (split)");
  static_assert(2 * IM::DEL_B + 16 * 2048 - IM::DEL_B <= STAGE, "d_lo address range stays inside the stage");
};

template <class C>
__global__ void __launch_bounds__(DwSplitGeom<C>::NT, 1) node_tc_dw_split_kernel(const unsigned char* __restrict__ images,
                                                                                 long long n_items,
                                                                                 float* __restrict__ grad_rep) {
  using G = DwSplitGeom<C>;
  using IM = ImgMap<C, true>;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + G::OFF_BAR);
  uint64_t* full = bars;
  uint64_t* empty = bars + G::NST;
  uint64_t* accf = bars + 2 * G::NST;       // accumulator b holds a finished block
  uint64_t* acce = bars + 2 * G::NST + 2;   // accumulator b has been drained
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + G::NBAR);
  const int tid = threadIdx.x, warp = tid / 32;
  const DwPair pr = dw_pair<C>(blockIdx.y);
  const int slab = blockIdx.x, nslab = gridDim.x;
  long long my_items = 0;
  if ((long long)slab < n_items) my_items = (n_items - slab + nslab - 1) / nslab;

  if (tid == 0) {
    for (int i = 0; i < G::NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&accf[i], 1); mbar_init(&acce[i], 128); }
    fence_barrier_init();
  }
  if (warp == 0) umma::tmem_alloc(tslot, 256);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;
  const int oDlo = pr.bytesD, oAhi = 2 * pr.bytesD, oAlo = 2 * pr.bytesD + pr.bytesA;

  if (my_items > 0) {
    if (warp == 4) {
      if (tid % 32 == 0) {   // TMA producer: four images per block
        for (long long i = 0; i < my_items; ++i) {
          const int s = (int)(i % G::NST);
          if (i >= G::NST) mbar_wait(&empty[s], (uint32_t)((i / G::NST) - 1) & 1);
          const unsigned char* blk = images + (size_t)(slab + i * nslab) * IM::STRIDE;
          unsigned char* st = smem_raw + s * G::STAGE;
          mbar_expect_tx(&full[s], 2 * (pr.bytesD + pr.bytesA));
          tma_load_1d(st, blk + pr.offD, pr.bytesD, &full[s]);
          tma_load_1d(st + oDlo, blk + IM::LO + pr.offD, pr.bytesD, &full[s]);
          tma_load_1d(st + oAhi, blk + pr.offA, pr.bytesA, &full[s]);
          tma_load_1d(st + oAlo, blk + IM::LO + pr.offA, pr.bytesA, &full[s]);
        }
      }
      __syncwarp();
    } else if (warp == 5) {
      if (tid % 32 == 0) {   // MMA issuer: per block  D = d_hi a_hi + d_hi a_lo + d_lo a_hi,  K = 128 trajectories
        const uint32_t idesc = umma::make_idesc_bf16(128, pr.N, true, true);
        for (long long i = 0; i < my_items; ++i) {
          const int s = (int)(i % G::NST), b = (int)(i & 1);
          if (i >= 2) mbar_wait(&acce[b], (uint32_t)((i >> 1) - 1) & 1);
          mbar_wait(&full[s], (uint32_t)(i / G::NST) & 1);
          umma::fence_after_sync();
          const uint32_t s0 = smem_u32(smem_raw + s * G::STAGE);
          const uint32_t dh = s0, dl = s0 + oDlo, ah = s0 + oAhi, al = s0 + oAlo;
          const uint32_t acc = tbase + b * 128;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma::mma_bf16(acc, umma::make_desc(dh + ks * 256, 128, 2048), umma::make_desc(ah + ks * 256, 128, 2048), idesc,
                           ks > 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma::mma_bf16(acc, umma::make_desc(dh + ks * 256, 128, 2048), umma::make_desc(al + ks * 256, 128, 2048), idesc, 1u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma::mma_bf16(acc, umma::make_desc(dl + ks * 256, 128, 2048), umma::make_desc(ah + ks * 256, 128, 2048), idesc, 1u);
          umma::commit(&empty[s]);
          umma::commit(&accf[b]);
        }
      }
      __syncwarp();
    } else {   // drain warps: row j of the result = feature j of the cotangent image
      float racc[G::NG8][8];
#pragma unroll
      for (int g = 0; g < G::NG8; ++g)
#pragma unroll
        for (int k = 0; k < 8; ++k) racc[g][k] = 0.f;
      const int ng8 = pr.N / 8;
      for (long long i = 0; i < my_items; ++i) {
        const int b = (int)(i & 1);
        mbar_wait(&accf[b], (uint32_t)(i >> 1) & 1);
        umma::fence_after_sync();
        const uint32_t trow = tbase + b * 128 + ((uint32_t)(warp * 32) << 16);
#pragma unroll
        for (int g = 0; g < G::NG8; ++g)
          if (g < ng8) {
            float v[8];
            umma::ld8f(trow + 8 * g, v);
            umma::wait_ld8(v);
#pragma unroll
            for (int k = 0; k < 8; ++k) racc[g][k] += v[k];
          }
        umma::fence_before_sync();
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&acce[b])) : "memory");
      }
      const int j = tid;
      float* grep = grad_rep + (size_t)(slab % NREP) * C::NGRAD;
      if (j < pr.rows) {
#pragma unroll
        for (int g = 0; g < G::NG8; ++g)
          if (g < ng8) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const int col = 8 * g + k;
              if (col < pr.cols) atomicAdd(grep + pr.goff_w + j * pr.ldw + col, racc[g][k]);
              else if (col == pr.cols) atomicAdd(grep + pr.goff_b + j, racc[g][k]);
            }
          }
      }
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 256);
}

struct TcBwdArgs {
  BwdArgs b;              // same meaning as node_bwd.cuh (grad_rep unused here)
  unsigned char* images;  // [n_tiles_in_chunk][n_steps * S][ImgMap::STRIDE] or null
  long long tile0;        // first tile of this chunk (rows tile0 * 128 ...)
  long long n_tiles;      // tiles in this chunk
};

template <class C, bool SPLIT = false>
__global__ void __launch_bounds__(TcBwdGeom<C, SPLIT>::NT, 1) node_tc_bwd_kernel(TcBwdArgs ta) {
  using T = TcCfg<C>;
  using G = TcBwdGeom<C, SPLIT>;
  using IM = ImgMap<C, SPLIT>;
  const BwdArgs& a = ta.b;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  float* sRes = reinterpret_cast<float*>(smem_raw + G::OFF_RES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + G::OFF_BAR);
  uint64_t* full = bars;
  uint64_t* empty = bars + G::NSLOT;
  uint64_t* done = bars + 2 * G::NSLOT;
  uint64_t* resbar = bars + 2 * G::NSLOT + G::G;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + G::NBAR);
  const int tid = threadIdx.x, warp = tid / 32;
  const Tableau tab = make_tableau(a.method);
  const int S = tab.S;

  const int n_cta = (int)gridDim.x;
  const long long n_pairs = (ta.n_tiles + G::G - 1) / G::G;
  long long my_iters = 0;
  if ((long long)blockIdx.x < n_pairs) my_iters = (n_pairs - blockIdx.x + n_cta - 1) / n_cta;
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
    if (cx.keeper) {
      mbar_expect_tx(resbar, C::RES_ALL * 4);
      tma_load_1d(sRes, a.packed, C::RES_ALL * 4, resbar);
      for (uint32_t p = 0; p + 1 < (uint32_t)G::NSLOT; ++p) ring_prefetch<G>(cx, p);
    }
    mbar_wait(resbar, 0);
    const long long n_se = (long long)a.n_steps * S;
#ifdef NLBAC_TC_PROF
    cx.pf_t = clock64(); cx.pf_epi = cx.pf_full = cx.pf_issue = cx.pf_done = cx.pf_n = 0;
    const long long pf_start = cx.pf_t;
#endif

    for (long long it = 0; it < my_iters; ++it) {
      const long long ltile = (it * n_cta + blockIdx.x) * G::G + cx.g;   // tile within the chunk
      const bool tile_live = ltile < ta.n_tiles;
      const long long b = (ta.tile0 + ltile) * T::ROWS + cx.lt;
      const bool live = tile_live && b < a.B;
      unsigned char* timg = (ta.images != nullptr && tile_live) ? ta.images + (size_t)ltile * n_se * IM::STRIDE : nullptr;
      float adj[C::NS], cst[C::NCONST], dcst[C::NCONST];
      float dks[4][C::NS];
#pragma unroll
      for (int d = 0; d < C::NS; ++d) adj[d] = 0.f;
#pragma unroll
      for (int d = 0; d < C::NCONST; ++d) { cst[d] = 0.f; dcst[d] = 0.f; }
      if (live) {
#pragma unroll
        for (int d = 0; d < C::NCONST; ++d) cst[d] = a.y0[b * C::D + C::NS + d];
      }
      int jo = a.n_t - 1;
      for (int step = a.n_steps - 1; step >= 0; --step) {
        const float dt = a.dts[step];
        float carry[C::NS];
#pragma unroll
        for (int d = 0; d < C::NS; ++d) carry[d] = 0.f;
        while (jo >= 1 && a.out_step[jo] == step) {
          if (live) {
            const float w = a.out_w[jo];
            const float* g = a.grad_out + ((long long)jo * a.B + b) * C::D;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) {
              const float gv = g[d];
              adj[d] = fmaf(w, gv, adj[d]);
              carry[d] = fmaf(1.f - w, gv, carry[d]);
            }
#pragma unroll
            for (int d = 0; d < C::NCONST; ++d) dcst[d] += g[C::NS + d];
          }
          --jo;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int d = 0; d < C::NS; ++d) dks[i][d] = dt * tab.csol[i] * adj[d];
        float ay[C::NS];
#pragma unroll
        for (int d = 0; d < C::NS; ++d) ay[d] = adj[d] + carry[d];
#pragma unroll 1
        for (int i = S - 1; i >= 0; --i) {
          float x[C::NS], dk[C::NS], dx[C::NS];
#pragma unroll
          for (int d = 0; d < C::NS; ++d) { x[d] = 0.f; dk[d] = dks[i][d]; }
          if (live) {
            const float* p = a.ckpt + (((long long)step * S + i) * a.B + b) * C::NS;
#pragma unroll
            for (int d = 0; d < C::NS; ++d) x[d] = p[d];
          } else {
#pragma unroll
            for (int d = 0; d < C::NS; ++d) dk[d] = 0.f;
          }
          unsigned char* img = timg ? timg + (size_t)((long long)step * S + i) * IM::STRIDE : nullptr;
          vjp_rhs_tc<C, SPLIT>(cx, img, x, cst, dk, dx, dcst);
#pragma unroll
          for (int d = 0; d < C::NS; ++d) ay[d] += dx[d];
#pragma unroll
          for (int j = 0; j < 3; ++j)
            if (j < i) {
              const float bij = dt * tab.beta[i][j];
#pragma unroll
              for (int d = 0; d < C::NS; ++d) dks[j][d] = fmaf(bij, dx[d], dks[j][d]);
            }
        }
#pragma unroll
        for (int d = 0; d < C::NS; ++d) adj[d] = ay[d];
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
#ifdef NLBAC_TC_PROF
    if (blockIdx.x == 0 && cx.lt == 0)
      printf("tc_prof pass A group %d: layers %lld  total %lld cyc  per layer: epilogue %lld  wait-weights %lld  issue %lld  "
             "wait-done %lld\n", cx.g, cx.pf_n, clock64() - pf_start, cx.pf_epi / cx.pf_n, cx.pf_full / cx.pf_n,
             cx.pf_issue / cx.pf_n, cx.pf_done / cx.pf_n);
#endif
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, G::TMEM_COLS);
}

}  // namespace nlbac
