// Device building blocks of the FP32-exact NODE kernels: the weight ring (TMA bulk copies +
// mbarriers), the register-tiled FP32 GEMM over a tile of trajectories, and the thin-layer helpers.
//
// Activation layout in shared memory: row = trajectory, RS floats per row (RS/4 odd), so a warp
// reading 32 consecutive rows with LDS.128 is bank-conflict free, and weight rows are broadcast.
#pragma once
#include <type_traits>

#include "node_cfg.cuh"

namespace nlbac {

// Chain loop (f_net, g_net) of the FP32 kernels rolled (1) or unrolled (0): rolled halves the code of the kernel
// (instruction-cache footprint: 55 -> 38 KB for the backward), unrolled keeps every per-chain size a compile-time
// constant.  Measured at 2^20 x 20 rk4 (round 2, same box A/B): backward 782 ms unrolled vs 843 ms rolled, forward equal
// -> unrolled.  (The tcgen05 3xTF32 sweep, 193 KB, is the kernel where rolling pays: node_tc_bwd.cuh.)
#ifndef NLBAC_ROLL_CHAINS_BWD
#define NLBAC_ROLL_CHAINS_BWD 0
#endif
#ifndef NLBAC_ROLL_CHAINS_FWD
#define NLBAC_ROLL_CHAINS_FWD 0
#endif
// k-loop of gemm_tile (4 k per iteration, 26 iterations at K = 104), per call site.  Measured at 2^20 x 20 rk4 (round 2,
// same box): backward sweep 791 ms rolled -> 747 (x2) -> 736 (x4); forward 170 ms rolled -> 176 (x2) -> 175 (x4): the
// backward's loads had nothing to overlap with at 2 warps per scheduler, the forward prefers the smaller loop body.
#ifndef NLBAC_GEMM_UNROLL_FWD
#define NLBAC_GEMM_UNROLL_FWD 1
#endif
#ifndef NLBAC_GEMM_UNROLL_BWD
#define NLBAC_GEMM_UNROLL_BWD 4
#endif
#ifndef NLBAC_DW_UNROLL
#define NLBAC_DW_UNROLL 8      // t-loop of dw_tile_reduce (one tile row per iteration): sweep 737 (x2) -> 731 (x4) -> 714 (x8) -> 717 ms (x16)
#endif
constexpr int NODE_DW_UNROLL = NLBAC_DW_UNROLL;
#ifndef NLBAC_THIN_UNROLL
#define NLBAC_THIN_UNROLL 16   // t-loops of the thin first / last layer dW and the bias sums in vjp_rhs_tile: sweep 714 (x2) -> 696 (x8) -> 692 ms (x16)
#endif
constexpr int NODE_THIN_UNROLL = NLBAC_THIN_UNROLL;
#ifndef NLBAC_THINOUT_UNROLL
#define NLBAC_THINOUT_UNROLL 4  // k-loop of thin_out (the H -> n_out last layer, one thread per trajectory): step 861 (x1) -> 855 ms (x4)
#endif
constexpr int NODE_THINOUT_UNROLL = NLBAC_THINOUT_UNROLL;
#ifndef NLBAC_DX0_UNROLL
#define NLBAC_DX0_UNROLL 4      // j-loop of the first-layer dx (owner threads) in vjp_rhs_tile: step 855 (x1) -> 847 ms (x4)
#endif
constexpr int NODE_DX0_UNROLL = NLBAC_DX0_UNROLL;
#ifndef NLBAC_GEMM_UNROLL_FWD_SMALL
#define NLBAC_GEMM_UNROLL_FWD_SMALL 4   // forward tiles with <= NODE_FWD_SMALL_TT rows per thread (small batches: latency-bound)
#endif
constexpr int NODE_GEMM_UNROLL_FWD_SMALL = NLBAC_GEMM_UNROLL_FWD_SMALL, NODE_FWD_SMALL_TT = 2;
constexpr int NODE_GEMM_UNROLL_FWD = NLBAC_GEMM_UNROLL_FWD, NODE_GEMM_UNROLL_BWD = NLBAC_GEMM_UNROLL_BWD;
constexpr bool NODE_ROLL_CHAINS_BWD = NLBAC_ROLL_CHAINS_BWD != 0, NODE_ROLL_CHAINS_FWD = NLBAC_ROLL_CHAINS_FWD != 0;

// ---------------------------------------------------------------------------------------------
// Weight ring: the hidden HxH matrices cycle through NSLOT shared-memory slots.  One thread
// issues cp.async.bulk (TMA, UBLKCP) for packet seq+NSLOT as soon as the CTA has finished reading
// packet seq; consumers wait on the slot's mbarrier (phase parity = (seq / NSLOT) & 1).
// If every packet of the cycle fits (NSLOT == CYCLE) they are loaded once and stay resident.
// ---------------------------------------------------------------------------------------------
template <int NSLOT, int CYCLE, int PKT_FLOATS>
struct WeightRing {
  static constexpr bool RESIDENT = (NSLOT == CYCLE);
  float* slots;
  uint64_t* full;
  uint32_t seq;    // packets consumed so far (CTA-uniform)
  uint32_t total;  // packets this CTA will consume in the whole launch

  template <class OffFn>
  __device__ __forceinline__ void start(const float* packed, OffFn off) {
    if (threadIdx.x == 0) {
      const uint32_t n = RESIDENT ? (uint32_t)NSLOT : (total < (uint32_t)NSLOT ? total : (uint32_t)NSLOT);
      for (uint32_t s = 0; s < n; ++s) {
        mbar_expect_tx(&full[s], PKT_FLOATS * 4);
        tma_load_1d(slots + s * PKT_FLOATS, packed + off(s), PKT_FLOATS * 4, &full[s]);
      }
    }
  }
  // wait for the current packet; returns its shared-memory address
  __device__ __forceinline__ const float* acquire() {
    const uint32_t slot = seq % NSLOT;
    if (!RESIDENT || seq < (uint32_t)NSLOT) mbar_wait(&full[slot], (seq / NSLOT) & 1);
    return slots + slot * PKT_FLOATS;
  }
  // call after the __syncthreads() that follows the last read of the current packet
  template <class OffFn>
  __device__ __forceinline__ void release(const float* packed, OffFn off) {
    if (!RESIDENT && threadIdx.x == 0) {
      const uint32_t nxt = seq + NSLOT;
      if (nxt < total) {
        const uint32_t slot = seq % NSLOT;
        fence_proxy_async();
        mbar_expect_tx(&full[slot], PKT_FLOATS * 4);
        tma_load_1d(slots + slot * PKT_FLOATS, packed + off(nxt), PKT_FLOATS * 4, &full[slot]);
      }
    }
    ++seq;
  }
};

// ---------------------------------------------------------------------------------------------
// Column ownership of a thread in a [rows][HP] tile.  8 column groups (ng = 0..7); group ng owns
//   4 consecutive columns at 32*v + 4*ng for v < NV4      (LDS.128 / STS.128, 16-byte aligned)
//   1 column at 32*NV4 + ng when HP % 32 == 8             (HP = 104: 13 columns per thread)
// so a warp of 32 row-threads x 1 column group reads weight rows as broadcasts, and 8 column groups
// x T/TT row-threads give a warp count that is a multiple of 4 (balanced over the 4 SM sub-partitions).
// ---------------------------------------------------------------------------------------------
template <int HP>
struct ColMap {
  static constexpr int NV4 = HP / 32;
  static constexpr int NREM = (HP % 32) / 8;
  static constexpr int NN = 4 * NV4 + NREM;
  static_assert(HP % 32 == 0 || HP % 32 == 8, "column map supports HP = 32m or 32m + 8");
  __device__ __forceinline__ static int col(int ng, int j) { return j < 4 * NV4 ? 32 * (j / 4) + 4 * ng + (j % 4) : 32 * NV4 + ng; }
};

constexpr int NODE_SMALL_TILE = 8;    // rows per CTA of the small-batch geometry (ColMapN)

// Narrow map for the small-batch geometry (8-row tiles, B <= ~1200): HP / 4 column groups of ONE float4 each, i.e. 4
// columns per thread instead of 13 -- the per-layer latency of the register-tiled GEMM is its k-loop over a thread's
// columns, and at B = 128 / 256 latency is all there is (4 / 8 CTAs with the wide map).  Same index formulas as ColMap
// (32 * v + 4 * ng with v = 0), so gemm_tile / store_relu / store_masked take it as a template argument.
template <int HP>
struct ColMapN {
  static constexpr int NV4 = 1;
  static constexpr int NREM = 0;
  static constexpr int NN = 4;
  static constexpr int NGRP = HP / 4;
  static_assert(HP % 4 == 0, "narrow column map: HP = 4m");
  __device__ __forceinline__ static int col(int ng, int j) { return 4 * ng + j; }
};

// acc[i][j] = bias[col j] + sum_{k<K} in[row_i][k] * W[k][col j],   row_i = tg + TG*i,  K % 4 == 0
// Inner product runs on packed FFMA2: the 4*NV4 vector columns as pairs, the remainder column scalar.
template <int TT, int TG, int HP, int UNR = NODE_GEMM_UNROLL_FWD, class CM = ColMap<HP>>
__device__ __forceinline__ void gemm_tile(const float* __restrict__ in, int ldi, const float* __restrict__ W,
                                          const float* __restrict__ bias, int K, int tg, int ng,
                                          float (&acc)[TT][CM::NN]) {
  constexpr int NP = 2 * CM::NV4;  // column pairs
  unsigned long long acc2[TT][NP];
  float accr[TT];
  {
    unsigned long long b2[NP];
#pragma unroll
    for (int p = 0; p < NP; ++p)
      b2[p] = bias != nullptr ? pack2(bias[CM::col(ng, 2 * p)], bias[CM::col(ng, 2 * p + 1)]) : 0ull;
    const float br = (CM::NREM && bias != nullptr) ? bias[CM::col(ng, CM::NN - 1)] : 0.f;
#pragma unroll
    for (int i = 0; i < TT; ++i) {
#pragma unroll
      for (int p = 0; p < NP; ++p) acc2[i][p] = b2[p];
      accr[i] = br;
    }
  }
  const float* inp = in + tg * ldi;
  const float* wp = W + 4 * ng;
#pragma unroll(UNR)
  for (int k = 0; k < K; k += 4) {
    float a[TT][4];
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const float4 v = *reinterpret_cast<const float4*>(inp + (size_t)(TG * i) * ldi + k);
      a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      unsigned long long w2[NP];
      const float* wr = wp + (k + kk) * HP;
#pragma unroll
      for (int v = 0; v < CM::NV4; ++v) {
        const float4 q = *reinterpret_cast<const float4*>(wr + 32 * v);
        w2[2 * v] = pack2(q.x, q.y);
        w2[2 * v + 1] = pack2(q.z, q.w);
      }
      float wrem = 0.f;
      if (CM::NREM) wrem = W[(k + kk) * HP + 32 * CM::NV4 + ng];
#pragma unroll
      for (int i = 0; i < TT; ++i) {
        const unsigned long long aa = pack2(a[i][kk], a[i][kk]);
#pragma unroll
        for (int p = 0; p < NP; ++p) acc2[i][p] = fma2(aa, w2[p], acc2[i][p]);
        if (CM::NREM) accr[i] = fmaf(a[i][kk], wrem, accr[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < TT; ++i) {
#pragma unroll
    for (int p = 0; p < NP; ++p) unpack2(acc2[i][p], acc[i][2 * p], acc[i][2 * p + 1]);
    if (CM::NREM) acc[i][CM::NN - 1] = accr[i];
  }
}

// out[row_i][col j] = relu(acc[i][j])
template <int TT, int TG, int HP, class CM = ColMap<HP>>
__device__ __forceinline__ void store_relu(float* __restrict__ out, int ldo, int tg, int ng,
                                           const float (&acc)[TT][CM::NN]) {
#pragma unroll
  for (int i = 0; i < TT; ++i) {
    float* o = out + (size_t)(tg + TG * i) * ldo;
#pragma unroll
    for (int v = 0; v < CM::NV4; ++v)
      *reinterpret_cast<float4*>(o + 32 * v + 4 * ng) =
          make_float4(fmaxf(acc[i][4 * v], 0.f), fmaxf(acc[i][4 * v + 1], 0.f), fmaxf(acc[i][4 * v + 2], 0.f),
                      fmaxf(acc[i][4 * v + 3], 0.f));
    if (CM::NREM) o[32 * CM::NV4 + ng] = fmaxf(acc[i][CM::NN - 1], 0.f);
  }
}

// out[row_i][col j] = acc[i][j] where out[row_i][col j] > 0 else 0   (ReLU backward, in place over the activation)
template <int TT, int TG, int HP, class CM = ColMap<HP>>
__device__ __forceinline__ void store_masked(float* out, int ldo, int tg, int ng, const float (&acc)[TT][CM::NN]) {
#pragma unroll
  for (int i = 0; i < TT; ++i) {
    float* o = out + (size_t)(tg + TG * i) * ldo;
#pragma unroll
    for (int v = 0; v < CM::NV4; ++v) {
      float4* p = reinterpret_cast<float4*>(o + 32 * v + 4 * ng);
      const float4 m = *p;
      *p = make_float4(m.x > 0.f ? acc[i][4 * v] : 0.f, m.y > 0.f ? acc[i][4 * v + 1] : 0.f,
                       m.z > 0.f ? acc[i][4 * v + 2] : 0.f, m.w > 0.f ? acc[i][4 * v + 3] : 0.f);
    }
    if (CM::NREM) {
      float* p = o + 32 * CM::NV4 + ng;
      *p = *p > 0.f ? acc[i][CM::NN - 1] : 0.f;
    }
  }
}

// Thin last layer, one thread per trajectory: o[0..OUTP) = bL + sum_{k<H} a[k] * WtL[k][0..OUTP)
template <int H, int OUTP>
__device__ __forceinline__ void thin_out(const float* __restrict__ arow, const float* __restrict__ WtL,
                                         const float* __restrict__ bL, float (&o)[OUTP]) {
#pragma unroll
  for (int j = 0; j < OUTP; ++j) o[j] = bL[j];
#pragma unroll(NODE_THINOUT_UNROLL)
  for (int k = 0; k < H; k += 4) {
    const float4 v = *reinterpret_cast<const float4*>(arow + k);
    const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float* w = WtL + (k + kk) * OUTP;
#pragma unroll
      for (int j = 0; j < OUTP; j += 4) {
        const float4 wv = *reinterpret_cast<const float4*>(w + j);
        o[j] = fmaf(a[kk], wv.x, o[j]); o[j + 1] = fmaf(a[kk], wv.y, o[j + 1]);
        o[j + 2] = fmaf(a[kk], wv.z, o[j + 2]); o[j + 3] = fmaf(a[kk], wv.w, o[j + 3]);
      }
    }
  }
}

}  // namespace nlbac
