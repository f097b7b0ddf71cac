// Device building blocks of the FP32-exact NODE kernels: the weight ring (TMA bulk copies +
// mbarriers), the register-tiled FP32 GEMM over a tile of trajectories, and the thin-layer helpers.
//
// Activation layout in shared memory: row = trajectory, RS floats per row (RS/4 odd), so a warp
// reading 32 consecutive rows with LDS.128 is bank-conflict free, and weight rows are broadcast.
#pragma once
#include "node_cfg.cuh"

namespace nlbac {

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
// acc[i][0..8) = bias[n0..n0+8) + sum_{k<K} in[row_i][k] * W[k][n0..n0+8),   row_i = tg + TG*i
// K % 4 == 0.  in: ld = ldi floats (rows 16-byte aligned), W: [K][ldw], n0 % 8 == 0.
// ---------------------------------------------------------------------------------------------
template <int TT, int TG>
__device__ __forceinline__ void gemm_tile(const float* __restrict__ in, int ldi, const float* __restrict__ W, int ldw,
                                          const float* __restrict__ bias, int K, int tg, int n0, float (&acc)[TT][8]) {
  if (bias != nullptr) {
    const float4 b0 = *reinterpret_cast<const float4*>(bias + n0);
    const float4 b1 = *reinterpret_cast<const float4*>(bias + n0 + 4);
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      acc[i][0] = b0.x; acc[i][1] = b0.y; acc[i][2] = b0.z; acc[i][3] = b0.w;
      acc[i][4] = b1.x; acc[i][5] = b1.y; acc[i][6] = b1.z; acc[i][7] = b1.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < TT; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  }
  const float* inp = in + tg * ldi;
  const float* wp = W + n0;
#pragma unroll 1
  for (int k = 0; k < K; k += 4) {
    float a[TT][4];
#pragma unroll
    for (int i = 0; i < TT; ++i) {
      const float4 v = *reinterpret_cast<const float4*>(inp + (size_t)(TG * i) * ldi + k);
      a[i][0] = v.x; a[i][1] = v.y; a[i][2] = v.z; a[i][3] = v.w;
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float4 w0 = *reinterpret_cast<const float4*>(wp + (k + kk) * ldw);
      const float4 w1 = *reinterpret_cast<const float4*>(wp + (k + kk) * ldw + 4);
#pragma unroll
      for (int i = 0; i < TT; ++i) {
        const float av = a[i][kk];
        acc[i][0] = fmaf(av, w0.x, acc[i][0]); acc[i][1] = fmaf(av, w0.y, acc[i][1]);
        acc[i][2] = fmaf(av, w0.z, acc[i][2]); acc[i][3] = fmaf(av, w0.w, acc[i][3]);
        acc[i][4] = fmaf(av, w1.x, acc[i][4]); acc[i][5] = fmaf(av, w1.y, acc[i][5]);
        acc[i][6] = fmaf(av, w1.z, acc[i][6]); acc[i][7] = fmaf(av, w1.w, acc[i][7]);
      }
    }
  }
}

// out[row_i][n0..n0+8) = relu(acc[i][:])
template <int TT, int TG>
__device__ __forceinline__ void store_relu(float* __restrict__ out, int ldo, int tg, int n0, const float (&acc)[TT][8]) {
#pragma unroll
  for (int i = 0; i < TT; ++i) {
    float* o = out + (size_t)(tg + TG * i) * ldo + n0;
    *reinterpret_cast<float4*>(o) =
        make_float4(fmaxf(acc[i][0], 0.f), fmaxf(acc[i][1], 0.f), fmaxf(acc[i][2], 0.f), fmaxf(acc[i][3], 0.f));
    *reinterpret_cast<float4*>(o + 4) =
        make_float4(fmaxf(acc[i][4], 0.f), fmaxf(acc[i][5], 0.f), fmaxf(acc[i][6], 0.f), fmaxf(acc[i][7], 0.f));
  }
}

// out[row_i][n0..) = acc[i][:] where mask[row_i][n0..] > 0 else 0   (ReLU backward; out may alias mask)
template <int TT, int TG>
__device__ __forceinline__ void store_masked(float* out, int ldo, int tg, int n0, const float (&acc)[TT][8]) {
#pragma unroll
  for (int i = 0; i < TT; ++i) {
    float* o = out + (size_t)(tg + TG * i) * ldo + n0;
    const float4 m0 = *reinterpret_cast<const float4*>(o);
    const float4 m1 = *reinterpret_cast<const float4*>(o + 4);
    *reinterpret_cast<float4*>(o) = make_float4(m0.x > 0.f ? acc[i][0] : 0.f, m0.y > 0.f ? acc[i][1] : 0.f,
                                                m0.z > 0.f ? acc[i][2] : 0.f, m0.w > 0.f ? acc[i][3] : 0.f);
    *reinterpret_cast<float4*>(o + 4) = make_float4(m1.x > 0.f ? acc[i][4] : 0.f, m1.y > 0.f ? acc[i][5] : 0.f,
                                                    m1.z > 0.f ? acc[i][6] : 0.f, m1.w > 0.f ? acc[i][7] : 0.f);
  }
}

// Thin last layer, one thread per trajectory: o[0..OUTP) = bL + sum_{k<H} a[k] * WtL[k][0..OUTP)
template <int H, int OUTP>
__device__ __forceinline__ void thin_out(const float* __restrict__ arow, const float* __restrict__ WtL,
                                         const float* __restrict__ bL, float (&o)[OUTP]) {
#pragma unroll
  for (int j = 0; j < OUTP; ++j) o[j] = bL[j];
#pragma unroll 1
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
