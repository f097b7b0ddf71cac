// tcgen05 operand-layout self-test and descriptor probe (the tensor-core NODE entry points live in node_abi.cu).
#include "../../include/nlbac.h"
#include "umma.cuh"
#include <cstring>
#include <cuda_bf16.h>

namespace nlbac {

// ------------------------------------------------------------------------------------------------
// Self-test of the two operand interpretations the tensor-core NODE kernels rely on (umma.cuh):
//   mode 0  kind::tf32  D[t][j] = sum_k X[t][k] W[j][k]   A K-major,  B K-major   (hidden layer; the dX product uses
//                                                                                  the same form on a transposed image)
//   mode 1  kind::f16   D[j][k] = sum_t X[t][j] Y[t][k]   A MN-major, B MN-major  (dW over bf16 row-owner images)
//   mode 2  kind::tf32  as mode 0 with A in TENSOR MEMORY, written by its row-owner threads with tcgen05.st
// kind::tf32 has no MN-major form without the 128-byte swizzle (measured: all-zero accumulators for every LBO/SBO
// assignment, tools/umma_sweep.py), which is why dX reads a transposed weight image and dW runs in bf16.
// X, Y: [128][104] row-major; W: [104][104] row-major; D: [128][112] row-major.
// ------------------------------------------------------------------------------------------------
constexpr int ST_ROWS = 128, ST_K = 104, ST_N = 112;
constexpr int ST_XCH = ST_ROWS * 16, ST_WCH = ST_N * 16;
constexpr int ST_SMEM = 32 * ST_XCH * 2 + 64;

__global__ void __launch_bounds__(128, 1) umma_selftest_kernel(int mode, const float* __restrict__ X,
                                                              const float* __restrict__ Y, float* __restrict__ D) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* bufA = sm;                    // 32 chunks x 2048 B
  unsigned char* bufB = sm + 32 * ST_XCH;      // up to 32 chunks x 2048 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 64 * ST_XCH);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(sm + 64 * ST_XCH + 16);
  const int tid = threadIdx.x, warp = tid / 32;
  for (int i = tid; i < 64 * ST_XCH / 4; i += 128) reinterpret_cast<float*>(sm)[i] = 0.f;
  __syncthreads();
  if (mode == 0 || mode == 2) {
    // A: row-owner fp32 image of X (chunk stride 2048); B: weight image, rows j (112 per chunk), columns k
    if (mode == 0)
      for (int c = 0; c < ST_K; ++c) *reinterpret_cast<float*>(bufA + umma::tile_off(tid, c, ST_XCH)) = X[tid * ST_K + c];
    if (tid < ST_K)
      for (int c = 0; c < ST_K; ++c) *reinterpret_cast<float*>(bufB + umma::tile_off(tid, c, ST_WCH)) = Y[tid * ST_K + c];
  } else {
    // both operands: row-owner bf16 images (8 features per 16-byte chunk)
    for (int c = 0; c < ST_K; ++c) {
      *reinterpret_cast<__nv_bfloat16*>(bufA + umma::tile_off16(tid, c, ST_XCH)) = __float2bfloat16(X[tid * ST_K + c]);
      *reinterpret_cast<__nv_bfloat16*>(bufB + umma::tile_off16(tid, c, ST_XCH)) = __float2bfloat16(Y[tid * ST_K + c]);
    }
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) umma::tmem_alloc(tslot, 256);
  fence_proxy_async();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;
  if (mode == 2) {   // A operand: row tid -> lane tid, columns 128 .. 128 + ST_K
    for (int c0 = 0; c0 < ST_K; c0 += 8) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = X[tid * ST_K + c0 + i];
      umma::st8(umma::tmem_addr(tbase, warp * 32, 128 + c0), v);
    }
    umma::wait_st();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
  }
  if (tid == 0) {
    const uint32_t a0 = smem_u32(bufA), b0 = smem_u32(bufB);
    if (mode == 2) {
      const uint32_t idesc = umma::make_idesc_tf32(128, ST_N, false, false);
      for (int ks = 0; ks < ST_K / 8; ++ks)
        umma::mma_tf32_ts(tbase, tbase + 128 + ks * 8, umma::make_desc(b0 + ks * 2 * ST_WCH, ST_WCH, 128), idesc, ks > 0);
    } else if (mode == 0) {
      const uint32_t idesc = umma::make_idesc_tf32(128, ST_N, false, false);
      for (int ks = 0; ks < ST_K / 8; ++ks)
        umma::mma_tf32(tbase, umma::make_desc(a0 + ks * 2 * ST_XCH, ST_XCH, 128),
                       umma::make_desc(b0 + ks * 2 * ST_WCH, ST_WCH, 128), idesc, ks > 0);
    } else {
      const uint32_t idesc = umma::make_idesc_bf16(128, ST_N, true, true);
      for (int ks = 0; ks < ST_ROWS / 16; ++ks)  // K = 16 rows (two 8-row core matrices, 256 B) per MMA
        umma::mma_bf16(tbase, umma::make_desc(a0 + ks * 256, 128, ST_XCH), umma::make_desc(b0 + ks * 256, 128, ST_XCH),
                       idesc, ks > 0);
    }
    umma::commit(bar);
  }
  mbar_wait(bar, 0);
  umma::fence_after_sync();
  for (int c0 = 0; c0 < ST_N; c0 += 8) {
    float v[8];
    umma::ld8(umma::tmem_addr(tbase, warp * 32, c0), v);
    umma::wait_ld();
#pragma unroll
    for (int i = 0; i < 8; ++i) D[tid * ST_N + c0 + i] = v[i];
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 256);
}

// Descriptor sweep used to establish the table above (tools/umma_sweep.py); diagnostic only.
struct ProbeArgs { int a_mn, b_mn, lboA, sboA, stepA, lboB, sboB, stepB, nsteps, b_is_weight; };
__global__ void __launch_bounds__(128, 1) umma_probe_kernel(ProbeArgs p, const float* __restrict__ X,
                                                           const float* __restrict__ Y, float* __restrict__ D) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* bufA = sm;
  unsigned char* bufB = sm + 32 * ST_XCH;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 64 * ST_XCH);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(sm + 64 * ST_XCH + 16);
  const int tid = threadIdx.x, warp = tid / 32;
  for (int i = tid; i < 64 * ST_XCH / 4; i += 128) reinterpret_cast<float*>(sm)[i] = 0.f;
  __syncthreads();
  if (p.b_is_weight == 2) {
    for (int c = 0; c < ST_K; ++c) {
      *reinterpret_cast<__nv_bfloat16*>(bufA + umma::tile_off16(tid, c, ST_XCH)) = __float2bfloat16(X[tid * ST_K + c]);
      *reinterpret_cast<__nv_bfloat16*>(bufB + umma::tile_off16(tid, c, ST_XCH)) = __float2bfloat16(Y[tid * ST_K + c]);
    }
  } else
  for (int c = 0; c < ST_K; ++c) *reinterpret_cast<float*>(bufA + umma::tile_off(tid, c, ST_XCH)) = X[tid * ST_K + c];
  if (p.b_is_weight == 2) {
  } else if (!p.b_is_weight) {
    for (int c = 0; c < ST_K; ++c) *reinterpret_cast<float*>(bufB + umma::tile_off(tid, c, ST_XCH)) = Y[tid * ST_K + c];
  } else if (tid < ST_K) {
    for (int c = 0; c < ST_K; ++c) *reinterpret_cast<float*>(bufB + umma::tile_off(tid, c, ST_WCH)) = Y[tid * ST_K + c];
  }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) umma::tmem_alloc(tslot, 128);
  fence_proxy_async();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;
  if (tid == 0) {
    const uint32_t a0 = smem_u32(bufA), b0 = smem_u32(bufB);
    if (p.b_is_weight == 2) {
      const uint32_t idesc = umma::make_idesc_bf16(128, ST_N, p.a_mn != 0, p.b_mn != 0);
      for (int ks = 0; ks < p.nsteps; ++ks)
        umma::mma_bf16(tbase, umma::make_desc(a0 + ks * p.stepA, p.lboA, p.sboA),
                       umma::make_desc(b0 + ks * p.stepB, p.lboB, p.sboB), idesc, ks > 0);
    } else {
    const uint32_t idesc = umma::make_idesc_tf32(128, ST_N, p.a_mn != 0, p.b_mn != 0);
    for (int ks = 0; ks < p.nsteps; ++ks)
      umma::mma_tf32(tbase, umma::make_desc(a0 + ks * p.stepA, p.lboA, p.sboA),
                     umma::make_desc(b0 + ks * p.stepB, p.lboB, p.sboB), idesc, ks > 0);
    }
    umma::commit(bar);
  }
  mbar_wait(bar, 0);
  umma::fence_after_sync();
  for (int c0 = 0; c0 < ST_N; c0 += 8) {
    float v[8];
    umma::ld8(umma::tmem_addr(tbase, warp * 32, c0), v);
    umma::wait_ld();
#pragma unroll
    for (int i = 0; i < 8; ++i) D[tid * ST_N + c0 + i] = v[i];
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 128);
}

// ------------------------------------------------------------------------------------------------
// Tensor-pipe peak probe: one CTA per SM, ONE thread issues tcgen05.mma back to back
// over operands that never leave shared / tensor memory (no registers, no HBM), K loop of 13 k-steps like a hidden
// layer.  Modes:
//   0  kind::tf32  SS  M=128 N=256   (the pipe's best case: 12 KB of operands per 128-cycle MMA)
//   1  kind::tf32  TS  M=128 N=256   (A operand in tensor memory)
//   2  kind::tf32  SS  M=128 N=112   (the NODE kernels' own tile: 7.6 KB per 56-cycle MMA -> shared-memory bound)
//   3  kind::tf32  TS  M=128 N=112   (the NODE kernels' tensor-memory group)
//   4  kind::f16 (bf16) SS M=128 N=256  (cross-check against MEASURED_PEAKS.json's cuBLAS figure)
// ------------------------------------------------------------------------------------------------
constexpr int PK_KSTEPS = 13;
constexpr int PK_SMEM = PK_KSTEPS * (4096 + 8192) + 64;
__global__ void __launch_bounds__(128, 1) umma_peak_kernel(int mode, int iters) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* bufA = sm;                          // 13 k-steps x [128 rows x 8 k] fp32 (4 KB each)
  unsigned char* bufB = sm + PK_KSTEPS * 4096;       // 13 k-steps x [256 rows x 8 k] fp32 (8 KB each)
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + PK_KSTEPS * (4096 + 8192));
  uint32_t* tslot = reinterpret_cast<uint32_t*>(sm + PK_KSTEPS * (4096 + 8192) + 16);
  const int tid = threadIdx.x, warp = tid / 32;
  uint32_t h = 0x9E3779B9u * (tid + 1) + blockIdx.x;
  for (int i = tid; i < PK_KSTEPS * (4096 + 8192) / 4; i += 128) {   // pseudo-random operands in [-1, 1)
    h = h * 1664525u + 1013904223u;
    reinterpret_cast<float*>(sm)[i] = __uint_as_float(0x3F800000u | (h >> 9)) * 2.f - 3.f;
  }
  if (mode == 4) {   // bf16 pairs: keep both halves finite
    for (int i = tid; i < PK_KSTEPS * (4096 + 8192) / 4; i += 128)
      reinterpret_cast<uint32_t*>(sm)[i] = (reinterpret_cast<uint32_t*>(sm)[i] & 0x0FFF0FFFu) | 0x30003000u;
  }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) umma::tmem_alloc(tslot, 512);
  fence_proxy_async();
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tbase = *tslot;
  const bool ts = mode == 1 || mode == 3;
  const int N = (mode == 2 || mode == 3) ? 112 : 256;
  const uint32_t top = (N == 112) ? 384u : 256u;     // A operand block in tensor memory (TS modes)
  if (ts) {
    for (int c0 = 0; c0 < 8 * PK_KSTEPS; c0 += 8) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 1e-3f * (float)((tid * 31 + c0 + i) % 97) - 0.05f;
      umma::st8(umma::tmem_addr(tbase, warp * 32, top + c0), v);
    }
    umma::wait_st();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
  }
  if (tid == 0) {
    const uint32_t a0 = smem_u32(bufA), b0 = smem_u32(bufB);
    const int nacc = (N == 112) ? 3 : (ts ? 1 : 2);
    const uint32_t astride = (N == 112) ? 128u : 256u;
    if (mode == 4) {
      const uint32_t idesc = umma::make_idesc_bf16(128, 256, false, false);   // K-major bf16: 16 k per MMA = 32 B = 2 chunks
      for (int it = 0; it < iters; ++it) {
        const uint32_t d = tbase + (uint32_t)(it % nacc) * astride;
#pragma unroll
        for (int ks = 0; ks < PK_KSTEPS; ++ks)
          umma::mma_bf16(d, umma::make_desc(a0 + ks * 4096, 2048, 128), umma::make_desc(b0 + ks * 8192, 4096, 128), idesc,
                         ks > 0 ? 1u : 0u);
      }
    } else {
      const uint32_t idesc = umma::make_idesc_tf32(128, N, false, false);
      for (int it = 0; it < iters; ++it) {
        const uint32_t d = tbase + (uint32_t)(it % nacc) * astride;
#pragma unroll
        for (int ks = 0; ks < PK_KSTEPS; ++ks) {
          const uint64_t bd = umma::make_desc(b0 + ks * 8192, 4096, 128);
          if (ts) umma::mma_tf32_ts(d, tbase + top + 8 * ks, bd, idesc, ks > 0 ? 1u : 0u);
          else umma::mma_tf32(d, umma::make_desc(a0 + ks * 4096, 2048, 128), bd, idesc, ks > 0 ? 1u : 0u);
        }
      }
    }
    umma::commit(bar);
  }
  mbar_wait(bar, 0);
  umma::fence_after_sync();
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, 512);
}

}  // namespace nlbac

using namespace nlbac;

extern "C" {

int nlbac_umma_selftest(int mode, const float* X, const float* Y, float* D, void* stream) {
  NLBAC_REQUIRE(mode >= 0 && mode <= 2 && X && Y && D, "bad arguments");
  NLBAC_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM));
  umma_selftest_kernel<<<1, 128, ST_SMEM, (cudaStream_t)stream>>>(mode, X, Y, D);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_umma_probe(const int* params10, const float* X, const float* Y, float* D, void* stream) {
  ProbeArgs p;
  memcpy(&p, params10, sizeof(p));
  NLBAC_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM));
  umma_probe_kernel<<<1, 128, ST_SMEM, (cudaStream_t)stream>>>(p, X, Y, D);
  NLBAC_CUDA(cudaGetLastError());
  return 0;
}

int nlbac_umma_peak_probe(int mode, int iters, double* flops, void* stream) {
  NLBAC_REQUIRE(mode >= 0 && mode <= 4 && iters > 0, "bad arguments");
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  NLBAC_CUDA(cudaFuncSetAttribute(umma_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PK_SMEM));
  umma_peak_kernel<<<sms, 128, PK_SMEM, (cudaStream_t)stream>>>(mode, iters);
  NLBAC_CUDA(cudaGetLastError());
  const int N = (mode == 2 || mode == 3) ? 112 : 256, K = (mode == 4) ? 16 : 8;
  if (flops) *flops = 2.0 * 128 * N * K * PK_KSTEPS * (double)iters * sms;
  return 0;
}

}  // extern "C"
