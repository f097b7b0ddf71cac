// K3/K4 — fused 2-hidden-layer ReLU MLP (in -> H -> H -> out), forward and backward, FP32.
//
// Replaces QNetwork / LyaNetwork / BarrierNetwork / GaussianPolicy.forward of the reference
// (U sac_cbf_clf/model.py:37-64, :67-83, :86-115; BU model.py:67-86) and autograd through them.
// The real minibatches are B = 128 / 256 rows: the whole forward is ONE launch
// of ceil(B/TM) CTAs that keep the row tile's activations in shared memory; weights (<= 0.27 MB)
// are staged through shared memory with cp.async (the 256 x 256 layer) or read through L1/L2.  The backward is one row-tile kernel (deltas, dX) + one batched
// A^T B kernel for the three weight gradients with the bias column sums folded in; no atomics,
// so gradients are bit-reproducible run to run.
#pragma once
#include "common.cuh"

namespace nlbac {

constexpr int MLP_TM_LARGE = 8;   // rows per CTA (large batches); small batches use MLP_TM_SMALL so that B = 128 / 256
constexpr int MLP_TM_SMALL = 2;   // still spreads over 64 / 128 of the 148 SMs instead of 16 / 32
constexpr long long MLP_SMALL_B = 148 * 8;   // below this many rows the 8-row tiling leaves SMs idle
constexpr int MLP_NT = 256;       // threads per CTA
constexpr int MLP_MAX_OUT = 8;    // output columns (twin-Q: 1, policy: 2*n_u <= 4)
constexpr int MLP_MAX_IN = 32;

struct MlpArgs {
  const float *W1, *b1, *W2, *b2, *W3, *b3;  // nn.Linear layout: W [out, in]
  const float* x;    // [B, n_in]
  float* out;        // [B, n_out]
  float* act;        // [2, B, H] post-ReLU activations (saved for the backward) or null
  long long B;
  int n_in, H, n_out;
  int stage;         // layer 2 reads W2 through a transposing shared-memory stage (host: H <= 512 and it fits)
};

constexpr int MLP_KC = 32;        // W2 columns per staged chunk (one coalesced 128-byte row segment per warp load)
#ifndef NLBAC_MLP_NBUF
#define NLBAC_MLP_NBUF 2
#endif
#ifndef NLBAC_MLP_ROTATE
#define NLBAC_MLP_ROTATE 0   // measured: 12.7 -> 13.5 us at B = 128 (the modulo costs more than lock-step L2 reads)
#endif
constexpr int MLP_NBUF = NLBAC_MLP_NBUF;   // chunks in flight: every CTA pulls all of W2 (256 KB at H = 256) from L2, latency-bound
__host__ __device__ inline size_t mlp_stage_floats(int H) { return (size_t)MLP_NBUF * H * (MLP_KC + 4); }

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// dynamic smem: x[TM][n_in] | h1[TM][H] | h2[TM][H] | (stage) W2 chunks [MLP_NBUF][H][MLP_KC + 4]
template <int MLP_TM>
__global__ void __launch_bounds__(MLP_NT) mlp_fwd_kernel(MlpArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int H = a.H, NI = a.n_in, NO = a.n_out;
  float* sx = sm;
  float* h1 = sx + MLP_TM * MLP_MAX_IN;
  float* h2 = h1 + MLP_TM * H;
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * MLP_TM;
  const int rows = (int)min((long long)MLP_TM, a.B - row0);
  for (int i = tid; i < MLP_TM * NI; i += MLP_NT) {
    const int r = i / NI, c = i - r * NI;
    sx[r * MLP_MAX_IN + c] = r < rows ? __ldg(a.x + (row0 + r) * NI + c) : 0.f;
  }
  __syncthreads();
  // layer 1: one hidden unit per thread, all rows of the tile
  for (int j = tid; j < H; j += MLP_NT) {
    float acc[MLP_TM];
    const float bj = __ldg(a.b1 + j);
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) acc[r] = bj;
    const float* w = a.W1 + (size_t)j * NI;
    for (int k = 0; k < NI; ++k) {
      const float wv = __ldg(w + k);
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[r] = fmaf(sx[r * MLP_MAX_IN + k], wv, acc[r]);
    }
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) h1[r * H + j] = fmaxf(acc[r], 0.f);
  }
  __syncthreads();
  if (a.stage) {
    // layer 2, staged: W2 is nn.Linear layout [out j][in k], so "thread j walks row j" makes every warp load touch 32
    // different 128-byte lines.  Here the 32-column chunk W2[:, k0 .. k0+32) is copied with 16-byte cp.async into a
    // row-major tile sW[j][32 + 4] (8 lanes per 128-byte row segment: coalesced; the 144-byte row stride keeps the
    // destination 16-byte aligned and makes thread j's LDS.128 reads conflict-free), MLP_NBUF chunks in flight.
    // ncu at B = 128 (profiles/r2_ncu_mlp_b128.txt): the kernel is instruction-latency bound (2 warps per scheduler),
    // so what counts is instructions per thread: 8 copies + 8 LDS.128 per chunk instead of 32 + 32.
    float* sW = h2 + MLP_TM * H;
    constexpr int WS = MLP_KC + 4;                 // row stride of the staged tile (floats)
    const int nchunk = H / MLP_KC;
    const int c_rot = NLBAC_MLP_ROTATE ? (int)(blockIdx.x % (unsigned)nchunk) : 0;
    auto issue = [&](int c) {
      float* dst = sW + (size_t)(c % MLP_NBUF) * H * WS + (tid & 7) * 4;
      const float* src = a.W2 + ((c + c_rot) % nchunk) * MLP_KC + (tid & 7) * 4;
      for (int j = tid >> 3; j < H; j += MLP_NT / 8) cp_async16(dst + j * WS, src + (size_t)j * H);
      cp_async_commit();
    };
    float acc[2][MLP_TM];
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int j = tid + m * MLP_NT;
      const float bj = j < H ? __ldg(a.b2 + j) : 0.f;
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[m][r] = bj;
    }
    // MLP_NBUF - 1 chunks in flight; one (possibly empty) commit group per iteration keeps the wait count constant
#pragma unroll
    for (int c = 0; c < MLP_NBUF - 1; ++c) {
      if (c < nchunk) issue(c);
      else cp_async_commit();
    }
    for (int c = 0; c < nchunk; ++c) {
      cp_async_wait<MLP_NBUF - 2>();
      __syncthreads();           // chunk c landed for every thread; everyone is done with chunk c - 1's buffer
      if (c + MLP_NBUF - 1 < nchunk) issue(c + MLP_NBUF - 1);
      else cp_async_commit();
      const float* w = sW + (size_t)(c % MLP_NBUF) * H * WS;
      const float* hc = h1 + ((c + c_rot) % nchunk) * MLP_KC;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int j = tid + m * MLP_NT;
        if (j < H) {
#pragma unroll
          for (int k4 = 0; k4 < MLP_KC / 4; ++k4) {
            const float4 wv = *reinterpret_cast<const float4*>(w + j * WS + 4 * k4);
#pragma unroll
            for (int r = 0; r < MLP_TM; ++r) {
              const float4 hv = *reinterpret_cast<const float4*>(hc + r * H + 4 * k4);
              acc[m][r] = fmaf(hv.x, wv.x, acc[m][r]);
              acc[m][r] = fmaf(hv.y, wv.y, acc[m][r]);
              acc[m][r] = fmaf(hv.z, wv.z, acc[m][r]);
              acc[m][r] = fmaf(hv.w, wv.w, acc[m][r]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int j = tid + m * MLP_NT;
      if (j < H) {
#pragma unroll
        for (int r = 0; r < MLP_TM; ++r) h2[r * H + j] = fmaxf(acc[m][r], 0.f);
      }
    }
  } else {
    // layer 2 (H > 512): thread j walks row j of W2 (each 128-byte line is fetched once and consumed from L1).  Measured at B = 128
    // (torch.profiler on a GraphedUpdate replay, round 2): 8-row tiles 18.7 us -> 2-row tiles 14.2 us; deeper unrolling
    // neutral; one WARP per output with coalesced row reads + shuffle reductions 25.7 us (32 dependent reductions per
    // warp cost more than the wavefronts they save) -- dropped.
    for (int j = tid; j < H; j += MLP_NT) {
      float acc[MLP_TM];
      const float bj = a.b2[j];
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[r] = bj;
      const float4* w = reinterpret_cast<const float4*>(a.W2 + (size_t)j * H);
#pragma unroll 8
      for (int k4 = 0; k4 < H / 4; ++k4) {
        const float4 wv = __ldg(w + k4);
#pragma unroll
        for (int r = 0; r < MLP_TM; ++r) {
          const float4 hv = *reinterpret_cast<const float4*>(h1 + r * H + 4 * k4);
          acc[r] = fmaf(hv.x, wv.x, acc[r]);
          acc[r] = fmaf(hv.y, wv.y, acc[r]);
          acc[r] = fmaf(hv.z, wv.z, acc[r]);
          acc[r] = fmaf(hv.w, wv.w, acc[r]);
        }
      }
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) h2[r * H + j] = fmaxf(acc[r], 0.f);
    }
  }
  __syncthreads();
  if (a.act != nullptr) {
    for (int i = tid; i < rows * H; i += MLP_NT) {
      const int r = i / H, c = i - r * H;
      a.act[(row0 + r) * H + c] = h1[i];
      a.act[((size_t)a.B + row0 + r) * H + c] = h2[i];
    }
  }
  // layer 3: one warp per (row, output) dot product
  const int warp = tid / 32, lane = tid % 32;
  for (int p = warp; p < rows * NO; p += MLP_NT / 32) {
    const int r = p / NO, o = p - r * NO;
    float s = 0.f;
    for (int k = lane; k < H; k += 32) s = fmaf(h2[r * H + k], __ldg(a.W3 + (size_t)o * H + k), s);
    s = warp_sum(s);
    if (lane == 0) a.out[(row0 + r) * NO + o] = s + __ldg(a.b3 + o);
  }
}

struct MlpBwdArgs {
  const float *W1, *W2, *W3;
  const float* act;       // [2, B, H]
  const float* grad_out;  // [B, n_out]
  float* d1;              // [B, H] pre-activation cotangent of layer 1 (workspace)
  float* d2;              // [B, H] pre-activation cotangent of layer 2 (workspace)
  float* grad_x;          // [B, n_in] or null
  long long B;
  int n_in, H, n_out;
  int stage;              // the d1 pass reads W2 through a shared-memory stage (host: H <= 512 and it fits)
};
#ifndef NLBAC_MLP_BWD_NBUF
#define NLBAC_MLP_BWD_NBUF 3
#endif
constexpr int MLP_BWD_NBUF = NLBAC_MLP_BWD_NBUF;
__host__ __device__ inline size_t mlp_bwd_stage_floats(int H) { return (size_t)MLP_BWD_NBUF * MLP_KC * H; }

// dynamic smem: g[TM][MAX_OUT] | d2[TM][H] | d1[TM][H] | (stage) W2 row blocks [MLP_BWD_NBUF][MLP_KC][H]
template <int MLP_TM>
__global__ void __launch_bounds__(MLP_NT) mlp_bwd_rows_kernel(MlpBwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int H = a.H, NI = a.n_in, NO = a.n_out;
  float* sg = sm;
  float* s2 = sg + MLP_TM * MLP_MAX_OUT;
  float* s1 = s2 + MLP_TM * H;
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * MLP_TM;
  const int rows = (int)min((long long)MLP_TM, a.B - row0);
  for (int i = tid; i < MLP_TM * NO; i += MLP_NT) {
    const int r = i / NO, o = i - r * NO;
    sg[r * MLP_MAX_OUT + o] = r < rows ? __ldg(a.grad_out + (row0 + r) * NO + o) : 0.f;
  }
  __syncthreads();
  const float* h1 = a.act;
  const float* h2 = a.act + (size_t)a.B * H;
  // d2[r][j] = relu'(h2) * sum_o g[r][o] W3[o][j]
  for (int j = tid; j < H; j += MLP_NT) {
    float acc[MLP_TM];
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) acc[r] = 0.f;
    for (int o = 0; o < NO; ++o) {
      const float wv = __ldg(a.W3 + (size_t)o * H + j);
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[r] = fmaf(sg[r * MLP_MAX_OUT + o], wv, acc[r]);
    }
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) {
      float v = 0.f;
      if (r < rows) {
        v = __ldg(h2 + (row0 + r) * H + j) > 0.f ? acc[r] : 0.f;
        a.d2[(row0 + r) * H + j] = v;
      }
      s2[r * H + j] = v;
    }
  }
  __syncthreads();
  // d1[r][k] = relu'(h1) * sum_j d2[r][j] W2[j][k]   (W2 rows read coalesced across threads k)
  if (a.stage) {
    // 256 dependent L2 round trips per thread at 2 warps per scheduler made this loop latency-bound (ncu at B = 128:
    // 16 long-scoreboard stall cycles per issue).  Staged: W2 rows j0 .. j0+32 are one contiguous 32 * H block --
    // 16-byte cp.async, MLP_BWD_NBUF blocks in flight, thread k reads sW[jj][k] (consecutive banks).
    float* sW = s1 + MLP_TM * H;
    const int nchunk = H / MLP_KC, per_chunk4 = MLP_KC * H / 4;
    auto issue = [&](int c) {
      float* dst = sW + (size_t)(c % MLP_BWD_NBUF) * MLP_KC * H;
      const float* src = a.W2 + (size_t)c * MLP_KC * H;
      for (int i = tid; i < per_chunk4; i += MLP_NT) cp_async16(dst + 4 * i, src + 4 * i);
      cp_async_commit();
    };
    float acc[2][MLP_TM];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[m][r] = 0.f;
#pragma unroll
    for (int c = 0; c < MLP_BWD_NBUF - 1; ++c) {
      if (c < nchunk) issue(c);
      else cp_async_commit();
    }
    for (int c = 0; c < nchunk; ++c) {
      cp_async_wait<MLP_BWD_NBUF - 2>();
      __syncthreads();
      if (c + MLP_BWD_NBUF - 1 < nchunk) issue(c + MLP_BWD_NBUF - 1);
      else cp_async_commit();
      const float* w = sW + (size_t)(c % MLP_BWD_NBUF) * MLP_KC * H;
#pragma unroll
      for (int j4 = 0; j4 < MLP_KC / 4; ++j4) {
        float4 dv[MLP_TM];
#pragma unroll
        for (int r = 0; r < MLP_TM; ++r) dv[r] = *reinterpret_cast<const float4*>(s2 + r * H + c * MLP_KC + 4 * j4);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const int k = tid + m * MLP_NT;
          if (k < H) {
            const float w0 = w[(4 * j4 + 0) * H + k], w1 = w[(4 * j4 + 1) * H + k];
            const float w2 = w[(4 * j4 + 2) * H + k], w3 = w[(4 * j4 + 3) * H + k];
#pragma unroll
            for (int r = 0; r < MLP_TM; ++r) {
              acc[m][r] = fmaf(dv[r].x, w0, acc[m][r]);
              acc[m][r] = fmaf(dv[r].y, w1, acc[m][r]);
              acc[m][r] = fmaf(dv[r].z, w2, acc[m][r]);
              acc[m][r] = fmaf(dv[r].w, w3, acc[m][r]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int k = tid + m * MLP_NT;
      if (k < H) {
#pragma unroll
        for (int r = 0; r < MLP_TM; ++r) {
          float v = 0.f;
          if (r < rows) {
            v = __ldg(h1 + (row0 + r) * H + k) > 0.f ? acc[m][r] : 0.f;
            a.d1[(row0 + r) * H + k] = v;
          }
          s1[r * H + k] = v;
        }
      }
    }
  } else {
  for (int k = tid; k < H; k += MLP_NT) {
    float acc[MLP_TM];
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) acc[r] = 0.f;
#pragma unroll 8
    for (int j = 0; j < H; ++j) {
      const float wv = __ldg(a.W2 + (size_t)j * H + k);
#pragma unroll
      for (int r = 0; r < MLP_TM; ++r) acc[r] = fmaf(s2[r * H + j], wv, acc[r]);
    }
#pragma unroll
    for (int r = 0; r < MLP_TM; ++r) {
      float v = 0.f;
      if (r < rows) {
        v = __ldg(h1 + (row0 + r) * H + k) > 0.f ? acc[r] : 0.f;
        a.d1[(row0 + r) * H + k] = v;
      }
      s1[r * H + k] = v;
    }
  }
  }
  __syncthreads();
  if (a.grad_x != nullptr) {
    const int warp = tid / 32, lane = tid % 32;
    for (int p = warp; p < rows * NI; p += MLP_NT / 32) {
      const int r = p / NI, i = p - r * NI;
      float s = 0.f;
      for (int k = lane; k < H; k += 32) s = fmaf(s1[r * H + k], __ldg(a.W1 + (size_t)k * NI + i), s);
      s = warp_sum(s);
      if (lane == 0) a.grad_x[(row0 + r) * NI + i] = s;
    }
  }
}

// C[M,N] = A^T B over the batch:  C[m][n] = sum_b A[b][m] * Bm[b][n];  three problems per launch (blockIdx.z)
struct AtbProb { const float* A; const float* Bm; float* C; int M, N; float* cs; };   // cs: column sums of A (bias gradient) or null
struct AtbArgs { AtbProb p[3]; long long B; };
// Small batches (B < MLP_SPLIT_MIN_B).  The batch is consumed 128 rows per step -- at B = 128 ONE global round trip per
// CTA instead of four dependent ones (ncu: 12 long-scoreboard stall cycles per issue with 32-row steps) -- and the n = 0
// CTAs of every M tile also reduce the column sums of A (the bias gradients), which removes the separate column-sum launch.
constexpr int ATB_BK = 128;
__global__ void __launch_bounds__(256) mlp_atb_kernel(AtbArgs a) {
  const AtbProb p = a.p[blockIdx.z];
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  if (m0 >= p.M || n0 >= p.N) return;
  __shared__ float sA[ATB_BK][33], sB[ATB_BK][33];
  const int tx = threadIdx.x % 32, ty = threadIdx.x / 32;  // 32 x 8; each thread 4 rows of the 32x32 tile
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float csum = 0.f;
  const bool am = m0 + tx < p.M, bn = n0 + tx < p.N;
  for (long long b0 = 0; b0 < a.B; b0 += ATB_BK) {
    // two phases: ALL global loads first, then the shared-memory stores, so that the 32 loads of a step are in flight
    // together instead of being scheduled load -> store -> load
    // (measured at B = 128 / 256, where this loop runs once / twice: fwd+bwd 39.4 -> 36.9 us / 49.3 -> 43.6 us)
    float av[ATB_BK / 8], bv[ATB_BK / 8];
#pragma unroll
    for (int i = 0; i < ATB_BK / 8; ++i) {
      const long long b = b0 + ty + 8 * i;
      av[i] = (b < a.B && am) ? __ldg(p.A + b * p.M + m0 + tx) : 0.f;
      bv[i] = (b < a.B && bn) ? __ldg(p.Bm + b * p.N + n0 + tx) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < ATB_BK / 8; ++i) {
      const int bb = ty + 8 * i;
      sA[bb][tx] = av[i];
      csum += av[i];
      sB[bb][tx] = bv[i];
    }
    __syncthreads();
#pragma unroll 8
    for (int bb = 0; bb < ATB_BK; ++bb) {
      const float bv = sB[bb][tx];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(sA[bb][ty + 8 * i], bv, acc[i]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty + 8 * i, n = n0 + tx;
    if (m < p.M && n < p.N) p.C[(size_t)m * p.N + n] = acc[i];
  }
  if (p.cs != nullptr && n0 == 0) {      // fixed-order reduction of the 8 partials: bit-reproducible
    sB[ty][tx] = csum;
    __syncthreads();
    if (ty == 0 && am) {
      float t = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) t += sB[q][tx];
      p.cs[m0 + tx] = t;
    }
  }
}

// db[c] = sum_b D[b][c] for the batch-split path (small batches: folded into mlp_atb_kernel)
struct ColSumProb { const float* D; float* out; int N; };
struct ColSumArgs { ColSumProb p[3]; long long B; };

// ------------------------------------------------------------------------------------------------
// Large-batch forms of the two reductions above (B >= MLP_SPLIT_MIN_B: the vectorised environments of SURVEY 8f.4 feed
// 10^4 .. 10^6-row batches).  The small-batch kernels loop over the WHOLE batch in 192 / 24 CTAs: at B = 2^20 ncu timed
// them at 129 ms / 32 ms (profiles/r2_ncu_tail_kernels.txt).  Here the batch is cut into chunks (blockIdx.z / nsplit), a
// CTA forms a 64 x 64 tile of C over its chunk with a 4 x 4 register tile per thread, and the chunks meet in a
// pre-zeroed output through one atomicAdd per element per chunk (fp32 sums in a chunk-dependent order: 1e-6-level
// differences from run to run, which is why the small-batch kernels, bit-reproducible, stay the default below it).
// ------------------------------------------------------------------------------------------------
constexpr long long MLP_SPLIT_MIN_B = 4096;
struct AtbSplitArgs { AtbProb p[3]; long long B; int nsplit; };
__global__ void __launch_bounds__(256) mlp_atb_split_kernel(AtbSplitArgs a) {
  const AtbProb p = a.p[blockIdx.z / a.nsplit];
  const int chunk = blockIdx.z % a.nsplit;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  if (m0 >= p.M || n0 >= p.N) return;
  const long long per = ((a.B + a.nsplit - 1) / a.nsplit + 31) / 32 * 32;
  const long long b_lo = (long long)chunk * per, b_hi = min(a.B, b_lo + per);
  if (b_lo >= b_hi) return;
  __shared__ __align__(16) float sA[32][64], sB[32][64];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;      // thread tile: rows m0 + 4 ty .., columns n0 + 4 tx ..
  const int lr = threadIdx.x / 8, lc = (threadIdx.x % 8) * 8;  // loader: batch row lr of the stage, 8 columns at lc
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // Operands of a 32-row step go through registers: all 16 global loads are issued together, and the NEXT step's
  // loads are in flight while this step's products are formed.
  float ra[8], rb[8];
  auto gload = [&](long long b0) {
    const long long b = b0 + lr;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      ra[c] = (b < b_hi && m0 + lc + c < p.M) ? __ldg(p.A + b * p.M + m0 + lc + c) : 0.f;
      rb[c] = (b < b_hi && n0 + lc + c < p.N) ? __ldg(p.Bm + b * p.N + n0 + lc + c) : 0.f;
    }
  };
  gload(b_lo);
  for (long long b0 = b_lo; b0 < b_hi; b0 += 32) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      sA[lr][lc + c] = ra[c];
      sB[lr][lc + c] = rb[c];
    }
    __syncthreads();
    if (b0 + 32 < b_hi) gload(b0 + 32);
#pragma unroll 8
    for (int bb = 0; bb < 32; ++bb) {
      const float4 av = *reinterpret_cast<const float4*>(&sA[bb][4 * ty]);
      const float4 bv = *reinterpret_cast<const float4*>(&sB[bb][4 * tx]);
      const float am[4] = {av.x, av.y, av.z, av.w}, bn[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(am[i], bn[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + 4 * ty + i, n = n0 + 4 * tx + j;
      if (m < p.M && n < p.N) atomicAdd(p.C + (size_t)m * p.N + n, acc[i][j]);
    }
}

struct ColSumSplitArgs { ColSumProb p[3]; long long B; int nsplit; };
__global__ void __launch_bounds__(256) mlp_colsum_split_kernel(ColSumSplitArgs a) {
  const ColSumProb p = a.p[blockIdx.y];
  const int c = blockIdx.x * 32 + threadIdx.x % 32;
  const int part = threadIdx.x / 32;
  const long long per = (a.B + a.nsplit - 1) / a.nsplit;
  const long long b_lo = (long long)blockIdx.z * per, b_hi = min(a.B, b_lo + per);
  __shared__ float red[8][33];
  float s = 0.f;
  if (c < p.N)
    for (long long b = b_lo + part; b < b_hi; b += 8) s += p.D[b * p.N + c];
  red[part][threadIdx.x % 32] = s;
  __syncthreads();
  if (part == 0 && c < p.N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(p.out + c, t);
  }
}

}  // namespace nlbac
