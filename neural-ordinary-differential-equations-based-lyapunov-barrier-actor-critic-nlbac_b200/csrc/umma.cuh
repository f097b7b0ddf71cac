// tcgen05 (5th-generation tensor core) primitives for sm_100a: TMEM allocation, shared-memory matrix
// descriptors, kind::tf32 MMA issue, commit -> mbarrier, TMEM -> register loads.
//
// Operand layout used throughout (no swizzle, "interleave" canonical form, 32-bit elements):
//   a tile X[r][c] (r = row held by its owner thread, c = feature) is stored as 16-byte chunks
//       byte(r, c) = (c / 4) * CHUNK + (r / 8) * 128 + (r % 8) * 16 + (c % 4) * 4,     CHUNK = ROWS * 16
//   so one "core matrix" (8 rows x 16 bytes) is 128 contiguous bytes.  As a K-major kind::tf32 operand
//   (rows = M/N, c = K): SBO = 128 (next 8 rows), LBO = CHUNK (next 4 k); one MMA consumes K = 8 -> advance the
//   start address by 2 * CHUNK per k-step.  kind::tf32 has NO MN-major form in this layout (it needs the 128-byte
//   swizzle; measured all-zero accumulators), so products that contract over the row index use 16-bit images:
//       byte16(r, c) = (c / 8) * CHUNK + (r / 8) * 128 + (r % 8) * 16 + (c % 8) * 2
//   read by kind::f16 as an MN-major operand (c = M/N, rows = K): LBO = 128 (next 8 k), SBO = CHUNK (next 8 m/n),
//   one MMA consumes K = 16 rows -> advance by 256 per k-step.
#pragma once
#include "common.cuh"

namespace nlbac {
namespace umma {

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 64-bit shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (Blackwell)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

// 32-bit instruction descriptor: D = F32, A = B = TF32, dense
__host__ __device__ constexpr uint32_t make_idesc_tf32(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// 32-bit instruction descriptor: D = F32, A = B = BF16, dense
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// byte offset of element (r, c) of a 16-bit tile: 16-byte chunks of 8 features
__host__ __device__ constexpr int tile_off16(int r, int c, int chunk_bytes) {
  return (c >> 3) * chunk_bytes + (r >> 3) * 128 + (r & 7) * 16 + (c & 7) * 2;
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// same with the A operand in TENSOR MEMORY: row r of A = TMEM lane r, element k of a 32-bit kind = column a_tmem + k
// (one MMA consumes K = 8 columns); the row-owner thread writes its activations there with tcgen05.st
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mbarrier arrives once all MMAs issued so far by this thread have completed (implies fence::before_thread_sync)
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// TMEM -> registers: the calling warp reads ITS 32 lanes (lane quadrant = warp_id % 4), 8 consecutive columns
__device__ __forceinline__ void ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// same load straight into float registers (a .b32 operand accepts .f32 registers): no move between the load and its
// wait, which matters once loads are software-pipelined
__device__ __forceinline__ void ld8f(uint32_t taddr, float (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void wait_ld8(float (&a)[8]) {   // wait::ld that names the registers of ONE earlier ld8f
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])
               :
               : "memory");
}
// wait::ld that NAMES the registers two earlier ld8 wrote: later reads of v0 / v1 cannot be scheduled above the wait
// (needed once loads are software-pipelined: another tcgen05.ld is already in flight when these values are consumed)
__device__ __forceinline__ void wait_ld16(float (&a)[8], float (&b)[8]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7]),
                 "+f"(b[0]), "+f"(b[1]), "+f"(b[2]), "+f"(b[3]), "+f"(b[4]), "+f"(b[5]), "+f"(b[6]), "+f"(b[7])
               :
               : "memory");
}

// registers -> TMEM: the calling warp writes ITS 32 lanes, 8 consecutive columns
__device__ __forceinline__ void st8(uint32_t taddr, const float (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
               "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
               "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
               "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
               : "memory");
}
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, int lane, int col) {
  return base + ((uint32_t)lane << 16) + (uint32_t)col;
}

// byte offset of element (r, c) of a tile stored in the layout described at the top
__host__ __device__ constexpr int tile_off(int r, int c, int chunk_bytes) {
  return (c >> 2) * chunk_bytes + (r >> 3) * 128 + (r & 7) * 16 + (c & 3) * 4;
}

}  // namespace umma
}  // namespace nlbac
