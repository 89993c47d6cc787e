// tcgen05 / TMEM / mbarrier primitives for sm_100a, hand-written PTX (no CUTLASS on the path).
//
// Operand staging: every tensor-core operand is an fp32 (tf32 container) matrix in shared memory in
// the UMMA canonical NO-swizzle K-major layout (core matrix = 8 x 16 bytes), in one of two forms:
//   "ns": column-wise, K = rows    -> Gram products (V^T A, V^T V) and the small B operands (W, X)
//   "rn": row-wise,    K = columns -> the A operand of the updates (A -= V W, Q = -V X)
// kind::tf32 has no usable MN-major form for these buffers (measured: such MMAs leave the
// accumulator untouched), so a reflector block is stored once in each form.
//
// hi/lo error correction: hi = tf32(x) rounded to nearest (cvt.rna), lo = x - hi (exact in
// fp32, read by the tensor core as tf32). A corrected product is three MMAs into one fp32
// TMEM accumulator: hi*hi + hi*lo + lo*hi. This is the split of reference
// src/tsqr.cu:546-549 / src/tcqr32x16.cu:923-926 fused into the operand store.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace tsqr_b200 {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// round to tf32 (10 mantissa bits), ties away from zero: bit-identical to cvt.rna.tf32.f32 for every finite
// input and for +-inf (the PTX instruction compiles to a 4-5 instruction sequence with NaN handling; the
// sign-magnitude add + mask is 2 ALU instructions; NaN inputs stay NaN or become inf, there are none here)
__device__ __forceinline__ float to_tf32_rna(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

// store x into the hi (and, when COR, lo) tile buffers
template <bool COR>
__device__ __forceinline__ void tile_store(char* hi, char* lo, uint32_t off, float x) {
  const float h = to_tf32_rna(x);
  *reinterpret_cast<float*>(hi + off) = h;
  if (COR) *reinterpret_cast<float*>(lo + off) = __fsub_rn(x, h);
}

// two values that are adjacent in the tile layout (8-byte aligned offset)
template <bool COR>
__device__ __forceinline__ void tile_store2(char* hi, char* lo, uint32_t off, float x0, float x1) {
  const float h0 = to_tf32_rna(x0), h1 = to_tf32_rna(x1);
  *reinterpret_cast<float2*>(hi + off) = make_float2(h0, h1);
  if (COR) *reinterpret_cast<float2*>(lo + off) = make_float2(__fsub_rn(x0, h0), __fsub_rn(x1, h1));
}

// ---- descriptors
// shared-memory matrix descriptor, version 1 (Blackwell); layout_type 0 = no swizzle
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                             uint32_t layout_type = 0) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version
  d |= (uint64_t)layout_type << 61;  // 0 = no swizzle, 2 = SWIZZLE_128B
  return d;
}

// instruction descriptor for kind::tf32, fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, bool a_mn_major, bool b_mn_major, bool neg_a) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((neg_a ? 1u : 0u) << 13) | ((a_mn_major ? 1u : 0u) << 15) |
         ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}

// ---- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Bounded spin: a wait that never completes (a lost arrive, a wrong transaction count) traps the kernel
// after ~2^26 probes instead of hanging the device.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok, spins = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (!ok && ++spins > (1u << 26)) __trap();
  } while (!ok);
}
// all previously issued tcgen05.mma of this thread arrive on `bar` when they complete
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- fences
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }


// D[tmem] (+)= A[tmem] * B[smem]: the A operand (M = 128 lanes x K = 8 consecutive 32-bit columns) is read
// straight from tensor memory, where the threads put it with tcgen05.st (row r of the tile = lane r)
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}

// plain arrive (release at CTA scope)
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- TMA (cp.async.bulk.tensor): one elected thread arms the mbarrier with the byte count of the box and
// issues the copy; the box lands in shared memory and completes the transaction on the barrier
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// box origin (c0 = index along the contiguous dimension, c1 = index along the strided one)
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* tmap, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tmap),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the shared-memory source of every committed store group has been read (the buffer may be rewritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// =========================================================================================
// "sw" tile buffer: K-major operand with the 128-byte hardware swizzle (layout_type 2), K = tile rows.
// One K block = 32 rows (128 bytes of a column); inside a K block, operand row m (= matrix column) is a
// 128-byte line whose 16-byte chunks are XOR-swizzled with m % 8; 8 lines form a 1024-byte atom
// (SBO = 1024). A K block holds kLines operand rows, so hi (lines [0, NP)) and lo (lines [NP, 2 NP))
// parts of one tile are ONE M = 128 operand. Thread-per-row stores of one column hit 32 distinct banks.
namespace sw {
constexpr uint32_t kLines = 128;                       // operand rows per K block
constexpr uint32_t kBlock = kLines * 128;              // 16384 bytes per K block
__host__ __device__ constexpr uint32_t bytes(int rows) { return (uint32_t)(rows / 32) * kBlock; }
__device__ __forceinline__ uint32_t off(int row, int line) {
  return (uint32_t)(row >> 5) * kBlock + (uint32_t)line * 128u + (uint32_t)((((row & 31) >> 2) ^ (line & 7)) << 4) +
         (uint32_t)((row & 3) << 2);
}
// D[128 x N] = sum over rows [0, krows) of OP[m][row] * OP[n][row], m < 128 lines, n < N lines
__device__ __forceinline__ void gram_issue(uint32_t d_tmem, uint32_t base, int N, int krows) {
  const uint32_t id = idesc_tf32(128, N, false, false, false);
  bool acc = false;
  for (int kb = 0; kb < krows / 32; ++kb) {
    const uint64_t d0 = smem_desc(base + (uint32_t)kb * kBlock, 16, 1024, 2);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const uint64_t d = d0 + (uint64_t)(kk * 2);   // +32 bytes in the 16-byte units of the address field
      mma_tf32(d_tmem, d, d, id, acc);
      acc = true;
    }
  }
}
}  // namespace sw

// small K-major no-swizzle B operand blocks with dense (128-byte) core matrices, written once per kernel:
// element (k, c) of a K x N block: (c/8)*SBO + (k/4)*128 + (c%8)*16 + (k%4)*4, SBO = (K/4)*128
namespace kb {
__host__ __device__ constexpr uint32_t sbo(int K) { return (uint32_t)(K / 4) * 128u; }
__host__ __device__ constexpr uint32_t bytes(int K, int N) { return (uint32_t)(N / 8) * sbo(K); }
__device__ __forceinline__ uint32_t off(int k, int c, int K) {
  return (uint32_t)(c >> 3) * sbo(K) + (uint32_t)(k >> 2) * 128u + (uint32_t)((c & 7) << 4) + (uint32_t)((k & 3) << 2);
}
}  // namespace kb

// ---- TMEM allocation (one full warp calls alloc / dealloc)
template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}

// ---- TMEM <-> registers, shape 32x32b: thread (warp w, lane l) owns TMEM lane 32*(w%4)+l and
// reads/writes consecutive 32-bit columns starting at the column of taddr.
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// the same load without the wait: issue several, then tmem_wait_ld() once
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr),
               "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
               "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
               "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
               : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// TMEM address of (lane base of this warp, column)
__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, int warp, int col) {
  return base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)col;
}

// =========================================================================================
// "ns" tile buffer: NO-swizzle canonical layout with a padded core-matrix stride. Core matrix =
// 8 columns x 4 rows (8 x 16 bytes, contiguous 128 B):
//     off(row, col) = (col/8)*SB + (row/4)*144 + (col%8)*16 + (row%4)*4,   SB = (rows/4)*144
// K-major view (K = rows, MN = cols): LBO = 144, SBO = SB. MN-major view (MN = rows, K = cols):
// SBO = 144, LBO = SB. The 144-byte stride makes thread-per-row scalar stores conflict free.
namespace ns {
constexpr uint32_t kCM = 144;
__host__ __device__ constexpr uint32_t sb(int rows) { return (uint32_t)(rows / 4) * kCM; }
__host__ __device__ constexpr uint32_t bytes(int rows, int cols) { return (uint32_t)(cols / 8) * sb(rows); }
__device__ __forceinline__ uint32_t off(int row, int col, int rows) {
  return (uint32_t)(col >> 3) * sb(rows) + (uint32_t)(row >> 2) * kCM + (uint32_t)((col & 7) << 4) + (uint32_t)((row & 3) << 2);
}
// column part of off(): off(row, col, rows) = off(row, 0, rows) + coloff(col, rows)
__host__ __device__ constexpr uint32_t coloff(int col, int rows) { return (uint32_t)(col >> 3) * sb(rows) + (uint32_t)((col & 7) << 4); }
// D[128 x N] (+)= A[:, a_c0 + m]^T B[:, b_c0 + n] over rows [0, krows)
template <bool COR>
__device__ __forceinline__ void gram_issue(uint32_t d_tmem, uint32_t a_hi, uint32_t a_lo, int a_rows, int a_c0,
                                           uint32_t b_hi, uint32_t b_lo, int b_rows, int b_c0, int N, int krows,
                                           bool accumulate) {
  const uint32_t id = idesc_tf32(128, N, false, false, false);
  const uint32_t asb = sb(a_rows), bsb = sb(b_rows);
  const uint32_t a0 = (uint32_t)(a_c0 >> 3) * asb, b0 = (uint32_t)(b_c0 >> 3) * bsb;
  // the start-address field is the low 14 bits (16-byte units): one K step of 8 rows = 2 core matrices
  uint64_t ah = smem_desc(a_hi + a0, kCM, asb), bh = smem_desc(b_hi + b0, kCM, bsb);
  uint64_t al = COR ? smem_desc(a_lo + a0, kCM, asb) : 0, bl = COR ? smem_desc(b_lo + b0, kCM, bsb) : 0;
  constexpr uint64_t kStep = (2 * kCM) >> 4;
  bool acc = accumulate;
#pragma unroll 4
  for (int k = 0; k < krows; k += 8) {
    mma_tf32(d_tmem, ah, bh, id, acc);
    acc = true;
    if (COR) {
      mma_tf32(d_tmem, ah, bl, id, true);
      mma_tf32(d_tmem, al, bh, id, true);
      al += kStep; bl += kStep;
    }
    ah += kStep; bh += kStep;
  }
}
}  // namespace ns

// "rn" buffer: 128 rows x cols stored row-wise without swizzle, a K-major operand with K = columns.
// Core matrix = 8 rows x 4 columns: off(row, col) = (col/4)*2048 + (row/8)*128 + (row%8)*16 + (col%4)*4
// (LBO = 2048 between K core matrices, SBO = 128 between 8-row groups). A thread that owns a row
// stores 4 consecutive columns as one float4; 8 consecutive lanes fill one 128-byte core matrix.
namespace rn {
__host__ __device__ constexpr uint32_t bytes(int cols) { return (uint32_t)cols * 512u; }
__device__ __forceinline__ uint32_t off4(int row, int col4) {  // col4 = col / 4
  return (uint32_t)col4 * 2048u + (uint32_t)(row >> 3) * 128u + (uint32_t)((row & 7) << 4);
}
template <bool COR>
__device__ __forceinline__ void store4(char* hi, char* lo, int row, int col4, float x0, float x1, float x2, float x3) {
  const uint32_t o = off4(row, col4);
  const float4 h = make_float4(to_tf32_rna(x0), to_tf32_rna(x1), to_tf32_rna(x2), to_tf32_rna(x3));
  *reinterpret_cast<float4*>(hi + o) = h;
  if (COR) *reinterpret_cast<float4*>(lo + o) = make_float4(__fsub_rn(x0, h.x), __fsub_rn(x1, h.y), __fsub_rn(x2, h.z), __fsub_rn(x3, h.w));
}
// D[128 rows x N] (+)= (neg ? -1 : 1) * A[:, a_c0 : a_c0+kcols] * B[b_k0 : b_k0+kcols, b_c0 : b_c0+N]
// A: rn buffer; B: ns buffer with b_rows rows (K-major view, K = its rows)
template <bool COR>
__device__ __forceinline__ void update_issue(uint32_t d_tmem, uint32_t a_hi, uint32_t a_lo, int a_c0, uint32_t b_hi,
                                             uint32_t b_lo, int b_rows, int b_c0, int b_k0, int N, int kcols, bool neg,
                                             bool accumulate) {
  const uint32_t id = idesc_tf32(128, N, false, false, neg);
  const uint32_t bsb = ns::sb(b_rows);
  const uint32_t ao = (uint32_t)(a_c0 >> 2) * 2048u;
  const uint32_t bo = (uint32_t)(b_c0 >> 3) * bsb + (uint32_t)(b_k0 >> 2) * ns::kCM;
  uint64_t ah = smem_desc(a_hi + ao, 2048, 128), bh = smem_desc(b_hi + bo, ns::kCM, bsb);
  uint64_t al = COR ? smem_desc(a_lo + ao, 2048, 128) : 0, bl = COR ? smem_desc(b_lo + bo, ns::kCM, bsb) : 0;
  constexpr uint64_t kStepA = 4096 >> 4, kStepB = (2 * ns::kCM) >> 4;
  bool acc = accumulate;
#pragma unroll 2
  for (int k = 0; k < kcols; k += 8) {
    mma_tf32(d_tmem, ah, bh, id, acc);
    acc = true;
    if (COR) {
      mma_tf32(d_tmem, ah, bl, id, true);
      mma_tf32(d_tmem, al, bh, id, true);
      al += kStepA; bl += kStepB;
    }
    ah += kStepA; bh += kStepB;
  }
}
}  // namespace rn


}  // namespace tc
}  // namespace tsqr_b200
