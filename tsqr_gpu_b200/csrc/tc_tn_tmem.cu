// BlockQR projection coefficients C = X^T Y for the corrected fp32 modes with the A OPERAND IN TENSOR MEMORY
// (reference src/blockqr.cu:92-116 / :272-291, first GEMM per panel). tc_tn_tma.cu stages both operands in shared memory and is
// bound by shared-memory bandwidth (128 KB of traffic per 16 KB block of X); here a block of X goes from its landing buffer
// straight into tensor memory, transposed: accumulator row = TMEM lane = column of X.
//   * the landing buffer is written by TMA with the 128-byte swizzle (boxes of 32 rows x 64 columns: one 128-byte line per
//     column and 32 rows), so that the thread that owns column c reads its 64 rows with 16 conflict-free LDS.128;
//   * two groups of 128 threads alternate over the blocks; in a group, thread L < 64 stores tf32(x) of column L, thread
//     64 + L stores x - tf32(x): the [X hi | X lo]^T operand (128 lanes x 64 K columns) by tcgen05.st;
//   * Y stays a shared-memory operand (K-major, [Y hi | Y lo] = 128 lines), converted by the 256 threads of both groups at
//     the start of a tile;
//   * D_b[128 x 128] += A_tmem . Y: 8 products of K = 8 per block; 3 accumulator blocks (192 columns of X) per launch:
//     512 TMEM columns = 3 x 128 (D) + 2 x 64 (A, one buffer per group).
// Shared-memory traffic per block: 16 (bulk write) + 32 (both halves read the block) + 32 (Y operand reads) = 80 KB.
#include <cuda.h>
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "tc05.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {
using namespace tc;

constexpr int kTile = 64;                    // rows per tile
constexpr int kXSlots = 2, kYSlots = 2;      // landing rings: kXSlots per lane group (a ring has ONE consumer group, so that
                                             // every phase of its barriers is observed in order), kYSlots for Y
constexpr uint32_t kSlot = 64u * 64u * 4u;   // 16 KB: one 64 x 64 fp32 block
constexpr uint32_t kOp = 2 * sw::kBlock;     // Y operand buffer: 2 K blocks x 128 lines
constexpr int kWarpMma = 8, kWarpTma = 9;   // warps 0-7: the two lane groups (together they also convert the Y block of a tile)
constexpr int kThreads = 10 * 32;
constexpr int kMaxBlk = 3;

struct TmParams {
  int kb, nb, nblk;
  long long M;
  int tiles, tiles_per_cta;
  float* Cp;
};

struct TmSmem {
  static constexpr uint32_t oBar = 0;
  static constexpr uint32_t oXland = 1024;
  static constexpr uint32_t oYland = oXland + 2 * kXSlots * kSlot;
  static constexpr uint32_t oYop = oYland + kYSlots * kSlot;
  static constexpr uint32_t kBytes = oYop + 2 * kOp;
};

__global__ void __launch_bounds__(kThreads, 1) tc_tn_tmem_kernel(const TmParams P, const __grid_constant__ CUtensorMap tmapX,
                                                                const __grid_constant__ CUtensorMap tmapY) {
  extern __shared__ __align__(1024) char sm[];
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + TmSmem::oBar);
  uint64_t* xfull = bars;                   // [2][kXSlots] X block landed in slot s of group g's ring
  uint64_t* xempty = xfull + 2 * kXSlots;   // [2][kXSlots] read by the 128 threads of the group
  uint64_t* yfull = xempty + 2 * kXSlots;   // [kYSlots]
  uint64_t* yempty = yfull + kYSlots;   // [kYSlots] read by the 256 threads of both groups
  uint64_t* bar_ra = yempty + kYSlots;  // [2] A buffer of group g written (128 arrivals)
  uint64_t* bar_ry = bar_ra + 2;        // [2] Y operand buffer written (256 arrivals)
  uint64_t* bar_mma = bar_ry + 2;       // [2] products that read A buffer g are complete
  uint64_t* bar_yfree = bar_mma + 2;    // [2] every product of the tile that used Y operand buffer s is complete
  uint64_t* bar_fin = bar_yfree + 2;    // every product of the CTA is complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + TmSmem::oBar + 256);
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  if (t == 0) {
    for (int i = 0; i < 2 * kXSlots; ++i) { mbar_init(xfull + i, 1); mbar_init(xempty + i, 128); }
    for (int i = 0; i < kYSlots; ++i) { mbar_init(yfull + i, 1); mbar_init(yempty + i, 256); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_ra + i, 128); mbar_init(bar_ry + i, 256); mbar_init(bar_mma + i, 1); mbar_init(bar_yfree + i, 1); }
    mbar_init(bar_fin, 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmapX);
    tma_prefetch_desc(&tmapY);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const uint32_t tmA = tm + 128u * kMaxBlk;   // A buffers: [tmA, tmA + 64) group 0, [tmA + 64, tmA + 128) group 1
  const int nblk = P.nblk;
  const int tile_beg = blockIdx.x * P.tiles_per_cta;
  const int tile_end = min(tile_beg + P.tiles_per_cta, P.tiles);
  const int ntl = max(tile_end - tile_beg, 0);
  const int nx = ntl * nblk;   // X blocks of this CTA, in order q = tile * nblk + block

  if (warp == kWarpTma) {
    // ---- producer: per tile the Y block, then its X blocks; each ring has its own slot counter
    if (lane == 0) {
      int ys = 0, xn = 0;
      for (int i = 0; i < ntl; ++i) {
        if (i >= kYSlots) mbar_wait(yempty + ys, (uint32_t)(i / kYSlots - 1) & 1u);
        mbar_arrive_expect_tx(yfull + ys, kSlot);
        tma_load_2d(sm + TmSmem::oYland + ys * kSlot, &tmapY, yfull + ys, (tile_beg + i) * kTile, 0);
        if (++ys == kYSlots) ys = 0;
        for (int b = 0; b < nblk; ++b, ++xn) {
          const int g = xn & 1, k = xn >> 1, xs = g * kXSlots + k % kXSlots;   // block xn = item k of group g
          if (k >= kXSlots) mbar_wait(xempty + xs, (uint32_t)(k / kXSlots - 1) & 1u);
          mbar_arrive_expect_tx(xfull + xs, kSlot);
          char* dst = sm + TmSmem::oXland + xs * kSlot;
          tma_load_2d(dst, &tmapX, xfull + xs, (tile_beg + i) * kTile, 64 * b);              // rows [0, 32): lines of K block 0
          tma_load_2d(dst + 8192, &tmapX, xfull + xs, (tile_beg + i) * kTile + 32, 64 * b);  // rows [32, 64)
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ---- tensor-core issuer
    if (lane == 0) {
      const uint32_t idesc = idesc_tf32(128, 128, false, false, false);
      for (int i = 0; i < ntl; ++i) {
        mbar_wait(bar_ry + (i & 1), (uint32_t)(i >> 1) & 1u);
        const uint32_t ya = smem_u32(sm + TmSmem::oYop + (i & 1) * kOp);
        for (int b = 0; b < nblk; ++b) {
          const int q = i * nblk + b, g = q & 1;
          mbar_wait(bar_ra + g, (uint32_t)(q >> 1) & 1u);
          fence_after_sync();
          const uint32_t d = tm + 128u * (uint32_t)b;
#pragma unroll
          for (int kblk = 0; kblk < 2; ++kblk) {
            const uint64_t dy = smem_desc(ya + (uint32_t)kblk * sw::kBlock, 16, 1024, 2);   // N = 128: lines [Y hi | Y lo]
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              mma_tf32_ts(d, tmA + 64u * (uint32_t)g + 32u * (uint32_t)kblk + 8u * (uint32_t)kk, dy + (uint64_t)(kk * 2), idesc,
                          i > 0 || kblk > 0 || kk > 0);
          }
          mma_commit(bar_mma + g);
        }
        mma_commit(bar_yfree + (i & 1));
      }
      mma_commit(bar_fin);
    }
  } else {
    // ---- lane groups: group g = warp / 4 takes the X blocks q with q % 2 == g; thread L of the group = TMEM lane L
    const int g = warp >> 2, L = t & 127, col = L & 63;
    const bool lo_half = L >= 64;
    const uint32_t tmAg = tmA + 64u * (uint32_t)g;
    // Y conversion (all 256 threads): thread = (row, 16-column group) of the 64 x 64 block
    const int yrow = t & 63, ycg = t >> 6;
    uint32_t offs[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) offs[e] = sw::off(yrow, e);
    int ys = 0;
    uint32_t ypar = 0;
    int done = 0, k = 0;   // this group's products known complete; its next item
    for (int i = 0; i < ntl; ++i) {
      {
        // the Y operand buffer is free once every product of tile i - 2 is complete (one commit per tile on
        // bar_yfree[i & 1]: its next completion needs this very conversion, so the awaited phase cannot be overtaken)
        if (i >= 2) mbar_wait(bar_yfree + (i & 1), (uint32_t)((i >> 1) - 1) & 1u);
        char* op = sm + TmSmem::oYop + (i & 1) * kOp + (16 * ycg) * 128;
        mbar_wait(yfull + ys, ypar);
        const float* land = reinterpret_cast<const float*>(sm + TmSmem::oYland + ys * kSlot) + (16 * ycg) * kTile + yrow;
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const float x = land[c * kTile];
          const float h = to_tf32_rna(x);
          *reinterpret_cast<float*>(op + (c & ~7) * 128 + offs[c & 7]) = h;
          *reinterpret_cast<float*>(op + 64 * 128 + (c & ~7) * 128 + offs[c & 7]) = __fsub_rn(x, h);
        }
        fence_async_smem();
        mbar_arrive(yempty + ys);
        mbar_arrive(bar_ry + (i & 1));
        if (++ys == kYSlots) { ys = 0; ypar ^= 1u; }
      }
      for (int b = 0; b < nblk; ++b) {
        if (((i * nblk + b) & 1) != g) continue;   // the other group's block
        const int slot = g * kXSlots + k % kXSlots;   // item k of this group's ring
        const uint32_t par = (uint32_t)(k / kXSlots) & 1u;
        mbar_wait(xfull + slot, par);
        const char* land = sm + TmSmem::oXland + slot * kSlot + col * 128;
        float v[64];
#pragma unroll
        for (int kblk = 0; kblk < 2; ++kblk)
#pragma unroll
          for (int j = 0; j < 8; ++j) {   // rows 32 kblk + 4 j .. + 3 of the column sit in chunk j ^ (col % 8) of its line
            const float4 x4 = *reinterpret_cast<const float4*>(land + kblk * 8192 + ((j ^ (col & 7)) << 4));
            const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float h = to_tf32_rna(xs[e]);
              v[32 * kblk + 4 * j + e] = lo_half ? __fsub_rn(xs[e], h) : h;
            }
          }
        fence_async_smem();   // the reads above precede the bulk write that reuses the slot
        mbar_arrive(xempty + slot);
        if (k >= 1) {   // the products of this group's previous block read the A buffer
          mbar_wait(bar_mma + g, (uint32_t)done & 1u);
          ++done;
          fence_after_sync();
        }
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 16) {
          float w[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) w[e] = v[c0 + e];
          tmem_st16(tmem_addr(tmAg + c0, warp, 0), w);
        }
        tmem_wait_st();
        fence_before_sync();
        mbar_arrive(bar_ra + g);
        ++k;
      }
    }
    // ---- all products (the issuer's last commit), then the accumulators: lane l < 64: row l of X hi^T [Y hi | Y lo] (the two halves are summed),
    // lane 64 + l: row l of X lo^T Y hi
    mbar_wait(bar_fin, 0);
    fence_after_sync();
    const int l = L & 63, half = L >> 6;
    float* out = P.Cp + ((size_t)blockIdx.x * 2 + half) * (size_t)P.kb * P.nb;
    for (int b = g; b < nblk; b += 2) {
      const int irow = 64 * b + l;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float a[16], w[16];
        tmem_ld16(tmem_addr(tm + 128 * b, warp, c0), a);
        tmem_ld16(tmem_addr(tm + 128 * b, warp, 64 + c0), w);
        if (irow < P.kb) {
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (c0 + e < P.nb) out[irow + (size_t)(c0 + e) * P.kb] = ntl == 0 ? 0.f : (half == 0 ? __fadd_rn(a[e], w[e]) : a[e]);
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeFn encode_fn() {
  static EncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<EncodeFn>(p);
  }();
  return fn;
}
// column-major M x n fp32 matrix; box = box_rows x 64 columns, optionally with the 128-byte swizzle (box_rows = 32)
bool f32_map(CUtensorMap* map, const void* A, long long lda, long long M, int n, int box_rows, bool swizzle) {
  EncodeFn fn = encode_fn();
  if (!fn || (reinterpret_cast<uintptr_t>(A) & 15u) != 0 || ((size_t)lda * 4) % 16 != 0 || M >= (1ll << 31)) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)M, (cuuint64_t)n};
  cuuint64_t gstr[1] = {(cuuint64_t)lda * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_rows, 64u};
  cuuint32_t estr[2] = {1u, 1u};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(A), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// C[k0 + i, c] (ld ldc) = sum over the partial slots, in slot order (bit-reproducible)
__global__ void __launch_bounds__(256) tm_sum_slots_kernel(const float* __restrict__ Cp, int slots, int kg, int nb, float* __restrict__ C, long long ldc) {
  // 64 elements per CTA x 4 slot groups: thread (e, sg) adds slots sg, sg + 4, ...; the four sums are added in fixed order
  __shared__ float part[4][64];
  const size_t count = (size_t)kg * nb;
  const int sg = threadIdx.x >> 6;
  const size_t e = (size_t)blockIdx.x * 64 + (threadIdx.x & 63);
  float acc = 0.f;
  if (e < count) {
#pragma unroll 8
    for (int sl = sg; sl < slots; sl += 4) acc += Cp[(size_t)sl * count + e];
  }
  part[sg][threadIdx.x & 63] = acc;
  __syncthreads();
  if (sg == 0 && e < count) {
    const int i = threadIdx.x;
    C[(e % kg) + (e / kg) * ldc] = (part[0][i] + part[1][i]) + (part[2][i] + part[3][i]);
  }
}

}  // namespace

// fp32 corrected modes only. false: nothing was launched (not applicable); the caller takes tc_tn_tma.cu.
bool launch_tc_tn_tmem(const void* X, long long ldx, int kb, const void* Y, long long ldy, int nb, long long M, int max_slabs, float* Cp,
                       float* C, cudaStream_t s) {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_TN_TMEM"); return !(e && e[0] == '0'); }();
  if (!on || kb < 1 || max_slabs < 1) return false;
  constexpr int kmax = 64 * kMaxBlk;
  const int ngroups = (kb + kmax - 1) / kmax;
  CUtensorMap mx[8], my;
  if (ngroups > 8 || !f32_map(&my, Y, ldy, M, nb, kTile, false)) return false;
  for (int gi = 0; gi < ngroups; ++gi)
    if (!f32_map(&mx[gi], static_cast<const char*>(X) + (size_t)gi * kmax * (size_t)ldx * 4, ldx, M, std::min(kmax, kb - gi * kmax), 32, true))
      return false;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(tc_tn_tmem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TmSmem::kBytes));
  for (int gi = 0; gi < ngroups; ++gi) {
    const int k0 = gi * kmax, kg = std::min(kmax, kb - k0);
    TmParams p{};
    p.kb = kg; p.nb = nb; p.nblk = (kg + 63) / 64; p.M = M; p.Cp = Cp;
    p.tiles = (int)((M + kTile - 1) / kTile);
    int grid = std::min(std::min(device_sm_count(), p.tiles), max_slabs);
    p.tiles_per_cta = (p.tiles + grid - 1) / grid;
    grid = (p.tiles + p.tiles_per_cta - 1) / p.tiles_per_cta;
    tc_tn_tmem_kernel<<<grid, kThreads, TmSmem::kBytes, s>>>(p, mx[gi], my);
    tm_sum_slots_kernel<<<(kg * nb + 63) / 64, 256, 0, s>>>(Cp, 2 * grid, kg, nb, C + k0, kb);
    note_launch(2);
  }
  TSQR_CUDA_CHECK(cudaGetLastError());
  return true;
}

}  // namespace tsqr_b200
