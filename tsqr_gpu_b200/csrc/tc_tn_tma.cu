// BlockQR projection coefficients C = X^T Y on sm_100a as a TMA-fed streaming kernel (the first of the two GEMMs per
// panel of reference src/blockqr.cu:92-116 / :272-291): X is m x kb (kb <= 512 per launch), Y is m x nb (nb <= 64).
// HBM-bound: X and Y are read once. One CTA per SM walks a contiguous range of 64-row tiles:
//   * TMA (cp.async.bulk.tensor) lands raw 64 x 64 blocks of Y and X in a 3-slot ring, decoupled from the compute;
//   * warp-specialised: one producer thread (TMA), one tensor-core issuer thread, 512 converter threads that only ever wait
//     on mbarriers (no CTA-wide barrier in the loop);
//   * the converter threads rewrite a landed block as the K-major tf32 operand (K = rows; 128-byte swizzle; hi in lines [0, 64),
//     lo = x - hi in lines [64, 128) for the corrected modes) -- conflict-free thread-per-row stores; 16 warps so that the
//     fence / barrier latencies of one warp hide behind the others;
//   * corrected modes: D_b[128 x 128] += [X_b hi | X_b lo]^T [Y hi | Y lo] as ONE M = 128, N = 128 product per K step (operand
//     bytes read from shared memory per product are what bounds this kernel), otherwise D_b[128 x 64] += X_b^T Y; one
//     accumulator block per 64 columns of X, kept in tensor memory over the whole row range (256 / 512 columns per launch);
//   * the two accumulator halves (hi^T Y, lo^T Y) leave as two partial slots per CTA, summed in fixed order by the caller.
// tc_project.cu keeps the thread-staged kernel for matrices TMA cannot address.
#include <cuda.h>
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "tc05.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {
using namespace tc;

constexpr int kThreads = 512;                   // converter threads (16 warps); + 1 issuer warp + 1 producer warp
constexpr int kConvWarps = kThreads / 32;
constexpr int kColsPerThread = 64 * 64 / kThreads;   // 8
constexpr int kTile = 64;                       // rows per tile = 2 K blocks of the sw layout
constexpr int kLand = 3;                        // landing slots
constexpr uint32_t kOp = 2 * sw::kBlock;        // one 64-row operand buffer: 128 lines (hi | lo), 32 KB

struct TnParams {
  int kb, nb, nblk;
  long long M;
  int tiles, tiles_per_cta;
  float* Cp;
};

template <typename TIo>
struct TnSmem {
  static constexpr uint32_t kSlot = (uint32_t)kTile * 64u * sizeof(TIo);   // raw block, [column][row]
  static constexpr uint32_t oBar = 0;
  static constexpr uint32_t oLand = 1024;
  static constexpr uint32_t oX = (oLand + kLand * kSlot + 1023) / 1024 * 1024;
  static constexpr uint32_t oY = oX + 2 * kOp;
  static constexpr uint32_t kBytes = oY + 2 * kOp;
};

template <bool COR, typename TIo>
__global__ void __launch_bounds__(kThreads + 64, 1) tc_tn_tma_kernel(const TnParams P, const __grid_constant__ CUtensorMap tmapX,
                                                                     const __grid_constant__ CUtensorMap tmapY) {
  using Sm = TnSmem<TIo>;
  constexpr int kDW = COR ? 128 : 64;   // accumulator columns per 64-column block of X
  extern __shared__ __align__(1024) char sm[];
  const int t = threadIdx.x, warp = t >> 5;
  uint64_t* bar_full = reinterpret_cast<uint64_t*>(sm + Sm::oBar);   // landing slot filled (TMA transaction bytes)
  uint64_t* bar_empty = bar_full + kLand;                            // landing slot read by all converter threads
  uint64_t* bar_mma = bar_empty + kLand;                             // [2] products that read X stage q & 1 are complete
  uint64_t* bar_rx = bar_mma + 2;                                    // [2] X operand stage written by all converter threads
  uint64_t* bar_ry = bar_rx + 2;                                     // [2] Y operand buffer written
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + Sm::oBar + 128);
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  if (t == 0) {
    for (int i = 0; i < kLand; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, kThreads); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_mma + i, 1); mbar_init(bar_rx + i, kThreads); mbar_init(bar_ry + i, kThreads); }
    fence_mbar_init();
    tma_prefetch_desc(&tmapX);
    tma_prefetch_desc(&tmapY);
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const int nblk = P.nblk, per = nblk + 1;   // items per tile: the Y block, then the X blocks
  const int tile_beg = blockIdx.x * P.tiles_per_cta;
  const int tile_end = min(tile_beg + P.tiles_per_cta, P.tiles);
  const int ntl = max(tile_end - tile_beg, 0);
  const int nitems = ntl * per;

  if (warp == kConvWarps + 1) {
    // ---- producer: one thread keeps kLand block loads in flight
    if (t == (kConvWarps + 1) * 32) {
      for (int j = 0; j < nitems; ++j) {
        const int slot = j % kLand, i = j / per, r = j % per;
        if (j >= kLand) mbar_wait(bar_empty + slot, (uint32_t)(j / kLand - 1) & 1u);
        mbar_arrive_expect_tx(bar_full + slot, Sm::kSlot);
        if (r == 0) tma_load_2d(sm + Sm::oLand + slot * Sm::kSlot, &tmapY, bar_full + slot, (tile_beg + i) * kTile, 0);
        else tma_load_2d(sm + Sm::oLand + slot * Sm::kSlot, &tmapX, bar_full + slot, (tile_beg + i) * kTile, 64 * (r - 1));
      }
    }
  } else if (warp == kConvWarps) {
    // ---- tensor-core issuer: one thread; product q reads X stage q & 1 and the Y buffer of its tile
    if (t == kConvWarps * 32) {
      const uint32_t idesc = idesc_tf32(128, kDW, false, false, false);
      for (int i = 0; i < ntl; ++i) {
        mbar_wait(bar_ry + (i & 1), (uint32_t)(i >> 1) & 1u);
        const uint32_t ya = smem_u32(sm + Sm::oY + (i & 1) * kOp);
        for (int b = 0; b < nblk; ++b) {
          const int q = i * nblk + b;
          mbar_wait(bar_rx + (q & 1), (uint32_t)(q >> 1) & 1u);
          fence_after_sync();
          const uint32_t xa = smem_u32(sm + Sm::oX + (q & 1) * kOp);
          const uint32_t d = tm + (uint32_t)kDW * (uint32_t)b;
#pragma unroll
          for (int kblk = 0; kblk < 2; ++kblk) {
            const uint64_t da = smem_desc(xa + (uint32_t)kblk * sw::kBlock, 16, 1024, 2);
            const uint64_t dy = smem_desc(ya + (uint32_t)kblk * sw::kBlock, 16, 1024, 2);   // N = 128: lines [Y hi | Y lo]
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              mma_tf32(d, da + (uint64_t)(kk * 2), dy + (uint64_t)(kk * 2), idesc, i > 0 || kblk > 0 || kk > 0);
          }
          mma_commit(bar_mma + (q & 1));
        }
      }
    }
  } else {
    // ---- converter threads: landed block -> K-major tf32 operand (hi, and lo for the corrected modes)
    int nw = 0;   // products [0, nw) are known complete (commit q arrives on bar_mma[q & 1], parity (q >> 1) & 1)
    auto wait_product = [&](int q) {
      while (nw <= q) {
        mbar_wait(bar_mma + (nw & 1), (uint32_t)(nw >> 1) & 1u);
        ++nw;
      }
    };
    const int row = t & 63, cg = t >> 6;   // this thread converts columns [8 cg, 8 cg + 8) of row `row`
    uint32_t offs[8];                       // swizzled chunk of the row inside a 128-byte line, by line % 8
#pragma unroll
    for (int e = 0; e < 8; ++e) offs[e] = sw::off(row, e);
    static_assert(kColsPerThread == 8, "one 8-line swizzle atom per thread");
    const uint32_t land_off = (uint32_t)((kColsPerThread * cg) * kTile + row) * (uint32_t)sizeof(TIo);
    const uint32_t dst_off = (uint32_t)(kColsPerThread * cg) * 128u;
    int slot = 0;            // landing ring position and the parity of its current use
    uint32_t par = 0;
    // one landed block -> operand buffer `op`; then the slot goes back to the producer and `ready` gets this thread's arrival
    auto convert = [&](char* op, uint64_t* ready) {
      mbar_wait(bar_full + slot, par);
      const TIo* land = reinterpret_cast<const TIo*>(sm + Sm::oLand + slot * Sm::kSlot + land_off);
      char* dst = op + dst_off;
#pragma unroll
      for (int c = 0; c < kColsPerThread; ++c) {
        const float x = io_plain(land + c * kTile);
        const float h = to_tf32_rna(x);
        *reinterpret_cast<float*>(dst + offs[c]) = h;
        if (COR) *reinterpret_cast<float*>(dst + 64 * 128 + offs[c]) = __fsub_rn(x, h);
      }
      fence_async_smem();   // landing-slot reads before the next bulk write; operand stores before the tensor core reads them
      mbar_arrive(bar_empty + slot);
      mbar_arrive(ready);
      if (++slot == kLand) { slot = 0; par ^= 1u; }
    };
    int q = 0;   // product index of the next X block
    for (int i = 0; i < ntl; ++i) {
      if (i >= 2) wait_product(q - nblk - 1);   // every product of tile i - 2 read this Y buffer
      convert(sm + Sm::oY + (i & 1) * kOp, bar_ry + (i & 1));
      for (int b = 0; b < nblk; ++b, ++q) {
        if (q >= 2) wait_product(q - 2);        // the product that read this X stage
        convert(sm + Sm::oX + (q & 1) * kOp, bar_rx + (q & 1));
      }
    }
    if (ntl > 0) wait_product(ntl * nblk - 1);
    fence_after_sync();
    // accumulator lane l < 64: row l of hi^T Y, lane 64 + l: row l of lo^T Y -> two partial slots; warp w takes the blocks
    // b = w / 4 (mod 4) (a warp reaches the 32 TMEM lanes of its index modulo 4)
    const int lane128 = (warp & 3) * 32 + (t & 31);
    const int l = lane128 & 63, half = lane128 >> 6;
    float* out = P.Cp + ((size_t)blockIdx.x * 2 + half) * (size_t)P.kb * P.nb;
    const bool zero = ntl == 0 || (!COR && half == 1);
    for (int b = warp >> 2; b < nblk; b += kConvWarps / 4) {
      const int irow = 64 * b + l;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16], w[16];
        tmem_ld16(tmem_addr(tm + kDW * b, warp, c0), v);
        if (COR) {   // accumulator columns [64, 128) of the hi lanes hold X hi^T Y lo
          tmem_ld16(tmem_addr(tm + kDW * b, warp, 64 + c0), w);
          if (half == 0) {
#pragma unroll
            for (int e = 0; e < 16; ++e) v[e] = __fadd_rn(v[e], w[e]);
          }
        }
        if (irow < P.kb) {
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (c0 + e < P.nb) out[irow + (size_t)(c0 + e) * P.kb] = zero ? 0.f : v[e];
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
// column-major M x n matrix, box = 64 rows x 64 columns, rows and columns out of range read as zero
bool block_map(CUtensorMap* map, const void* A, long long lda, long long M, int n, IoType io) {
  EncodeFn fn = encode_fn();
  const size_t es = io == IoType::f16 ? 2 : 4;
  if (!fn || (reinterpret_cast<uintptr_t>(A) & 15u) != 0 || ((size_t)lda * es) % 16 != 0 || M >= (1ll << 31)) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)M, (cuuint64_t)n};
  cuuint64_t gstr[1] = {(cuuint64_t)lda * es};
  cuuint32_t box[2] = {(cuuint32_t)kTile, 64u};
  cuuint32_t estr[2] = {1u, 1u};
  return fn(map, io == IoType::f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(A), gdim,
            gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// C[k0 + i, c] (ld ldc) = sum over the partial slots, in slot order (bit-reproducible)
__global__ void __launch_bounds__(256) tn_sum_slots_kernel(const float* __restrict__ Cp, int slots, int kg, int nb, float* __restrict__ C, long long ldc) {
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

template <bool COR, typename TIo>
bool launch_t(IoType io, const void* X, long long ldx, int kb, const void* Y, long long ldy, int nb, long long M, int max_slabs,
              float* Cp, float* C, cudaStream_t s) {
  constexpr int kmax = COR ? 256 : 512;   // 512 TMEM columns / accumulator columns per block
  const int ngroups = (kb + kmax - 1) / kmax;
  const size_t es = io == IoType::f16 ? 2 : 4;
  CUtensorMap mx[4], my;
  if (ngroups > 4 || !block_map(&my, Y, ldy, M, nb, io)) return false;
  for (int gi = 0; gi < ngroups; ++gi)   // every map before the first launch: all or nothing
    if (!block_map(&mx[gi], static_cast<const char*>(X) + (size_t)gi * kmax * (size_t)ldx * es, ldx, M, std::min(kmax, kb - gi * kmax), io))
      return false;
  auto kern = tc_tn_tma_kernel<COR, TIo>;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TnSmem<TIo>::kBytes));
  for (int gi = 0; gi < ngroups; ++gi) {
    const int k0 = gi * kmax, kg = std::min(kmax, kb - k0);
    TnParams p{};
    p.kb = kg; p.nb = nb; p.nblk = (kg + 63) / 64; p.M = M; p.Cp = Cp;
    p.tiles = (int)((M + kTile - 1) / kTile);
    int grid = std::min(std::min(device_sm_count(), p.tiles), max_slabs);
    p.tiles_per_cta = (p.tiles + grid - 1) / grid;
    grid = (p.tiles + p.tiles_per_cta - 1) / p.tiles_per_cta;   // no CTA without a tile
    kern<<<grid, kThreads + 64, TnSmem<TIo>::kBytes, s>>>(p, mx[gi], my);
    tn_sum_slots_kernel<<<(kg * nb + 63) / 64, 256, 0, s>>>(Cp, 2 * grid, kg, nb, C + k0, kb);
    note_launch(2);
  }
  TSQR_CUDA_CHECK(cudaGetLastError());
  return true;
}

}  // namespace

// C (kb x nb, ld kb) = X^T Y; Cp: room for 2 x max_slabs partial slots of kb x nb floats. false: nothing was launched (TMA
// cannot address X or Y, or TSQR_B200_TMA=0) and the caller runs its thread-staged kernel.
// tc_tn_tmem.cu: the corrected fp32 modes with the A operand in tensor memory
bool launch_tc_tn_tmem(const void* X, long long ldx, int kb, const void* Y, long long ldy, int nb, long long M, int max_slabs, float* Cp,
                       float* C, cudaStream_t s);

bool launch_tc_tn_tma(bool corrected, IoType io, const void* X, long long ldx, int kb, const void* Y, long long ldy, int nb,
                      long long M, int max_slabs, float* Cp, float* C, cudaStream_t s) {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_TMA"); return !(e && e[0] == '0'); }();
  if (!on || kb < 1 || max_slabs < 1) return false;
  if (corrected && io == IoType::f32 && launch_tc_tn_tmem(X, ldx, kb, Y, ldy, nb, M, max_slabs, Cp, C, s)) return true;
  if (io == IoType::f32)
    return corrected ? launch_t<true, float>(io, X, ldx, kb, Y, ldy, nb, M, max_slabs, Cp, C, s)
                     : launch_t<false, float>(io, X, ldx, kb, Y, ldy, nb, M, max_slabs, Cp, C, s);
  return corrected ? launch_t<true, __half>(io, X, ldx, kb, Y, ldy, nb, M, max_slabs, Cp, C, s)
                   : launch_t<false, __half>(io, X, ldx, kb, Y, ldy, nb, M, max_slabs, Cp, C, s);
}

}  // namespace tsqr_b200
