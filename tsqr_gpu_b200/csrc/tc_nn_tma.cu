// BlockQR projection update Y <- Y - X C on sm_100a as a TMA-fed streaming kernel (the second of the two GEMMs per panel
// of reference src/blockqr.cu:92-116 / :272-291): X is m x kb (columns already factored), C is kb x nb (nb <= 64),
// Y the current panel. HBM-bound: X is read once, Y read and written once per launch.
//   * four independent tile pipelines per CTA (one CTA per SM), a tile = 128 rows, thread t = tile row t = TMEM lane t;
//   * X arrives as 128 x 16 panel boxes (cp.async.bulk.tensor) through a 3-slot ring per pipeline, so global-memory
//     latency is decoupled from the staging;
//   * the row's thread moves its 16 values into TENSOR MEMORY as the A operand (tf32 hi, and lo for the corrected
//     modes; double-buffered), so no transposing shared-memory staging is needed;
//   * C (hi / lo) stays in shared memory for the whole kernel as the K-major B operand; D += X_p C[16p:16p+16, :] with
//     tcgen05.mma kind::tf32 (1 or 3 products), fp32 accumulator in TMEM;
//   * the Y row is loaded behind the tile's last products and combined with the accumulator row.
// kb is cut into groups that fit the shared-memory image of C (256 columns corrected, 512 otherwise).
// tc_project.cu keeps the thread-staged kernel for matrices TMA cannot address (ld * element size not a multiple of 16).
#include <cuda.h>
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "tc05.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {
using namespace tc;

constexpr int kWg = 4;          // tile pipelines per CTA
constexpr int kSlots = 3;       // landing slots per pipeline
constexpr int kTmemPerWg = 128; // D [0, 64), A hi [64, 80) / [80, 96), A lo [96, 112) / [112, 128)

struct NnParams {
  const float* C;   // kb_all x nb, ld ldc; this launch uses rows [0, kb)
  long long ldc;
  int kb, kp;       // columns of X in this launch, padded to a multiple of 16
  void* Y;
  long long ldy;
  int nb;
  long long M;
  int tiles, tiles_per_cta;
};

template <typename TIo>
struct NnSmem {
  static constexpr uint32_t kSlot = 128u * 16u * sizeof(TIo);
  static constexpr uint32_t kBarBytes = 512;   // kWg x (2 kSlots + 2) mbarriers, then the TMEM slot in the last 8 bytes
  static constexpr uint32_t oBar = 0;
  static constexpr uint32_t oRing = 1024;
  static constexpr uint32_t oC = oRing + kWg * kSlots * kSlot;
  static uint32_t bytes(int kp, bool cor) { return oC + (cor ? 2u : 1u) * kb::bytes(kp, 64); }
};

template <bool COR, typename TIo>
__global__ void __launch_bounds__(128 * kWg, 1) tc_nn_tma_kernel(const NnParams P, const __grid_constant__ CUtensorMap tmapX) {
  using Sm = NnSmem<TIo>;
  constexpr int S = kSlots;
  extern __shared__ __align__(1024) char sm[];
  const int tid = threadIdx.x, wg = tid >> 7, t = tid & 127, warp = (tid >> 5) & 3;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + Sm::oBar) + wg * (2 * S + 2);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + S;
  uint64_t* bar_mma = bars + 2 * S;   // [2]: one per A-operand buffer
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + Sm::oBar + Sm::kBarBytes - 8);
  char* ring = sm + Sm::oRing + wg * S * Sm::kSlot;
  char* CHi = sm + Sm::oC;
  char* CLo = CHi + kb::bytes(P.kp, 64);
  const int npan = P.kp / 16;

  if (tid < 32) tmem_alloc<512>(tmem_slot);
  if (t == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, 128); }
    mbar_init(bar_mma, 1);
    mbar_init(bar_mma + 1, 1);
    fence_mbar_init();
    if (wg == 0) tma_prefetch_desc(&tmapX);
  }
  for (int idx = tid; idx < P.kp * 64; idx += 128 * kWg) {   // C as the K-major B operand, zero padded
    const int k = idx % P.kp, c = idx / P.kp;
    const float v = (k < P.kb && c < P.nb) ? P.C[k + (long long)c * P.ldc] : 0.f;
    tile_store<COR>(CHi, CLo, kb::off(k, c, P.kp), v);
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot + (uint32_t)(kTmemPerWg * wg);
  const uint32_t tmD = tm, tmAh = tm + 64, tmAl = tm + 96;

  const int tile_beg = blockIdx.x * P.tiles_per_cta;
  const int tile_end = min(tile_beg + P.tiles_per_cta, P.tiles);
  const int ntl_cta = max(tile_end - tile_beg, 0);
  const int ntl = ntl_cta > wg ? (ntl_cta - wg + kWg - 1) / kWg : 0;
  auto tile_of = [&](int i) { return tile_beg + wg + kWg * i; };
  const int npanels = ntl * npan;
  // producer state (thread 0 of the pipeline): the next panel to load, as (tile index, panel, slot)
  int ld_it = 0, ld_p = 0, ld_slot = 0, ld_left = npanels;
  auto issue_next = [&]() {
    mbar_arrive_expect_tx(bar_full + ld_slot, Sm::kSlot);
    tma_load_2d(ring + ld_slot * Sm::kSlot, &tmapX, bar_full + ld_slot, tile_of(ld_it) * 128, 16 * ld_p);
    if (++ld_p == npan) { ld_p = 0; ++ld_it; }
    if (++ld_slot == S) ld_slot = 0;
    --ld_left;
  };
  if (t == 0)
    for (int i = 0; i < S && ld_left > 0; ++i) issue_next();
  const uint64_t dCh = smem_desc(smem_u32(CHi), 128, kb::sbo(P.kp)), dCl = smem_desc(smem_u32(CLo), 128, kb::sbo(P.kp));
  const uint32_t id = idesc_tf32(128, 64, false, false, false);
  TIo* Y = reinterpret_cast<TIo*>(P.Y);
  int nw = 0;   // products [0, nw) of this pipeline are known complete (commit g arrives on bar_mma[g & 1], parity (g >> 1) & 1)
  auto wait_product = [&](int q) {
    while (nw <= q) {
      mbar_wait(bar_mma + (nw & 1), (uint32_t)(nw >> 1) & 1u);
      ++nw;
    }
  };
  int g = 0, slot = 0;   // panel counter of the pipeline, its landing slot and the parity of the slot's current use
  uint32_t par = 0;

  for (int it = 0; it < ntl; ++it) {
    const int tile = tile_of(it);
    const long long row = (long long)tile * 128 + t;
    const bool valid = row < P.M;
    uint64_t dp = 0;   // descriptor offset of C rows [16 p, 16 p + 16)
    for (int p = 0; p < npan; ++p, ++g, dp += 32) {
      const int ab = g & 1;
      float xh[16], xl[16];
      mbar_wait(bar_full + slot, par);
      {
        const TIo* land = reinterpret_cast<const TIo*>(ring + slot * Sm::kSlot);
#pragma unroll
        for (int i = 0; i < 16; ++i) {   // the split consumes the loaded value: every load has returned before the release
          const float x = io_plain(land + i * 128 + t);
          xh[i] = to_tf32_rna(x);
          xl[i] = __fsub_rn(x, xh[i]);
        }
      }
      fence_async_smem();   // generic-proxy reads before the async-proxy write that reuses the slot
      mbar_arrive(bar_empty + slot);
      if (t == 0 && ld_left > 0) {
        mbar_wait(bar_empty + slot, par);
        issue_next();
      }
      if (g >= 2) {   // the products that read this A buffer two panels ago
        wait_product(g - 2);
        fence_after_sync();
      }
      tmem_st16(tmem_addr(tmAh + 16 * ab, warp, 0), xh);
      if (COR) tmem_st16(tmem_addr(tmAl + 16 * ab, warp, 0), xl);
      tmem_wait_st();
      fence_before_sync();
      named_bar_sync(1 + wg, 128);
      if (t == 0) {
        fence_after_sync();
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint64_t off = dp + (uint64_t)(kk * 16);
          mma_tf32_ts(tmD, tmAh + 16 * ab + 8 * kk, dCh + off, id, p > 0 || kk > 0);
          if (COR) {
            mma_tf32_ts(tmD, tmAh + 16 * ab + 8 * kk, dCl + off, id, true);
            mma_tf32_ts(tmD, tmAl + 16 * ab + 8 * kk, dCh + off, id, true);
          }
        }
        mma_commit(bar_mma + ab);
      }
      if (++slot == S) { slot = 0; par ^= 1u; }
    }
    // ---- the Y row (its loads fly while the last products finish), all products of the tile, then Y_row -= D_row
    float yv[64];
    {
      const TIo* yp = Y + row;
#pragma unroll
      for (int c = 0; c < 64; ++c) yv[c] = (valid && c < P.nb) ? io_plain(yp + (long long)c * P.ldy) : 0.f;
    }
    wait_product(g - 1);
    fence_after_sync();
    {
      TIo* yp = Y + row;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float d[16];
        tmem_ld16(tmem_addr(tmD, warp, c0), d);
#pragma unroll
        for (int q = 0; q < 16; ++q)
          if (valid && c0 + q < P.nb) io_store(yp + (long long)(c0 + q) * P.ldy, __fsub_rn(yv[c0 + q], d[q]));
      }
    }
    fence_before_sync();   // the accumulator reads above precede the next tile's first product
    named_bar_sync(1 + wg, 128);
  }
  fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc<512>(*tmem_slot);
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
// column-major M x n matrix: dimension 0 = rows (contiguous), dimension 1 = columns; box = 128 rows x 16 columns,
// rows and columns out of range read as zero
bool panel_map(CUtensorMap* map, const void* X, long long ldx, long long M, int n, IoType io) {
  EncodeFn fn = encode_fn();
  const size_t es = io == IoType::f16 ? 2 : 4;
  if (!fn || (reinterpret_cast<uintptr_t>(X) & 15u) != 0 || ((size_t)ldx * es) % 16 != 0 || M >= (1ll << 31)) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)M, (cuuint64_t)n};
  cuuint64_t gstr[1] = {(cuuint64_t)ldx * es};
  cuuint32_t box[2] = {128u, 16u};
  cuuint32_t estr[2] = {1u, 1u};
  return fn(map, io == IoType::f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(X), gdim,
            gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <bool COR, typename TIo>
bool launch_t(IoType io, const void* X, long long ldx, int kb, const float* C, void* Y, long long ldy, int nb, long long M, cudaStream_t s) {
  const int kmax = COR ? 256 : 512;
  auto kern = tc_nn_tma_kernel<COR, TIo>;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NnSmem<TIo>::bytes(kmax, COR)));
  const int tiles = (int)((M + 127) / 128);
  const int grid = std::min(device_sm_count(), tiles);
  CUtensorMap maps[8];
  const int ngroups = (kb + kmax - 1) / kmax;
  if (ngroups > 8) return false;
  const size_t es = io == IoType::f16 ? 2 : 4;
  for (int gi = 0; gi < ngroups; ++gi) {   // every map before the first launch: all or nothing
    const int k0 = gi * kmax, kg = std::min(kmax, kb - k0);
    if (!panel_map(&maps[gi], static_cast<const char*>(X) + (size_t)k0 * (size_t)ldx * es, ldx, M, kg, io)) return false;
  }
  for (int gi = 0; gi < ngroups; ++gi) {
    const int k0 = gi * kmax, kg = std::min(kmax, kb - k0);
    NnParams p{};
    p.C = C + k0; p.ldc = kb; p.kb = kg; p.kp = (kg + 15) / 16 * 16;
    p.Y = Y; p.ldy = ldy; p.nb = nb; p.M = M; p.tiles = tiles; p.tiles_per_cta = (tiles + grid - 1) / grid;
    kern<<<grid, 128 * kWg, NnSmem<TIo>::bytes(p.kp, COR), s>>>(p, maps[gi]);
    note_launch();
  }
  TSQR_CUDA_CHECK(cudaGetLastError());
  return true;
}

}  // namespace

// false: nothing was launched (TMA cannot address X, or TSQR_B200_TMA=0); the caller runs its thread-staged kernel
bool launch_tc_nn_tma(bool corrected, IoType io, const void* X, long long ldx, int kb, const float* C, void* Y, long long ldy, int nb,
                      long long M, cudaStream_t s) {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_TMA"); return !(e && e[0] == '0'); }();
  if (!on || kb < 1) return false;
  if (io == IoType::f32) return corrected ? launch_t<true, float>(io, X, ldx, kb, C, Y, ldy, nb, M, s) : launch_t<false, float>(io, X, ldx, kb, C, Y, ldy, nb, M, s);
  return corrected ? launch_t<true, __half>(io, X, ldx, kb, C, Y, ldy, nb, M, s) : launch_t<false, __half>(io, X, ldx, kb, C, Y, ldy, nb, M, s);
}

}  // namespace tsqr_b200
