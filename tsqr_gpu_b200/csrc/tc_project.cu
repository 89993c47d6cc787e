// BlockQR projection step on the tensor cores (tcgen05 kind::tf32, accumulators in TMEM):
//     C = Q1^T Y        (kb x nb, kb = columns already factored, nb <= 64 = current panel)
//     Y <- Y - Q1 C
// i.e. the two cuBLAS GEMMs per panel of reference src/blockqr.cu:92-116 (and :272-291 for the
// re-orthogonalisation pass). Both are tall-skinny and HBM-bound (32 flop per byte of Q1): each
// kernel streams Q1 once in 128-row tiles, converts a tile to the tf32 hi/lo operand layout in
// shared memory while the previous tile's MMAs run (double buffer, one mbarrier per buffer), and
// keeps every partial result in TMEM until a row slab / tile group is finished.
//   tn kernel: one CTA per row slab; D_b[128 x 64] += [Q1_b hi | Q1_b lo]^T Y for every 64-column
//              block b of Q1 (accumulator lanes 0..63 = hi^T Y, 64..127 = lo^T Y, both products with
//              Y hi and Y lo); the two halves leave as two partial slots, summed in fixed order.
//   nn kernel: one CTA per group of 8 row tiles; D_g[128 x 64] -= Q1[g, b] C_b over the blocks b,
//              then Y[g] += D_g.
#include "common.cuh"
#include "tc05.cuh"
#include "driver.h"

namespace tsqr_b200 {
// tc_nn_tma.cu: the TMA-fed form of the update; false = not applicable, nothing launched
bool launch_tc_nn_tma(bool corrected, IoType io, const void* X, long long ldx, int kb, const float* C, void* Y, long long ldy, int nb,
                      long long M, cudaStream_t s);
// tc_tn_tma.cu: the TMA-fed form of the coefficient product (C complete on return); false = not applicable, nothing launched
bool launch_tc_tn_tma(bool corrected, IoType io, const void* X, long long ldx, int kb, const void* Y, long long ldy, int nb,
                      long long M, int max_slabs, float* Cp, float* C, cudaStream_t s);
namespace {
using namespace tc;

constexpr int kT = 128;
constexpr uint32_t kTileB = ns::bytes(128, 64);   // 36864: one 128 x 64 ns buffer (hi or lo)
constexpr uint32_t kRnB = rn::bytes(64);          // 32768: one 128 x 64 rn buffer
constexpr uint32_t kCbB = ns::bytes(64, 64);      // 18432: one 64 x 64 ns buffer

struct Pending {  // one outstanding tcgen05.commit per buffer at most
  uint32_t issued[2] = {0, 0}, waited[2] = {0, 0};
};
__device__ __forceinline__ void wait_buf(Pending& p, uint64_t* bars, int b) {
  if (p.issued[b] > p.waited[b]) {
    mbar_wait(bars + b, p.waited[b] & 1);
    p.waited[b]++;
    fence_after_sync();
  }
}

// ---------------------------------------------------------------- C partials = Q1^T Y
template <bool COR, typename T>
__global__ void __launch_bounds__(kT, 1) tc_tn_kernel(const T* __restrict__ X, long long ldx, int kb, const T* __restrict__ Y,
                                                      long long ldy, int nb, long long M, long long rows_per_slab,
                                                      float* __restrict__ Cp) {
  extern __shared__ __align__(128) char sm[];
  char* Qb[2] = {sm, sm + 2 * kTileB};          // each: hi | lo
  char* Yhi = sm + 4 * kTileB;
  char* Ylo = Yhi + kTileB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + 6 * kTileB);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + 6 * kTileB + 16);
  const int t = threadIdx.x, warp = t >> 5;
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  if (t == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); fence_mbar_init(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const int nblk = (kb + 63) / 64;
  const long long rbeg = (long long)blockIdx.x * rows_per_slab;
  const long long rend = rbeg + rows_per_slab < M ? rbeg + rows_per_slab : M;
  const uint32_t rowoff = ns::off(t, 0, 128);
  Pending pend;
  int use = 0;
  bool first = true;
  for (long long r0 = rbeg; r0 < rend; r0 += kT) {
    const long long row = r0 + t;
    const bool valid = row < rend;
    wait_buf(pend, bars, 0);  // the previous tile's products still read Y
    wait_buf(pend, bars, 1);
    {  // all 64 loads are issued before the first use: 32 KB in flight per SM
      float x[64];
      const T* yp = Y + row;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        x[c] = (valid && c < nb) ? io_load(yp) : 0.f;
        yp += ldy;
      }
#pragma unroll
      for (int c = 0; c < 64; ++c) tile_store<COR>(Yhi + rowoff, Ylo + rowoff, ns::coloff(c, 128), x[c]);
    }
    for (int b = 0; b < nblk; ++b, ++use) {
      const int bf = use & 1;
      wait_buf(pend, bars, bf);
      {
        float x[64];
        const T* xp = X + row + (long long)(64 * b) * ldx;
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          x[c] = (valid && 64 * b + c < kb) ? io_load(xp) : 0.f;
          xp += ldx;
        }
#pragma unroll
        for (int c = 0; c < 64; ++c) tile_store<COR>(Qb[bf] + rowoff, Qb[bf] + kTileB + rowoff, ns::coloff(c, 128), x[c]);
      }
      fence_async_smem();
      fence_before_sync();
      __syncthreads();
      if (t == 0) {
        fence_after_sync();
        const uint32_t qa = smem_u32(Qb[bf]);
        // A = [hi | lo] (128 "columns"), B = Y hi, then B = Y lo
        ns::gram_issue<false>(tm + 64 * b, qa, 0, 128, 0, smem_u32(Yhi), 0, 128, 0, 64, 128, !first);
        if (COR) ns::gram_issue<false>(tm + 64 * b, qa, 0, 128, 0, smem_u32(Ylo), 0, 128, 0, 64, 128, true);
        mma_commit(bars + bf);
      }
      pend.issued[bf]++;
    }
    first = false;
  }
  wait_buf(pend, bars, 0);
  wait_buf(pend, bars, 1);
  // accumulator lane l < 64: row l of hi^T Y, lane 64 + l: row l of lo^T Y -> two partial slots
  {
    const int l = t & 63, half = t >> 6;
    float* out = Cp + ((size_t)blockIdx.x * 2 + half) * (size_t)kb * nb;
    for (int b = 0; b < nblk; ++b) {
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_addr(tm + 64 * b, warp, c0), v);
        const int i = 64 * b + l;
        if (i < kb) {
#pragma unroll
          for (int q = 0; q < 16; ++q)
            if (c0 + q < nb) out[i + (size_t)(c0 + q) * kb] = (first || (!COR && half == 1)) ? 0.f : v[q];
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

// ---------------------------------------------------------------- Y -= Q1 C
template <bool COR, typename T>
__global__ void __launch_bounds__(kT, 1) tc_nn_kernel(const T* __restrict__ X, long long ldx, int kb, const float* __restrict__ C,
                                                      T* Y, long long ldy, int nb, long long M, int ngroups) {
  extern __shared__ __align__(128) char sm[];
  char* Qr[2] = {sm, sm + 2 * kRnB};            // each: hi | lo
  char* Chi = sm + 4 * kRnB;
  char* Clo = Chi + kCbB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + 4 * kRnB + 2 * kCbB);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + 4 * kRnB + 2 * kCbB + 16);
  const int t = threadIdx.x, warp = t >> 5;
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  if (t == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); fence_mbar_init(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const int nblk = (kb + 63) / 64;
  Pending pend;
  int use = 0;
  for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    const long long g0 = (long long)grp * 8 * kT;
    const int ntile = (int)((M - g0 + kT - 1) / kT < 8 ? (M - g0 + kT - 1) / kT : 8);
    for (int b = 0; b < nblk; ++b) {
      wait_buf(pend, bars, 0);  // products of the previous block still read C_b
      wait_buf(pend, bars, 1);
      {  // C rows [64b, 64b+64) as the K-major B operand: ns buffer with 64 rows (K), 64 columns (N)
        const int k = t & 63, jh = t >> 6;
#pragma unroll 8
        for (int jj = 0; jj < 32; ++jj) {
          const int j = 32 * jh + jj, i = 64 * b + j;
          tile_store<COR>(Chi, Clo, ns::off(j, k, 64), (i < kb && k < nb) ? C[i + (size_t)k * kb] : 0.f);
        }
      }
      for (int g = 0; g < ntile; ++g, ++use) {
        const int bf = use & 1;
        wait_buf(pend, bars, bf);
        const long long row = g0 + (long long)g * kT + t;
        const bool valid = row < M;
        {
          float x[64];
          const T* xp = X + row + (long long)(64 * b) * ldx;
#pragma unroll
          for (int c = 0; c < 64; ++c) {
            x[c] = (valid && 64 * b + c < kb) ? io_load(xp) : 0.f;
            xp += ldx;
          }
#pragma unroll
          for (int c4 = 0; c4 < 16; ++c4) rn::store4<COR>(Qr[bf], Qr[bf] + kRnB, t, c4, x[4 * c4], x[4 * c4 + 1], x[4 * c4 + 2], x[4 * c4 + 3]);
        }
        fence_async_smem();
        fence_before_sync();
        __syncthreads();
        if (t == 0) {
          fence_after_sync();
          rn::update_issue<COR>(tm + 64 * g, smem_u32(Qr[bf]), smem_u32(Qr[bf] + kRnB), 0, smem_u32(Chi), smem_u32(Clo), 64, 0,
                                0, 64, 64, true, b > 0);
          mma_commit(bars + bf);
        }
        pend.issued[bf]++;
      }
    }
    wait_buf(pend, bars, 0);
    wait_buf(pend, bars, 1);
    for (int g = 0; g < ntile; ++g) {
      const long long row = g0 + (long long)g * kT + t;
      const bool valid = row < M;
      T* yp = Y + row;
#pragma unroll
      for (int c0 = 0; c0 < 64; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_addr(tm + 64 * g, warp, c0), v);
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          if (valid && c0 + q < nb) io_store(yp, io_plain(yp) + v[q]);
          yp += ldy;
        }
      }
    }
    fence_before_sync();
    __syncthreads();
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

constexpr uint32_t kTnSmem = 6 * kTileB + 64;
constexpr uint32_t kNnSmem = 4 * kRnB + 2 * kCbB + 64;

template <bool COR, typename T>
void launch_typed(const T* X, long long ldx, int kb, T* Y, long long ldy, int nb, long long M, int nslabs,
                  long long rows_per_slab, float* Cp, float* C, void (*reduce)(const float*, int, size_t, float*, cudaStream_t),
                  cudaStream_t s) {
  auto tn = tc_tn_kernel<COR, T>;
  auto nn = tc_nn_kernel<COR, T>;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(tn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTnSmem));
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(nn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kNnSmem));
  const IoType io = sizeof(T) == 2 ? IoType::f16 : IoType::f32;
  if (!launch_tc_tn_tma(COR, io, X, ldx, kb, Y, ldy, nb, M, nslabs, Cp, C, s)) {
    tn<<<nslabs, kT, kTnSmem, s>>>(X, ldx, kb, Y, ldy, nb, M, rows_per_slab, Cp);
    note_launch();
    reduce(Cp, 2 * nslabs, (size_t)kb * nb, C, s);
  }
  if (!launch_tc_nn_tma(COR, io, X, ldx, kb, C, Y, ldy, nb, M, s)) {
    const int ngroups = (int)((M + 8 * kT - 1) / (8 * kT));
    const int grid = std::min(ngroups, device_sm_count());
    nn<<<grid, kT, kNnSmem, s>>>(X, ldx, kb, C, Y, ldy, nb, M, ngroups);
    note_launch();
  }
  TSQR_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

void launch_tc_project(bool corrected, IoType io, const void* X, long long ldx, int kb, void* Y, long long ldy, int nb,
                       long long M, int nslabs, long long rows_per_slab, float* Cp, float* C,
                       void (*reduce)(const float*, int, size_t, float*, cudaStream_t), cudaStream_t s) {
  if (io == IoType::f32) {
    if (corrected) launch_typed<true, float>((const float*)X, ldx, kb, (float*)Y, ldy, nb, M, nslabs, rows_per_slab, Cp, C, reduce, s);
    else launch_typed<false, float>((const float*)X, ldx, kb, (float*)Y, ldy, nb, M, nslabs, rows_per_slab, Cp, C, reduce, s);
  } else {
    if (corrected) launch_typed<true, __half>((const __half*)X, ldx, kb, (__half*)Y, ldy, nb, M, nslabs, rows_per_slab, Cp, C, reduce, s);
    else launch_typed<false, __half>((const __half*)X, ldx, kb, (__half*)Y, ldy, nb, M, nslabs, rows_per_slab, Cp, C, reduce, s);
  }
}

}  // namespace tsqr_b200
