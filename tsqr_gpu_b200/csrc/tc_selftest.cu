// Known-answer self test of the tcgen05 primitives in tc05.cuh: every descriptor form the
// TSQR kernels use (column-wise Gram operands, row-wise update operands, column-group offsets, operand
// negation, accumulation onto tcgen05.st-initialised TMEM, 1-term and 3-term products) is
// exercised on caller-provided matrices so that the host can compare against plain fp32/fp64.
#include "common.cuh"
#include "tc05.cuh"
#include "../../include/tsqr_b200.h"

namespace tsqr_b200 {
namespace {

constexpr int kRows = 128, kCols = 64;
constexpr uint32_t kNsTile = tc::ns::bytes(128, 64);  // 36864
constexpr uint32_t kNsX = tc::ns::bytes(64, 64);      // 18432
constexpr uint32_t kSmemNs = 4 * kNsTile + 2 * kNsX + 40960 + 1024;
constexpr uint32_t kSmemRk = 2 * 32768 + 2 * kNsX + 1024;

// VAR 0: Gram products on column-wise "ns" buffers -> D1, D3
// VAR 1: update products, A operand in the row-wise "rn" buffer, B in an "ns" buffer -> D2, D4
template <bool COR, int VAR>
__global__ void __launch_bounds__(128) selftest_kernel(const float* A, const float* B, const float* X, float* D1,
                                                       float* D2, float* D3, float* D4) {
  extern __shared__ char smem_raw[];
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5;

  if (warp == 0) tc::tmem_alloc<256>(&tmem_base_s);
  if (t == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  char *Ahi, *Alo, *Bhi = nullptr, *Blo = nullptr, *Xhi, *Xlo;
  if (VAR == 0) {
    Ahi = base; Alo = Ahi + kNsTile; Bhi = Alo + kNsTile; Blo = Bhi + kNsTile; Xhi = Blo + kNsTile; Xlo = Xhi + kNsX;
    for (int i = t; i < 40960 / 4; i += 128) reinterpret_cast<float*>(Xlo + kNsX)[i] = 0.f;
    for (int c = 0; c < kCols; ++c) {
      const uint32_t off = tc::ns::off(t, c, 128);
      tc::tile_store<COR>(Ahi, Alo, off, A[t + c * kRows]);
      tc::tile_store<COR>(Bhi, Blo, off, B[t + c * kRows]);
    }
    if (t < 64)
      for (int c = 0; c < 64; ++c) tc::tile_store<COR>(Xhi, Xlo, tc::ns::off(t, c, 64), X[t + c * 64]);
  } else {
    Ahi = base; Alo = Ahi + 32768; Xhi = Alo + 32768; Xlo = Xhi + kNsX;
    for (int c4 = 0; c4 < kCols / 4; ++c4)
      tc::rn::store4<COR>(Ahi, Alo, t, c4, A[t + (4 * c4) * kRows], A[t + (4 * c4 + 1) * kRows], A[t + (4 * c4 + 2) * kRows],
                          A[t + (4 * c4 + 3) * kRows]);
    if (t < 64)
      for (int c = 0; c < 64; ++c) tc::tile_store<COR>(Xhi, Xlo, tc::ns::off(t, c, 64), X[t + c * 64]);
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tm = tmem_base_s;
  const uint32_t dD1 = tm, dD2 = tm + 64, dD3 = tm + 128, dD4 = tm + 160;

  if (VAR == 1) {
    // D4 starts as B[:, 0:48] (tcgen05.st), then accumulates an update product
    for (int c0 = 0; c0 < 48; c0 += 16) {
      float v[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = B[t + (c0 + i) * kRows];
      tc::tmem_st16(tc::tmem_addr(dD4, warp, c0), v);
    }
    tc::tmem_wait_st();
  }
  tc::fence_before_sync();
  __syncthreads();
  if (t == 0) {
    tc::fence_after_sync();
    const uint32_t ah = tc::smem_u32(Ahi), al = tc::smem_u32(Alo), xh = tc::smem_u32(Xhi), xl = tc::smem_u32(Xlo);
    if (VAR == 0) {
      const uint32_t bh = tc::smem_u32(Bhi), bl = tc::smem_u32(Blo);
      tc::ns::gram_issue<COR>(dD1, ah, al, 128, 0, bh, bl, 128, 0, 64, 128, false);
      tc::ns::gram_issue<COR>(dD3, ah, al, 128, 16, bh, bl, 128, 48, 16, 128, false);
    } else {
      tc::rn::update_issue<COR>(dD2, ah, al, 0, xh, xl, 64, 0, 0, 64, 64, true, false);
      tc::rn::update_issue<COR>(dD4, ah, al, 16, xh, xl, 64, 16, 16, 48, 16, false, true);
    }
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  float v[16];
  for (int c0 = 0; c0 < 64; c0 += 16) {
    tc::tmem_ld16(tc::tmem_addr(VAR == 0 ? dD1 : dD2, warp, c0), v);
#pragma unroll
    for (int i = 0; i < 16; ++i) (VAR == 0 ? D1 : D2)[t + (c0 + i) * kRows] = v[i];
  }
  if (VAR == 0) {
    tc::tmem_ld16(tc::tmem_addr(dD3, warp, 0), v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D3[t + i * kRows] = v[i];
  } else {
    for (int c0 = 0; c0 < 48; c0 += 16) {
      tc::tmem_ld16(tc::tmem_addr(dD4, warp, c0), v);
#pragma unroll
      for (int i = 0; i < 16; ++i) D4[t + (c0 + i) * kRows] = v[i];
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(tm);
}

template <bool COR, int VAR>
void launch_selftest(const float* A, const float* B, const float* X, float* D1, float* D2, float* D3, float* D4,
                     cudaStream_t s) {
  const uint32_t smem = VAR == 0 ? kSmemNs : kSmemRk;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(selftest_kernel<COR, VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  selftest_kernel<COR, VAR><<<1, 128, smem, s>>>(A, B, X, D1, D2, D3, D4);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

}  // namespace
}  // namespace tsqr_b200

// A, B: 128 x 64 col-major (ld 128); X: 64 x 64 col-major (ld 64), all device fp32.
// D1 = A^T B (rows 0..63 valid) | D2 = -A X | D3[m][n] = sum_r A[r][16+m] B[r][48+n] (m < 48)
// D4[r][n] = B[r][n] + sum_{j<16} A[r][16+j] X[16+j][16+n] (n < 48); D* are 128-row col-major.
// variant 0 computes D1 and D3 (Gram form), variant 1 computes D2 and D4 (update form).
extern "C" int tsqr_b200_selftest_mma(int corrected, int variant, const float* A, const float* B, const float* X,
                                      float* D1, float* D2, float* D3, float* D4, void* stream) {
  using namespace tsqr_b200;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  try {
    if (variant == 0) {
      if (corrected) launch_selftest<true, 0>(A, B, X, D1, D2, D3, D4, s);
      else launch_selftest<false, 0>(A, B, X, D1, D2, D3, D4, s);
    } else {
      if (corrected) launch_selftest<true, 1>(A, B, X, D1, D2, D3, D4, s);
      else launch_selftest<false, 1>(A, B, X, D1, D2, D3, D4, s);
    }
  } catch (const std::exception&) {
    return TSQR_B200_ERR_RUNTIME;
  }
  return 0;
}
