// Known-answer self test of the tcgen05 primitives in tc05.cuh: every descriptor form the
// TSQR kernels use (K-major Gram view, MN-major update view, column-group offsets, operand
// negation, accumulation onto tcgen05.st-initialised TMEM, 1-term and 3-term products) is
// exercised on caller-provided matrices so that the host can compare against plain fp32/fp64.
#include "common.cuh"
#include "tc05.cuh"
#include "../../include/tsqr_b200.h"

namespace tsqr_b200 {
namespace {

constexpr int kRows = 128, kCols = 64;
constexpr uint32_t kTileBytes = kRows * kCols * 4;  // 32 KiB
constexpr uint32_t kXBytes = 64 * 64 * 4;           // 16 KiB
constexpr uint32_t kSmem = 4 * kTileBytes + 2 * kXBytes + 16384 + 1024;

template <bool COR>
__global__ void __launch_bounds__(128) selftest_kernel(const float* A, const float* B, const float* X, float* D1,
                                                       float* D2, float* D3, float* D4) {
  extern __shared__ char smem_raw[];
  char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  char* Ahi = base;
  char* Alo = Ahi + kTileBytes;
  char* Bhi = Alo + kTileBytes;
  char* Blo = Bhi + kTileBytes;
  char* Xhi = Blo + kTileBytes;
  char* Xlo = Xhi + kXBytes;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5;

  if (warp == 0) tc::tmem_alloc<256>(&tmem_base_s);
  if (t == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  // zero the slack the padded-M Gram reads run into
  for (int i = t; i < 16384 / 4; i += 128) reinterpret_cast<float*>(Xlo + kXBytes)[i] = 0.f;
  // thread-per-row operand staging
  for (int c = 0; c < kCols; ++c) {
    const uint32_t off = tc::tile_off(t, c, kCols);
    tc::tile_store<COR>(Ahi, Alo, off, A[t + c * kRows]);
    tc::tile_store<COR>(Bhi, Blo, off, B[t + c * kRows]);
  }
  if (t < 64)
    for (int c = 0; c < 64; ++c) tc::tile_store<COR>(Xhi, Xlo, tc::tile_off(t, c, 64), X[t + c * 64]);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tm = tmem_base_s;
  const uint32_t dD1 = tm, dD2 = tm + 64, dD3 = tm + 128, dD4 = tm + 160;

  // D4 starts as B[:, 0:48] (tcgen05.st), then accumulates an update product
  for (int c0 = 0; c0 < 48; c0 += 16) {
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = B[t + (c0 + i) * kRows];
    tc::tmem_st16(tc::tmem_addr(dD4, warp, c0), v);
  }
  tc::tmem_wait_st();
  tc::fence_before_sync();
  __syncthreads();
  if (t == 0) {
    tc::fence_after_sync();
    const uint32_t ah = tc::smem_u32(Ahi), al = tc::smem_u32(Alo), bh = tc::smem_u32(Bhi), bl = tc::smem_u32(Blo);
    const uint32_t xh = tc::smem_u32(Xhi), xl = tc::smem_u32(Xlo);
    tc::gram_issue<COR>(dD1, ah, al, kCols, 0, bh, bl, kCols, 0, 64, 128, false);
    tc::update_issue<COR>(dD2, ah, al, kCols, 0, xh, xl, 64, 0, 0, 64, 64, true, false);
    tc::gram_issue<COR>(dD3, ah, al, kCols, 16, bh, bl, kCols, 48, 16, 128, false);
    tc::update_issue<COR>(dD4, ah, al, kCols, 16, xh, xl, 64, 16, 16, 48, 16, false, true);
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  float v[16];
  for (int c0 = 0; c0 < 64; c0 += 16) {
    tc::tmem_ld16(tc::tmem_addr(dD1, warp, c0), v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D1[t + (c0 + i) * kRows] = v[i];
    tc::tmem_ld16(tc::tmem_addr(dD2, warp, c0), v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D2[t + (c0 + i) * kRows] = v[i];
  }
  tc::tmem_ld16(tc::tmem_addr(dD3, warp, 0), v);
#pragma unroll
  for (int i = 0; i < 16; ++i) D3[t + i * kRows] = v[i];
  for (int c0 = 0; c0 < 48; c0 += 16) {
    tc::tmem_ld16(tc::tmem_addr(dD4, warp, c0), v);
#pragma unroll
    for (int i = 0; i < 16; ++i) D4[t + (c0 + i) * kRows] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc<256>(tm);
}

}  // namespace
}  // namespace tsqr_b200

// A, B: 128 x 64 col-major (ld 128); X: 64 x 64 col-major (ld 64), all device fp32.
// D1 = A^T B (rows 0..63 valid) | D2 = -A X | D3[m][n] = sum_r A[r][16+m] B[r][48+n] (m < 48)
// D4[r][n] = B[r][n] + sum_{j<16} A[r][16+j] X[16+j][16+n] (n < 48); D* are 128-row col-major.
extern "C" int tsqr_b200_selftest_mma(int corrected, const float* A, const float* B, const float* X, float* D1,
                                      float* D2, float* D3, float* D4, void* stream) {
  using namespace tsqr_b200;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  try {
    if (corrected) {
      TSQR_CUDA_CHECK(cudaFuncSetAttribute(selftest_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
      selftest_kernel<true><<<1, 128, kSmem, s>>>(A, B, X, D1, D2, D3, D4);
    } else {
      TSQR_CUDA_CHECK(cudaFuncSetAttribute(selftest_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
      selftest_kernel<false><<<1, 128, kSmem, s>>>(A, B, X, D1, D2, D3, D4);
    }
    note_launch();
    TSQR_CUDA_CHECK(cudaGetLastError());
  } catch (const std::exception&) {
    return TSQR_B200_ERR_RUNTIME;
  }
  return 0;
}
