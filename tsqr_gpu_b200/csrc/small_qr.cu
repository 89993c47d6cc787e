// Batched QR of small row blocks (<= 32 rows x <= 16 columns): one warp per block, one
// matrix row per lane, the whole block in registers. Householder reflectors stay implicit
// (v, tau); Q is formed by applying them to the leading columns of the identity.
// Stands in for reference src/tcqr32x16.cu:1471-1634 (qr32x16[_batched] kernels), which
// build a dense 32x32 H per column step instead.
#include "common.cuh"
#include "small_qr.h"

namespace tsqr_b200 {
namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}

template <typename QT, typename RT, typename AT>
__global__ void __launch_bounds__(128) small_qr_kernel(QT* q, long long ldq, RT* r, long long ldr, const AT* a,
                                                       long long lda, unsigned m_total, unsigned n,
                                                       unsigned long long batch, const unsigned* starts) {
  const int lane = threadIdx.x & 31;
  const unsigned long long b = (unsigned long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (b >= batch) return;
  const unsigned r0 = starts ? starts[b] : 0u;
  const unsigned r1 = starts ? starts[b + 1] : m_total;
  const int rows = (int)(r1 - r0) > 32 ? 32 : (int)(r1 - r0);
  float x[16], tau[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) x[c] = (lane < rows && c < (int)n) ? io_load(a + r0 + lane + (long long)c * lda) : 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    tau[j] = 0.f;
    if (j < (int)n) {  // warp-uniform
      const float xj = lane >= j ? x[j] : 0.f;
      const float nrm2 = warp_sum(xj * xj);
      const float alpha = __shfl_sync(kFullMask, x[j], j);
      float v = 0.f, beta = alpha;
      if (nrm2 > 0.f) {
        beta = -copysignf(sqrtf(nrm2), alpha);
        tau[j] = (beta - alpha) / beta;
        const float rden = 1.f / (alpha - beta);
        v = lane > j ? x[j] * rden : (lane == j ? 1.f : 0.f);
      }
#pragma unroll
      for (int k = j + 1; k < 16; ++k) {
        const float s = warp_sum(v * x[k]);
        x[k] -= tau[j] * s * v;
      }
      if (lane == j) x[j] = beta;
      else if (lane > j) x[j] = v;
    }
  }
  // R block b: rows [n*b, n*b+n)
  if (lane < (int)n) {
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (k < (int)n) io_store(r + (long long)n * b + lane + (long long)k * ldr, lane <= k ? x[k] : 0.f);
  }
  // Q = H_0 ... H_{n-1} [I; 0]
  float qv[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) qv[c] = (lane == c) ? 1.f : 0.f;
#pragma unroll
  for (int j = 15; j >= 0; --j) {
    if (j < (int)n) {
      const float v = lane > j ? x[j] : (lane == j ? 1.f : 0.f);
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const float s = warp_sum(v * qv[c]);
        qv[c] -= tau[j] * s * v;
      }
    }
  }
  if (lane < rows) {
#pragma unroll
    for (int c = 0; c < 16; ++c)
      if (c < (int)n) io_store(q + r0 + lane + (long long)c * ldq, qv[c]);
  }
}

}  // namespace

template <typename QT, typename RT, typename AT>
void small_qr_launch(QT* q, size_t ldq, RT* r, size_t ldr, const AT* a, size_t lda, unsigned m, unsigned n,
                     size_t batch, const unsigned* starts, cudaStream_t s) {
  if (batch == 0 || n == 0) return;
  if (n > 16) throw std::runtime_error("tsqr_b200: qr32x16 needs n <= 16");
  const unsigned blocks = (unsigned)((batch + 3) / 4);
  small_qr_kernel<QT, RT, AT><<<blocks, 128, 0, s>>>(q, (long long)ldq, r, (long long)ldr, a, (long long)lda, m, n,
                                                     (unsigned long long)batch, starts);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

template void small_qr_launch<float, float, float>(float*, size_t, float*, size_t, const float*, size_t, unsigned,
                                                   unsigned, size_t, const unsigned*, cudaStream_t);
template void small_qr_launch<__half, float, float>(__half*, size_t, float*, size_t, const float*, size_t, unsigned,
                                                    unsigned, size_t, const unsigned*, cudaStream_t);
template void small_qr_launch<__half, __half, __half>(__half*, size_t, __half*, size_t, const __half*, size_t,
                                                      unsigned, unsigned, size_t, const unsigned*, cudaStream_t);

}  // namespace tsqr_b200
