// On-device validation metrics in fp64, any n <= m:
//   orthogonality  |Q^T Q - I|_F / sqrt(n)      replaces reference src/validation.cu:43-80 (check_orthogonality16: fp64
//                                               cuBLAS GEMM of the whole Q^T Q, then a host loop)
//   residual       |A - Q R|_F / |A|_F          replaces reference src/test.cu:146-165 (cuBLAS GEMM, D2H of m x n floats
//                                               and a float accumulation on the host, which loses accuracy at 2^30 elements)
// Both stream their matrices once per 64-column block pair, accumulate in fp64 on the CUDA cores (4 x 4 register blocks
// per thread) and finish on the device; nothing is copied to the host except the two numbers the caller asks for.
#include "../../include/tsqr_b200.h"
#include "common.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {

constexpr int kVB = 64;       // column block
constexpr int kVRows = 32;    // rows staged per step
constexpr int kVSlab = 8192;  // rows per CTA

// G[bi, bj] += Q[slab, bi]^T Q[slab, bj] for one 64 x 64 block pair (bi <= bj) and one row slab; 256 threads, thread
// (ty, tx) owns the 4 x 4 sub-block rows 4 ty.., columns 4 tx..
template <typename T>
__global__ void __launch_bounds__(256) gram64_kernel(const T* __restrict__ Q, long long ldq, long long m, int n, double* __restrict__ G) {
  __shared__ float sa[kVRows][kVB + 1], sb[kVRows][kVB + 1];
  const int nb = (n + kVB - 1) / kVB;
  int pair = blockIdx.y, bi = 0;
  while (pair >= nb - bi) { pair -= nb - bi; ++bi; }
  const int bj = bi + pair;
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const long long r0 = (long long)blockIdx.x * kVSlab, r1 = min(m, r0 + kVSlab);
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  for (long long r = r0; r < r1; r += kVRows) {
    for (int e = t; e < kVRows * kVB; e += 256) {
      const int rr = e % kVRows, c = e / kVRows;
      const long long row = r + rr;
      const int ca = bi * kVB + c, cb = bj * kVB + c;
      sa[rr][c] = (row < r1 && ca < n) ? io_load(Q + row + (long long)ca * ldq) : 0.f;
      sb[rr][c] = (row < r1 && cb < n) ? io_load(Q + row + (long long)cb * ldq) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int rr = 0; rr < kVRows; ++rr) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = (double)sa[rr][4 * ty + i]; b[i] = (double)sb[rr][4 * tx + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gi = bi * kVB + 4 * ty + i, gj = bj * kVB + 4 * tx + j;
      if (gi < n && gj < n) atomicAdd(G + (size_t)gi * n + gj, acc[i][j]);
    }
}

// sum over the upper block triangle of (G - I)^2, off-diagonal blocks counted twice -> out = sqrt(sum / n)
__global__ void __launch_bounds__(256) orth_finish_kernel(const double* __restrict__ G, int n, double* __restrict__ out) {
  __shared__ double red[256];
  double s = 0.0;
  for (size_t e = threadIdx.x; e < (size_t)n * n; e += 256) {
    const int i = (int)(e / n), j = (int)(e % n);
    const int bi = i / kVB, bj = j / kVB;
    if (bi > bj) continue;
    const double d = G[e] - (i == j ? 1.0 : 0.0);
    s += (bi == bj ? 1.0 : 2.0) * d * d;
  }
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sqrt(red[0] / (double)n);
}

// sums[0] += |A - Q R|^2, sums[1] += |A|^2 over one 64-row tile and one column block bj of the product
template <typename T>
__global__ void __launch_bounds__(256) resid64_kernel(const T* __restrict__ A, long long lda, const T* __restrict__ Q, long long ldq,
                                                     const T* __restrict__ R, long long ldr, long long m, int n,
                                                     double* __restrict__ sums) {
  __shared__ float sq[kVB][kVB + 1], sr[kVB][kVB + 1];   // Q tile (rows x k), R block (k x columns)
  __shared__ double red[2][256];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const long long r0 = (long long)blockIdx.x * kVB;
  const int bj = blockIdx.y;
  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
  for (int bk = 0; bk <= bj; ++bk) {   // R is upper triangular: blocks above and on the diagonal only
    for (int e = t; e < kVB * kVB; e += 256) {
      const int rr = e % kVB, c = e / kVB;
      const long long row = r0 + rr;
      const int k = bk * kVB + c;
      sq[rr][c] = (row < m && k < n) ? io_load(Q + row + (long long)k * ldq) : 0.f;
      const int kr = bk * kVB + rr, cj = bj * kVB + c;
      sr[rr][c] = (kr < n && cj < n && kr <= cj) ? io_load(R + kr + (long long)cj * ldr) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < kVB; ++k) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = (double)sq[4 * ty + i][k]; b[i] = (double)sr[k][4 * tx + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  double num = 0.0, den = 0.0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long row = r0 + 4 * ty + i;
      const int c = bj * kVB + 4 * tx + j;
      if (row < m && c < n) {
        const double a = (double)io_load(A + row + (long long)c * lda);
        const double d = a - acc[i][j];
        num += d * d;
        den += a * a;
      }
    }
  red[0][t] = num; red[1][t] = den;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (t < o) { red[0][t] += red[0][t + o]; red[1][t] += red[1][t + o]; }
    __syncthreads();
  }
  if (t == 0) { atomicAdd(sums, red[0][0]); atomicAdd(sums + 1, red[1][0]); }
}
__global__ void resid_finish_kernel(const double* __restrict__ sums, double* __restrict__ out) {
  out[0] = sums[1] > 0.0 ? sqrt(sums[0] / sums[1]) : 0.0;
}

template <typename T>
void orth_typed(const T* q, long long ldq, long long m, int n, double* out, double* ws, cudaStream_t s) {
  TSQR_CUDA_CHECK(cudaMemsetAsync(ws, 0, (size_t)n * n * sizeof(double), s));
  const int nb = (n + kVB - 1) / kVB;
  dim3 grid((unsigned)((m + kVSlab - 1) / kVSlab), (unsigned)(nb * (nb + 1) / 2));
  gram64_kernel<T><<<grid, 256, 0, s>>>(q, ldq, m, n, ws);
  orth_finish_kernel<<<1, 256, 0, s>>>(ws, n, out);
  note_launch(2);
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <typename T>
void resid_typed(const T* a, long long lda, const T* q, long long ldq, const T* r, long long ldr, long long m, int n, double* out,
                 double* ws, cudaStream_t s) {
  TSQR_CUDA_CHECK(cudaMemsetAsync(ws, 0, 2 * sizeof(double), s));
  dim3 grid((unsigned)((m + kVB - 1) / kVB), (unsigned)((n + kVB - 1) / kVB));
  resid64_kernel<T><<<grid, 256, 0, s>>>(a, lda, q, ldq, r, ldr, m, n, ws);
  resid_finish_kernel<<<1, 1, 0, s>>>(ws, out);
  note_launch(2);
  TSQR_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

size_t validation_workspace_bytes(size_t n) { return n * n * sizeof(double) + 64; }
void orthogonality(IoType io, const void* q, long long ldq, long long m, int n, double* out, void* ws, cudaStream_t s) {
  if (io == IoType::f32) orth_typed<float>((const float*)q, ldq, m, n, out, (double*)ws, s);
  else orth_typed<__half>((const __half*)q, ldq, m, n, out, (double*)ws, s);
}
void residual(IoType io, const void* a, long long lda, const void* q, long long ldq, const void* r, long long ldr, long long m, int n,
              double* out, void* ws, cudaStream_t s) {
  if (io == IoType::f32) resid_typed<float>((const float*)a, lda, (const float*)q, ldq, (const float*)r, ldr, m, n, out, (double*)ws, s);
  else resid_typed<__half>((const __half*)a, lda, (const __half*)q, ldq, (const __half*)r, ldr, m, n, out, (double*)ws, s);
}

}  // namespace tsqr_b200
