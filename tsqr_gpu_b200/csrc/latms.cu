// Conditioned test matrices on the device: A = U diag(s) V^T, m x n fp32, any n <= m, with
// U (m x n) and V (n x n) the orthogonal factors of THIS library's QR of counter-based standard normals and
// s_k = cond^(-k / (n - 1)) (geometric from 1 down to 1 / cond: LAPACK slatms MODE = 3; singular values as reference
// src/test_cond.cu:31-50). Replaces mtk::utils::latms (reference src/latms.cu:9-121: cuRAND MT19937 normals + cuSOLVER
// geqrf/orgqr + cuBLAS gemm); no cuRAND, cuSOLVER or cuBLAS here. The singular values are known by construction, so the
// gesvd of the reference's get_cond (src/latms.cu:129-170) has no counterpart.
#include <algorithm>
#include "common.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {   // splitmix64 finaliser
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// element (row, col) of the normal matrix of stream `seed`: a pure function of its coordinates (Box-Muller on two
// hashed uniforms), so the matrix does not depend on the launch shape
__device__ __forceinline__ float normal_at(unsigned long long seed, unsigned long long row, unsigned long long col) {
  const unsigned long long h = mix64(mix64(seed ^ (col * 0xD1342543DE82EF95ull)) + row);
  const float u1 = ((float)(unsigned)(h >> 40) + 1.0f) * (1.0f / 16777217.0f);       // (0, 1)
  const float u2 = (float)(unsigned)((h >> 8) & 0xFFFFFFu) * (1.0f / 16777216.0f);   // [0, 1)
  return sqrtf(-2.0f * __logf(u1)) * __cosf(6.28318530717958647692f * u2);
}
__global__ void normal_fill_kernel(float* __restrict__ a, long long lda, long long m, int n, unsigned long long seed) {
  const long long total = m * n;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long row = e % m, col = e / m;
    a[row + col * lda] = normal_at(seed, (unsigned long long)row, (unsigned long long)col);
  }
}
// A[row, :] <- (A[row, :] diag(s)) V^T, in place, one row per thread; W[k][j] = s_k V[j][k] in shared memory
__global__ void __launch_bounds__(128) scale_rotate_kernel(float* __restrict__ a, long long lda, long long m, int n,
                                                          const float* __restrict__ V, int ldv, float cond) {
  __shared__ float W[64][65];
  for (int e = threadIdx.x; e < 64 * 64; e += 128) {   // all of W: the padded entries multiply the zero-padded row
    const int k = e / 64, j = e % 64;
    const float sk = n > 1 ? exp2f(-log2f(cond) * (float)k / (float)(n - 1)) : 1.f;
    W[k][j] = (k < n && j < n) ? sk * V[j + (long long)k * ldv] : 0.f;
  }
  __syncthreads();
  const long long row = blockIdx.x * 128ll + threadIdx.x;
  if (row >= m) return;
  float u[64];
#pragma unroll
  for (int k = 0; k < 64; ++k) u[k] = k < n ? a[row + (long long)k * lda] : 0.f;
  for (int j = 0; j < n; ++j) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int k = 0; k < 64; k += 2) { s0 = __fmaf_rn(u[k], W[k][j], s0); s1 = __fmaf_rn(u[k + 1], W[k + 1][j], s1); }
    a[row + (long long)j * lda] = s0 + s1;
  }
}

// n > 64: A[rows, 64 jb ..] = sum over 64-column blocks kb of U[rows, kb] W[kb, jb], W[k][j] = s_k V[j][k]; one row per
// thread, out of place (U lives in the workspace)
__global__ void __launch_bounds__(128) scale_rotate_wide_kernel(const float* __restrict__ U, long long ldu, float* __restrict__ a,
                                                               long long lda, long long m, int n, const float* __restrict__ V, int ldv,
                                                               float cond) {
  __shared__ float W[64][65];
  const long long row = blockIdx.x * 128ll + threadIdx.x;
  const int j0 = blockIdx.y * 64;
  float acc[64];
#pragma unroll
  for (int j = 0; j < 64; ++j) acc[j] = 0.f;
  for (int k0 = 0; k0 < n; k0 += 64) {
    __syncthreads();
    for (int e = threadIdx.x; e < 64 * 64; e += 128) {
      const int k = k0 + e / 64, j = j0 + e % 64;
      const float sk = exp2f(-log2f(cond) * (float)k / (float)(n - 1));
      W[e / 64][e % 64] = (k < n && j < n) ? sk * V[j + (long long)k * ldv] : 0.f;
    }
    __syncthreads();
    if (row < m) {
      float u[64];
#pragma unroll
      for (int k = 0; k < 64; ++k) u[k] = k0 + k < n ? U[row + (long long)(k0 + k) * ldu] : 0.f;
#pragma unroll 4
      for (int j = 0; j < 64; ++j) {
        float s0 = acc[j], s1 = 0.f;
#pragma unroll
        for (int k = 0; k < 64; k += 2) { s0 = __fmaf_rn(u[k], W[k][j], s0); s1 = __fmaf_rn(u[k + 1], W[k + 1][j], s1); }
        acc[j] = s0 + s1;
      }
    }
  }
  if (row < m) {
#pragma unroll
    for (int j = 0; j < 64; ++j)
      if (j0 + j < n) a[row + (long long)(j0 + j) * lda] = acc[j];
  }
}

struct LatmsWs { size_t v_off, vq_off, r_off, u_off, qr_off, total; };
LatmsWs carve(long long m, int n) {
  LatmsWs w{};
  size_t off = 0;
  auto take = [&](size_t f) { size_t o = off; off = (off + f + 63) / 64 * 64; return o; };
  const bool wide = n > kMaxN;   // BlockQR does not work in place: U and the Q factor of V get their own buffers
  w.v_off = take((size_t)n * n);
  w.vq_off = take(wide ? (size_t)n * n : 0);
  w.r_off = take((size_t)n * n);
  w.u_off = take(wide ? (size_t)m * n : 0);
  w.qr_off = take(std::max(qr_workspace_bytes(m, n, wide), qr_workspace_bytes(n, n, wide)) / sizeof(float) + 64);
  w.total = off;
  return w;
}

}  // namespace

size_t latms_workspace_bytes(long long m, int n) { return carve(m, n).total * sizeof(float) + 256; }

void latms(float* a, long long lda, long long m, int n, float cond, unsigned long long seed, float* ws, size_t ws_bytes,
           cudaStream_t s) {
  if (n < 1 || m < n) throw std::runtime_error("tsqr_b200: latms takes 1 <= n <= m");
  if (!(cond >= 1.f)) throw std::runtime_error("tsqr_b200: latms takes cond >= 1");
  const LatmsWs w = carve(m, n);
  if (w.total * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
  float* V = ws + w.v_off;
  float* R = ws + w.r_off;
  float* qws = ws + w.qr_off;
  const size_t qws_bytes = (w.total - w.qr_off) * sizeof(float);
  const int blocks = (int)std::min<long long>((m * n + 255) / 256, 148ll * 16);
  normal_fill_kernel<<<blocks, 256, 0, s>>>(a, lda, m, n, seed);
  normal_fill_kernel<<<(int)std::min<long long>(((long long)n * n + 255) / 256, 1024), 256, 0, s>>>(V, n, n, n, seed ^ 0xA5A5A5A5DEADBEEFull);
  note_launch(2);
  TSQR_CUDA_CHECK(cudaGetLastError());
  if (n <= kMaxN) {
    // U over the normals, in place (q may alias a); V likewise
    if (qr_full(fp32_tc_cor, false, a, lda, R, n, a, lda, m, n, qws, qws_bytes, s) != 0) throw std::runtime_error("tsqr_b200: latms U factorisation failed");
    if (qr_full(fp32_notc, false, V, n, R, n, V, n, n, n, qws, qws_bytes, s) != 0) throw std::runtime_error("tsqr_b200: latms V factorisation failed");
    scale_rotate_kernel<<<(unsigned)((m + 127) / 128), 128, 0, s>>>(a, lda, m, n, V, n, cond);
  } else {
    // BlockQR with re-orthogonalisation: U into the workspace (a is overwritten by the factorisation), Vq beside V
    float* U = ws + w.u_off;
    float* Vq = ws + w.vq_off;
    if (qr_full(fp32_tc_cor, true, U, m, R, n, a, lda, m, n, qws, qws_bytes, s) != 0) throw std::runtime_error("tsqr_b200: latms U factorisation failed");
    if (qr_full(fp32_notc, true, Vq, n, R, n, V, n, n, n, qws, qws_bytes, s) != 0) throw std::runtime_error("tsqr_b200: latms V factorisation failed");
    dim3 grid((unsigned)((m + 127) / 128), (unsigned)((n + 63) / 64));
    scale_rotate_wide_kernel<<<grid, 128, 0, s>>>(U, m, a, lda, m, n, Vq, n, cond);
  }
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tsqr_b200
