// BlockQR for n > 64: left-looking block Gram-Schmidt over 64-column panels, each panel
// factored by the native TSQR; optional second projection + second TSQR per panel
// (re-orthogonalisation). Mirrors the *algorithm* of reference src/blockqr.cu:45-178 and
// :180-390 (panel width 64 instead of 16, hand-written GEMMs instead of cuBLAS).
#include "driver.h"
#include "common.cuh"
#include <algorithm>

namespace tsqr_b200 {
namespace {

constexpr int kNB = kMaxN;     // panel width
constexpr int kSlabRows = 32;  // rows staged per step in the T-N product
constexpr int kTcProjectMaxK = 512;   // columns of Q1 one tc_tn_kernel launch can hold (8 accumulator blocks of 64 TMEM columns)

// Cp[slab][i + k*kb] = sum over the slab's rows of X[row, i] * Y[row, k]   (i < kb, k < nb)
template <typename T>
__global__ void __launch_bounds__(256) gemm_tn_partial(const T* __restrict__ X, long long ldx, int kb,
                                                       const T* __restrict__ Y, long long ldy, int nb,
                                                       long long M, long long rows_per_slab, float* __restrict__ Cp) {
  __shared__ float sX[64][kSlabRows + 1];
  __shared__ float sY[64][kSlabRows + 1];
  const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
  const int i0 = blockIdx.y * 64;
  const long long rbeg = (long long)blockIdx.x * rows_per_slab;
  long long rend = rbeg + rows_per_slab;
  if (rend > M) rend = M;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int lr = t & 31, lc = t >> 5;  // loader: row within step, column group (8 columns each)
  for (long long r0 = rbeg; r0 < rend; r0 += kSlabRows) {
    const long long row = r0 + lr;
    const bool rok = row < rend;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const int col = lc * 8 + c;
      sX[col][lr] = (rok && i0 + col < kb) ? io_load(X + row + (long long)(i0 + col) * ldx) : 0.f;
      sY[col][lr] = (rok && col < nb) ? io_load(Y + row + (long long)col * ldy) : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int rr = 0; rr < kSlabRows; ++rr) {
      float xv[4], yv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) xv[i] = sX[ty * 4 + i][rr];
#pragma unroll
      for (int j = 0; j < 4; ++j) yv[j] = sY[tx * 4 + j][rr];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xv[i], yv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* out = Cp + (size_t)blockIdx.x * kb * nb;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ci = i0 + ty * 4 + i, ck = tx * 4 + j;
      if (ci < kb && ck < nb) out[ci + (size_t)ck * kb] = acc[i][j];
    }
}

// C[i] = sum_slab Cp[slab][i] in fixed order (deterministic)
__global__ void reduce_partials(const float* __restrict__ Cp, int nslabs, size_t count, float* __restrict__ C) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int sl = 0; sl < nslabs; ++sl) s += Cp[(size_t)sl * count + i];
    C[i] = s;
  }
}

// Y[row, 0:nb] -= X[row, 0:kb] * C[kb x nb]   (C column-major, ld kb), one row per thread
template <typename T>
__global__ void __launch_bounds__(128) gemm_nn_update(const T* __restrict__ X, long long ldx, int kb,
                                                      const float* __restrict__ C, T* Y, long long ldy, int nb,
                                                      long long M) {
  __shared__ __align__(16) float sC[16][kNB];
  const int t = threadIdx.x;
  const long long row = (long long)blockIdx.x * 128 + t;
  const bool rok = row < M;
  float acc[kNB];
#pragma unroll
  for (int k = 0; k < kNB; ++k) acc[k] = 0.f;
  for (int c0 = 0; c0 < kb; c0 += 16) {
    __syncthreads();
    for (int i = t; i < 16 * kNB; i += 128) {
      const int c = i / kNB, k = i % kNB;
      sC[c][k] = (c0 + c < kb && k < nb) ? C[(c0 + c) + (size_t)k * kb] : 0.f;
    }
    __syncthreads();
    float xv[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) xv[c] = (rok && c0 + c < kb) ? io_load(X + row + (long long)(c0 + c) * ldx) : 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
#pragma unroll
      for (int k4 = 0; k4 < kNB / 4; ++k4) {
        const float4 w = *reinterpret_cast<const float4*>(&sC[c][k4 * 4]);
        acc[k4 * 4 + 0] = fmaf(xv[c], w.x, acc[k4 * 4 + 0]);
        acc[k4 * 4 + 1] = fmaf(xv[c], w.y, acc[k4 * 4 + 1]);
        acc[k4 * 4 + 2] = fmaf(xv[c], w.z, acc[k4 * 4 + 2]);
        acc[k4 * 4 + 3] = fmaf(xv[c], w.w, acc[k4 * 4 + 3]);
      }
    }
  }
  if (rok) {
#pragma unroll
    for (int k = 0; k < kNB; ++k)
      if (k < nb) {
        T* p = Y + row + (long long)k * ldy;
        io_store(p, io_plain(p) - acc[k]);
      }
  }
}

// C[i + k*ldc] (+)= sum_d X[i + d*ldx] * Y[d + k*ldy]   (small, one output per thread)
__global__ void small_mm(const float* X, int ldx, const float* Y, int ldy, float* C, int ldc, int rows, int inner,
                         int cols, int accumulate) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * cols) return;
  const int i = idx % rows, k = idx / rows;
  float s = 0.f;
  for (int d = 0; d < inner; ++d) s = fmaf(X[i + (size_t)d * ldx], Y[d + (size_t)k * ldy], s);
  C[i + (size_t)k * ldc] = accumulate ? C[i + (size_t)k * ldc] + s : s;
}

__global__ void copy_block(const float* src, int lds, float* dst, int ldd, int rows, int cols) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * cols) return;
  const int i = idx % rows, k = idx / rows;
  dst[i + (size_t)k * ldd] = src[i + (size_t)k * lds];
}

// user R[0:kb+nb, c0:c0+nb] <- Rcol (fp32, ld = ldrc); strictly-lower part of the diagonal block zeroed
template <typename T>
__global__ void store_rcol(const float* Rcol, int ldrc, int kb, int nb, T* R, long long ldr, int c0) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int rows = kb + nb;
  if (idx >= rows * nb) return;
  const int i = idx % rows, k = idx / rows;
  const float v = (i <= kb + k) ? Rcol[i + (size_t)k * ldrc] : 0.f;
  io_store(R + i + (long long)(c0 + k) * ldr, v);
}

struct QrWs {
  size_t tsqr_off, cp_off, c_off, s_off, rcol_off, w_off, cg_off, total;
  int nslabs;
  long long rows_per_slab;
};

QrWs carve(long long m, long long n, bool reorth) {
  QrWs w{};
  const int nbmax = (int)std::min<long long>(n, kNB);
  size_t off = 0;
  auto take = [&](size_t f) { size_t o = off; off = (off + f + 63) / 64 * 64; return o; };
  w.tsqr_off = take(plan_floats_any(m, nbmax));
  if (n > kNB) {
    const long long kbmax = (n - 1) / kNB * kNB;
    w.nslabs = (int)std::min<long long>(2LL * device_sm_count(), (m + 1023) / 1024);
    if (w.nslabs < 1) w.nslabs = 1;
    w.rows_per_slab = ((m + w.nslabs - 1) / w.nslabs + kSlabRows - 1) / kSlabRows * kSlabRows;
    w.nslabs = (int)((m + w.rows_per_slab - 1) / w.rows_per_slab);
    w.cp_off = take((size_t)2 * w.nslabs * kbmax * kNB);  // the tensor-core product writes two slots per slab
    w.c_off = take((size_t)kbmax * kNB);
    w.s_off = take(reorth ? (size_t)kbmax * kNB : 0);
    w.rcol_off = take((size_t)n * kNB);
    w.w_off = take((size_t)kNB * kNB * 2);
    w.cg_off = take(kbmax > kTcProjectMaxK ? (size_t)kTcProjectMaxK * kNB : 0);
  }
  w.total = off;
  return w;
}

void reduce_partials_launch(const float* Cp, int nslots, size_t count, float* C, cudaStream_t s) {
  reduce_partials<<<(unsigned)std::min<size_t>((count + 255) / 256, 1024), 256, 0, s>>>(Cp, nslots, count, C);
  note_launch();
}

template <typename T>
void project(int mode, const T* Q1, long long ldq, int kb, T* Y, long long ldy, int nb, long long m, const QrWs& w, float* ws,
             float* Cout, cudaStream_t s) {
  if (use_tc(mode, m)) {  // tcgen05 path: both GEMMs on the tensor cores
    const IoType io = sizeof(T) == 2 ? IoType::f16 : IoType::f32;
    if (kb <= kTcProjectMaxK) {
      launch_tc_project(mode_is_cor(mode), io, Q1, ldq, kb, Y, ldy, nb, m, w.nslabs, w.rows_per_slab, ws + w.cp_off, Cout,
                        reduce_partials_launch, s);
      return;
    }
    // The tn kernel keeps one 64-column accumulator block per 64 columns of Q1 in its 512 TMEM columns: wider
    // Q1 goes through in groups of 512 columns, each group projected out of Y before the next one is measured
    // (block Gram-Schmidt across the groups; A2 = sum_g Q1_g C_g + Y' holds exactly as before).
    float* Cg = ws + w.cg_off;
    for (int g0 = 0; g0 < kb; g0 += kTcProjectMaxK) {
      const int kg = std::min(kTcProjectMaxK, kb - g0);
      launch_tc_project(mode_is_cor(mode), io, Q1 + (long long)g0 * ldq, ldq, kg, Y, ldy, nb, m, w.nslabs, w.rows_per_slab,
                        ws + w.cp_off, Cg, reduce_partials_launch, s);
      copy_block<<<(kg * nb + 255) / 256, 256, 0, s>>>(Cg, kg, Cout + g0, kb, kg, nb);
      note_launch();
    }
    TSQR_CUDA_CHECK(cudaGetLastError());
    return;
  }
  dim3 grid(w.nslabs, (kb + 63) / 64);
  gemm_tn_partial<T><<<grid, 256, 0, s>>>(Q1, ldq, kb, Y, ldy, nb, m, w.rows_per_slab, ws + w.cp_off);
  note_launch();
  const size_t count = (size_t)kb * nb;
  reduce_partials<<<(unsigned)std::min<size_t>((count + 255) / 256, 1024), 256, 0, s>>>(ws + w.cp_off, w.nslabs, count, Cout);
  note_launch();
  gemm_nn_update<T><<<(unsigned)((m + 127) / 128), 128, 0, s>>>(Q1, ldq, kb, Cout, Y, ldy, nb, m);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

template <typename T>
int qr_typed(int mode, bool reorth, T* q, long long ldq, T* r, long long ldr, T* a, long long lda, long long m,
             long long n, float* ws, size_t ws_bytes, cudaStream_t s) {
  const QrWs w = carve(m, n, reorth);
  if (w.total * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
  if (n <= kNB) {
    tsqr_full(mode, q, ldq, r, ldr, a, lda, m, (int)n, ws + w.tsqr_off, ws_bytes - w.tsqr_off * sizeof(float), s);
    return 0;
  }
  float* tws = ws + w.tsqr_off;
  float* Rcol = ws + w.rcol_off;
  const int ldrc = (int)n;
  for (long long c0 = 0; c0 < n; c0 += kNB) {
    const int nb = (int)std::min<long long>(kNB, n - c0);
    const int kb = (int)c0;
    T* A2 = a + c0 * lda;
    T* Q2 = q + c0 * ldq;
    if (kb > 0) project<T>(mode, q, ldq, kb, A2, lda, nb, m, w, ws, ws + w.c_off, s);
    Plan plan = make_plan_for(mode, m, nb);
    const float* r22 = tsqr_reduce(plan, mode, A2, lda, Q2, ldq, tws, s);
    if (kb > 0) copy_block<<<(kb * nb + 255) / 256, 256, 0, s>>>(ws + w.c_off, kb, Rcol, ldrc, kb, nb);
    note_launch();
    copy_block<<<(nb * nb + 255) / 256, 256, 0, s>>>(r22, nb, Rcol + kb, ldrc, nb, nb);
    note_launch();
    tsqr_build(plan, mode, A2, lda, Q2, ldq, nullptr, 0, tws, s);
    if (reorth && kb > 0) {
      // second projection and second TSQR of the panel (CGS2), then fix up R
      project<T>(mode, q, ldq, kb, Q2, ldq, nb, m, w, ws, ws + w.s_off, s);
      const float* wfac = tsqr_reduce(plan, mode, Q2, ldq, Q2, ldq, tws, s);
      float* Wm = ws + w.w_off;
      float* R22old = Wm + kNB * kNB;
      copy_block<<<(nb * nb + 255) / 256, 256, 0, s>>>(wfac, nb, Wm, nb, nb, nb);
      note_launch();
      copy_block<<<(nb * nb + 255) / 256, 256, 0, s>>>(Rcol + kb, ldrc, R22old, nb, nb, nb);
      note_launch();
      tsqr_build(plan, mode, Q2, ldq, Q2, ldq, nullptr, 0, tws, s);
      small_mm<<<(kb * nb + 255) / 256, 256, 0, s>>>(ws + w.s_off, kb, R22old, nb, Rcol, ldrc, kb, nb, nb, 1);
      note_launch();
      small_mm<<<(nb * nb + 255) / 256, 256, 0, s>>>(Wm, nb, R22old, nb, Rcol + kb, ldrc, nb, nb, nb, 0);
      note_launch();
    }
    store_rcol<T><<<((kb + nb) * nb + 255) / 256, 256, 0, s>>>(Rcol, ldrc, kb, nb, r, ldr, (int)c0);
    note_launch();
    TSQR_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

}  // namespace

size_t qr_workspace_bytes(long long m, long long n, bool reorth) {
  return carve(m, n, reorth).total * sizeof(float) + 256;
}

int qr_full(int mode, bool reorth, void* q, long long ldq, void* r, long long ldr, void* a, long long lda,
            long long m, long long n, float* ws, size_t ws_bytes, cudaStream_t s) {
  if (n > m || m <= 0 || n <= 0) return 1;  // reference src/blockqr.cu:407-409
  if (mode_io(mode) == IoType::f32)
    return qr_typed<float>(mode, reorth, (float*)q, ldq, (float*)r, ldr, (float*)a, lda, m, n, ws, ws_bytes, s);
  return qr_typed<__half>(mode, reorth, (__half*)q, ldq, (__half*)r, ldr, (__half*)a, lda, m, n, ws, ws_bytes, s);
}

}  // namespace tsqr_b200
