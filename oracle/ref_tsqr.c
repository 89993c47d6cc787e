/* ORACLE -- TEST INFRASTRUCTURE ONLY. Never linked into, imported by or called from the
 * product (tsqr_gpu_b200/, libtsqr_b200.so). Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may use it, and only as the checker.
 *
 * Plain-C (CPU, fp32) restatement of the reference's fp32_notc algorithm, written from the
 * behaviour of enp1s0/tsqr-gpu, function by function:
 *   ref_qr32x16      <- src/tcqr32x16.cu:1373-1469 (qr32x16_core) with the generic make_h
 *                       (:117-137) and update_qr (:464-496); tile load/store semantics of
 *                       :1503-1531 and src/matrix_copy.cuh:95-162
 *   ref_tsqr16       <- src/tsqr.cu:39-44 (row split), :1064-1279 (tsqr16_geq32: leaf level,
 *                       binary tree of 2n x n tiles, backward products :143-204, :591-656),
 *                       :1281-1310 (m <= 32 single tile)
 *   ref_blockqr      <- src/blockqr.cu:45-178 (block_qr_core) and :180-390 (re-orthogonalised
 *                       variant); the cuBLAS gemms of :92-116, :230-332 are plain fp32 loops
 *
 * PARITY PINNING: the reference ships no golden vectors, fixed seeds or expected values
 * (SURVEY.md section 8c), and it has no CPU path, so this restatement cannot be pinned
 * against reference outputs inside the CPU container. It is pinned (a) against host LAPACK
 * sgeqrf/sorgqr through the reference's own metric definitions (tests/test_oracle.py) and
 * (b) on the GPU box against the reference's own sm_100 build (oracle/_ref, see
 * oracle/build_ref.sh) by tests/test_reference_gpu.py.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define TM 32 /* tile rows   (FRAGMENT_DIM_M) */
#define TN 16 /* tile cols   (FRAGMENT_DIM_N) */

static float sign_f(float x) { return signbit(x) ? -1.0f : 1.0f; } /* cutf::math::sign, +0 -> +1 */

/* QR of one rows x n tile (rows <= 32, n <= 16), explicit 32x32 Householder matrix per step.
 * q: rows x n (ldq), r: n x n (ldr), a: rows x n (lda); all column-major. */
void ref_qr32x16(float* q, size_t ldq, float* r, size_t ldr, const float* a, size_t lda, unsigned rows, unsigned n) {
  float Q[TM * TM], R[TM * TN], H[TM * TM], u[TM], T[TM * TM];
  memset(R, 0, sizeof(R));
  for (unsigned c = 0; c < n; ++c)
    for (unsigned i = 0; i < rows; ++i) R[c * TM + i] = a[c * lda + i]; /* zero padded load */
  memset(Q, 0, sizeof(Q));
  for (unsigned i = 0; i < TM; ++i) Q[i * TM + i] = 1.0f;
  for (unsigned k = 0; k < n; ++k) {
    for (unsigned i = 0; i < TM; ++i) u[i] = (i >= k && i < rows) ? R[k * TM + i] : 0.0f;
    float s = 0.0f;
    for (unsigned i = 0; i < rows; ++i) s += u[i] * u[i];
    const float norm_u0 = sqrtf(s);
    u[k] += sign_f(u[k]) * norm_u0;
    float norm2_u1 = 0.0f;
    for (unsigned i = 0; i < rows; ++i) norm2_u1 += u[i] * u[i];
    /* H = I - 2 u u^T / |u|^2  (0/0 -> NaN for a zero column, as in the reference) */
    for (unsigned x = 0; x < TM; ++x)
      for (unsigned y = 0; y < TM; ++y) {
        float t = (x == y) ? 1.0f : 0.0f;
        if (x < rows && y < rows) t -= (2.0f * u[y] / norm2_u1) * u[x];
        H[x * TM + y] = t;
      }
    /* Q <- H Q (32x32x32), R <- H R (32x32x16) */
    for (unsigned c = 0; c < TM; ++c)
      for (unsigned i = 0; i < TM; ++i) {
        float acc = 0.0f;
        for (unsigned l = 0; l < TM; ++l) acc = fmaf(H[l * TM + i], Q[c * TM + l], acc);
        T[c * TM + i] = acc;
      }
    memcpy(Q, T, sizeof(Q));
    for (unsigned c = 0; c < TN; ++c)
      for (unsigned i = 0; i < TM; ++i) {
        float acc = 0.0f;
        for (unsigned l = 0; l < TM; ++l) acc = fmaf(H[l * TM + i], R[c * TM + l], acc);
        T[c * TM + i] = acc;
      }
    memcpy(R, T, sizeof(float) * TM * TN);
  }
  /* accumulated Q is H_{n-1}...H_0; the tile's Q factor is its transpose, first n columns */
  for (unsigned c = 0; c < n; ++c)
    for (unsigned i = 0; i < rows; ++i) q[c * ldq + i] = Q[i * TM + c];
  for (unsigned c = 0; c < n; ++c)
    for (unsigned i = 0; i < n; ++i) r[c * ldr + i] = R[c * TM + i]; /* full n x n block */
}

size_t ref_batch_size_log2(size_t m) {
  unsigned l = (unsigned)ceilf(log2f((float)m));
  if (l < 5u) l = 5u;
  return l - 5u;
}

/* TSQR of an m x n matrix, n <= 16. */
void ref_tsqr16(float* q, size_t ldq, float* r, size_t ldr, const float* a, size_t lda, size_t m, size_t n) {
  if (m <= 32) {
    ref_qr32x16(q, ldq, r, ldr, a, lda, (unsigned)m, (unsigned)n);
    return;
  }
  const size_t L = ref_batch_size_log2(m), batch = (size_t)1 << L;
  /* leaf level: Q_leaf stacked in an m x n array, R's stacked in (n*batch) x n */
  float* qleaf = (float*)malloc(sizeof(float) * m * n);
  float** qlev = (float**)calloc(L + 1, sizeof(float*)); /* qlev[k]: (2n*2^k) x n, k < L */
  float* rcur = (float*)malloc(sizeof(float) * n * batch * n);
  size_t ldrc = n * batch;
  for (size_t i = 0; i < batch; ++i) {
    const size_t r0 = m * i / batch, r1 = m * (i + 1) / batch;
    ref_qr32x16(qleaf + r0, m, rcur + n * i, ldrc, a + r0, lda, (unsigned)(r1 - r0), (unsigned)n);
  }
  for (size_t k = L; k-- > 0;) { /* level k has 2^k tiles of 2n rows */
    const size_t tiles = (size_t)1 << k;
    float* rnext = (float*)malloc(sizeof(float) * n * tiles * n);
    qlev[k] = (float*)malloc(sizeof(float) * 2 * n * tiles * n);
    for (size_t i = 0; i < tiles; ++i) {
      if (k == 0)
        ref_qr32x16(qlev[k], 2 * n, r, ldr, rcur, ldrc, (unsigned)(2 * n), (unsigned)n);
      else
        ref_qr32x16(qlev[k] + 2 * n * i, 2 * n * tiles, rnext + n * i, n * tiles, rcur + 2 * n * i, ldrc,
                    (unsigned)(2 * n), (unsigned)n);
    }
    free(rcur);
    rcur = rnext;
    ldrc = n * tiles;
  }
  free(rcur);
  /* backward: level k tile i <- tile * B, B = rows [n*i, n*i+n) of level k-1 */
  float tmp[TM * TN];
  for (size_t k = 1; k < L; ++k) {
    const size_t tiles = (size_t)1 << k, ldk = 2 * n * tiles, ldp = n * tiles;
    for (size_t i = 0; i < tiles; ++i) {
      float* ac = qlev[k] + 2 * n * i;
      const float* b = qlev[k - 1] + n * i;
      for (size_t c = 0; c < n; ++c)
        for (size_t x = 0; x < 2 * n; ++x) {
          float acc = 0.0f;
          for (size_t l = 0; l < n; ++l) acc = fmaf(ac[l * ldk + x], b[c * ldp + l], acc);
          tmp[c * TM + x] = acc;
        }
      for (size_t c = 0; c < n; ++c)
        for (size_t x = 0; x < 2 * n; ++x) ac[c * ldk + x] = tmp[c * TM + x];
    }
  }
  for (size_t i = 0; i < batch; ++i) { /* final: Q rows of leaf i = Q_leaf_i * B_i */
    const size_t r0 = m * i / batch, r1 = m * (i + 1) / batch;
    const float* b = qlev[L - 1] + n * i;
    const size_t ldb = n * batch;
    for (size_t c = 0; c < n; ++c)
      for (size_t x = r0; x < r1; ++x) {
        float acc = 0.0f;
        for (size_t l = 0; l < n; ++l) acc = fmaf(qleaf[l * m + x], b[c * ldb + l], acc);
        q[c * ldq + x] = acc;
      }
  }
  for (size_t k = 0; k < L; ++k) free(qlev[k]);
  free(qlev);
  free(qleaf);
}

/* C (p x w, ldc) = X[:, 0:p]^T Y[:, 0:w] (+= if acc) */
static void gemm_tn(float* c, size_t ldc, const float* x, size_t ldx, const float* y, size_t ldy, size_t m, size_t p,
                    size_t w, int accumulate) {
  for (size_t j = 0; j < w; ++j)
    for (size_t i = 0; i < p; ++i) {
      float s = 0.0f;
      for (size_t l = 0; l < m; ++l) s = fmaf(x[i * ldx + l], y[j * ldy + l], s);
      c[j * ldc + i] = accumulate ? c[j * ldc + i] + s : s;
    }
}
/* Y[:, 0:w] -= X[:, 0:p] C (p x w) */
static void gemm_nn_sub(float* y, size_t ldy, const float* x, size_t ldx, const float* c, size_t ldc, size_t m,
                        size_t p, size_t w) {
  for (size_t j = 0; j < w; ++j)
    for (size_t l = 0; l < p; ++l) {
      const float cv = c[j * ldc + l];
      for (size_t i = 0; i < m; ++i) y[j * ldy + i] = fmaf(-x[l * ldx + i], cv, y[j * ldy + i]);
    }
}

/* BlockQR, 16-column panels; a is overwritten when n > 16. returns 0, or 1 for bad sizes. */
int ref_blockqr(float* q, size_t ldq, float* r, size_t ldr, float* a, size_t lda, size_t m, size_t n, int reorth) {
  if (n > m || m == 0 || n == 0) return 1;
  const size_t nblocks = (n + TN - 1) / TN;
  float* s2 = reorth ? (float*)malloc(sizeof(float) * m * TN) : NULL; /* stored with ld = m in the reference */
  float r2[TN * TN], w[TN * TN], t[TN * TN];
  for (size_t b = 0; b < nblocks; ++b) {
    const size_t c0 = b * TN, wd = (n - c0 < TN) ? n - c0 : TN;
    float* a2 = a + c0 * lda;
    float* q2 = q + c0 * ldq;
    if (b == 0 || !reorth) {
      if (b > 0) {
        gemm_tn(r + c0 * ldr, ldr, q, ldq, a2, lda, m, c0, wd, 0);
        gemm_nn_sub(a2, lda, q, ldq, r + c0 * ldr, ldr, m, c0, wd);
      }
      ref_tsqr16(q2, ldq, r + c0 * ldr + c0, ldr, a2, lda, m, wd);
      continue;
    }
    gemm_tn(r + c0 * ldr, ldr, q, ldq, a2, lda, m, c0, wd, 0);
    gemm_nn_sub(a2, lda, q, ldq, r + c0 * ldr, ldr, m, c0, wd);
    ref_tsqr16(q2, ldq, r2, TN, a2, lda, m, wd);
    gemm_tn(s2, m, q, ldq, q2, ldq, m, c0, wd, 0);
    gemm_nn_sub(q2, ldq, q, ldq, s2, m, m, c0, wd);
    {
      float* qtmp = (float*)malloc(sizeof(float) * m * wd);
      for (size_t j = 0; j < wd; ++j) memcpy(qtmp + j * m, q2 + j * ldq, sizeof(float) * m);
      ref_tsqr16(q2, ldq, w, TN, qtmp, m, m, wd);
      free(qtmp);
    }
    /* R12 += S2 R2 ; R22 = W R2 */
    for (size_t j = 0; j < wd; ++j)
      for (size_t i = 0; i < c0; ++i) {
        float s = 0.0f;
        for (size_t l = 0; l < wd; ++l) s = fmaf(s2[l * m + i], r2[j * TN + l], s);
        r[(c0 + j) * ldr + i] += s;
      }
    for (size_t j = 0; j < wd; ++j)
      for (size_t i = 0; i < wd; ++i) {
        float s = 0.0f;
        for (size_t l = 0; l < wd; ++l) s = fmaf(w[l * TN + i], r2[j * TN + l], s);
        t[j * TN + i] = s;
      }
    for (size_t j = 0; j < wd; ++j)
      for (size_t i = 0; i < wd; ++i) r[(c0 + j) * ldr + c0 + i] = t[j * TN + i];
  }
  free(s2);
  return 0;
}
