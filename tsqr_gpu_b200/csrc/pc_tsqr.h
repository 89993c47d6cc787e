// Preconditioned-Cholesky TSQR ("PC" path): the fast path of every *_tc_* compute mode on tall matrices.
//
//   sample : R_pre = R factor of a Householder QR of h = 8 NP rows of A (h / 16 clusters of 16 rows, evenly spread)
//   pass 1 : Z = A R_pre^-1 (blocked forward substitution, never an explicit inverse), G = Z^T Z,
//            accumulated per CTA;  A is only read
//   finish : G = sum of the CTA partials (fp64), R_z = chol(G), certificate
//            est = |G|_F |G^-1|_F / n  (est = 1 for G ~ c I; n est bounds cond_2(G)),
//            R = R_z R_pre, C = R_z^-1
//   pass 2 : Z again (same code, same bits), Q = Z (C B)
// A tree node of the reference (QR of two stacked R factors, src/tsqr.cu:1115-1172) becomes a sum of
// Gram matrices: no levels, no sequential Householder chain per tile, 3 M N s bytes of HBM traffic.
// Z = A R_pre^-1 has cond(Z) ~ (1 + sqrt(NP/h)) / (1 - sqrt(NP/h)) for rows that are statistically alike,
// whatever cond(A) is, so chol(Z^T Z) is accurate; when the sample does not represent the matrix (graded or
// permuted rows, rank deficiency) the certificate fails ON THE DEVICE and the Householder passes
// (cs_reduce / tc_build), which are always enqueued behind it, do the work instead.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include "ts_pass.h"

namespace tsqr_b200 {

constexpr int kPcTileRows = 128;
constexpr long long kPcMinRows = 16384;
constexpr float kPcMaxEst = 6.0f;      // certificate threshold (see pc_finish_kernel)

inline int pc_sample_rows(int np) { return 8 * np; }

// float offsets of the PC region inside the workspace (all 64-float aligned)
struct PcLayout {
  int np = 0, grid = 0, parts = 0;   // CTAs of a pass (one per SM), fp64 Gram partials (3 per CTA)
  long long tiles = 0;
  int tiles_per_cta = 0;
  size_t rpre = 0;     // NP x NP fp32, ld NP
  size_t status = 0;   // int[16]: [0] status, [1] block counter, [2] est (float bits)
  size_t rout = 0;     // n x n fp32, ld n  (the R of this call, from whichever path produced it)
  size_t cmat = 0;     // NP x NP fp32, ld NP: C of pass 2
  size_t rzinv = 0;    // NP x NP fp32, ld NP: R_z^-1
  size_t cimg = 0;     // staging copy of the constant-memory image of R_pre's diagonal blocks
  size_t gpart = 0;    // [parts][NP][128] doubles
  size_t gsum = 0;     // [2 NP][NP] doubles
  size_t total = 0;
};
PcLayout pc_layout(long long m, int n);

struct PcCall {
  const void* A; long long lda; long long M; int n;
  void* Q; long long ldq;
  IoType io;
  bool cor;
  bool notc;           // *_notc modes: both passes on the CUDA cores in plain fp32 (pc_gram_notc_kernel / pc_apply_notc_kernel)
  float* ws;           // base of the PC region
  PcLayout L;
};
// enqueue sample + pass 1 + finish: afterwards ws + L.rout holds R (if status == kPcOk)
void pc_reduce(const PcCall& c, cudaStream_t s);
// enqueue (C = R_z^-1 B) + pass 2; both return at once on the device unless status == kPcOk
void pc_build(const PcCall& c, const float* B, long long ldb, cudaStream_t s);

// the stages one by one (row-sharded driver): sample | adopt an R_pre written into ws + L.rpre by somebody else |
// pass 1 | per-rank sums into ws + L.gsum ((2 NP) x NP doubles) | factor ws + L.gsum -> R, C, status. pc_build is pass 2.
void pc_stage_sample(const PcCall& c, cudaStream_t s);
void pc_stage_adopt_rpre(const PcCall& c, cudaStream_t s);
void pc_stage_pass1(const PcCall& c, cudaStream_t s);
void pc_stage_partial_sums(const PcCall& c, cudaStream_t s);
void pc_stage_factor(const PcCall& c, cudaStream_t s);

// tests: pass 1 with R_pre = rpre (gsum_out receives the 2 NP x NP fp64 sums: rows [0, NP) = Z_hi^T Z_hi,
// rows [NP, 2 NP) = Z_lo^T Z_hi), then pass 2 with C = cmat. All pointers are device memory.
void pc_probe(const PcCall& c, const float* rpre, const float* cmat, double* gsum_out, bool use_tma, cudaStream_t s);

}  // namespace tsqr_b200
