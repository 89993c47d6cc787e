// Level-pass description shared by the host driver and the TSQR kernels.
//
// One "level" of the TSQR is a tall matrix (column-major, rows x n) that is cut into
// `nchunks` contiguous row chunks, one CTA per chunk. A chunk is walked tile by tile
// (tile = 128*RPT rows) as a flat tree: the running N x N factor R is stacked on top of
// each tile and the stacked matrix [R; A_tile] is reduced by a blocked Householder QR.
//
//   reduce pass : A -> one R per chunk (written into the stacked matrix S of the next
//                 level) plus one packed checkpoint of R per tile (the running R *before*
//                 that tile), so that the build pass can redo any tile bit for bit.
//   build pass  : walks the tiles of a chunk in reverse, refactors [R_ckpt; A_tile] to get
//                 the reflectors (Y, T) again and emits Q_tile = -Y (T B), B <- (I - T) B,
//                 where B (n x n) starts as this chunk's block of the parent level's Q.
//
// This replaces reference src/tsqr.cu:1064-1279 (leaf batch + binary R tree + backward
// products, 2L+1 launches) and src/tcqr32x16.cu:1373-1469 (explicit-H leaf QR).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace tsqr_b200 {

constexpr int kPanel = 16;        // Householder panel width (columns held in registers)
constexpr int kPassThreads = 128; // threads per CTA, one (or RPT) matrix rows per thread
constexpr int kMaxN = 64;         // widest matrix the TSQR kernels take natively

enum class IoType : int { f32 = 0, f16 = 1 };

struct PassParams {
  const void* A;        // level input, column-major
  long long lda;
  long long M;          // rows of the level input
  int n;                // true column count (<= kMaxN)
  int np;               // n rounded up to a multiple of kPanel
  int nchunks;
  long long rows_per_chunk;  // multiple of the tile height
  int tiles_per_chunk;       // ceil(rows_per_chunk / tile height)
  float* ckpt;               // [nchunks][tiles_per_chunk][np*(np+1)/2] packed upper R
  // reduce only
  float* S;             // stacked R output, (nchunks*n) x n column-major
  long long lds;
  // build only
  const float* Bsrc;    // parent Q: chunk c uses rows [c*n, c*n+n); nullptr => identity
  long long ldb;
  void* Q;              // level output, column-major, same row count as A
  long long ldq;
  int emu;              // 0: fp32 products; 1 / 2 / 3: tf32_tc_nocor_emu / tf32_tc_cor_emu / mixed_tc_cor_emu operand rounding
  const int* skip_ok;   // when set: the kernel returns at once if *skip_ok == kPcOk (the PC path did the work)
};

// rows-per-thread used for a given padded width
inline int pass_rpt(int np) { return np <= 16 ? 4 : (np <= 32 ? 2 : 1); }
inline int pass_tile_rows(int np) { return kPassThreads * pass_rpt(np); }
inline size_t ckpt_floats(int np) { return (size_t)np * (np + 1) / 2; }

// SIMT (CUDA-core fp32) implementation; io may be fp32 or fp16, arithmetic is fp32.
void launch_pass_simt(const PassParams& p, bool build, IoType in, IoType out, cudaStream_t s);
int pass_simt_max_ctas_per_sm(int np, bool build);


// ---------------------------------------------------------------------------------------------
// Tensor-core (tcgen05 / TMEM) level passes, used by every *_tc_* compute mode.
//
// A chunk of rows [cbeg, cend) is: a HEAD block (its first NP rows, plain Householder QR) that
// seeds the running R, followed by 128-row TILES reduced one after the other as [R ; A_tile]
// with structured reflectors [e_j ; v_j] (flat tree).
//   reduce: A -> R per chunk (stacked into S) and the reflector tails V, written into the Q
//           buffer at the rows they belong to (head: strictly lower part of V0).
//   build : walks a chunk's tiles in reverse; per tile S = V^T V (tensor cores), solves
//           T^-1 X = B with T^-1 = striu(S) + diag((1+S_jj)/2), Q_tile = -V X (tensor cores),
//           B <- B - X; finally Q_head = H_0 .. H_{NP-1} B. Q overwrites V in place.
struct TcParams {
  const void* A;
  long long lda;
  long long M;               // rows of the level input
  int n;                     // true column count (<= NP)
  int nchunks;
  long long rows_per_chunk;  // the last chunk takes every remaining row
  void* Q;                   // V (reduce: out, build: in) then Q (build: out)
  long long ldq;
  float* S;                  // reduce: stacked R, (nchunks*n) x n column-major
  long long lds;
  const float* Bsrc;         // build: parent Q, chunk c uses rows [c*n, c*n+n); nullptr => identity
  long long ldb;
  const int* skip_ok;        // when set: the kernel returns at once if *skip_ok == kPcOk (the PC path did the work)
};
constexpr int kPcOk = 0x600D;    // status word of the preconditioned-Cholesky path: certificate passed
constexpr int kTcTileRows = 128;
inline int tc_np(int n) { return n <= 16 ? 16 : (n <= 32 ? 32 : 64); }
// reduce pass of the Householder tensor-core path: CUDA-core sweep in the (column, row-segment) layout
void launch_cs_reduce(const TcParams& p, IoType in, IoType q, cudaStream_t s);
int cs_reduce_ctas_per_sm(int np);
void launch_tc_build(const TcParams& p, bool corrected, IoType q, cudaStream_t s);

}  // namespace tsqr_b200
