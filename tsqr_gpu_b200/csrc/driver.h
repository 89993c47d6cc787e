// Host driver: level planning, workspace carving and the TSQR / BlockQR call sequences.
// Everything here is stream-ordered; no host synchronisation happens inside a call
// (the reference synchronises 2L+3 times per 16-column panel, src/tsqr.cu:1112-1260).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <vector>
#include "ts_pass.h"
#include "pc_tsqr.h"

namespace tsqr_b200 {

// numeric enum shared by mtk::qr / mtk::tsqr / mtk::tcqr (reference src/tsqr.hpp:9-20)
enum Mode : int {
  fp16_notc = 0, fp16_tc_nocor = 1, fp32_notc = 2, fp32_tc_cor = 3, fp32_tc_nocor = 4,
  mixed_tc_cor_emu = 5, tf32_tc_cor = 6, tf32_tc_cor_emu = 7, tf32_tc_nocor = 8, tf32_tc_nocor_emu = 9,
  kNumModes = 10
};
inline bool mode_valid(int m) { return m >= 0 && m < kNumModes; }
inline IoType mode_io(int m) { return (m == fp16_notc || m == fp16_tc_nocor) ? IoType::f16 : IoType::f32; }
inline size_t io_bytes(IoType t) { return t == IoType::f16 ? 2 : 4; }

struct Level {
  long long rows;
  int nchunks;
  long long rows_per_chunk;
  int tiles_per_chunk;
  size_t ckpt_off;   // float offsets into the workspace
  size_t S_off;      // stacked R written by this level's reduce: (nchunks*n) x n
  size_t Qup_off;    // Q of the next level (same shape as S), the B source of this level
};

struct Plan {
  bool tc = false;  // tensor-core passes (TcParams chunk format) instead of the CUDA-core passes
  bool pc = false;  // preconditioned-Cholesky fast path first (pc_tsqr.h); the tensor-core Householder levels
                    // below are enqueued behind it and return at once unless its certificate failed
  size_t pc_off = 0;   // float offset of the PC region in the workspace
  PcLayout pcl;
  long long m;
  int n, np;
  std::vector<Level> levels;
  size_t total_floats;
};

int device_sm_count();

// Optional per-launch timing of the level passes (CUDA events on the launching stream, read
// back after the caller synchronised). label = level * 2 + (build ? 1 : 0) for the Householder levels,
// 100 = PC sample, 101 = PC pass 1, 102 = PC finish, 103 = PC pass 2 (with its C product).
struct ProfScope {
  cudaStream_t s; int label; cudaEvent_t beg = nullptr, end = nullptr; bool on;
  ProfScope(cudaStream_t st, int lb);
  ~ProfScope();
};
void profile_enable(bool on);
int profile_read(int* labels, float* ms, int cap);  // returns #records copied, clears them
// levels for an m x n TSQR (n <= kMaxN). max_chunks0 <= 0 picks 1 CTA wave per SM x occupancy.
Plan make_plan(long long m, int n, int max_chunks0 = 0, bool tc = false);
// the *_tc_* modes run on the tensor-core passes when the matrix is tall enough
inline bool mode_is_tc(int m) { return m == fp16_tc_nocor || m == fp32_tc_cor || m == fp32_tc_nocor || m == tf32_tc_cor || m == tf32_tc_nocor; }
// operand rounding emulated by the CUDA-core passes in the three _emu modes (reference src/matmul.hpp:26-209)
inline int mode_emu(int m) { return m == tf32_tc_nocor_emu ? 1 : (m == tf32_tc_cor_emu ? 2 : (m == mixed_tc_cor_emu ? 3 : 0)); }
inline bool mode_is_cor(int m) { return m == fp32_tc_cor || m == tf32_tc_cor; }
constexpr long long kTcMinRows = 1024;
inline bool use_tc(int mode, long long m) { return mode_is_tc(mode) && m >= kTcMinRows; }
bool pc_enabled();   // TSQR_B200_PC=0 switches the fast path off (Householder passes only)
bool pc_notc_enabled();   // TSQR_B200_PC_NOTC=0 keeps the *_notc modes on the Householder passes
// the two plain CUDA-core modes take the fast path with CUDA-core passes (the _emu modes keep their emulated rounding)
inline bool mode_is_notc_fast(int m) { return m == fp32_notc || m == fp16_notc; }
inline bool use_pc(int mode, long long m) {
  return m >= kPcMinRows && pc_enabled() && (mode_is_tc(mode) || (mode_is_notc_fast(mode) && pc_notc_enabled()));
}
Plan make_plan_for(int mode, long long m, int n);
// workspace floats that cover every mode's plan
size_t plan_floats_any(long long m, int n);
size_t tsqr_workspace_bytes(long long m, int n);

// A (m x n) -> R_local (n x n fp32, column-major ld n, upper triangular, inside the
// workspace at the returned pointer) and the checkpoints needed by tsqr_build.
// The tensor-core plan also writes the reflector tails into q (same shape as a; q may alias a).
const float* tsqr_reduce(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq, float* ws,
                         cudaStream_t s);
// Q = Q_local * B, with B (n x n fp32, column-major ldb) or identity when B == nullptr.
void tsqr_build(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq,
                const float* B, long long ldb, float* ws, cudaStream_t s);
// copy the upper triangle of an n x n fp32 matrix into the user's R (io type), zero below.
void store_r(const float* r32, int n, void* r, long long ldr, IoType t, cudaStream_t s);

// Full single-GPU TSQR, n <= kMaxN.
void tsqr_full(int mode, void* q, long long ldq, void* r, long long ldr, const void* a, long long lda,
               long long m, int n, float* ws, size_t ws_bytes, cudaStream_t s);

// Y <- Y - X (X^T Y) on the tensor cores (tc_project.cu); C receives X^T Y (kb x nb, ld kb), Cp holds
// 2 * nslabs partial products, `reduce` sums partial slots in fixed order.
void launch_tc_project(bool corrected, IoType io, const void* X, long long ldx, int kb, void* Y, long long ldy, int nb,
                       long long M, int nslabs, long long rows_per_slab, float* Cp, float* C,
                       void (*reduce)(const float*, int, size_t, float*, cudaStream_t), cudaStream_t s);

// BlockQR for any n <= m (panels of kMaxN columns); overwrites A when n > kMaxN
// (reference src/blockqr.cu:45-178, :180-390).
size_t qr_workspace_bytes(long long m, long long n, bool reorth);
int qr_full(int mode, bool reorth, void* q, long long ldq, void* r, long long ldr, void* a, long long lda,
            long long m, long long n, float* ws, size_t ws_bytes, cudaStream_t s);

// fp64 validation metrics on the device (validate.cu); `out` is one double in device memory
size_t validation_workspace_bytes(size_t n);
void orthogonality(IoType io, const void* q, long long ldq, long long m, int n, double* out, void* ws, cudaStream_t s);
void residual(IoType io, const void* a, long long lda, const void* q, long long ldq, const void* r, long long ldr, long long m, int n,
              double* out, void* ws, cudaStream_t s);
// conditioned test matrices on the device (latms.cu)
size_t latms_workspace_bytes(long long m, int n);
void latms(float* a, long long lda, long long m, int n, float cond, unsigned long long seed, float* ws, size_t ws_bytes, cudaStream_t s);

}  // namespace tsqr_b200
