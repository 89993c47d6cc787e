/* tsqr_b200 -- thin C ABI over the B200-native tall-skinny QR library.
 *
 * The reference (enp1s0/tsqr-gpu) has no C ABI or FFI: its boundary is the C++ template
 * API in include/tsqr/{tsqr,blockqr,tcqr}.hpp (reference src/tsqr.hpp:108-140,
 * src/blockqr.hpp:142-175, src/tcqr.hpp:22-37). Each entry point below names the reference
 * interface it stands in for. All matrices are column-major DEVICE memory (or pinned host
 * memory mapped into the device), leading dimensions are in elements. `mode` is the
 * numeric value of mtk::qr::compute_mode (reference src/tsqr.hpp:9-20):
 *   0 fp16_notc  1 fp16_tc_nocor  2 fp32_notc  3 fp32_tc_cor  4 fp32_tc_nocor
 *   5 mixed_tc_cor_emu  6 tf32_tc_cor  7 tf32_tc_cor_emu  8 tf32_tc_nocor  9 tf32_tc_nocor_emu
 * Modes 0 and 1 take/return IEEE half elements, all others float.
 *
 * Return codes: 0 success, 1 invalid matrix size (reference error_invalid_matrix_size),
 * negative = failure inside the library (see tsqr_b200_last_error()). Calls are ordered on
 * `stream` (a cudaStream_t passed as void*; NULL = default stream) and do not block the host.
 */
#ifndef TSQR_B200_H
#define TSQR_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TSQR_B200_OK 0
#define TSQR_B200_INVALID_SIZE 1
#define TSQR_B200_ERR_ARGUMENT (-1)
#define TSQR_B200_ERR_RUNTIME (-2)
#define TSQR_B200_MAX_TSQR_N 64

int tsqr_b200_version(void);
const char* tsqr_b200_last_error(void);
/* bytes per matrix element for `mode` (2 or 4); 0 if mode is invalid */
size_t tsqr_b200_io_bytes(int mode);

/* 1 when a TSQR of m rows in `mode` runs on the tcgen05 tensor-core passes (the *_tc_* modes of
 * reference src/tsqr.hpp:9-20 on tall enough matrices), 0 when it runs on the CUDA-core passes */
int tsqr_b200_mode_uses_tensor_cores(int mode, size_t m);

/* workspace size in bytes -- replaces mtk::qr::get_working_{q,r,l}_size
 * (reference src/blockqr.cu:34-42) and mtk::qr::buffer::allocate (src/blockqr.hpp:76-95). */
size_t tsqr_b200_qr_workspace_bytes(size_t m, size_t n, int reorthogonalize);
/* replaces mtk::tsqr::get_working_{q,r}_size (reference src/tsqr.cu:47-60) */
size_t tsqr_b200_tsqr_workspace_bytes(size_t m, size_t n);

/* QR of A (m x n, any n <= m): Q (m x n), R (n x n upper block-triangular part written).
 * A is overwritten when n > 64. Replaces mtk::qr::qr<mode, Reorthogonalize>
 * (reference src/blockqr.hpp:142-175, src/blockqr.cu:394-433). */
int tsqr_b200_qr(int mode, int reorthogonalize,
                 void* q, size_t ldq, void* r, size_t ldr, void* a, size_t lda,
                 size_t m, size_t n, void* workspace, size_t workspace_bytes, void* stream);

/* TSQR of A (m x n, n <= 64), A is not modified. Replaces mtk::tsqr::tsqr16<mode>
 * (reference src/tsqr.hpp:108-140, src/tsqr.cu:1281-1310). */
int tsqr_b200_tsqr(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                   size_t m, size_t n, void* workspace, size_t workspace_bytes, void* stream);

/* Row-sharded (multi-GPU) TSQR in three steps; the reference is single-GPU only, these
 * split tsqr16 (reference src/tsqr.cu:1064-1279) at its R-tree root:
 *   1. reduce : local slab A_p (m_p x n) -> R_p (n x n float, ld n, upper triangular)
 *   2. (caller) gather all R_p into an (P*n) x n float matrix, run tsqr_b200_tsqr on it
 *      (mode 2) to get the final R and the (P*n) x n root factor; rows [p*n, p*n+n) of
 *      the root factor are B_p
 *   3. build  : Q_p = Qlocal_p * B_p (B_p n x n float, ld ldb; NULL = identity)
 * The same workspace AND the same q (m_p x n, element type of `mode`) must be passed to reduce
 * and build, untouched in between: the tensor-core modes keep the slab's reflectors in q. */
int tsqr_b200_tsqr_reduce(int mode, const void* a, size_t lda, size_t m, size_t n,
                          float* r_local, void* q, size_t ldq, void* workspace, size_t workspace_bytes, void* stream);
int tsqr_b200_tsqr_build(int mode, void* q, size_t ldq, const void* a, size_t lda, size_t m, size_t n,
                         const float* b, size_t ldb, void* workspace, size_t workspace_bytes, void* stream);

/* Row-sharded multi-GPU TSQR inside the library (one process or thread per GPU; NCCL on the caller's stream; the
 * reference is single-GPU, this splits its tsqr16, reference src/tsqr.cu:1064-1279, across ranks). Rank p passes its
 * own slab a (m_local x n, n <= 64) and receives its slab of Q and the full n x n R (identical on every rank).
 *   tsqr_b200_nccl_unique_id : 128 bytes of ncclUniqueId, made on one rank and handed to the others by the caller
 *   tsqr_b200_sharded_init   : collective; creates the communicator (current device = this rank's GPU); NULL on failure
 *   tsqr_b200_sharded_attach : for callers that already own an ncclComm_t (passed as void*); not destroyed by _destroy
 *   tsqr_b200_tsqr_sharded   : collective, every rank with the same mode and n. Tall slabs (>= 16384 rows on every rank)
 *       in a *_tc_* mode: one ncclBroadcast of the 16 KB preconditioner + one ncclAllReduce of the 64 KB fp64 Gram
 *       sums, no root factorisation; otherwise one ncclAllGather of the R factors + a redundant root factorisation.
 *       Synchronises `stream` once per call (certificate read-back); NCCL is resolved at run time (libnccl.so.2). */
int tsqr_b200_nccl_unique_id(void* out128);
void* tsqr_b200_sharded_init(const void* id128, int nranks, int rank);
void* tsqr_b200_sharded_attach(void* nccl_comm, int nranks, int rank);
void tsqr_b200_sharded_destroy(void* handle);
size_t tsqr_b200_sharded_workspace_bytes(size_t m_local, size_t n, int nranks);
int tsqr_b200_tsqr_sharded(void* handle, int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                           size_t m_local, size_t n, void* workspace, size_t workspace_bytes, void* stream);

/* TSQR (n <= 64) of a matrix in HOST memory, overlapped with its own transfers (csrc/host_pipeline.cu). The reference's
 * drivers stage the whole matrix with blocking copies around mtk::qr::qr (reference src/test.cu:118-127, :141-145); here
 * the rows are cut into `nslabs` slabs, slab k is reduced while slab k+1 arrives and copied out while slab k+1 is built.
 *   create : allocates the device copies of A and Q, the per-slab workspaces, two copy streams and the events on the
 *            current device; NULL on failure (tsqr_b200_last_error)
 *   run    : a_host, q_host column-major m x n of the mode's element type (lda, ldq >= m), r_host n x n float (ldr >= n,
 *            the full R, zeros below the diagonal). Pinned host memory is needed for the overlap. Asynchronous: ordered
 *            after `stream`'s work, and `stream` waits for the last copy -- synchronise it before reading the host
 *            buffers. One run at a time per handle, on the device it was created on. */
void* tsqr_b200_host_pipeline_create(int mode, size_t m, size_t n, int nslabs);
int tsqr_b200_host_pipeline_slabs(void* handle);
int tsqr_b200_host_pipeline_run(void* handle, const void* a_host, size_t lda, void* q_host, size_t ldq, float* r_host,
                                size_t ldr, void* stream);
void tsqr_b200_host_pipeline_destroy(void* handle);

/* Batched QR of small row blocks (<= 32 rows x <= 16 columns each). Replaces
 * mtk::tcqr::qr32x16_batched / qr32x16 (reference src/tcqr.hpp:22-37,
 * src/tcqr32x16.cu:1584-1634). q/a element type follows `mode`; r is float unless the
 * mode is a half mode. a_start_position has batch_size+1 entries (device memory). */
int tsqr_b200_qr32x16_batched(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                              unsigned m, unsigned n, size_t batch_size, const unsigned* a_start_position,
                              void* stream);

/* Optional per-launch timing of the TSQR level passes with CUDA events on the caller's stream
 * (the analogue of the reference's MEASURE_QR_TIME / PROFILE_BREAKDOWN compile-time switches,
 * reference src/tsqr.cu:21-26, src/blockqr.cu:6-7). label = level * 2 + (1 for the Q-build
 * pass, 0 for the R-reduce pass). Read after synchronising the stream. */
void tsqr_b200_profile_enable(int on);
int tsqr_b200_profile_read(int* labels, float* ms, int capacity);

/* Known-answer self test of the tcgen05/TMEM operand layouts used by the tensor-core kernels
 * (no reference counterpart; the reference's analogue is src/wmma_extension/test/correctness).
 * A, B: 128 x 64 col-major fp32 (ld 128), X: 64 x 64 (ld 64); D1..D4: 128 x 64 outputs. */
int tsqr_b200_selftest_mma(int corrected, int variant, const float* A, const float* B, const float* X, float* D1,
                           float* D2, float* D3, float* D4, void* stream);

/* The *_tc_* modes factor matrices of >= 16384 rows with the preconditioned-Cholesky passes (csrc/pc_tsqr.h)
 * and fall back, on the device, to the Householder passes when the certificate of that path fails.
 * tsqr_b200_pc_status: which of the two produced the last result held in `workspace` (the one passed to
 * tsqr_b200_tsqr / tsqr_reduce for an m x n matrix in `mode`): 1 = fast path, 0 = Householder fallback,
 * -1 = this (mode, m) never takes the fast path. Blocks on `stream`. *est (optional) receives the
 * certificate value |G|_F |G^-1|_F / n. No reference counterpart. */
int tsqr_b200_pc_status(int mode, size_t m, size_t n, const void* workspace, void* stream, float* est);
/* Test probe of the two passes with caller-supplied factors (device pointers): pass 1 with R_pre = rpre
 * (NP x NP float, ld NP, NP = n rounded up to 16/32/64) -> gsum (2 NP x NP doubles, row-major: rows [0, NP) =
 * Z_hi^T Z_hi, rows [NP, 2 NP) = Z_lo^T Z_hi, Z = A rpre^-1); pass 2 with C = c -> q = Z c. */
int tsqr_b200_pc_probe(int mode, const void* a, size_t lda, size_t m, size_t n, const float* rpre, const float* c,
                       void* q, size_t ldq, double* gsum, int use_tma, void* workspace, size_t workspace_bytes,
                       void* stream);

/* Validation metrics in fp64 ON THE DEVICE, any n <= m; element type of the matrices follows `mode`. `result` is one
 * double in device memory, `workspace` at least tsqr_b200_validation_workspace_bytes(n) bytes of device memory.
 *   orthogonality = |Q^T Q - I|_F / sqrt(n)   replaces mtk::validation::check_orthogonality16 (reference src/validation.cu:43-80)
 *   residual      = |A - Q R|_F / |A|_F       replaces the residual of the reference's test driver (reference src/test.cu:146-165);
 *                   only the upper triangle of r is read */
size_t tsqr_b200_validation_workspace_bytes(size_t n);
int tsqr_b200_orthogonality(int mode, const void* q, size_t ldq, size_t m, size_t n, double* result, void* workspace,
                            size_t workspace_bytes, void* stream);
int tsqr_b200_residual(int mode, const void* a, size_t lda, const void* q, size_t ldq, const void* r, size_t ldr, size_t m, size_t n,
                       double* result, void* workspace, size_t workspace_bytes, void* stream);

/* Conditioned test matrix on the device: a (m x n float, any n <= m) = U diag(s) V^T with U, V orthogonal factors of this
 * library's QR of counter-based normals (a pure function of `seed`) and s_k = cond^(-k/(n-1)). Replaces
 * mtk::utils::latms (reference src/latms.cu:9-121; singular values of src/test_cond.cu:31-50) without cuRAND/cuSOLVER. */
size_t tsqr_b200_latms_workspace_bytes(size_t m, size_t n);
int tsqr_b200_latms(float* a, size_t lda, size_t m, size_t n, float cond, unsigned long long seed, void* workspace,
                    size_t workspace_bytes, void* stream);

/* TSQR of A (m x n, n <= 64) for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). No counterpart in the reference.
 * Same result as tsqr_b200_tsqr, but the workspace is a file-scope static __device__ array of
 * TSQR_B200_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until the process
 * exits); the call allocates and frees nothing. A (mode, m, n) whose level plan needs more than that is rejected
 * with TSQR_B200_ERR_ARGUMENT (tsqr_b200_last_error() names both numbers): 64 MiB admits every *_tc_* mode at any m
 * (their workspace is bounded by the SM count: 40 MB at n = 64) and the CUDA-core modes (*_notc, *_emu; one R checkpoint
 * per 128-row tile) up to about 2^18 rows at n = 64, 2^24 rows at n = 16. ALL calls on a device share the one array: calls that overlap on the GPU (different streams, different
 * threads) race -- serialise them. The array keeps the previous call's contents; nothing in it needs clearing. */
#define TSQR_B200_STATIC_SCRATCH_BYTES (64u << 20)
int tsqr_b200_tsqr_static(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                          size_t m, size_t n, void* stream);

/* how many kernels this library has launched in this process (bench bookkeeping) */
unsigned long long tsqr_b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif
