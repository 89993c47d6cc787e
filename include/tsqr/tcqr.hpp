// tsqr_b200 -- drop-in replacement for the `mtk::tcqr` interface of enp1s0/tsqr-gpu
// (reference src/tcqr.hpp:1-41): QR of one (or a batch of) small row blocks,
// at most 32 rows x 16 columns each, column-major.
#pragma once
#include <cstddef>
#include <cuda_runtime_api.h>

namespace mtk {
namespace tcqr {

enum compute_mode {
	fp16_notc         = 0,
	fp16_tc_nocor     = 1,
	fp32_notc         = 2,
	fp32_tc_cor       = 3,
	fp32_tc_nocor     = 4,
	mixed_tc_cor_emu  = 5,
	tf32_tc_cor       = 6,
	tf32_tc_cor_emu   = 7,
	tf32_tc_nocor     = 8,
	tf32_tc_nocor_emu = 9,
};

// q (m x n), r (n x n) <- QR of a (m x n), m <= 32, n <= 16
template <compute_mode mode, class Q_T, class R_T, class A_T>
void qr32x16(
		Q_T* const q, const std::size_t ldq,
		R_T* const r, const std::size_t ldr,
		const A_T* const a, const std::size_t lda,
		const unsigned int m, const unsigned int n,
		cudaStream_t const cuda_stream = nullptr);

// block i = rows [a_start_position[i], a_start_position[i+1]) of a (m rows in total);
// its Q goes to the same rows of q, its R to rows [n*i, n*i+n) of r.
template <compute_mode mode, class Q_T, class R_T, class A_T>
void qr32x16_batched(
		Q_T *const q, const std::size_t ldq,
		R_T *const r, const std::size_t ldr,
		const A_T *const a, const std::size_t lda,
		const unsigned int m, const unsigned int n,
		const std::size_t batch_size, const unsigned* a_start_position,
		cudaStream_t const cuda_stream = nullptr);

} // namespace tcqr
} // namespace mtk
