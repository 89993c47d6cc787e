// tsqr_b200 -- drop-in replacement for the `mtk::qr` interface of enp1s0/tsqr-gpu
// (reference src/blockqr.hpp:1-178): QR of an m x n column-major matrix, any n <= m.
// n <= 64 runs as a single native TSQR (A untouched); wider matrices run as a block
// algorithm over 64-column panels and overwrite A, as the reference does for n > 16.
//
// The cuBLAS handle is only a stream carrier here: the library never calls a cuBLAS
// compute routine. A null handle means the default stream.
#pragma once
#include <cstddef>
#include <cstdlib>
#include <stdexcept>
#include <cublas_v2.h>
#include "tsqr.hpp"

namespace mtk {
namespace qr {

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

// kept (spelling included) because user code may name it; the native panel is 64 wide
constexpr std::size_t tsqr_colmun_size = 16;

using state_t = int;
const state_t success_factorization     = 0;
const state_t error_invalid_matrix_size = 1;

// the two enums are value-identical by construction
template <mtk::qr::compute_mode mode>
constexpr mtk::tsqr::compute_mode get_tsqr_compute_mode() {
	return static_cast<mtk::tsqr::compute_mode>(static_cast<int>(mode));
}

template <mtk::qr::compute_mode mode>
struct get_working_q_type { using type = typename mtk::tsqr::get_working_q_type<get_tsqr_compute_mode<mode>()>::type; };
template <mtk::qr::compute_mode mode>
struct get_working_r_type { using type = typename mtk::tsqr::get_working_r_type<get_tsqr_compute_mode<mode>()>::type; };
template <mtk::qr::compute_mode mode>
struct get_io_type { using type = typename mtk::tsqr::get_io_type<get_tsqr_compute_mode<mode>()>::type; };

std::size_t get_working_q_size(const std::size_t m, const std::size_t n);
std::size_t get_working_r_size(const std::size_t m, const std::size_t n);
std::size_t get_working_l_size(const std::size_t m);

template <mtk::qr::compute_mode mode, bool Reorthogonalize>
struct buffer {
	using wq_t = typename get_working_q_type<mode>::type;
	using wr_t = typename get_working_r_type<mode>::type;
	using io_t = typename get_io_type<mode>::type;

	wq_t* dwq;
	wr_t* dwr;
	io_t* dw_reorth_r;
	unsigned* dl;
	unsigned* hl;
	std::size_t total_memory_size;

	buffer() : dwq(nullptr), dwr(nullptr), dw_reorth_r(nullptr), dl(nullptr), hl(nullptr), total_memory_size(0lu) {}
	~buffer() { destroy(); }

	void allocate(const std::size_t m, const std::size_t n) { reserve(m, n, false); }
	void allocate_host(const std::size_t m, const std::size_t n) { reserve(m, n, true); }

	void destroy() {
		cudaFree(dwq); cudaFree(dwr); cudaFree(dw_reorth_r); cudaFree(dl); cudaFreeHost(hl);
		dwq = nullptr; dwr = nullptr; dw_reorth_r = nullptr; dl = nullptr; hl = nullptr;
	}
	void destroy_host() {
		cudaFreeHost(dwq); cudaFreeHost(dwr); cudaFreeHost(dw_reorth_r); cudaFreeHost(dl); cudaFreeHost(hl);
		dwq = nullptr; dwr = nullptr; dw_reorth_r = nullptr; dl = nullptr; hl = nullptr;
	}
	std::size_t get_device_memory_size() const { return total_memory_size; }

private:
	void reserve(const std::size_t m, const std::size_t n, const bool pinned_host) {
		if (dwq != nullptr || dwr != nullptr || dl != nullptr || hl != nullptr) {
			throw std::runtime_error("The buffer has been already allocated");
		}
		const std::size_t q_bytes = sizeof(wq_t) * get_working_q_size(m, n);
		const std::size_t r_bytes = sizeof(wr_t) * get_working_r_size(m, n);
		const std::size_t l_bytes = sizeof(unsigned) * get_working_l_size(m);
		// same element count as the reference (src/blockqr.hpp:91); unused by this library
		const std::size_t o_bytes = Reorthogonalize
			? sizeof(io_t) * (tsqr_colmun_size * tsqr_colmun_size * 2 + m * tsqr_colmun_size) : 0lu;
		void** const slots[3] = {reinterpret_cast<void**>(&dwq), reinterpret_cast<void**>(&dwr), reinterpret_cast<void**>(&dl)};
		const std::size_t bytes[3] = {q_bytes, r_bytes, l_bytes};
		for (int i = 0; i < 3; i++) {
			if (pinned_host) cudaMallocHost(slots[i], bytes[i]);
			else             cudaMalloc(slots[i], bytes[i]);
		}
		if (Reorthogonalize) {
			if (pinned_host) cudaMallocHost(reinterpret_cast<void**>(&dw_reorth_r), o_bytes);
			else             cudaMalloc(reinterpret_cast<void**>(&dw_reorth_r), o_bytes);
		}
		cudaMallocHost(reinterpret_cast<void**>(&hl), l_bytes);
		total_memory_size = q_bytes + r_bytes + l_bytes + o_bytes;
	}
};

// returns success_factorization, or error_invalid_matrix_size when n > m, m == 0 or n == 0
template <mtk::qr::compute_mode mode, bool Reorthogonalize>
state_t qr(
		typename mtk::qr::get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
		typename mtk::qr::get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
		typename mtk::qr::get_io_type<mode>::type* const a_ptr, const std::size_t lda,
		const std::size_t m, const std::size_t n,
		typename mtk::qr::get_working_q_type<mode>::type* const wq_ptr,
		typename mtk::qr::get_working_r_type<mode>::type* const wr_ptr,
		typename mtk::qr::get_io_type<mode>::type* const reorth_r_ptr,
		unsigned* const d_wl_ptr,
		unsigned* const h_wl_ptr,
		cublasHandle_t const main_cublas_handle);

template <mtk::qr::compute_mode mode, bool Reorthogonalize>
inline state_t qr(
		typename mtk::qr::get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
		typename mtk::qr::get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
		typename mtk::qr::get_io_type<mode>::type* const a_ptr, const std::size_t lda,
		const std::size_t m, const std::size_t n,
		buffer<mode, Reorthogonalize>& bf,
		cublasHandle_t const main_cublas_handle) {
	return qr<mode, Reorthogonalize>(q_ptr, ldq, r_ptr, ldr, a_ptr, lda, m, n,
			bf.dwq, bf.dwr, bf.dw_reorth_r, bf.dl, bf.hl, main_cublas_handle);
}

} // namespace qr
} // namespace mtk
