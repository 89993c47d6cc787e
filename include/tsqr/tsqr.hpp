// tsqr_b200 -- drop-in replacement for the `mtk::tsqr` interface of enp1s0/tsqr-gpu
// (reference src/tsqr.hpp:1-140). Same namespace, enumerator values, traits, buffer
// fields and function signatures, so code written against the reference recompiles and
// links against libtsqr_b200 unchanged. The implementation behind it is new (sm_100a).
//
// Differences a caller can observe:
//  * tsqr16 accepts any n <= 64 (the reference is undefined for n > 16); `tsqr` is an
//    alias with the width-general name.
//  * the work buffers are sized by this library's own get_working_*_size(); only `dwq`
//    carries real workspace, the other pointers are kept for ABI compatibility.
//  * calls are stream-ordered and do not block the host; `hl` is never written.
#pragma once
#include <cstddef>
#include <stdexcept>
#include <cuda_fp16.h>
#include <cuda_runtime_api.h>

namespace mtk {
namespace tsqr {

// Enumerator values are part of the link contract (they appear in mangled names).
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

std::size_t get_batch_size_log2(const std::size_t m);
std::size_t get_batch_size(const std::size_t m);

namespace detail {
template <compute_mode mode> struct is_half_io    { static constexpr bool value = (mode == fp16_notc || mode == fp16_tc_nocor); };
template <compute_mode mode> struct is_half_workq { static constexpr bool value = is_half_io<mode>::value || mode == fp32_tc_nocor; };
template <bool use_half> struct pick_type        { using type = float; };
template <>              struct pick_type<true>  { using type = half; };
} // namespace detail

// element types of the user matrices and of the two work arrays, per mode
template <compute_mode mode> struct get_io_type        { using type = typename detail::pick_type<detail::is_half_io<mode>::value>::type; };
template <compute_mode mode> struct get_working_q_type { using type = typename detail::pick_type<detail::is_half_workq<mode>::value>::type; };
template <compute_mode mode> struct get_working_r_type { using type = typename detail::pick_type<detail::is_half_io<mode>::value>::type; };

// element counts of the work arrays (all real workspace lives in the `q` array)
std::size_t get_working_q_size(const std::size_t m, const std::size_t n);
std::size_t get_working_r_size(const std::size_t m, const std::size_t n);
inline std::size_t get_working_l_size(const std::size_t m) { return get_batch_size(m) + 1lu; }

template <compute_mode mode>
struct buffer {
	using wq_t = typename get_working_q_type<mode>::type;
	using wr_t = typename get_working_r_type<mode>::type;

	wq_t* dwq;
	wr_t* dwr;
	unsigned* dl;
	unsigned* hl;
	std::size_t total_memory_size;

	buffer() : dwq(nullptr), dwr(nullptr), dl(nullptr), hl(nullptr), total_memory_size(0lu) {}
	~buffer() { destroy(); }

	void allocate(const std::size_t m, const std::size_t n) { reserve(m, n, false); }
	void allocate_host(const std::size_t m, const std::size_t n) { reserve(m, n, true); }

	void destroy() {
		cudaFree(dwq); cudaFree(dwr); cudaFree(dl); cudaFreeHost(hl);
		dwq = nullptr; dwr = nullptr; dl = nullptr; hl = nullptr;
	}
	void destroy_host() {
		cudaFreeHost(dwq); cudaFreeHost(dwr); cudaFreeHost(dl); cudaFreeHost(hl);
		dwq = nullptr; dwr = nullptr; dl = nullptr; hl = nullptr;
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
		if (pinned_host) {
			cudaMallocHost(reinterpret_cast<void**>(&dwq), q_bytes);
			cudaMallocHost(reinterpret_cast<void**>(&dwr), r_bytes);
			cudaMallocHost(reinterpret_cast<void**>(&dl), l_bytes);
		} else {
			cudaMalloc(reinterpret_cast<void**>(&dwq), q_bytes);
			cudaMalloc(reinterpret_cast<void**>(&dwr), r_bytes);
			cudaMalloc(reinterpret_cast<void**>(&dl), l_bytes);
		}
		cudaMallocHost(reinterpret_cast<void**>(&hl), l_bytes);
		total_memory_size = q_bytes + r_bytes + l_bytes;
	}
};

// Q (m x n), R (n x n, upper triangle) <- QR of A (m x n); column-major, n <= 64.
template <compute_mode mode>
void tsqr16(
		typename get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
		typename get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
		const typename get_io_type<mode>::type* const a_ptr, const std::size_t lda,
		const std::size_t m,
		const std::size_t n,
		typename get_working_q_type<mode>::type* const working_q_ptr,
		typename get_working_r_type<mode>::type* const working_r_ptr,
		unsigned* const d_working_l_ptr,
		unsigned* const h_working_l_ptr,
		cudaStream_t const cuda_stream = nullptr);

template <compute_mode mode>
inline void tsqr16(
		typename get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
		typename get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
		const typename get_io_type<mode>::type* const a_ptr, const std::size_t lda,
		const std::size_t m, const std::size_t n,
		buffer<mode>& buffer,
		cudaStream_t const cuda_stream) {
	tsqr16<mode>(q_ptr, ldq, r_ptr, ldr, a_ptr, lda, m, n,
			buffer.dwq, buffer.dwr, buffer.dl, buffer.hl, cuda_stream);
}

// width-general name used by the north-star API description
template <compute_mode mode>
inline void tsqr(
		typename get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
		typename get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
		const typename get_io_type<mode>::type* const a_ptr, const std::size_t lda,
		const std::size_t m, const std::size_t n,
		buffer<mode>& buffer,
		cudaStream_t const cuda_stream = nullptr) {
	tsqr16<mode>(q_ptr, ldq, r_ptr, ldr, a_ptr, lda, m, n, buffer, cuda_stream);
}

} // namespace tsqr
} // namespace mtk
