// ORACLE -- TEST INFRASTRUCTURE ONLY. C entry points around the UNMODIFIED reference
// (enp1s0/tsqr-gpu, compiled from /root/reference/src where it lies, for sm_100) so that the
// GPU parity tests and bench.py --impl reference can call mtk::qr::qr<mode, reorth> of the
// reference on the same device buffers as the new library. Built by oracle/build_ref.sh into
// oracle/_ref/libref_tcqr.so; never linked into the product.
#include <stdexcept>     // the reference's headers use std::runtime_error without including this
#include <cstdio>
#include <cublas_v2.h>
#include <blockqr.hpp>   // the reference's header (-I /root/reference/src)

namespace {
cublasHandle_t g_handle = nullptr;

template <mtk::qr::compute_mode mode, bool reorth>
int run(void* q, size_t ldq, void* r, size_t ldr, void* a, size_t lda, size_t m, size_t n, cudaStream_t stream) {
  using io_t = typename mtk::qr::get_io_type<mode>::type;
  static mtk::qr::buffer<mode, reorth> buf;   // one per instantiation, re-allocated when the shape changes
  static size_t cur_m = 0, cur_n = 0;
  if (cur_m != m || cur_n != n) {
    if (cur_m != 0) buf.destroy();
    buf.allocate(m, n);
    cur_m = m; cur_n = n;
  }
  if (!g_handle) cublasCreate(&g_handle);
  cublasSetStream(g_handle, stream);
  return mtk::qr::qr<mode, reorth>((io_t*)q, ldq, (io_t*)r, ldr, (io_t*)a, lda, m, n, buf, g_handle);
}

template <bool reorth>
int dispatch(int mode, void* q, size_t ldq, void* r, size_t ldr, void* a, size_t lda, size_t m, size_t n, cudaStream_t s) {
  using namespace mtk::qr;
  switch (mode) {
    case 0: return run<fp16_notc, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 1: return run<fp16_tc_nocor, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 2: return run<fp32_notc, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 3: return run<fp32_tc_cor, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 4: return run<fp32_tc_nocor, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 5: return run<mixed_tc_cor_emu, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 6: return run<tf32_tc_cor, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 7: return run<tf32_tc_cor_emu, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 8: return run<tf32_tc_nocor, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
    case 9: return run<tf32_tc_nocor_emu, reorth>(q, ldq, r, ldr, a, lda, m, n, s);
  }
  return -1;
}
}  // namespace

extern "C" int ref_gpu_qr(int mode, int reorth, void* q, size_t ldq, void* r, size_t ldr, void* a, size_t lda,
                          size_t m, size_t n, void* stream) {
  try {
    return reorth ? dispatch<true>(mode, q, ldq, r, ldr, a, lda, m, n, (cudaStream_t)stream)
                  : dispatch<false>(mode, q, ldq, r, ldr, a, lda, m, n, (cudaStream_t)stream);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "ref_gpu_qr: %s\n", e.what());
    return -2;
  }
}
