// Public surface of libtsqr_b200: the C ABI (include/tsqr_b200.h) and the explicit
// instantiations of the mtk:: templates the reference's static library exports
// (reference src/blockqr.cu:436-456, src/tsqr.cu:1313-1323, src/tcqr32x16.cu:1608-1646).
#include "../../include/tsqr_b200.h"
#include "../../include/tsqr/blockqr.hpp"
#include "../../include/tsqr/tcqr.hpp"
#include "common.cuh"
#include "driver.h"
#include "small_qr.h"
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <mutex>

namespace tsqr_b200 {
static std::atomic<unsigned long long> g_launches{0};
void note_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static thread_local char g_err[512] = "";
static void set_err(const char* msg) {
  std::strncpy(g_err, msg, sizeof(g_err) - 1);
  g_err[sizeof(g_err) - 1] = 0;
}

void set_error_message(const char* msg) { set_err(msg); }

// The reference takes its stream from the cuBLAS handle (src/blockqr.cu:58-59). cuBLAS is
// not linked: cublasGetStream_v2 is looked up at run time, in the process first (the
// caller that owns a handle has the library loaded), then by soname.
static cudaStream_t stream_of(cublasHandle_t h) {
  if (h == nullptr) return nullptr;
  using fn_t = int (*)(cublasHandle_t, cudaStream_t*);
  static fn_t fn = [] {
    void* sym = dlsym(RTLD_DEFAULT, "cublasGetStream_v2");
    if (!sym) {
      void* lib = dlopen("libcublas.so.12", RTLD_NOW | RTLD_GLOBAL);
      if (!lib) lib = dlopen("libcublas.so", RTLD_NOW | RTLD_GLOBAL);
      if (lib) sym = dlsym(lib, "cublasGetStream_v2");
    }
    return reinterpret_cast<fn_t>(sym);
  }();
  if (!fn) throw std::runtime_error("tsqr_b200: a cublasHandle_t was given but cublasGetStream_v2 cannot be resolved");
  cudaStream_t s = nullptr;
  if (fn(h, &s) != 0) throw std::runtime_error("tsqr_b200: cublasGetStream failed");
  return s;
}

template <typename F>
static int guarded(F&& f) {
  try {
    return f();
  } catch (const std::exception& e) {
    set_err(e.what());
    return TSQR_B200_ERR_RUNTIME;
  } catch (...) {
    set_err("unknown exception");
    return TSQR_B200_ERR_RUNTIME;
  }
}
}  // namespace tsqr_b200

using namespace tsqr_b200;

// ------------------------------------------------------------------ C ABI
extern "C" {

int tsqr_b200_version(void) { return 100; }
const char* tsqr_b200_last_error(void) { return g_err; }
size_t tsqr_b200_io_bytes(int mode) { return mode_valid(mode) ? io_bytes(mode_io(mode)) : 0; }
unsigned long long tsqr_b200_launch_count(void) { return g_launches.load(); }
void tsqr_b200_profile_enable(int on) { profile_enable(on != 0); }
int tsqr_b200_profile_read(int* labels, float* ms, int cap) { return profile_read(labels, ms, cap); }

int tsqr_b200_mode_uses_tensor_cores(int mode, size_t m) { return mode_valid(mode) && use_tc(mode, (long long)m) ? 1 : 0; }

size_t tsqr_b200_qr_workspace_bytes(size_t m, size_t n, int reorth) {
  if (m == 0 || n == 0 || n > m) return 256;
  return qr_workspace_bytes((long long)m, (long long)n, reorth != 0);
}
size_t tsqr_b200_tsqr_workspace_bytes(size_t m, size_t n) {
  if (m == 0 || n == 0 || n > (size_t)kMaxN) return 256;
  return tsqr_workspace_bytes((long long)m, (int)n);
}

int tsqr_b200_qr(int mode, int reorth, void* q, size_t ldq, void* r, size_t ldr, void* a, size_t lda, size_t m,
                 size_t n, void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (n > m || m == 0 || n == 0) return TSQR_B200_INVALID_SIZE;
  if (!q || !r || !a || !ws) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    return qr_full(mode, reorth != 0, q, (long long)ldq, r, (long long)ldr, a, (long long)lda, (long long)m,
                   (long long)n, (float*)ws, ws_bytes, (cudaStream_t)stream);
  });
}

int tsqr_b200_tsqr(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda, size_t m, size_t n,
                   void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > (size_t)kMaxN) return TSQR_B200_INVALID_SIZE;
  if (!q || !r || !a || !ws) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    tsqr_full(mode, q, (long long)ldq, r, (long long)ldr, a, (long long)lda, (long long)m, (int)n, (float*)ws,
              ws_bytes, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_tsqr_reduce(int mode, const void* a, size_t lda, size_t m, size_t n, float* r_local, void* q, size_t ldq,
                          void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > (size_t)kMaxN) return TSQR_B200_INVALID_SIZE;
  if (!a || !ws || !r_local || !q) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    Plan plan = make_plan_for(mode, (long long)m, (int)n);
    if (plan.total_floats * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
    const float* r32 = tsqr_reduce(plan, mode, a, (long long)lda, q, (long long)ldq, (float*)ws, (cudaStream_t)stream);
    store_r(r32, (int)n, r_local, (long long)n, IoType::f32, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_tsqr_build(int mode, void* q, size_t ldq, const void* a, size_t lda, size_t m, size_t n, const float* b,
                         size_t ldb, void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > (size_t)kMaxN) return TSQR_B200_INVALID_SIZE;
  if (!a || !ws || !q) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    Plan plan = make_plan_for(mode, (long long)m, (int)n);
    if (plan.total_floats * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
    tsqr_build(plan, mode, a, (long long)lda, q, (long long)ldq, b, (long long)ldb, (float*)ws, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_pc_probe(int mode, const void* a, size_t lda, size_t m, size_t n, const float* rpre, const float* c,
                       void* q, size_t ldq, double* gsum, int use_tma, void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode) || !(mode_is_tc(mode) || mode_is_notc_fast(mode))) { set_err("the probe takes a *_tc_* or *_notc mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > (size_t)kMaxN) return TSQR_B200_INVALID_SIZE;
  return guarded([&] {
    PcCall call{};
    call.A = a; call.lda = (long long)lda; call.M = (long long)m; call.n = (int)n;
    call.Q = q; call.ldq = (long long)ldq;
    call.io = mode_io(mode); call.cor = mode_is_cor(mode); call.notc = !mode_is_tc(mode);
    call.ws = (float*)ws; call.L = pc_layout((long long)m, (int)n);
    if (call.L.total * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
    pc_probe(call, rpre, c, gsum, use_tma != 0, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_pc_status(int mode, size_t m, size_t n, const void* ws, void* stream, float* est) {
  if (!mode_valid(mode) || m == 0 || n == 0 || n > (size_t)kMaxN) return -1;
  if (!use_pc(mode, (long long)m)) return -1;
  return guarded([&] {
    Plan plan = make_plan_for(mode, (long long)m, (int)n);
    int st[16] = {0};
    TSQR_CUDA_CHECK(cudaMemcpyAsync(st, (const float*)ws + plan.pc_off + plan.pcl.status, sizeof(st), cudaMemcpyDeviceToHost,
                                    (cudaStream_t)stream));
    TSQR_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    if (est) std::memcpy(est, &st[2], sizeof(float));
    if (std::getenv("TSQR_B200_PC_PHASES"))   // SM-clock cycles of the single-CTA phases (tuning aid)
      std::fprintf(stderr, "tsqr_b200 pc phases: sample gather %d sweep %d | finish chol %d inverse %d rest %d (copies %d certificate %d)\n",
                   st[8], st[9], st[4], st[5], st[6], st[10], st[11]);
    return st[0] == kPcOk ? 1 : 0;
  });
}

size_t tsqr_b200_validation_workspace_bytes(size_t n) { return validation_workspace_bytes(n); }

int tsqr_b200_orthogonality(int mode, const void* q, size_t ldq, size_t m, size_t n, double* result, void* ws, size_t ws_bytes,
                            void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > m) return TSQR_B200_INVALID_SIZE;
  if (!q || !result || !ws || ws_bytes < validation_workspace_bytes(n)) { set_err("null pointer or workspace too small"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    orthogonality(mode_io(mode), q, (long long)ldq, (long long)m, (int)n, result, ws, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_residual(int mode, const void* a, size_t lda, const void* q, size_t ldq, const void* r, size_t ldr, size_t m, size_t n,
                       double* result, void* ws, size_t ws_bytes, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > m) return TSQR_B200_INVALID_SIZE;
  if (!a || !q || !r || !result || !ws || ws_bytes < 64) { set_err("null pointer or workspace too small"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    residual(mode_io(mode), a, (long long)lda, q, (long long)ldq, r, (long long)ldr, (long long)m, (int)n, result, ws, (cudaStream_t)stream);
    return 0;
  });
}

size_t tsqr_b200_latms_workspace_bytes(size_t m, size_t n) {
  if (m == 0 || n == 0 || n > m) return 256;
  return latms_workspace_bytes((long long)m, (int)n);
}

int tsqr_b200_latms(float* a, size_t lda, size_t m, size_t n, float cond, unsigned long long seed, void* ws, size_t ws_bytes,
                    void* stream) {
  if (m == 0 || n == 0 || n > m) return TSQR_B200_INVALID_SIZE;
  if (!a || !ws) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  return guarded([&] {
    latms(a, (long long)lda, (long long)m, (int)n, cond, seed, (float*)ws, ws_bytes, (cudaStream_t)stream);
    return 0;
  });
}

int tsqr_b200_qr32x16_batched(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                              unsigned m, unsigned n, size_t batch, const unsigned* starts, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (n == 0 || n > 16 || m == 0) return TSQR_B200_INVALID_SIZE;
  return guarded([&] {
    if (mode_io(mode) == IoType::f16)
      small_qr_launch<__half, __half, __half>((__half*)q, ldq, (__half*)r, ldr, (const __half*)a, lda, m, n, batch,
                                              starts, (cudaStream_t)stream);
    else
      small_qr_launch<float, float, float>((float*)q, ldq, (float*)r, ldr, (const float*)a, lda, m, n, batch, starts,
                                           (cudaStream_t)stream);
    return 0;
  });
}

}  // extern "C"

// ------------------------------------------------------------------ TSQR without any allocation
// Workspace of tsqr_b200_tsqr_static: a file-scope array in device memory. Every GPU that loads the library holds
// these TSQR_B200_STATIC_SCRATCH_BYTES until the process exits. In HOST code the name below is only a stand-in for
// the symbol, not a device address: the address is looked up with cudaGetSymbolAddress once per device (with that
// device current) and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_static_scratch[TSQR_B200_STATIC_SCRATCH_BYTES];

namespace {
float* static_scratch_address() {
  static std::mutex mu;
  static void* ptrs[64] = {};
  int dev = 0;
  TSQR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) throw std::runtime_error("tsqr_b200: device index out of range for the static scratch table");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    TSQR_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) throw std::runtime_error("tsqr_b200: bad static scratch address");
    ptrs[dev] = p;
  }
  return reinterpret_cast<float*>(ptrs[dev]);
}
}  // namespace

extern "C" int tsqr_b200_tsqr_static(int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda, size_t m,
                                     size_t n, void* stream) {
  if (!mode_valid(mode)) { set_err("invalid compute mode"); return TSQR_B200_ERR_ARGUMENT; }
  if (m == 0 || n == 0 || n > (size_t)kMaxN || n > m) return TSQR_B200_INVALID_SIZE;
  if (!q || !r || !a) { set_err("null pointer argument"); return TSQR_B200_ERR_ARGUMENT; }
  const size_t need = make_plan_for(mode, (long long)m, (int)n).total_floats * sizeof(float);   // this mode's plan only
  if (need > (size_t)TSQR_B200_STATIC_SCRATCH_BYTES) {
    char msg[256];
    std::snprintf(msg, sizeof(msg), "%zu x %zu in mode %d needs %zu bytes of workspace, the static scratch array holds %zu: use "
                  "tsqr_b200_tsqr with a caller-provided workspace", m, n, mode, need, (size_t)TSQR_B200_STATIC_SCRATCH_BYTES);
    set_err(msg);
    return TSQR_B200_ERR_ARGUMENT;
  }
  return guarded([&] {
    // Nothing in the workspace has to start at zero: the first kernel of a call resets the status words it uses and
    // every other word is written before it is read, so the previous call's contents are harmless.
    tsqr_full(mode, q, (long long)ldq, r, (long long)ldr, a, (long long)lda, (long long)m, (int)n, static_scratch_address(),
              (size_t)TSQR_B200_STATIC_SCRATCH_BYTES, (cudaStream_t)stream);
    return 0;
  });
}

// ------------------------------------------------------------------ mtk:: C++ API
namespace {
// The reference's entries return after a stream synchronisation (src/blockqr.cu:140, src/tsqr.cu:1260) and its own
// driver times them with a wall clock and no sync (src/test.cu:299-309). The mtk:: entries keep that contract;
// TSQR_B200_ASYNC=1 opts out (the C ABI is always asynchronous).
bool cxx_blocking() {
  static const bool b = [] { const char* e = std::getenv("TSQR_B200_ASYNC"); return !(e && e[0] == '1'); }();
  return b;
}
// All workspace lives in the `wq` array. Its element count is sized for 2-byte elements
// (see get_working_q_size), so the byte capacity is at least count * 2 for every mode.
size_t wq_capacity_bytes_qr(size_t m, size_t n) { return mtk::qr::get_working_q_size(m, n) * 2; }
size_t wq_capacity_bytes_tsqr(size_t m, size_t n) { return mtk::tsqr::get_working_q_size(m, n) * 2; }
}  // namespace

// batch sizes are kept with the reference's meaning (leaves of <= 32 rows,
// reference src/tsqr.cu:39-44) because get_working_l_size() is defined through them.
std::size_t mtk::tsqr::get_batch_size_log2(const std::size_t m) {
  std::size_t lg = 0;
  while ((std::size_t(1) << lg) < m) ++lg;
  return lg > 5 ? lg - 5 : 0;
}
std::size_t mtk::tsqr::get_batch_size(const std::size_t m) { return std::size_t(1) << get_batch_size_log2(m); }

std::size_t mtk::tsqr::get_working_q_size(const std::size_t m, const std::size_t n) {
  return (tsqr_b200_tsqr_workspace_bytes(m, n) + 1) / 2;
}
std::size_t mtk::tsqr::get_working_r_size(const std::size_t, const std::size_t n) { return n * n + 16; }

std::size_t mtk::qr::get_working_q_size(const std::size_t m, const std::size_t n) {
  return (tsqr_b200_qr_workspace_bytes(m, n, 1) + 1) / 2;
}
std::size_t mtk::qr::get_working_r_size(const std::size_t, const std::size_t n) {
  const std::size_t w = n < 64 ? n : 64;
  return w * w + 16;
}
std::size_t mtk::qr::get_working_l_size(const std::size_t m) { return mtk::tsqr::get_working_l_size(m); }

template <mtk::tsqr::compute_mode mode>
void mtk::tsqr::tsqr16(typename mtk::tsqr::get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
                       typename mtk::tsqr::get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
                       const typename mtk::tsqr::get_io_type<mode>::type* const a_ptr, const std::size_t lda,
                       const std::size_t m, const std::size_t n,
                       typename get_working_q_type<mode>::type* const working_q_ptr,
                       typename get_working_r_type<mode>::type* const, unsigned* const, unsigned* const,
                       cudaStream_t const cuda_stream) {
  if (m == 0 || n == 0) return;
  if (n > (std::size_t)kMaxN) throw std::runtime_error("tsqr_b200: tsqr16 takes n <= 64");
  tsqr_full((int)mode, q_ptr, (long long)ldq, r_ptr, (long long)ldr, a_ptr, (long long)lda, (long long)m, (int)n,
            reinterpret_cast<float*>(working_q_ptr), wq_capacity_bytes_tsqr(m, n), cuda_stream);
  if (cxx_blocking()) TSQR_CUDA_CHECK(cudaStreamSynchronize(cuda_stream));
}

template <mtk::qr::compute_mode mode, bool Reorthogonalize>
mtk::qr::state_t mtk::qr::qr(typename mtk::qr::get_io_type<mode>::type* const q_ptr, const std::size_t ldq,
                             typename mtk::qr::get_io_type<mode>::type* const r_ptr, const std::size_t ldr,
                             typename mtk::qr::get_io_type<mode>::type* const a_ptr, const std::size_t lda,
                             const std::size_t m, const std::size_t n,
                             typename mtk::qr::get_working_q_type<mode>::type* const wq_ptr,
                             typename mtk::qr::get_working_r_type<mode>::type* const,
                             typename mtk::qr::get_io_type<mode>::type* const, unsigned* const, unsigned* const,
                             cublasHandle_t const main_cublas_handle) {
  if (n > m || m == 0 || n == 0) return mtk::qr::error_invalid_matrix_size;
  cudaStream_t const stream = stream_of(main_cublas_handle);
  const int st = qr_full((int)mode, Reorthogonalize, q_ptr, (long long)ldq, r_ptr, (long long)ldr, a_ptr, (long long)lda,
                         (long long)m, (long long)n, reinterpret_cast<float*>(wq_ptr), wq_capacity_bytes_qr(m, n), stream);
  if (cxx_blocking()) TSQR_CUDA_CHECK(cudaStreamSynchronize(stream));
  return st;
}

template <mtk::tcqr::compute_mode mode, class Q_T, class R_T, class A_T>
void mtk::tcqr::qr32x16(Q_T* const q, const std::size_t ldq, R_T* const r, const std::size_t ldr, const A_T* const a,
                        const std::size_t lda, const unsigned int m, const unsigned int n,
                        cudaStream_t const cuda_stream) {
  small_qr_launch<Q_T, R_T, A_T>(q, ldq, r, ldr, a, lda, m, n, 1, nullptr, cuda_stream);
}
template <mtk::tcqr::compute_mode mode, class Q_T, class R_T, class A_T>
void mtk::tcqr::qr32x16_batched(Q_T* const q, const std::size_t ldq, R_T* const r, const std::size_t ldr,
                                const A_T* const a, const std::size_t lda, const unsigned int m, const unsigned int n,
                                const std::size_t batch_size, const unsigned* a_start_position,
                                cudaStream_t const cuda_stream) {
  small_qr_launch<Q_T, R_T, A_T>(q, ldq, r, ldr, a, lda, m, n, batch_size, a_start_position, cuda_stream);
}

// ---- explicit instantiations: the symbol set of the reference's libtcqr.a ----
#define TSQR_B200_FOR_EACH_MODE(X) \
  X(fp16_notc) X(fp16_tc_nocor) X(fp32_notc) X(fp32_tc_cor) X(fp32_tc_nocor) X(mixed_tc_cor_emu) \
  X(tf32_tc_cor) X(tf32_tc_cor_emu) X(tf32_tc_nocor) X(tf32_tc_nocor_emu)

#define TSQR_INST(md)                                                                                             \
  template void mtk::tsqr::tsqr16<mtk::tsqr::compute_mode::md>(                                                   \
      mtk::tsqr::get_io_type<mtk::tsqr::compute_mode::md>::type* const, const std::size_t,                        \
      mtk::tsqr::get_io_type<mtk::tsqr::compute_mode::md>::type* const, const std::size_t,                        \
      const mtk::tsqr::get_io_type<mtk::tsqr::compute_mode::md>::type* const, const std::size_t, const std::size_t, \
      const std::size_t, mtk::tsqr::get_working_q_type<mtk::tsqr::compute_mode::md>::type* const,                 \
      mtk::tsqr::get_working_r_type<mtk::tsqr::compute_mode::md>::type* const, unsigned* const, unsigned* const,  \
      cudaStream_t const);
TSQR_B200_FOR_EACH_MODE(TSQR_INST)

#define QR_INST2(md, re)                                                                                       \
  template mtk::qr::state_t mtk::qr::qr<mtk::qr::compute_mode::md, re>(                                        \
      mtk::qr::get_io_type<mtk::qr::compute_mode::md>::type* const, const std::size_t,                         \
      mtk::qr::get_io_type<mtk::qr::compute_mode::md>::type* const, const std::size_t,                         \
      mtk::qr::get_io_type<mtk::qr::compute_mode::md>::type* const, const std::size_t, const std::size_t,      \
      const std::size_t, mtk::qr::get_working_q_type<mtk::qr::compute_mode::md>::type* const,                  \
      mtk::qr::get_working_r_type<mtk::qr::compute_mode::md>::type* const,                                     \
      mtk::qr::get_io_type<mtk::qr::compute_mode::md>::type* const, unsigned* const, unsigned* const,          \
      cublasHandle_t const);
#define QR_INST(md) QR_INST2(md, true) QR_INST2(md, false)
TSQR_B200_FOR_EACH_MODE(QR_INST)

#define TCQR_INST(md, QT, RT, AT)                                                                               \
  template void mtk::tcqr::qr32x16<mtk::tcqr::compute_mode::md, QT, RT, AT>(                                    \
      QT* const, const std::size_t, RT* const, const std::size_t, const AT* const, const std::size_t,           \
      const unsigned int, const unsigned int, cudaStream_t const);                                              \
  template void mtk::tcqr::qr32x16_batched<mtk::tcqr::compute_mode::md, QT, RT, AT>(                            \
      QT* const, const std::size_t, RT* const, const std::size_t, const AT* const, const std::size_t,           \
      const unsigned int, const unsigned int, const std::size_t, const unsigned*, cudaStream_t const);
TCQR_INST(fp32_tc_cor, float, float, float)
TCQR_INST(fp32_notc, float, float, float)
TCQR_INST(fp32_tc_nocor, half, float, float)
TCQR_INST(fp32_tc_nocor, float, float, float)
TCQR_INST(fp16_notc, half, half, half)
TCQR_INST(fp16_tc_nocor, half, half, half)
TCQR_INST(tf32_tc_nocor, float, float, float)
TCQR_INST(tf32_tc_cor, float, float, float)
TCQR_INST(tf32_tc_nocor_emu, float, float, float)
TCQR_INST(tf32_tc_cor_emu, float, float, float)
TCQR_INST(mixed_tc_cor_emu, float, float, float)
