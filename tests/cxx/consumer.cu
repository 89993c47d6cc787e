// A C++ consumer of the drop-in headers, written the way a user of enp1s0/tsqr-gpu writes one: the README
// sample (reference README.md:60-103, typos fixed), the buffer& overloads (reference src/blockqr.hpp:155-175,
// src/tsqr.hpp:125-140), allocate_host / destroy_host (src/blockqr.hpp:107-126), the double-allocate throw
// (src/blockqr.hpp:77-79), get_device_memory_size and the size error code (src/blockqr.cu:407-409).
// It links against libtsqr_b200.so only through the mtk:: symbols the reference's libtcqr.a exports, and it
// relies on the reference's host-blocking semantics (src/blockqr.cu:140, src/tsqr.cu:1260): the results are
// read right after the call returns, and the stream must be idle by then.
#include <blockqr.hpp>
#include <tsqr.hpp>
#include <cmath>
#include <cstdio>
#include <stdexcept>
#include <vector>

namespace {
int failures = 0;
#define CHECK(cond, ...) do { if (!(cond)) { std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); ++failures; } } while (0)

std::vector<float> random_matrix(std::size_t m, std::size_t n, unsigned seed) {
  std::vector<float> a(m * n);
  unsigned long long s = 0x9E3779B97F4A7C15ull ^ seed;
  for (auto& v : a) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    v = (float)((double)(s >> 11) / 9007199254740992.0 * 2.0 - 1.0);   // uniform(-1, 1), reference src/test.cu:97
  }
  return a;
}

// orthogonality |Q^T Q - I|_F / sqrt(n) (reference src/validation.cu:43-80) and residual |QR - A|_F / |A|_F
// (reference src/test.cu:146-165), in double on the host
void metrics(const std::vector<float>& a, const std::vector<float>& q, const std::vector<float>& r, std::size_t m,
             std::size_t n, double& orth, double& res) {
  double so = 0.0;
  for (std::size_t i = 0; i < n; i++)
    for (std::size_t j = 0; j < n; j++) {
      double s = 0.0;
      for (std::size_t k = 0; k < m; k++) s += (double)q[k + i * m] * (double)q[k + j * m];
      s -= (i == j) ? 1.0 : 0.0;
      so += s * s;
    }
  orth = std::sqrt(so / (double)n);
  double sr = 0.0, sa = 0.0;
  for (std::size_t j = 0; j < n; j++)
    for (std::size_t k = 0; k < m; k++) {
      double s = 0.0;
      for (std::size_t i = 0; i <= j; i++) s += (double)q[k + i * m] * (double)r[i + j * n];
      const double d = s - (double)a[k + j * m];
      sr += d * d;
      sa += (double)a[k + j * m] * (double)a[k + j * m];
    }
  res = std::sqrt(sr / sa);
}

template <mtk::qr::compute_mode mode, bool Reorth>
void run_qr_sample(std::size_t M, std::size_t N, bool host_buffer, const char* what) {
  using compute_t = typename mtk::qr::get_io_type<mode>::type;
  static_assert(sizeof(compute_t) == 4, "this consumer checks the float modes");
  const std::vector<float> a = random_matrix(M, N, (unsigned)(M + N));
  compute_t *d_a, *d_r, *d_q;
  cudaMalloc((void**)&d_a, sizeof(compute_t) * M * N);
  cudaMalloc((void**)&d_r, sizeof(compute_t) * N * N);
  cudaMalloc((void**)&d_q, sizeof(compute_t) * M * N);
  cudaMemcpy(d_a, a.data(), sizeof(float) * M * N, cudaMemcpyHostToDevice);
  cudaMemset(d_r, 0, sizeof(compute_t) * N * N);   // the caller zeroes R (reference src/test.cu:129)

  mtk::qr::buffer<mode, Reorth> buffer;
  if (host_buffer) buffer.allocate_host(M, N); else buffer.allocate(M, N);
  CHECK(buffer.get_device_memory_size() > 0, "get_device_memory_size");
  bool thrown = false;
  try { buffer.allocate(M, N); } catch (const std::runtime_error&) { thrown = true; }
  CHECK(thrown, "a second allocate() must throw std::runtime_error");

  cublasHandle_t cublas_handle;
  cublasCreate(&cublas_handle);
  cudaStream_t stream;
  cudaStreamCreate(&stream);
  cublasSetStream(cublas_handle, stream);

  const mtk::qr::state_t st = mtk::qr::qr<mode, Reorth>(d_q, M, d_r, N, d_a, M, M, N, buffer, cublas_handle);
  CHECK(st == mtk::qr::success_factorization, "%s: state %d", what, st);
  CHECK(cudaStreamQuery(stream) == cudaSuccess, "%s: mtk::qr::qr must be host-blocking (stream still busy)", what);

  std::vector<float> q(M * N), r(N * N);
  cudaMemcpy(q.data(), d_q, sizeof(float) * M * N, cudaMemcpyDeviceToHost);
  cudaMemcpy(r.data(), d_r, sizeof(float) * N * N, cudaMemcpyDeviceToHost);
  double orth, res;
  metrics(a, q, r, M, N, orth, res);
  std::printf("%s: %zu x %zu orthogonality %.3e residual %.3e\n", what, M, N, orth, res);
  CHECK(orth < 5e-6 && res < 5e-6, "%s: accuracy", what);
  for (std::size_t j = 0; j < N; j++)
    for (std::size_t i = j + 1; i < N; i++) CHECK(r[i + j * N] == 0.0f, "%s: R below the diagonal", what);

  CHECK((mtk::qr::qr<mode, Reorth>(d_q, M, d_r, N, d_a, M, N - 1, N, buffer, cublas_handle)) == mtk::qr::error_invalid_matrix_size,
        "n > m must return error_invalid_matrix_size");
  if (host_buffer) buffer.destroy_host(); else buffer.destroy();
  CHECK(buffer.dwq == nullptr, "destroy");
  cublasDestroy(cublas_handle);
  cudaStreamDestroy(stream);
  cudaFree(d_a); cudaFree(d_r); cudaFree(d_q);
}

template <mtk::tsqr::compute_mode mode>
void run_tsqr(std::size_t M, std::size_t N, bool generic_entry) {
  const std::vector<float> a = random_matrix(M, N, 7u + (unsigned)N);
  float *d_a, *d_r, *d_q;
  cudaMalloc((void**)&d_a, sizeof(float) * M * N);
  cudaMalloc((void**)&d_r, sizeof(float) * N * N);
  cudaMalloc((void**)&d_q, sizeof(float) * M * N);
  cudaMemcpy(d_a, a.data(), sizeof(float) * M * N, cudaMemcpyHostToDevice);
  cudaMemset(d_r, 0, sizeof(float) * N * N);
  mtk::tsqr::buffer<mode> buffer;
  buffer.allocate(M, N);
  cudaStream_t stream;
  cudaStreamCreate(&stream);
  if (generic_entry) mtk::tsqr::tsqr<mode>(d_q, M, d_r, N, d_a, M, M, N, buffer, stream);
  else mtk::tsqr::tsqr16<mode>(d_q, M, d_r, N, d_a, M, M, N, buffer, stream);
  CHECK(cudaStreamQuery(stream) == cudaSuccess, "tsqr16 must be host-blocking");
  std::vector<float> q(M * N), r(N * N);
  cudaMemcpy(q.data(), d_q, sizeof(float) * M * N, cudaMemcpyDeviceToHost);
  cudaMemcpy(r.data(), d_r, sizeof(float) * N * N, cudaMemcpyDeviceToHost);
  double orth, res;
  metrics(a, q, r, M, N, orth, res);
  std::printf("tsqr%s<%d>: %zu x %zu orthogonality %.3e residual %.3e\n", generic_entry ? "" : "16", (int)mode, M, N, orth, res);
  CHECK(orth < 5e-6 && res < 5e-6, "tsqr accuracy");
  CHECK(buffer.get_device_memory_size() > 0, "tsqr buffer size");
  buffer.destroy();
  cudaStreamDestroy(stream);
  cudaFree(d_a); cudaFree(d_r); cudaFree(d_q);
}
}  // namespace

int main(int argc, char** argv) {
  if (argc > 1) {   // "link-only" invocation of the CPU test: the templates above were instantiated and linked
    std::printf("linked: working_q_size(9211, 51) = %zu\n", mtk::qr::get_working_q_size(9211, 51));
    return 0;
  }
  run_qr_sample<mtk::qr::fp32_tc_cor, false>(9211, 51, false, "README sample (fp32_tc_cor)");
  run_qr_sample<mtk::qr::fp32_tc_cor, true>(20000, 130, false, "BlockQR + reorth (fp32_tc_cor)");
  run_qr_sample<mtk::qr::fp32_notc, true>(4096, 32, true, "allocate_host (fp32_notc)");
  run_qr_sample<mtk::qr::tf32_tc_cor, false>(40000, 64, false, "fast path (tf32_tc_cor)");
  run_tsqr<mtk::tsqr::fp32_notc>(16384, 16, false);
  run_tsqr<mtk::tsqr::fp32_tc_cor>(30000, 64, true);
  if (failures) { std::printf("%d failure(s)\n", failures); return 1; }
  std::printf("consumer OK\n");
  return 0;
}
