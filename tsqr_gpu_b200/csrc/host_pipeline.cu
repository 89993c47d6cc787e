// TSQR (n <= 64) of a matrix that lives in HOST memory, overlapped with its own transfers: the caller side of the hot
// path (SURVEY.md section 8 "next": the reference's drivers stage whole matrices with blocking copies before and after
// mtk::qr::qr, src/test.cu:118-127 and :141-145). The rows are cut into slabs; slab k is copied in on a copy stream and
// reduced as soon as it has arrived, the slab R factors are stacked and factored once (the root step of the row-sharded
// form), then every slab is built with its block of the root Q and copied out while the next slab is being built.
// C ABI: tsqr_b200_host_pipeline_{create,run,destroy}; tsqr_gpu_b200.HostPipeline is the Python twin.
#include <algorithm>
#include <cstring>
#include <utility>
#include <vector>
#include "../../include/tsqr_b200.h"
#include "common.cuh"
#include "driver.h"

namespace tsqr_b200 {
void set_error_message(const char* msg);   // api.cu

namespace {

// gathered[k][c][i] (S blocks of n x n, column-major) -> stacked (S n) x n, column-major, ld S n
__global__ void stack_blocks_kernel(const float* __restrict__ g, int S, int n, float* __restrict__ st) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < S * n * n; e += gridDim.x * blockDim.x) {
    const int i = e % n, c = (e / n) % n, k = e / (n * n);
    st[(size_t)k * n + i + (size_t)c * S * n] = g[e];
  }
}

struct HostPipe {
  int mode = 0, n = 0, device = 0;
  long long m = 0;
  size_t esz = 4;
  std::vector<std::pair<long long, long long>> bounds;   // [lo, hi) rows of each slab
  std::vector<Plan> plans;
  std::vector<float*> ws;
  char* a = nullptr;      // m x n, ld m, io type
  char* q = nullptr;
  float* gath = nullptr;  // S blocks of n x n
  float* stack = nullptr; // (S n) x n
  float* qroot = nullptr; // (S n) x n
  float* rroot = nullptr; // n x n
  float* rootws = nullptr;
  size_t rootws_bytes = 0;
  cudaStream_t h2d = nullptr, d2h = nullptr;
  std::vector<cudaEvent_t> ev_in, ev_out;
  cudaEvent_t ev_start = nullptr, ev_root = nullptr, ev_done = nullptr;

  ~HostPipe() {
    for (float* w : ws) cudaFree(w);
    cudaFree(a); cudaFree(q); cudaFree(gath); cudaFree(stack); cudaFree(qroot); cudaFree(rroot); cudaFree(rootws);
    for (cudaEvent_t e : ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : ev_out) cudaEventDestroy(e);
    if (ev_start) cudaEventDestroy(ev_start);
    if (ev_root) cudaEventDestroy(ev_root);
    if (ev_done) cudaEventDestroy(ev_done);
    if (h2d) cudaStreamDestroy(h2d);
    if (d2h) cudaStreamDestroy(d2h);
  }
};

// contiguous slabs covering [0, m): heights a multiple of the 128-row tile (except the last), at least 1024 rows and
// at least n rows each; a short tail is folded into the previous slab
std::vector<std::pair<long long, long long>> slab_bounds(long long m, int n, int nslabs) {
  long long ms = (m + std::max(1, nslabs) - 1) / std::max(1, nslabs);
  ms = std::max<long long>(std::max<long long>(1024, n), (ms + 127) / 128 * 128);
  std::vector<std::pair<long long, long long>> b;
  for (long long lo = 0; lo < m; lo += ms) b.emplace_back(lo, std::min(m, lo + ms));
  if (b.size() > 1 && b.back().second - b.back().first < std::max<long long>(n, 1024)) {
    const long long lo = b[b.size() - 2].first;
    b.pop_back();
    b.back() = {lo, m};
  }
  return b;
}

template <typename T>
void dev_alloc(T** p, size_t bytes) { TSQR_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(p), std::max<size_t>(bytes, 256))); }

}  // namespace
}  // namespace tsqr_b200

using namespace tsqr_b200;

extern "C" {

void* tsqr_b200_host_pipeline_create(int mode, size_t m, size_t n, int nslabs) {
  HostPipe* hp = nullptr;
  try {
    if (!mode_valid(mode)) throw std::runtime_error("tsqr_b200: unknown compute mode");
    if (n == 0 || n > (size_t)kMaxN || n > m) throw std::runtime_error("tsqr_b200: the host pipeline takes 0 < n <= 64, n <= m");
    hp = new HostPipe();
    hp->mode = mode; hp->m = (long long)m; hp->n = (int)n; hp->esz = io_bytes(mode_io(mode));
    TSQR_CUDA_CHECK(cudaGetDevice(&hp->device));
    hp->bounds = slab_bounds(hp->m, hp->n, nslabs);
    const size_t S = hp->bounds.size();
    for (auto& b : hp->bounds) {
      hp->plans.push_back(make_plan_for(mode, b.second - b.first, hp->n));
      float* w = nullptr;
      dev_alloc(&w, hp->plans.back().total_floats * sizeof(float) + 256);
      hp->ws.push_back(w);
    }
    dev_alloc(&hp->a, m * n * hp->esz);
    dev_alloc(&hp->q, m * n * hp->esz);
    dev_alloc(&hp->gath, S * n * n * sizeof(float));
    dev_alloc(&hp->stack, S * n * n * sizeof(float));
    dev_alloc(&hp->qroot, S * n * n * sizeof(float));
    dev_alloc(&hp->rroot, n * n * sizeof(float));
    hp->rootws_bytes = tsqr_workspace_bytes((long long)(S * n), hp->n);
    dev_alloc(&hp->rootws, hp->rootws_bytes);
    TSQR_CUDA_CHECK(cudaStreamCreateWithFlags(&hp->h2d, cudaStreamNonBlocking));
    TSQR_CUDA_CHECK(cudaStreamCreateWithFlags(&hp->d2h, cudaStreamNonBlocking));
    auto mk = [](cudaEvent_t* e) { TSQR_CUDA_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming)); };
    hp->ev_in.resize(S); hp->ev_out.resize(S);
    for (size_t k = 0; k < S; ++k) { mk(&hp->ev_in[k]); mk(&hp->ev_out[k]); }
    mk(&hp->ev_start); mk(&hp->ev_root); mk(&hp->ev_done);
    return hp;
  } catch (const std::exception& e) {
    set_error_message(e.what());
    delete hp;
    return nullptr;
  }
}

void tsqr_b200_host_pipeline_destroy(void* handle) { delete static_cast<HostPipe*>(handle); }

int tsqr_b200_host_pipeline_slabs(void* handle) { return handle ? (int)static_cast<HostPipe*>(handle)->bounds.size() : 0; }

int tsqr_b200_host_pipeline_run(void* handle, const void* a_host, size_t lda, void* q_host, size_t ldq, float* r_host, size_t ldr,
                                void* stream) {
  auto* hp = static_cast<HostPipe*>(handle);
  if (!hp || !a_host || !q_host || !r_host) { set_error_message("tsqr_b200: null argument"); return TSQR_B200_ERR_ARGUMENT; }
  if (lda < (size_t)hp->m || ldq < (size_t)hp->m || ldr < (size_t)hp->n) {
    set_error_message("tsqr_b200: leading dimension smaller than the matrix");
    return TSQR_B200_ERR_ARGUMENT;
  }
  try {
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int n = hp->n, S = (int)hp->bounds.size();
    const long long m = hp->m;
    const size_t esz = hp->esz;
    TSQR_CUDA_CHECK(cudaEventRecord(hp->ev_start, s));
    TSQR_CUDA_CHECK(cudaStreamWaitEvent(hp->h2d, hp->ev_start, 0));
    TSQR_CUDA_CHECK(cudaStreamWaitEvent(hp->d2h, hp->ev_start, 0));
    for (int k = 0; k < S; ++k) {   // one strided copy per slab: n columns of (hi - lo) elements
      const long long lo = hp->bounds[k].first, rows = hp->bounds[k].second - lo;
      TSQR_CUDA_CHECK(cudaMemcpy2DAsync(hp->a + lo * esz, m * esz, static_cast<const char*>(a_host) + lo * esz, lda * esz,
                                        rows * esz, n, cudaMemcpyHostToDevice, hp->h2d));
      TSQR_CUDA_CHECK(cudaEventRecord(hp->ev_in[k], hp->h2d));
    }
    for (int k = 0; k < S; ++k) {
      const long long lo = hp->bounds[k].first;
      TSQR_CUDA_CHECK(cudaStreamWaitEvent(s, hp->ev_in[k], 0));
      const float* r32 = tsqr_reduce(hp->plans[k], hp->mode, hp->a + lo * esz, m, hp->q + lo * esz, m, hp->ws[k], s);
      store_r(r32, n, S > 1 ? hp->gath + (size_t)k * n * n : hp->rroot, n, IoType::f32, s);   // upper triangle, zeros below
    }
    if (S > 1) {
      stack_blocks_kernel<<<(S * n * n + 255) / 256, 256, 0, s>>>(hp->gath, S, n, hp->stack);
      note_launch();
      TSQR_CUDA_CHECK(cudaGetLastError());
      tsqr_full(fp32_notc, hp->qroot, (long long)S * n, hp->rroot, n, hp->stack, (long long)S * n, (long long)S * n, n, hp->rootws,
                hp->rootws_bytes, s);
    }
    TSQR_CUDA_CHECK(cudaEventRecord(hp->ev_root, s));
    for (int k = 0; k < S; ++k) {
      const long long lo = hp->bounds[k].first, rows = hp->bounds[k].second - lo;
      tsqr_build(hp->plans[k], hp->mode, hp->a + lo * esz, m, hp->q + lo * esz, m, S > 1 ? hp->qroot + (size_t)k * n : nullptr,
                 (long long)S * n, hp->ws[k], s);
      TSQR_CUDA_CHECK(cudaEventRecord(hp->ev_out[k], s));
      TSQR_CUDA_CHECK(cudaStreamWaitEvent(hp->d2h, hp->ev_out[k], 0));
      TSQR_CUDA_CHECK(cudaMemcpy2DAsync(static_cast<char*>(q_host) + lo * esz, ldq * esz, hp->q + lo * esz, m * esz, rows * esz, n,
                                        cudaMemcpyDeviceToHost, hp->d2h));
    }
    TSQR_CUDA_CHECK(cudaStreamWaitEvent(hp->d2h, hp->ev_root, 0));
    TSQR_CUDA_CHECK(cudaMemcpy2DAsync(r_host, ldr * sizeof(float), hp->rroot, (size_t)n * sizeof(float), (size_t)n * sizeof(float), n,
                                      cudaMemcpyDeviceToHost, hp->d2h));
    TSQR_CUDA_CHECK(cudaEventRecord(hp->ev_done, hp->d2h));
    TSQR_CUDA_CHECK(cudaStreamWaitEvent(s, hp->ev_done, 0));
    return 0;
  } catch (const std::exception& e) {
    set_error_message(e.what());
    return TSQR_B200_ERR_RUNTIME;
  }
}

}  // extern "C"
