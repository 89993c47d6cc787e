// Row-sharded multi-GPU TSQR inside the library: one process (or thread) per GPU, rank p owns rows
// [p M/P, (p+1) M/P) of A and of Q; NCCL on the caller's stream does the exchange (BASELINE.json
// north_star item 4). The reference is single-GPU; this splits its tsqr16 (reference src/tsqr.cu:1064-1279) across ranks.
//
// Tall slabs in a tensor-core mode (every rank >= 16384 rows), the preconditioned-Cholesky form:
//     rank 0: R_pre from a sample of its slab  ->  ncclBroadcast (NP x NP floats, 16 KB)
//     all   : pass 1 with the common R_pre     ->  ncclAllReduce of the (2 NP) x NP fp64 Gram sums (64 KB)
//     all   : the same Cholesky factor, certificate, R and C on every rank (bit-identical input), pass 2.
// No root QR, no bulk data on NVLink, two latency-bound collectives. If the (global) certificate fails, or for the
// other modes / short slabs, the general form runs: local R-reduce -> ncclAllGather of the R factors -> every rank
// factors the stacked (P n) x n matrix redundantly with the same kernel (identical bits) -> local Q-build.
// The certificate is read back once per call (one stream synchronisation): the two forms use different collectives,
// so the choice has to be made on the host, identically on every rank.
// NCCL is not linked: its entry points are looked up in the process (the caller that owns a communicator, or torch,
// has the library loaded), then by soname.
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include "../../include/tsqr_b200.h"
#include "common.cuh"
#include "driver.h"

namespace tsqr_b200 {
namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};
const NcclApi& nccl() {
  static NcclApi api = [] {
    NcclApi a;
    void* lib = RTLD_DEFAULT;
    if (!dlsym(RTLD_DEFAULT, "ncclAllReduce")) {
      lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
      if (!lib) return a;
    }
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(dlsym(lib, "ncclBroadcast"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(lib, "ncclAllReduce"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(lib, "ncclAllGather"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.Broadcast && a.AllReduce && a.AllGather;
    return a;
  }();
  if (!api.ok) throw std::runtime_error("tsqr_b200: NCCL (libnccl.so.2) cannot be resolved");
  return api;
}
void nccl_check(ncclResult_t r, const char* what) {
  if (r != ncclSuccess)
    throw std::runtime_error(std::string("tsqr_b200 NCCL error in ") + what + ": " + (nccl().GetErrorString ? nccl().GetErrorString(r) : "?"));
}

struct Shard {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;
  bool owns = false;
  int* h_flag = nullptr;                                        // pinned: certificate / agreement read-backs
  std::map<std::tuple<int, long long, int>, bool> pc_agreed;    // (mode, m_local, n) -> every rank takes the PC form
};

// gathered[p][k][i] (P blocks of n x n, column-major) -> stacked (P n) x n, column-major, ld P n
__global__ void restack_kernel(const float* __restrict__ g, int P, int n, float* __restrict__ st) {
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < P * n * n; e += gridDim.x * blockDim.x) {
    const int i = e % n, k = (e / n) % n, p = e / (n * n);
    st[(size_t)p * n + i + (size_t)k * P * n] = g[e];
  }
}

struct ShWs { size_t local, gath, stack, qroot, rroot, rootws, flag, total; };
ShWs carve(int mode, long long m, int n, int P) {
  ShWs w{};
  size_t off = 0;
  auto take = [&](size_t f) { size_t o = off; off = (off + f + 63) / 64 * 64; return o; };
  w.local = take(plan_floats_any(m, n) + 64);
  w.gath = take((size_t)P * n * n);
  w.stack = take((size_t)P * n * n);
  w.qroot = take((size_t)P * n * n);
  w.rroot = take((size_t)n * n);
  w.rootws = take(tsqr_workspace_bytes((long long)P * n, n) / sizeof(float) + 64);
  w.flag = take(64);
  w.total = off;
  return w;
}

PcCall pc_call_of(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq, float* ws) {
  PcCall c{};
  c.A = a; c.lda = lda; c.M = plan.m; c.n = plan.n;
  c.Q = q; c.ldq = ldq;
  c.io = mode_io(mode);
  c.cor = mode_is_cor(mode);
  c.notc = !mode_is_tc(mode);
  c.ws = ws + plan.pc_off;
  c.L = plan.pcl;
  return c;
}

void general_form(Shard& sh, const Plan& plan, int mode, void* q, long long ldq, void* r, long long ldr, const void* a, long long lda,
                  int n, float* ws, const ShWs& w, cudaStream_t s) {
  const int P = sh.nranks;
  float* lws = ws + w.local;
  const float* r32 = tsqr_reduce(plan, mode, a, lda, q, ldq, lws, s);
  nccl_check(nccl().AllGather(r32, ws + w.gath, (size_t)n * n, ncclFloat32, sh.comm, s), "ncclAllGather");
  restack_kernel<<<(P * n * n + 255) / 256, 256, 0, s>>>(ws + w.gath, P, n, ws + w.stack);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
  tsqr_full(fp32_notc, ws + w.qroot, (long long)P * n, ws + w.rroot, n, ws + w.stack, (long long)P * n, (long long)P * n, n,
            ws + w.rootws, (w.flag - w.rootws) * sizeof(float), s);
  store_r(ws + w.rroot, n, r, ldr, mode_io(mode), s);
  tsqr_build(plan, mode, a, lda, q, ldq, ws + w.qroot + (size_t)sh.rank * n, (long long)P * n, lws, s);
}

}  // namespace
}  // namespace tsqr_b200

using namespace tsqr_b200;

extern "C" {

int tsqr_b200_nccl_unique_id(void* out128) {
  try {
    ncclUniqueId id;
    nccl_check(nccl().GetUniqueId(&id), "ncclGetUniqueId");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    std::memcpy(out128, &id, 128);
    return 0;
  } catch (const std::exception& e) { set_error_message(e.what()); return TSQR_B200_ERR_RUNTIME; }
}

void* tsqr_b200_sharded_init(const void* id128, int nranks, int rank) {
  try {
    auto* sh = new Shard();
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    nccl_check(nccl().CommInitRank(&sh->comm, nranks, id, rank), "ncclCommInitRank");
    sh->rank = rank; sh->nranks = nranks; sh->owns = true;
    TSQR_CUDA_CHECK(cudaMallocHost(reinterpret_cast<void**>(&sh->h_flag), 64));
    return sh;
  } catch (const std::exception& e) { set_error_message(e.what()); return nullptr; }
}

void* tsqr_b200_sharded_attach(void* nccl_comm, int nranks, int rank) {
  try {
    auto* sh = new Shard();
    sh->comm = reinterpret_cast<ncclComm_t>(nccl_comm);
    sh->rank = rank; sh->nranks = nranks; sh->owns = false;
    TSQR_CUDA_CHECK(cudaMallocHost(reinterpret_cast<void**>(&sh->h_flag), 64));
    return sh;
  } catch (const std::exception& e) { set_error_message(e.what()); return nullptr; }
}

void tsqr_b200_sharded_destroy(void* handle) {
  auto* sh = reinterpret_cast<Shard*>(handle);
  if (!sh) return;
  if (sh->owns && sh->comm) nccl().CommDestroy(sh->comm);
  if (sh->h_flag) cudaFreeHost(sh->h_flag);
  delete sh;
}

size_t tsqr_b200_sharded_workspace_bytes(size_t m_local, size_t n, int nranks) {
  if (m_local == 0 || n == 0 || n > (size_t)kMaxN || nranks < 1) return 256;
  return carve(0, (long long)m_local, (int)n, nranks).total * sizeof(float) + 256;
}

int tsqr_b200_tsqr_sharded(void* handle, int mode, void* q, size_t ldq, void* r, size_t ldr, const void* a, size_t lda,
                           size_t m_local, size_t n_, void* workspace, size_t ws_bytes, void* stream) {
  auto* sh = reinterpret_cast<Shard*>(handle);
  if (!sh || !mode_valid(mode) || !q || !r || !a || !workspace) return TSQR_B200_ERR_ARGUMENT;
  if (m_local == 0 || n_ == 0 || n_ > (size_t)kMaxN) return TSQR_B200_INVALID_SIZE;
  try {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int n = (int)n_, P = sh->nranks;
    const long long m = (long long)m_local;
    float* ws = reinterpret_cast<float*>(workspace);
    const ShWs w = carve(mode, m, n, P);
    if (w.total * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
    Plan plan = make_plan_for(mode, m, n);
    int* d_flag = reinterpret_cast<int*>(ws + w.flag);
    // every rank must take the same form: agree once per (mode, m_local, n)
    const auto key = std::make_tuple(mode, m, n);
    auto it = sh->pc_agreed.find(key);
    if (it == sh->pc_agreed.end()) {
      sh->h_flag[0] = plan.pc ? 1 : 0;
      TSQR_CUDA_CHECK(cudaMemcpyAsync(d_flag, sh->h_flag, sizeof(int), cudaMemcpyHostToDevice, s));
      if (P > 1) nccl_check(nccl().AllReduce(d_flag, d_flag, 1, ncclInt32, ncclMin, sh->comm, s), "ncclAllReduce");
      TSQR_CUDA_CHECK(cudaMemcpyAsync(sh->h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
      TSQR_CUDA_CHECK(cudaStreamSynchronize(s));
      it = sh->pc_agreed.emplace(key, sh->h_flag[0] == 1).first;
    }
    bool done = false;
    if (it->second) {
      const PcCall c = pc_call_of(plan, mode, a, (long long)lda, q, (long long)ldq, ws + w.local);
      const size_t np2 = (size_t)c.L.np * c.L.np;
      if (sh->rank == 0) pc_stage_sample(c, s);
      if (P > 1) nccl_check(nccl().Broadcast(c.ws + c.L.rpre, c.ws + c.L.rpre, np2, ncclFloat32, 0, sh->comm, s), "ncclBroadcast");
      pc_stage_adopt_rpre(c, s);
      pc_stage_pass1(c, s);
      pc_stage_partial_sums(c, s);
      double* gsum = reinterpret_cast<double*>(c.ws + c.L.gsum);
      if (P > 1) nccl_check(nccl().AllReduce(gsum, gsum, (c.cor ? 2 : 1) * np2, ncclFloat64, ncclSum, sh->comm, s), "ncclAllReduce");
      pc_stage_factor(c, s);
      TSQR_CUDA_CHECK(cudaMemcpyAsync(sh->h_flag, c.ws + c.L.status, sizeof(int), cudaMemcpyDeviceToHost, s));
      TSQR_CUDA_CHECK(cudaStreamSynchronize(s));   // the certificate: identical on every rank (same Gram matrix)
      if (sh->h_flag[0] == kPcOk) {
        {
          ProfScope prof(s, 103);
          pc_build(c, nullptr, 0, s);
        }
        store_r(c.ws + c.L.rout, n, r, (long long)ldr, mode_io(mode), s);
        done = true;
      }
    }
    if (!done) {
      Plan hplan = plan;
      hplan.pc = false;   // Householder / CUDA-core passes only
      general_form(*sh, hplan, mode, q, (long long)ldq, r, (long long)ldr, a, (long long)lda, n, ws, w, s);
    }
    return 0;
  } catch (const std::exception& e) {
    set_error_message(e.what());
    return TSQR_B200_ERR_RUNTIME;
  }
}

}  // extern "C"
