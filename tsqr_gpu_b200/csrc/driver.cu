#include "driver.h"
#include "common.cuh"
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <mutex>

namespace tsqr_b200 {

int device_sm_count() {
  static int cached[64];
  static std::once_flag once;
  std::call_once(once, [] { for (int& c : cached) c = 0; });
  int dev = 0;
  // size queries must also work on a host without a device (planning is pure host logic)
  if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return 148; }
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) { (void)cudaGetLastError(); sms = 0; }
    cached[dev] = sms > 0 ? sms : 148;
  }
  return cached[dev];
}

namespace {
struct ProfRec { int label; cudaEvent_t beg, end; };
std::atomic<bool> g_prof_on{false};
std::vector<ProfRec> g_prof;
std::vector<cudaEvent_t> g_prof_pool;   // events are recycled: none is created inside a timed region twice
std::mutex g_prof_mu;
cudaEvent_t prof_event() {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  if (!g_prof_pool.empty()) { cudaEvent_t e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace
ProfScope::ProfScope(cudaStream_t st, int lb) : s(st), label(lb), on(g_prof_on.load(std::memory_order_relaxed)) {
  if (on) { beg = prof_event(); end = prof_event(); cudaEventRecord(beg, s); }
}
ProfScope::~ProfScope() {
  if (on) {
    cudaEventRecord(end, s);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (g_prof.size() < 65536) g_prof.push_back({label, beg, end});
    else { g_prof_pool.push_back(beg); g_prof_pool.push_back(end); }
  }
}

void profile_enable(bool on) { g_prof_on.store(on); }
int profile_read(int* labels, float* ms, int cap) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int n = 0;
  for (auto& r : g_prof) {
    float t = 0.f;
    cudaEventSynchronize(r.end);
    cudaEventElapsedTime(&t, r.beg, r.end);
    if (n < cap) { labels[n] = r.label; ms[n] = t; ++n; }
    g_prof_pool.push_back(r.beg); g_prof_pool.push_back(r.end);
  }
  g_prof.clear();
  return n;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static Plan make_plan_tc(long long m, int n, int max_chunks0) {
  Plan plan;
  plan.tc = true;
  plan.m = m;
  plan.n = n;
  plan.np = tc_np(n);
  const int NP = plan.np, H = kTcTileRows;
  size_t off = 0;
  long long rows = m;
  int level = 0;
  while (true) {
    Level L{};
    L.rows = rows;
    long long tiles_pc;
    if (rows <= NP + 8LL * H) {
      tiles_pc = (rows > NP) ? (rows - NP + H - 1) / H : 0;
    } else if (level == 0) {
      const long long want = max_chunks0 > 0 ? max_chunks0 : (long long)cs_reduce_ctas_per_sm(NP) * device_sm_count();
      const long long per = (rows + want - 1) / want;
      tiles_pc = std::max<long long>(4, (per - NP + H - 1) / H);
    } else {
      tiles_pc = 4;  // short chains keep the upper levels shallow in time
    }
    L.rows_per_chunk = NP + tiles_pc * H;
    L.tiles_per_chunk = (int)tiles_pc;
    L.nchunks = (int)std::max<long long>(1, (rows + L.rows_per_chunk - 1) / L.rows_per_chunk);
    L.ckpt_off = off;
    L.S_off = off;
    off = align_up(off + (size_t)L.nchunks * n * n, 64);
    L.Qup_off = off;
    off = align_up(off + (size_t)L.nchunks * n * n, 64);
    plan.levels.push_back(L);
    if (L.nchunks == 1) break;
    rows = (long long)L.nchunks * n;
    ++level;
  }
  plan.total_floats = off;
  return plan;
}

Plan make_plan(long long m, int n, int max_chunks0, bool tc) {
  if (tc) return make_plan_tc(m, n, max_chunks0);
  Plan plan;
  plan.m = m;
  plan.n = n;
  plan.np = (n + kPanel - 1) / kPanel * kPanel;
  const int H = pass_tile_rows(plan.np);
  const int sms = max_chunks0 > 0 ? 1 : device_sm_count();
  size_t off = 0;
  long long rows = m;
  int level = 0;
  while (true) {
    Level L;
    L.rows = rows;
    const long long ntiles = (rows + H - 1) / H;
    long long want;
    if (level == 0) {
      want = max_chunks0 > 0 ? max_chunks0
                             : (long long)sms * pass_simt_max_ctas_per_sm(plan.np, true);
    } else {
      // upper levels: short chains (4 tiles per CTA) keep the tree shallow in time
      want = (ntiles + 3) / 4;
    }
    if (ntiles <= 4) want = 1;
    long long tiles_pc = (ntiles + want - 1) / want;
    if (tiles_pc < 1) tiles_pc = 1;
    L.rows_per_chunk = tiles_pc * H;
    L.tiles_per_chunk = (int)tiles_pc;
    L.nchunks = (int)((rows + L.rows_per_chunk - 1) / L.rows_per_chunk);
    L.ckpt_off = off;
    off = align_up(off + (size_t)L.nchunks * L.tiles_per_chunk * ckpt_floats(plan.np), 64);
    L.S_off = off;
    off = align_up(off + (size_t)L.nchunks * n * n, 64);
    L.Qup_off = off;
    off = align_up(off + (size_t)L.nchunks * n * n, 64);
    plan.levels.push_back(L);
    if (L.nchunks == 1) break;
    rows = (long long)L.nchunks * n;
    ++level;
  }
  plan.total_floats = off;
  return plan;
}

bool pc_enabled() {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_PC"); return !(e && e[0] == '0'); }();
  return on;
}

bool pc_notc_enabled() {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_PC_NOTC"); return !(e && e[0] == '0'); }();
  return on;
}

Plan make_plan_for(int mode, long long m, int n) {
  Plan plan = make_plan(m, n, 0, use_tc(mode, m));
  if (use_pc(mode, m)) {
    plan.pc = true;
    plan.pc_off = align_up(plan.total_floats, 64);
    plan.pcl = pc_layout(m, n);
    plan.total_floats = plan.pc_off + plan.pcl.total;
  }
  return plan;
}

size_t plan_floats_any(long long m, int n) {
  size_t tcf = make_plan(m, n, 0, true).total_floats, ccf = make_plan(m, n, 0, false).total_floats;
  if (m >= kPcMinRows) {
    const size_t pc = pc_layout(m, n).total;
    tcf = align_up(tcf, 64) + pc;
    ccf = align_up(ccf, 64) + pc;
  }
  return std::max(ccf, tcf);
}

size_t tsqr_workspace_bytes(long long m, int n) { return plan_floats_any(m, n) * sizeof(float) + 256; }

static TcParams tc_level_params(const Plan& plan, size_t li, const void* a, long long lda, void* q, long long ldq, float* ws) {
  const Level& L = plan.levels[li];
  TcParams p{};
  p.n = plan.n;
  p.M = L.rows;
  p.nchunks = L.nchunks;
  p.rows_per_chunk = L.rows_per_chunk;
  if (li == 0) {
    p.A = a; p.lda = lda; p.Q = q; p.ldq = ldq;
  } else {
    p.A = ws + plan.levels[li - 1].S_off; p.lda = L.rows;
    p.Q = ws + plan.levels[li - 1].Qup_off; p.ldq = L.rows;
  }
  p.S = ws + L.S_off;
  p.lds = (long long)L.nchunks * plan.n;
  return p;
}

static PassParams level_params(const Plan& plan, size_t li, const void* a, long long lda, float* ws) {
  const Level& L = plan.levels[li];
  PassParams p{};
  p.n = plan.n;
  p.np = plan.np;
  p.M = L.rows;
  p.nchunks = L.nchunks;
  p.rows_per_chunk = L.rows_per_chunk;
  p.tiles_per_chunk = L.tiles_per_chunk;
  p.ckpt = ws + L.ckpt_off;
  if (li == 0) {
    p.A = a;
    p.lda = lda;
  } else {
    p.A = ws + plan.levels[li - 1].S_off;
    p.lda = L.rows;
  }
  p.S = ws + L.S_off;
  p.lds = (long long)L.nchunks * plan.n;
  return p;
}

static PcCall pc_call(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq, float* ws) {
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

const float* tsqr_reduce(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq, float* ws,
                         cudaStream_t s) {
  if (plan.tc) {
    const int* skip = nullptr;
    float* rout = nullptr;
    if (plan.pc) {   // fast path first; the levels below return at once when its certificate passed
      pc_reduce(pc_call(plan, mode, a, lda, q, ldq, ws), s);
      skip = reinterpret_cast<const int*>(ws + plan.pc_off + plan.pcl.status);
      rout = ws + plan.pc_off + plan.pcl.rout;
    }
    for (size_t li = 0; li < plan.levels.size(); ++li) {
      TcParams p = tc_level_params(plan, li, a, lda, q, ldq, ws);
      p.skip_ok = skip;
      if (rout && li + 1 == plan.levels.size()) { p.S = rout; p.lds = plan.n; }   // one home for R, whoever wrote it
      ProfScope prof(s, (int)li * 2);
      const IoType io = li == 0 ? mode_io(mode) : IoType::f32;
      launch_cs_reduce(p, io, io, s);
    }
    return rout ? rout : ws + plan.levels.back().S_off;
  }
  const int* skip = nullptr;
  float* rout = nullptr;
  if (plan.pc) {   // *_notc fast path (CUDA-core passes); the levels below return at once when its certificate passed
    pc_reduce(pc_call(plan, mode, a, lda, q, ldq, ws), s);
    skip = reinterpret_cast<const int*>(ws + plan.pc_off + plan.pcl.status);
    rout = ws + plan.pc_off + plan.pcl.rout;
  }
  for (size_t li = 0; li < plan.levels.size(); ++li) {
    PassParams p = level_params(plan, li, a, lda, ws);
    p.emu = mode_emu(mode);
    p.skip_ok = skip;
    if (rout && li + 1 == plan.levels.size()) { p.S = rout; p.lds = plan.n; }   // one home for R, whoever wrote it
    ProfScope prof(s, (int)li * 2);
    launch_pass_simt(p, false, li == 0 ? mode_io(mode) : IoType::f32, IoType::f32, s);
  }
  return rout ? rout : ws + plan.levels.back().S_off;
}

void tsqr_build(const Plan& plan, int mode, const void* a, long long lda, void* q, long long ldq,
                const float* B, long long ldb, float* ws, cudaStream_t s) {
  if (plan.tc) {
    const int* skip = nullptr;
    if (plan.pc) {
      ProfScope prof(s, 103);
      pc_build(pc_call(plan, mode, a, lda, q, ldq, ws), B, ldb, s);
      skip = reinterpret_cast<const int*>(ws + plan.pc_off + plan.pcl.status);
    }
    for (size_t li = plan.levels.size(); li-- > 0;) {
      const Level& L = plan.levels[li];
      TcParams p = tc_level_params(plan, li, a, lda, q, ldq, ws);
      p.skip_ok = skip;
      ProfScope prof(s, (int)li * 2 + 1);
      if (li + 1 == plan.levels.size()) { p.Bsrc = B; p.ldb = ldb; }
      else { p.Bsrc = ws + L.Qup_off; p.ldb = (long long)L.nchunks * plan.n; }
      launch_tc_build(p, mode_is_cor(mode), li == 0 ? mode_io(mode) : IoType::f32, s);
    }
    return;
  }
  const int* skip = nullptr;
  if (plan.pc) {
    ProfScope prof(s, 103);
    pc_build(pc_call(plan, mode, a, lda, q, ldq, ws), B, ldb, s);
    skip = reinterpret_cast<const int*>(ws + plan.pc_off + plan.pcl.status);
  }
  for (size_t li = plan.levels.size(); li-- > 0;) {
    const Level& L = plan.levels[li];
    PassParams p = level_params(plan, li, a, lda, ws);
    p.emu = mode_emu(mode);
    p.skip_ok = skip;
    ProfScope prof(s, (int)li * 2 + 1);
    if (li + 1 == plan.levels.size()) {
      p.Bsrc = B;
      p.ldb = ldb;
    } else {
      p.Bsrc = ws + L.Qup_off;
      p.ldb = (long long)L.nchunks * plan.n;
    }
    if (li == 0) {
      p.Q = q;
      p.ldq = ldq;
      launch_pass_simt(p, true, mode_io(mode), mode_io(mode), s);
    } else {
      p.Q = ws + plan.levels[li - 1].Qup_off;
      p.ldq = L.rows;
      launch_pass_simt(p, true, IoType::f32, IoType::f32, s);
    }
  }
}

namespace {
template <typename T>
__global__ void store_r_kernel(const float* __restrict__ r32, int n, T* r, long long ldr) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n * n; i += gridDim.x * blockDim.x) {
    const int j = i % n, k = i / n;
    io_store(r + j + (long long)k * ldr, j <= k ? r32[j + (long long)k * n] : 0.f);
  }
}
}  // namespace

void store_r(const float* r32, int n, void* r, long long ldr, IoType t, cudaStream_t s) {
  const int blocks = std::min(32, (n * n + 255) / 256);
  if (t == IoType::f32) store_r_kernel<float><<<blocks, 256, 0, s>>>(r32, n, (float*)r, ldr);
  else store_r_kernel<__half><<<blocks, 256, 0, s>>>(r32, n, (__half*)r, ldr);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

void tsqr_full(int mode, void* q, long long ldq, void* r, long long ldr, const void* a, long long lda,
               long long m, int n, float* ws, size_t ws_bytes, cudaStream_t s) {
  Plan plan = make_plan_for(mode, m, n);
  if (plan.total_floats * sizeof(float) > ws_bytes) throw std::runtime_error("tsqr_b200: workspace too small");
  const float* r32 = tsqr_reduce(plan, mode, a, lda, q, ldq, ws, s);
  store_r(r32, n, r, ldr, mode_io(mode), s);
  tsqr_build(plan, mode, a, lda, q, ldq, nullptr, 0, ws, s);
}

}  // namespace tsqr_b200
