// Tensor-core TSQR level passes for sm_100a (tcgen05.mma kind::tf32, accumulators in TMEM).
// See ts_pass.h (TcParams) for the chunk format and tc05.cuh for the operand layouts.
//
// What runs where:
//   CUDA cores : the 16-column Householder panel sweep (column norms and reflector dot
//                products = per-thread partial sums over a 16-row segment + one warp shuffle +
//                a 4-warp combine), the 16x16 / NPxNP triangular solves, the NP x NP head QR.
//   tensor core: every genuine GEMM of the build pass -- S = V^T V and Q_tile = -V X -- with the hi/lo split of the
//                *_cor modes applied when an operand is stored to shared memory (3 MMAs per
//                product into one fp32 TMEM accumulator).
// Replaces reference src/tcqr32x16.cu:1373-1469 (explicit-H leaf QR), src/tsqr.cu:143-1061
// (backward WMMA kernels) and the per-level launch/sync loop of src/tsqr.cu:1064-1279.
#include "common.cuh"
#include "tc05.cuh"
#include "ts_pass.h"

namespace tsqr_b200 {
namespace {
using namespace tc;

constexpr int kT = 128;          // threads per CTA = rows per tile
constexpr int kPS = 132;         // panel buffer column stride (floats), 16 B aligned, bank-skewed
constexpr unsigned kFull = 0xffffffffu;

struct HhScal { float beta, tau, rden; };
// Householder scalars for pivot alpha and tail norm^2 g: reflector [1 ; x * rden], R_jj = beta,
// tau = (beta - alpha) / beta = 1 + |alpha| / norm. A zero column gives v = 0, tau = 2
// (= 2 / (1 + |v|^2), the rule the build pass uses). sqrt = rsqrt.approx + one Newton step
// (correctly rounded to within 1 ulp), reciprocals = rcp.approx (1 ulp): the reflectors are
// stored, never recomputed, so no bit-reproducibility is needed here.
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ HhScal hh_scalars(float alpha, float g) {
  HhScal h;
  const float nx2 = __fmaf_rn(alpha, alpha, g);
  if (nx2 > 1e-36f) {
    const float y = rsqrt_approx(nx2);
    const float n0 = __fmul_rn(nx2, y);
    const float norm = __fmaf_rn(__fmul_rn(0.5f, y), __fmaf_rn(-n0, n0, nx2), n0);
    h.beta = -copysignf(norm, alpha);
    h.rden = rcp_approx(__fsub_rn(alpha, h.beta));
    h.tau = __fmaf_rn(fabsf(alpha), rcp_approx(norm), 1.f);
  } else {
    h.beta = 0.f; h.rden = 0.f; h.tau = 2.f;
  }
  return h;
}

// optional per-phase cycle counters of the build pass (make PHASES=1; tools/phase_clocks.py)
#ifdef TSQR_B200_PHASES
__device__ unsigned long long g_phase_clk[16];
#define PH_INIT() unsigned long long ph_acc[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; unsigned long long ph_s = 0; unsigned long long ph_t = clock64()
#define PH(i) do { const unsigned long long ph_now = clock64(); ph_acc[i] += ph_now - ph_t; ph_t = ph_now; } while (0)
#define PHS0() do { ph_s = clock64(); } while (0)
#define PHS(i) do { const unsigned long long ph_now = clock64(); ph_acc[i] += ph_now - ph_s; ph_s = ph_now; } while (0)
#define PH_FINI() do { if (threadIdx.x == 0 && blockIdx.x == 0) for (int i = 0; i < 14; ++i) g_phase_clk[i] = ph_acc[i]; } while (0)
#else
#define PH_INIT() do {} while (0)
#define PH(i) do {} while (0)
#define PH_FINI() do {} while (0)
#define PHS0() do {} while (0)
#define PHS(i) do {} while (0)
#endif
#include "tc_pass_build.inc"

}  // namespace

void launch_tc_build(const TcParams& p, bool corrected, IoType q, cudaStream_t s) {
  if (q == IoType::f32) {
    if (corrected) launch_build_np<true, float>(p, s);
    else launch_build_np<false, float>(p, s);
  } else {
    if (corrected) launch_build_np<true, __half>(p, s);
    else launch_build_np<false, __half>(p, s);
  }
}

}  // namespace tsqr_b200

#ifdef TSQR_B200_PHASES
extern "C" int tsqr_b200_debug_phases(unsigned long long* out) {
  return (int)cudaMemcpyFromSymbol(out, tsqr_b200::g_phase_clk, 14 * sizeof(unsigned long long));
}
#endif
