// R-reduce pass on CUDA cores in the (column, row-segment) layout -- the default reduce pass of the
// tensor-core path (same chunk format as tc_pass.cu: head QR, then [R ; tile] steps; the reflector
// tails V go to the Q buffer, the chunk's R to the stacked matrix S).
//
// Thread t of the 128 owns rows [16s, 16s+16), s = t / 16, of the columns c = t % 16 + 16k
// (k < NP/16) of a 128 x NP tile, in registers. A Householder step j is
//   * the pivot column's row segment read from shared memory (published by a lane of the same
//     half warp, so only a __syncwarp orders it),
//   * per column one 16-term partial dot product per thread, one warp shuffle, a 4-warp combine
//     through shared memory (the only block barrier of the step),
//   * the Householder scalars (rsqrt / rcp), and the rank-1 update of the thread's columns.
// Measured on B200: this right-looking, unblocked sweep needs fewer instructions per tile than
// the blocked tcgen05 variant (tc_reduce_kernel, kept for comparison: TSQR_B200_REDUCE=tc) once
// that one's operand conversions, TMEM round trips and single-thread MMA issue are counted, and
// its 51 KB of shared memory allow 4 CTAs per SM instead of 2. The build pass (S = V^T V,
// Q = -V X) stays on the tensor cores.
// Replaces reference src/tcqr32x16.cu:1373-1469 (qr32x16_core with explicit H) and the R-tree
// levels of src/tsqr.cu:1102-1172.
#include "common.cuh"
#include "tc05.cuh"
#include "ts_pass.h"

namespace tsqr_b200 {
namespace {

constexpr int kT = 128;
constexpr int kPS = 132;  // tile buffer column stride (floats): 16 B aligned, bank-skewed
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

struct HhScal { float beta, tau, rden; };
// reflector [1 ; x * rden], R_jj = beta, tau = 1 + |alpha| / norm; zero column: v = 0, tau = 2
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

template <int NP>
struct CsSmem {
  static constexpr int LD = NP + 1;
  static constexpr uint32_t oTile = 0;                         // NP x kPS floats
  static constexpr uint32_t oR = oTile + NP * kPS * 4;         // NP x LD floats, R[j * LD + c]
  static constexpr uint32_t oPart = oR + NP * LD * 4;          // 2 x NP x 4 floats
  static constexpr uint32_t oRden = oPart + 2 * NP * 4 * 4;    // NP floats
  static constexpr uint32_t kBytes = oRden + NP * 4 + 16;
};


// the 16 Householder steps of panel PJ (columns [16 PJ, 16 PJ + 16)) applied to every column
// >= 16 PJ of the tile. KP = NP / 16 column groups per thread; a[k][i] packs rows 2i, 2i+1.
template <int PJ, int KP>
__device__ __forceinline__ void sweep_panel(f32x2 (&a)[KP][8], float* tileS, float* Rs, float* part, float* rdenv,
                                            int LD, int t) {
  const int c16 = t & 15, s = t >> 4, lane = t & 31, warp = t >> 5;
#pragma unroll 1
  for (int jl = 0; jl < 16; ++jl) {
    const int j = 16 * PJ + jl;
    __syncwarp();  // column j of this row segment was published by a lane of this half warp
    f32x2 vj[8];
    {
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + j * kPS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const ulonglong2 v = src[i]; vj[2 * i] = v.x; vj[2 * i + 1] = v.y; }
    }
    // R entries of row j are read before the barrier and rewritten after it
    const float alpha = Rs[j * LD + j];
    float rjc[KP];
#pragma unroll
    for (int k = PJ; k < KP; ++k) rjc[k] = Rs[j * LD + c16 + 16 * k];
    float* pb = part + (jl & 1) * (KP * 64);
#pragma unroll
    for (int k = PJ; k < KP; ++k) {
      f32x2 q0 = 0ull, q1 = 0ull;
#pragma unroll
      for (int i = 0; i < 8; i += 2) { q0 = fma2(vj[i], a[k][i], q0); q1 = fma2(vj[i + 1], a[k][i + 1], q1); }
      float p0, p1, p2, p3;
      unpack2(q0, p0, p1);
      unpack2(q1, p2, p3);
      float p = __fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3));
      p = __fadd_rn(p, __shfl_xor_sync(kFull, p, 16));
      if (lane < 16) pb[(c16 + 16 * k) * 4 + warp] = p;
    }
    __syncthreads();
    const float4 gj4 = *reinterpret_cast<const float4*>(pb + j * 4);
    const HhScal h = hh_scalars(alpha, __fadd_rn(__fadd_rn(gj4.x, gj4.y), __fadd_rn(gj4.z, gj4.w)));
#pragma unroll
    for (int k = PJ; k < KP; ++k) {
      const int c = c16 + 16 * k;
      const float4 g4 = *reinterpret_cast<const float4*>(pb + c * 4);
      const float g = __fadd_rn(__fadd_rn(g4.x, g4.y), __fadd_rn(g4.z, g4.w));
      const float coef = __fmul_rn(h.tau, __fmaf_rn(g, h.rden, rjc[k]));
      const bool active = (k > PJ) || (c16 > jl);
      const float ncv = active ? -__fmul_rn(coef, h.rden) : 0.f;
      const f32x2 ncv2 = pack2(ncv, ncv);
#pragma unroll
      for (int i = 0; i < 8; ++i) a[k][i] = fma2(ncv2, vj[i], a[k][i]);
      if (s == 0 && active) Rs[j * LD + c] = __fsub_rn(rjc[k], coef);
    }
    if (c16 == jl + 1) {  // publish the next pivot column of this panel (raw, unscaled)
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + (j + 1) * kPS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_ulonglong2(a[PJ][2 * i], a[PJ][2 * i + 1]);
    }
    if (s == 0 && c16 == jl) { Rs[j * LD + j] = h.beta; rdenv[j] = h.rden; }
  }
  if (PJ + 1 < KP && c16 == 0) {  // first pivot column of the next panel
    constexpr int KN = PJ + 1 < KP ? PJ + 1 : PJ;
    ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + (16 * (PJ + 1)) * kPS + 16 * s);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_ulonglong2(a[KN][2 * i], a[KN][2 * i + 1]);
  }
}

template <int PJ, int KP>
struct SweepAll {
  static __device__ __forceinline__ void run(f32x2 (&a)[KP][8], float* tileS, float* Rs, float* part, float* rdenv, int LD, int t) {
    sweep_panel<PJ, KP>(a, tileS, Rs, part, rdenv, LD, t);
    SweepAll<PJ + 1, KP>::run(a, tileS, Rs, part, rdenv, LD, t);
  }
};
template <int KP>
struct SweepAll<KP, KP> {
  static __device__ __forceinline__ void run(f32x2 (&)[KP][8], float*, float*, float*, float*, int, int) {}
};

template <int NP, int MINB, typename TIn, typename TQ>
__global__ void __launch_bounds__(kT, MINB) cs_reduce_kernel(const TcParams P) {
  using L = CsSmem<NP>;
  constexpr int LD = L::LD, KP = NP / 16;
  extern __shared__ __align__(128) char sm[];
  float* tileS = reinterpret_cast<float*>(sm + L::oTile);
  float* Rs = reinterpret_cast<float*>(sm + L::oR);
  float* part = reinterpret_cast<float*>(sm + L::oPart);
  float* rdenv = reinterpret_cast<float*>(sm + L::oRden);
  const int t = threadIdx.x, c16 = t & 15, s = t >> 4;
  const int n = P.n;
  const TIn* A = reinterpret_cast<const TIn*>(P.A);
  TQ* Q = reinterpret_cast<TQ*>(P.Q);
  const int chunk = blockIdx.x;
  const long long cbeg = (long long)chunk * P.rows_per_chunk;
  const long long cend = (chunk == P.nchunks - 1) ? P.M : cbeg + P.rows_per_chunk;

  // ---------------- head: plain Householder QR of the first NP rows (in Rs, column c at Hs[c*LD]) ----
  {
    float* Hs = tileS;  // Hs[c * LD + r]
    for (int idx = t; idx < NP * NP; idx += kT) {
      const int r = idx % NP, c = idx / NP;
      const long long row = cbeg + r;
      Hs[c * LD + r] = (row < cend && c < n) ? io_load(A + row + (long long)c * P.lda) : 0.f;
    }
    __syncthreads();
    for (int j = 0; j < NP; ++j) {
      const float* hj = Hs + j * LD;
      float g = 0.f;
      for (int r = j + 1; r < NP; ++r) g = __fmaf_rn(hj[r], hj[r], g);
      const HhScal h = hh_scalars(hj[j], g);
      if (t > j && t < NP) {
        float* hc = Hs + t * LD;
        float dot = 0.f;
        for (int r = j + 1; r < NP; ++r) dot = __fmaf_rn(hj[r], hc[r], dot);
        const float coef = __fmul_rn(h.tau, __fmaf_rn(dot, h.rden, hc[j]));
        const float cv = __fmul_rn(coef, h.rden);
        hc[j] = __fsub_rn(hc[j], coef);
        for (int r = j + 1; r < NP; ++r) hc[r] = __fmaf_rn(-cv, hj[r], hc[r]);
      }
      __syncthreads();
      if (t == j) {
        float* hw = Hs + j * LD;
        for (int r = j + 1; r < NP; ++r) hw[r] = __fmul_rn(hw[r], h.rden);
        hw[j] = h.beta;
      }
    }
    __syncthreads();
    for (int idx = t; idx < NP * NP; idx += kT) {  // R0 (upper) -> Rs
      const int c = idx % NP, j = idx / NP;
      Rs[j * LD + c] = (j <= c) ? Hs[c * LD + j] : 0.f;
    }
    if (t < NP && cbeg + t < cend) {  // V0 (strictly lower) -> Q head rows
      for (int j = 0; j < n; ++j) io_store(Q + cbeg + t + (long long)j * P.ldq, t > j ? Hs[j * LD + t] : 0.f);
    }
    __syncthreads();
  }

  // ---------------- tiles ----------------
  const long long t0row = cbeg + NP;
  const int ntiles = cend > t0row ? (int)((cend - t0row + kT - 1) / kT) : 0;
  for (int tile = 0; tile < ntiles; ++tile) {
    const long long row = t0row + (long long)tile * kT + t;
    const bool valid = row < cend;
    {
      const TIn* ap = A + row;
#pragma unroll 8
      for (int c = 0; c < NP; ++c) {
        tileS[c * kPS + t] = (valid && c < n) ? io_load(ap) : 0.f;
        ap += P.lda;
      }
    }
    if (tile + 1 < ntiles) {  // pull the next tile towards L2: one 128-byte line per warp and column
      const int lane = t & 31;
      const long long r0 = row - lane + kT;
      for (int c = lane; c < n; c += 32)
        if (r0 < cend) tc::prefetch_l2(A + r0 + (long long)c * P.lda);
    }
    __syncthreads();
    f32x2 a[KP][8];
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + (c16 + 16 * k) * kPS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const ulonglong2 v = src[i]; a[k][2 * i] = v.x; a[k][2 * i + 1] = v.y; }
    }
    SweepAll<0, KP>::run(a, tileS, Rs, part, rdenv, LD, t);
    __syncthreads();
    if (valid) {  // V = x_j * rden_j, every column was published (raw) when it became the pivot
      TQ* qp = Q + row;
#pragma unroll 8
      for (int c = 0; c < NP; ++c) {
        if (c < n) io_store(qp, __fmul_rn(tileS[c * kPS + t], rdenv[c]));
        qp += P.ldq;
      }
    }
    __syncthreads();
  }

  // ---------------- chunk result: R -> stacked S ----------------
  for (int idx = t; idx < n * n; idx += kT) {
    const int j = idx % n, c = idx / n;
    P.S[(long long)chunk * n + j + (long long)c * P.lds] = (j <= c) ? Rs[j * LD + c] : 0.f;
  }
}

template <int NP, typename TIn, typename TQ>
void launch_cs_t(const TcParams& p, cudaStream_t s) {
  constexpr int MINB = NP == 64 ? 4 : 6;
  auto kern = cs_reduce_kernel<NP, MINB, TIn, TQ>;
  const int bytes = (int)CsSmem<NP>::kBytes;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  kern<<<p.nchunks, kT, bytes, s>>>(p);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <typename TIn, typename TQ>
void launch_cs_np(const TcParams& p, cudaStream_t s) {
  switch (tc_np(p.n)) {
    case 16: launch_cs_t<16, TIn, TQ>(p, s); break;
    case 32: launch_cs_t<32, TIn, TQ>(p, s); break;
    default: launch_cs_t<64, TIn, TQ>(p, s); break;
  }
}

}  // namespace

void launch_cs_reduce(const TcParams& p, IoType in, IoType q, cudaStream_t s) {
  if (in != q) throw std::runtime_error("tsqr_b200: reduce pass needs matching A and Q element types");
  if (in == IoType::f32) launch_cs_np<float, float>(p, s);
  else launch_cs_np<__half, __half>(p, s);
}
int cs_reduce_ctas_per_sm(int np) { return np == 64 ? 4 : 6; }

}  // namespace tsqr_b200
