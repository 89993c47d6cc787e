// R-reduce pass on CUDA cores in the (column, row-segment) layout -- the default reduce pass of the
// tensor-core path (same chunk format as tc_pass.cu: head QR, then [R ; tile] steps; the reflector
// tails V go to the Q buffer, the chunk's R to the stacked matrix S).
//
// Thread t of the 16*RS owns rows [RPT s, RPT s + RPT), s = t / 16, RPT = 128 / RS, of the columns
// c = t % 16 + 16k (k < NP/16) of a 128 x NP tile, in registers. Two shapes are built: RS = 8
// (128 threads, 16 rows per thread) and RS = 4 (64 threads, 32 rows per thread: the per-column
// reduction/scalar work of a step is shared by twice as many FFMA2s). A Householder step j is
//   * the pivot column's row segment read from shared memory (published by a lane of the same
//     half warp, so only a __syncwarp orders it),
//   * per column one RPT-term partial dot product per thread, one warp shuffle, an RS/2-warp combine
//     through shared memory (the only block barrier of the step),
//   * the Householder scalars (rsqrt / rcp), and the rank-1 update of the thread's columns.
// Its 51 KB of shared memory allow 4 CTAs per SM. The build pass (S = V^T V, Q = -V X) runs on the tensor cores.
// (A blocked variant with the trailing update on tcgen05 was measured in round 1 -- 15.0 ms against 7.8 ms at 2^24 x 64,
// profiles/r1_tc_reduce_full.* -- and removed in round 2, when the preconditioned-Cholesky passes of pc_tsqr.cu
// became the fast path and this kernel its fallback.)
// Replaces reference src/tcqr32x16.cu:1373-1469 (qr32x16_core with explicit H) and the R-tree
// levels of src/tsqr.cu:1102-1172.
#include <cstdlib>
#include <string>
#include "common.cuh"
#include "tc05.cuh"
#include "ts_pass.h"

namespace tsqr_b200 {
namespace {

constexpr int kTileRows = 128;
constexpr int kPS = 132;  // tile buffer column stride (floats): 16 B aligned, bank-skewed
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// raw element loads: the conversion of fp16 input is kept apart from the load so that a whole batch of
// loads is in flight before the first use (with __half2float(__ldg(p)) the compiler reuses one register
// for load + convert and serialises the batch)
__device__ __forceinline__ float raw_load(const float* p) { return *p; }   // plain load: A may alias Q, which this kernel writes
__device__ __forceinline__ unsigned short raw_load(const __half* p) {
  unsigned short v; asm volatile("ld.global.u16 %0, [%1];" : "=h"(v) : "l"(p)); return v;
}
__device__ __forceinline__ float raw_cvt(float v) { return v; }
__device__ __forceinline__ float raw_cvt(unsigned short v) { return __half2float(__ushort_as_half(v)); }
template <typename T> struct RawOf { typedef float type; };
template <> struct RawOf<__half> { typedef unsigned short type; };

struct HhScal { float beta, tau, rden; };
// reflector [1 ; x * rden], R_jj = beta, tau = 1 + |alpha| / norm; zero column: v = 0, tau = 2
__device__ __forceinline__ HhScal hh_scalars(float alpha, float g) {
  HhScal h;
  const float nx2 = __fmaf_rn(alpha, alpha, g);
  const bool ok = nx2 > 1e-36f;              // branch-free: the zero column is a select at the end
  const float nx2s = ok ? nx2 : 1.f;
  const float y = rsqrt_approx(nx2s);
  const float n0 = __fmul_rn(nx2s, y);
  const float norm = __fmaf_rn(__fmul_rn(0.5f, y), __fmaf_rn(-n0, n0, nx2s), n0);
  const float beta = -copysignf(norm, alpha);
  const float rden = rcp_approx(__fsub_rn(alpha, beta));
  const float tau = __fmaf_rn(fabsf(alpha), rcp_approx(norm), 1.f);
  h.beta = ok ? beta : 0.f;
  h.rden = ok ? rden : 0.f;
  h.tau = ok ? tau : 2.f;
  return h;
}

template <int NP, int RS, int CL>
struct CsSmem {
  static constexpr int LD = NP + 1, NW = RS * CL / 32;
  static constexpr uint32_t oTile = 0;                         // NP x kPS floats
  static constexpr uint32_t oR = oTile + NP * kPS * 4;         // NP x LD floats, R[j * LD + c]
  static constexpr uint32_t oPart = oR + NP * LD * 4;          // 2 x NP x NW floats
  static constexpr uint32_t oRden = oPart + 2 * NP * NW * 4;   // NP floats
  static constexpr uint32_t kBytes = oRden + NP * 4 + 16;
};


// the 16 Householder steps of panel PJ (columns [16 PJ, 16 PJ + 16)) applied to every column
// >= 16 PJ of the tile. KP = NP / 16 column groups per thread; a[k][i] packs rows 2i, 2i+1.
template <int NW> struct PartVec;
template <> struct PartVec<4> { static __device__ __forceinline__ float sum(const float* p) {
  const float4 v = *reinterpret_cast<const float4*>(p); return __fadd_rn(__fadd_rn(v.x, v.y), __fadd_rn(v.z, v.w)); } };
template <> struct PartVec<2> { static __device__ __forceinline__ float sum(const float* p) {
  const float2 v = *reinterpret_cast<const float2*>(p); return __fadd_rn(v.x, v.y); } };

template <int PJ, int NP, int RS, int CL>
__device__ __forceinline__ void sweep_panel(f32x2 (&a)[NP / CL][64 / RS], float* tileS, float* Rs, float* part, float* rdenv,
                                            int LD, int t) {
  constexpr int KP = NP / CL, RPT = kTileRows / RS, NI = RPT / 2, NW = RS * CL / 32;
  constexpr int K0 = (16 * PJ) / CL;        // first column group that still has live columns
  constexpr int JR0 = 16 * PJ - CL * K0;    // lane of this panel's first pivot inside group K0
  const int cl = t & (CL - 1), s = t / CL, lane = t & 31, warp = t >> 5;
#pragma unroll 1
  for (int jl = 0; jl < 16; ++jl) {
    const int j = 16 * PJ + jl, jr = JR0 + jl;
    __syncwarp();  // column j of this row segment was published by a lane of this warp
    f32x2 vj[NI];
    {
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + j * kPS + RPT * s);
#pragma unroll
      for (int i = 0; i < NI / 2; ++i) { const ulonglong2 v = src[i]; vj[2 * i] = v.x; vj[2 * i + 1] = v.y; }
    }
    // R entries of row j are read before the barrier and rewritten after it
    const float alpha = Rs[j * LD + j];
    float rjc[KP];
#pragma unroll
    for (int k = K0; k < KP; ++k) rjc[k] = Rs[j * LD + cl + CL * k];
    float* pb = part + (jl & 1) * (NP * NW);
#pragma unroll
    for (int k = K0; k < KP; ++k) {
      f32x2 q0 = 0ull, q1 = 0ull;
#pragma unroll
      for (int i = 0; i < NI; i += 2) { q0 = fma2(vj[i], a[k][i], q0); q1 = fma2(vj[i + 1], a[k][i + 1], q1); }
      float p0, p1, p2, p3;
      unpack2(q0, p0, p1);
      unpack2(q1, p2, p3);
      float p = __fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3));
      if (CL == 16) p = __fadd_rn(p, __shfl_xor_sync(kFull, p, 16));
      if (CL == 32 || lane < 16) pb[(cl + CL * k) * NW + warp] = p;
    }
    __syncthreads();
    const HhScal h = hh_scalars(alpha, PartVec<NW>::sum(pb + j * NW));
#pragma unroll
    for (int k = K0; k < KP; ++k) {
      const int c = cl + CL * k;
      const float g = PartVec<NW>::sum(pb + c * NW);
      const float coef = __fmul_rn(h.tau, __fmaf_rn(g, h.rden, rjc[k]));
      const bool active = (k > K0) || (cl > jr);
      const float ncv = active ? -__fmul_rn(coef, h.rden) : 0.f;
      const f32x2 ncv2 = pack2(ncv, ncv);
#pragma unroll
      for (int i = 0; i < NI; ++i) a[k][i] = fma2(ncv2, vj[i], a[k][i]);
      if (s == 0 && active) Rs[j * LD + c] = __fsub_rn(rjc[k], coef);
    }
    if (jl < 15 && cl == jr + 1) {  // publish the next pivot column of this panel (raw, unscaled)
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + (j + 1) * kPS + RPT * s);
#pragma unroll
      for (int i = 0; i < NI / 2; ++i) dst[i] = make_ulonglong2(a[K0][2 * i], a[K0][2 * i + 1]);
    }
    if (s == 0 && cl == jr) { Rs[j * LD + j] = h.beta; rdenv[j] = h.rden; }
  }
  if (16 * (PJ + 1) < NP) {  // first pivot column of the next panel
    constexpr int JN = 16 * (PJ + 1) < NP ? 16 * (PJ + 1) : 0, K1 = JN / CL, L1 = JN % CL;
    if (cl == L1) {
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + JN * kPS + RPT * s);
#pragma unroll
      for (int i = 0; i < NI / 2; ++i) dst[i] = make_ulonglong2(a[K1][2 * i], a[K1][2 * i + 1]);
    }
  }
}

template <int PJ, int NP, int RS, int CL>
struct SweepAll {
  static __device__ __forceinline__ void run(f32x2 (&a)[NP / CL][64 / RS], float* tileS, float* Rs, float* part, float* rdenv, int LD, int t) {
    sweep_panel<PJ, NP, RS, CL>(a, tileS, Rs, part, rdenv, LD, t);
    SweepAll<PJ + 1, NP, RS, CL>::run(a, tileS, Rs, part, rdenv, LD, t);
  }
};
template <int NP, int RS, int CL>
struct SweepAll<NP / 16, NP, RS, CL> {
  static __device__ __forceinline__ void run(f32x2 (&)[NP / CL][64 / RS], float*, float*, float*, float*, int, int) {}
};

// Step J of the head QR with column c in the registers of thread c (compile-time row indices).
template <int J, int NP>
struct HeadSteps {
  static __device__ __forceinline__ void run(float (&h)[NP], float* pv, int t) {
    float* pj = pv + (J & 1) * NP;
    if (t == J) {
#pragma unroll
      for (int r = J; r < NP; ++r) pj[r] = h[r];
    }
    __syncthreads();
    constexpr int R0 = J & ~3;
    // two passes over the shared pivot vector (dots, then update) keep only h[] live in registers
    float g0 = 0.f, g1 = 0.f, d0 = 0.f, d1 = 0.f, alpha = 0.f;
#pragma unroll
    for (int r = R0; r < NP; r += 4) {
      const float4 v = *reinterpret_cast<const float4*>(pj + r);
      const float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (r + i == J) alpha = x[i];
        if (r + i > J) {
          if (i & 1) { g1 = __fmaf_rn(x[i], x[i], g1); d1 = __fmaf_rn(x[i], h[r + i], d1); }
          else { g0 = __fmaf_rn(x[i], x[i], g0); d0 = __fmaf_rn(x[i], h[r + i], d0); }
        }
      }
    }
    const HhScal hs = hh_scalars(alpha, __fadd_rn(g0, g1));
    if (t > J) {
      const float coef = __fmul_rn(hs.tau, __fmaf_rn(__fadd_rn(d0, d1), hs.rden, h[J]));
      const float ncv = -__fmul_rn(coef, hs.rden);
      h[J] = __fsub_rn(h[J], coef);
#pragma unroll
      for (int r = R0; r < NP; r += 4) {
        const float4 v = *reinterpret_cast<const float4*>(pj + r);
        const float x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (r + i > J) h[r + i] = __fmaf_rn(ncv, x[i], h[r + i]);
      }
    } else if (t == J) {
#pragma unroll
      for (int r = J + 1; r < NP; ++r) h[r] = __fmul_rn(h[r], hs.rden);
      h[J] = hs.beta;
    }
    HeadSteps<J + 1, NP>::run(h, pv, t);
  }
};
template <int NP>
struct HeadSteps<NP, NP> {
  static __device__ __forceinline__ void run(float (&)[NP], float*, int) {}
};

template <int NP, int RS, int CL, int MINB, typename TIn, typename TQ>
__global__ void __launch_bounds__(CL * RS, MINB) cs_reduce_kernel(const TcParams P) {
  using L = CsSmem<NP, RS, CL>;
  constexpr int LD = L::LD, KP = NP / CL, kT = CL * RS, RPT = kTileRows / RS, NI = RPT / 2, RT = kTileRows / kT;
  extern __shared__ __align__(128) char sm[];
  if (P.skip_ok && *P.skip_ok == kPcOk) return;
  float* tileS = reinterpret_cast<float*>(sm + L::oTile);
  float* Rs = reinterpret_cast<float*>(sm + L::oR);
  float* part = reinterpret_cast<float*>(sm + L::oPart);
  float* rdenv = reinterpret_cast<float*>(sm + L::oRden);
  const int t = threadIdx.x, cl = t & (CL - 1), s = t / CL;
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
      Hs[c * LD + r] = (row < cend && c < n) ? io_plain(A + row + (long long)c * P.lda) : 0.f;
    }
    __syncthreads();
    // Column c lives in the registers of thread c (fully unrolled steps, compile-time row indices);
    // the pivot column goes through a double-buffered shared vector, one barrier per step, and
    // every thread derives the Householder scalars from it redundantly.
    {
      float* pv = part;  // 2 x NP floats (the partial-sum buffers are idle during the head)
      static_assert(L::NW >= 1, "pivot buffer does not fit");
      const int cc = t < NP ? t : NP - 1;
      float h[NP];
#pragma unroll
      for (int r = 0; r < NP; ++r) h[r] = Hs[cc * LD + r];
      HeadSteps<0, NP>::run(h, pv, t);
      __syncthreads();
      if (t < NP) {
#pragma unroll
        for (int r = 0; r < NP; ++r) Hs[t * LD + r] = h[r];
      }
    }
    __syncthreads();
    for (int idx = t; idx < NP * NP; idx += kT) {  // R0 (upper) -> Rs
      const int c = idx % NP, j = idx / NP;
      Rs[j * LD + c] = (j <= c) ? Hs[c * LD + j] : 0.f;
    }
    for (int r = t; r < NP; r += kT) {  // V0 (strictly lower) -> Q head rows
      if (cbeg + r < cend)
        for (int j = 0; j < n; ++j) io_store(Q + cbeg + r + (long long)j * P.ldq, r > j ? Hs[j * LD + r] : 0.f);
    }
    __syncthreads();
  }

  // ---------------- tiles ----------------
  const long long t0row = cbeg + NP;
  const int ntiles = cend > t0row ? (int)((cend - t0row + kTileRows - 1) / kTileRows) : 0;
  // Tile i+1 is loaded while tile i's reflectors are stored: a thread overwrites exactly the shared-memory
  // words it has just read for the V store, so the two are fused column batch by column batch (the loads are
  // in flight behind the batch's stores) with no barrier in between. STORE = false is the prologue (tile 0).
  constexpr int CB = (NP < 64 && RT == 1) ? 8 : 16;  // columns per batch (8 where the register cap is 85)
  auto store_load = [&](int tile_st, int tile_ld) {
    const long long row_st = t0row + (long long)tile_st * kTileRows + t;  // this thread: rows row + kT rr
    const long long row_ld = t0row + (long long)tile_ld * kTileRows + t;
    const bool do_st = tile_st >= 0, do_ld = tile_ld < ntiles;
#pragma unroll 1
    for (int c0 = 0; c0 < NP; c0 += CB) {
      typename RawOf<TIn>::type v[CB][RT];
      if (do_ld) {
        const TIn* ap = A + row_ld + (long long)c0 * P.lda;
#pragma unroll
        for (int c = 0; c < CB; ++c) {
#pragma unroll
          for (int rr = 0; rr < RT; ++rr) {
            v[c][rr] = 0;
            if (row_ld + kT * rr < cend && c0 + c < n) v[c][rr] = raw_load(ap + (long long)c * P.lda + kT * rr);
          }
        }
      }
      if (do_st) {  // V = x_j * rden_j, every column was published (raw) when it became the pivot
        TQ* qp = Q + row_st + (long long)c0 * P.ldq;
#pragma unroll
        for (int c = 0; c < CB; ++c) {
#pragma unroll
          for (int rr = 0; rr < RT; ++rr)
            if (c0 + c < n && row_st + kT * rr < cend)
              io_store(qp + (long long)c * P.ldq + kT * rr, __fmul_rn(tileS[(c0 + c) * kPS + t + kT * rr], rdenv[c0 + c]));
        }
      }
      if (do_ld) {
#pragma unroll
        for (int c = 0; c < CB; ++c) {
#pragma unroll
          for (int rr = 0; rr < RT; ++rr) tileS[(c0 + c) * kPS + t + kT * rr] = raw_cvt(v[c][rr]);
        }
      }
    }
  };
  if (ntiles > 0) store_load(-1, 0);
  for (int tile = 0; tile < ntiles; ++tile) {
    if (tile + 1 < ntiles) {  // pull the next tile towards L2: one 128-byte line per warp and column
      const int lane = t & 31;
      const long long row = t0row + (long long)tile * kTileRows + t;
#pragma unroll
      for (int rr = 0; rr < RT; ++rr) {
        const long long r0 = row - lane + kTileRows + kT * rr;
        for (int c = lane; c < n; c += 32)
          if (r0 < cend) tc::prefetch_l2(A + r0 + (long long)c * P.lda);
      }
    }
    __syncthreads();
    f32x2 a[KP][NI];
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + (cl + CL * k) * kPS + RPT * s);
#pragma unroll
      for (int i = 0; i < NI / 2; ++i) { const ulonglong2 v = src[i]; a[k][2 * i] = v.x; a[k][2 * i + 1] = v.y; }
    }
    SweepAll<0, NP, RS, CL>::run(a, tileS, Rs, part, rdenv, LD, t);
    __syncthreads();
    store_load(tile, tile + 1);
  }
  __syncthreads();

  // ---------------- chunk result: R -> stacked S ----------------
  for (int idx = t; idx < n * n; idx += kT) {
    const int j = idx % n, c = idx / n;
    P.S[(long long)chunk * n + j + (long long)c * P.lds] = (j <= c) ? Rs[j * LD + c] : 0.f;
  }
}

// Tile shapes (RS row segments x CL column lanes): NP = 64 runs 4 x 16 (64 threads, 32 rows per thread), the
// narrower widths 8 x 16 (128 threads). Measured at 2^24 x 64 on B200: 4x16 7.7 ms, 8x16 8.9 ms, 4x32 (128
// threads, a warp = one 32-row segment, lane = column, no shuffle, 16 warps per SM at 128 registers) 10.1 ms.
// TSQR_B200_CS_SHAPE=8x16 | 4x16 | 4x32 forces one (4x32 needs NP = 64).
int cs_shape(int np) {
  static const int forced = [] {
    const char* e = std::getenv("TSQR_B200_CS_SHAPE");
    if (!e) return 0;
    const std::string v(e);
    return v == "8x16" ? 816 : (v == "4x16" ? 416 : (v == "4x32" ? 432 : 0));
  }();
  if (forced == 432 && np != 64) return 816;
  return forced ? forced : (np == 64 ? 416 : 816);
}

template <int NP, int RS, int CL, int MINB, typename TIn, typename TQ>
void launch_cs_t(const TcParams& p, cudaStream_t s) {
  auto kern = cs_reduce_kernel<NP, RS, CL, MINB, TIn, TQ>;
  const int bytes = (int)CsSmem<NP, RS, CL>::kBytes;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  kern<<<p.nchunks, CL * RS, bytes, s>>>(p);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <typename TIn, typename TQ>
void launch_cs_shape(const TcParams& p, cudaStream_t s) {
  const int np = tc_np(p.n), shape = cs_shape(np);
  if (np == 64) {
    if (shape == 432) launch_cs_t<64, 4, 32, 4, TIn, TQ>(p, s);
    else if (shape == 416) launch_cs_t<64, 4, 16, 4, TIn, TQ>(p, s);
    else launch_cs_t<64, 8, 16, 4, TIn, TQ>(p, s);
  } else if (np == 32) {
    if (shape == 416) launch_cs_t<32, 4, 16, 5, TIn, TQ>(p, s);
    else launch_cs_t<32, 8, 16, 6, TIn, TQ>(p, s);
  } else {
    if (shape == 416) launch_cs_t<16, 4, 16, 8, TIn, TQ>(p, s);
    else launch_cs_t<16, 8, 16, 6, TIn, TQ>(p, s);
  }
}

}  // namespace

void launch_cs_reduce(const TcParams& p, IoType in, IoType q, cudaStream_t s) {
  if (in != q) throw std::runtime_error("tsqr_b200: reduce pass needs matching A and Q element types");
  if (in == IoType::f32) launch_cs_shape<float, float>(p, s);
  else launch_cs_shape<__half, __half>(p, s);
}
int cs_reduce_ctas_per_sm(int np) {
  const int shape = cs_shape(np);
  if (np == 64) return 4;
  if (np == 32) return shape == 416 ? 5 : 6;
  return shape == 416 ? 8 : 6;
}

}  // namespace tsqr_b200
