// Tensor-core TSQR level passes for sm_100a (tcgen05.mma kind::tf32, accumulators in TMEM).
// See ts_pass.h (TcParams) for the chunk format and tc05.cuh for the operand layouts.
//
// What runs where:
//   CUDA cores : the 16-column Householder panel sweep (column norms and reflector dot
//                products = per-thread partial sums over a 16-row segment + one warp shuffle +
//                a 4-warp combine), the 16x16 / NPxNP triangular solves, the NP x NP head QR.
//   tensor core: every genuine GEMM -- G = A_trail^T V_p and A_trail -= V_p W (reduce pass),
//                S = V^T V and Q_tile = -V X (build pass) -- with the hi/lo split of the
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

// ---- reduce-pass shared memory map (bytes) ----------------------------------------------
constexpr uint32_t rMisc = 0;                       // Rd, Rn, Sp (16x17 each), tau, part, bar, tmem slot
constexpr uint32_t rPbuf = 4096;                    // 16 x kPS floats
constexpr uint32_t rAtHi = rPbuf + 16 * kPS * 4;    // 12544: trailing columns 16..63, ns layout
constexpr uint32_t kAtBytes = ns::bytes(128, 48);   // 27648
constexpr uint32_t rAtLo = rAtHi + kAtBytes;
constexpr uint32_t rVcHi = rAtLo + kAtBytes;        // panel reflectors, column layout (Gram operand)
constexpr uint32_t kVcBytes = ns::bytes(128, 16);   // 9216
constexpr uint32_t rVcLo = rVcHi + kVcBytes;
constexpr uint32_t rVrHi = rVcLo + kVcBytes;        // panel reflectors, row layout (update operand)
constexpr uint32_t kVrBytes = rn::bytes(16);        // 8192
constexpr uint32_t rVrLo = rVrHi + kVrBytes;
constexpr uint32_t rWHi = rVrLo + kVrBytes;         // W (16 x 64), ns layout with 16 rows
constexpr uint32_t kWBytes = ns::bytes(16, 64);     // 4608
constexpr uint32_t rWLo = rWHi + kWBytes;
constexpr uint32_t kReduceSmem = rWLo + kWBytes;    // 111872
// The Gram product addresses the trailing buffer by ABSOLUTE column (accumulator row m =
// column m), i.e. from a virtual base two column groups before rAt*, and with M = 128 it reads
// 16 column groups: both ends must stay inside this CTA's shared memory.
static_assert(rAtHi >= 2 * ns::sb(128), "virtual column 0 of the trailing buffer must not be negative");
static_assert(rAtLo - 2 * ns::sb(128) + 15 * ns::sb(128) + 31 * ns::kCM + 128 <= kReduceSmem, "Gram over-read leaves the CTA");

// One 16-column panel sweep of [R_diag ; panel]. Thread (c = t%16, s = t/16) owns rows
// [16s, 16s+16) of panel column c in registers. Columns stay UNSCALED (x_j, not v_j = x_j * rden_j):
// rdenv[j] is applied when the reflector block is written out, off the critical path.
// Out: Rn (new diagonal block of R), Sp[i*16+l] = v_i . v_l (i < l), tau, rdenv.
__device__ __forceinline__ void sweep16(float* Pbuf, const float* Rd, float* Rn, float* Sp, float* tau, float* rdenv,
                                        float* part, int t) {
  const int c = t & 15, s = t >> 4, lane = t & 31, warp = t >> 5;
  float a[16];
  float my_rden = 0.f;
  {
    const float4* src = reinterpret_cast<const float4*>(Pbuf + c * kPS + 16 * s);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float4 v = src[i]; a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w; }
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    __syncwarp();  // column j of this row segment was published by a lane of this half warp
    float vj[16];
    {
      const float4* src = reinterpret_cast<const float4*>(Pbuf + j * kPS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float4 v = src[i]; vj[4 * i] = v.x; vj[4 * i + 1] = v.y; vj[4 * i + 2] = v.z; vj[4 * i + 3] = v.w; }
    }
    const float alpha = Rd[j * 17 + j];
    const float rjc = Rd[j * 17 + c];
    float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i += 4) {
      p0 = __fmaf_rn(vj[i], a[i], p0); p1 = __fmaf_rn(vj[i + 1], a[i + 1], p1);
      p2 = __fmaf_rn(vj[i + 2], a[i + 2], p2); p3 = __fmaf_rn(vj[i + 3], a[i + 3], p3);
    }
    float p = __fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3));
    p = __fadd_rn(p, __shfl_xor_sync(kFull, p, 16));
    float* pb = part + (j & 1) * 64;
    if (lane < 16) pb[c * 4 + warp] = p;
    __syncthreads();
    const float4 gc4 = *reinterpret_cast<const float4*>(pb + c * 4);
    const float4 gj4 = *reinterpret_cast<const float4*>(pb + j * 4);
    const float gc = __fadd_rn(__fadd_rn(gc4.x, gc4.y), __fadd_rn(gc4.z, gc4.w));
    const float gj = __fadd_rn(__fadd_rn(gj4.x, gj4.y), __fadd_rn(gj4.z, gj4.w));
    const HhScal h = hh_scalars(alpha, gj);
    // branch-free: lanes with c <= j run the update with a zero coefficient
    const float coef = __fmul_rn(h.tau, __fmaf_rn(gc, h.rden, rjc));
    const float cv = (c > j) ? __fmul_rn(coef, h.rden) : 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = __fmaf_rn(-cv, vj[i], a[i]);
    if (c == j + 1) {
      float4* dst = reinterpret_cast<float4*>(Pbuf + c * kPS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(a[4 * i], a[4 * i + 1], a[4 * i + 2], a[4 * i + 3]);
    }
    if (c == j) my_rden = h.rden;
    if (s == 0) {
      if (c > j) Rn[j * 17 + c] = __fsub_rn(rjc, coef);
      else if (c == j) { Rn[j * 17 + j] = h.beta; tau[j] = h.tau; rdenv[j] = h.rden; }
      else Sp[c * 16 + j] = __fmul_rn(__fmul_rn(gc, my_rden), h.rden);  // v_c . v_j
    }
  }
  __syncthreads();
}

template <int NP, bool COR, typename TIn, typename TQ>
__global__ void __launch_bounds__(kT, 2) tc_reduce_kernel(const TcParams P) {
  extern __shared__ __align__(128) char sm[];
  if (P.skip_ok && *P.skip_ok == kPcOk) return;
  float* Rd = reinterpret_cast<float*>(sm + rMisc);
  float* Rn = Rd + 272;
  float* Sp = Rn + 272;
  float* tau = Sp + 272;
  float* rdenv = tau + 16;
  float* part = rdenv + 16;  // 128 floats
  float* Tp = Rd;            // 16 x 16, written after the sweep is done with Rd
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 3968);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + 3976);
  float* Pbuf = reinterpret_cast<float*>(sm + rPbuf);
  char* AtHi = sm + rAtHi; char* AtLo = sm + rAtLo;
  char* VcHi = sm + rVcHi; char* VcLo = sm + rVcLo;
  char* VrHi = sm + rVrHi; char* VrLo = sm + rVrLo;
  char* WHi = sm + rWHi;   char* WLo = sm + rWLo;
  const int t = threadIdx.x, warp = t >> 5;
  const int n = P.n;
  const TIn* A = reinterpret_cast<const TIn*>(P.A);
  TQ* Q = reinterpret_cast<TQ*>(P.Q);

  if (warp == 0) tmem_alloc<256>(tmem_slot);
  if (t == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot;
  const uint32_t dA = tm, dG = tm + 64, dR = tm + 128;  // A tile | Gram result | R (lane = column)
  uint32_t phase = 0;

  const int chunk = blockIdx.x;
  const long long cbeg = (long long)chunk * P.rows_per_chunk;
  const long long cend = (chunk == P.nchunks - 1) ? P.M : cbeg + P.rows_per_chunk;

  // ---------------- head: plain Householder QR of the first NP rows ----------------
  {
    constexpr int LD = NP + 1;
    float* Hs = reinterpret_cast<float*>(AtHi);  // Hs[c * LD + r]
    for (int idx = t; idx < NP * NP; idx += kT) {
      const int r = idx % NP, c = idx / NP;
      const long long row = cbeg + r;
      Hs[c * LD + r] = (row < cend && c < n) ? io_plain(A + row + (long long)c * P.lda) : 0.f;
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
    if (warp * 32 < NP) {  // R0 -> TMEM, lane = column
#pragma unroll
      for (int j0 = 0; j0 < NP; j0 += 16) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = (t < NP && j0 + i <= t) ? Hs[t * LD + j0 + i] : 0.f;
        tmem_st16(tmem_addr(dR, warp, j0), v);
      }
    }
    if (t < NP && cbeg + t < cend) {  // V0 (strictly lower) -> Q head rows
      for (int j = 0; j < n; ++j) io_store(Q + cbeg + t + (long long)j * P.ldq, t > j ? Hs[j * LD + t] : 0.f);
    }
    tmem_wait_st();
    __syncthreads();
  }

  // ---------------- tiles ----------------
  const long long t0row = cbeg + NP;
  const int ntiles = cend > t0row ? (int)((cend - t0row + kT - 1) / kT) : 0;
  const uint32_t at_hi_v = smem_u32(AtHi) - 2 * ns::sb(128), at_lo_v = smem_u32(AtLo) - 2 * ns::sb(128);
  for (int tile = 0; tile < ntiles; ++tile) {
    const long long row = t0row + (long long)tile * kT + t;
    const bool valid = row < cend;
    const uint32_t rowoff = ns::off(t, 0, 128);  // row part of an ns offset; columns add compile-time terms
#pragma unroll
    for (int c0 = 0; c0 < NP; c0 += 16) {
      float a[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = (valid && c0 + i < n) ? io_plain(A + row + (long long)(c0 + i) * P.lda) : 0.f;
      if (c0 == 0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) Pbuf[i * kPS + t] = a[i];
      } else {
        tmem_st16(tmem_addr(dA, warp, c0), a);
#pragma unroll
        for (int i = 0; i < 16; ++i) tile_store<COR>(AtHi + rowoff, AtLo + rowoff, ns::coloff(c0 - 16 + i, 128), a[i]);
      }
    }
    if (tile + 1 < ntiles) {  // pull the next tile towards L2: one 128-byte line per warp and column
      const int lane = t & 31;
      const long long r0 = row - lane + kT;
      for (int c = lane; c < n; c += 32)
        if (r0 < cend) prefetch_l2(A + r0 + (long long)c * P.lda);
    }
    if (NP > 16) tmem_wait_st();

#pragma unroll 1
    for (int c0 = 0; c0 < NP; c0 += 16) {
      {  // diagonal block of R for this panel: thread t in [c0, c0+16) owns column t
        float rv[16];
        tmem_ld16(tmem_addr(dR, warp, c0), rv);
        if (t >= c0 && t < c0 + 16) {
#pragma unroll
          for (int i = 0; i < 16; ++i) Rd[i * 17 + (t - c0)] = rv[i];
        }
      }
      __syncthreads();
      sweep16(Pbuf, Rd, Rn, Sp, tau, rdenv, part, t);
      {
        float rv[16];
        tmem_ld16(tmem_addr(dR, warp, c0), rv);
        if (t >= c0 && t < c0 + 16) {
#pragma unroll
          for (int i = 0; i < 16; ++i) rv[i] = (i <= t - c0) ? Rn[i * 17 + (t - c0)] : 0.f;
        }
        tmem_st16(tmem_addr(dR, warp, c0), rv);
      }
      // reflector block (scaled here): both operand layouts, then the global V store while the
      // Gram product runs
      float vout[16];
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        const float4 rd4 = *reinterpret_cast<const float4*>(rdenv + 4 * i4);
        vout[4 * i4 + 0] = __fmul_rn(Pbuf[(4 * i4 + 0) * kPS + t], rd4.x);
        vout[4 * i4 + 1] = __fmul_rn(Pbuf[(4 * i4 + 1) * kPS + t], rd4.y);
        vout[4 * i4 + 2] = __fmul_rn(Pbuf[(4 * i4 + 2) * kPS + t], rd4.z);
        vout[4 * i4 + 3] = __fmul_rn(Pbuf[(4 * i4 + 3) * kPS + t], rd4.w);
        if (c0 + 16 < NP) {
#pragma unroll
          for (int q = 0; q < 4; ++q) tile_store<COR>(VcHi + rowoff, VcLo + rowoff, ns::coloff(4 * i4 + q, 128), vout[4 * i4 + q]);
          rn::store4<COR>(VrHi, VrLo, t, i4, vout[4 * i4], vout[4 * i4 + 1], vout[4 * i4 + 2], vout[4 * i4 + 3]);
        }
      }
      tmem_wait_st();
      if (c0 + 16 < NP) {
        fence_async_smem();
        fence_before_sync();
        __syncthreads();
        if (t == 0) {
          fence_after_sync();
          ns::gram_issue<COR>(dG, at_hi_v, at_lo_v, 128, 0, smem_u32(VcHi), smem_u32(VcLo), 128, 0, 16, 128, false);
          mma_commit(bar);
        }
      }
      if (valid) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (c0 + i < n) io_store(Q + row + (long long)(c0 + i) * P.ldq, vout[i]);
      }
      if (c0 + 16 < NP) {
        // while the Gram product runs: T_p = (striu(Sp) + diag(1/tau))^-1, one row per thread
        // (larft recurrence, a row depends only on its own earlier entries)
        if (warp == 1 && (t & 31) < 16) {
          const int i = t & 31;
          float trow[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float acc = 0.f;
#pragma unroll
            for (int l = 0; l < j; ++l)
              if (l >= i) acc = __fmaf_rn(trow[l], Sp[l * 16 + j], acc);
            trow[j] = (j < i) ? 0.f : (j == i ? tau[j] : __fmul_rn(-tau[j], acc));
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) Tp[i * 16 + j] = trow[j];
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        fence_after_sync();
        __syncthreads();
        {  // W = T_p^T (R[J, c] + G[c, :]) for the trailing column c = t; R[J, c] -= W
          float g[16], rr[16];
          tmem_ld16(tmem_addr(dG, warp, 0), g);
          tmem_ld16(tmem_addr(dR, warp, c0), rr);
          if (t >= c0 + 16 && t < NP) {
            float w[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) w[i] = 0.f;
#pragma unroll
            for (int l = 0; l < 16; ++l) {
              const float rhs = __fadd_rn(rr[l], g[l]);
              const float4* tr = reinterpret_cast<const float4*>(Tp + l * 16);
              float tl[16];
#pragma unroll
              for (int q = 0; q < 4; ++q) { const float4 x = tr[q]; tl[4 * q] = x.x; tl[4 * q + 1] = x.y; tl[4 * q + 2] = x.z; tl[4 * q + 3] = x.w; }
#pragma unroll
              for (int i = l; i < 16; ++i) w[i] = __fmaf_rn(tl[i], rhs, w[i]);
            }
            char* wh = WHi + ns::off(0, t, 16);
            char* wl = WLo + ns::off(0, t, 16);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              rr[i] = __fsub_rn(rr[i], w[i]);
              tile_store<COR>(wh, wl, (uint32_t)((i >> 2) * ns::kCM + (i & 3) * 4), w[i]);
            }
          }
          tmem_st16(tmem_addr(dR, warp, c0), rr);
        }
        tmem_wait_st();
        fence_async_smem();
        fence_before_sync();
        __syncthreads();
        if (t == 0) {
          fence_after_sync();
          rn::update_issue<COR>(dA + c0 + 16, smem_u32(VrHi), smem_u32(VrLo), 0, smem_u32(WHi), smem_u32(WLo), 16,
                                c0 + 16, 0, NP - c0 - 16, 16, true, true);
          mma_commit(bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        fence_after_sync();
        {
          float a[16];
          tmem_ld16(tmem_addr(dA, warp, c0 + 16), a);
#pragma unroll
          for (int i = 0; i < 16; ++i) Pbuf[i * kPS + t] = a[i];
        }
        for (int cc = c0 + 32; cc < NP; cc += 16) {
          float a[16];
          tmem_ld16(tmem_addr(dA, warp, cc), a);
#pragma unroll
          for (int i = 0; i < 16; ++i) tile_store<COR>(AtHi + rowoff + ns::coloff(cc - 16, 128), AtLo + rowoff + ns::coloff(cc - 16, 128), ns::coloff(i, 128), a[i]);
        }
      }
    }
    fence_before_sync();
    __syncthreads();
  }

  // ---------------- chunk result: R -> stacked S ----------------
  if (warp * 32 < NP) {
#pragma unroll
    for (int j0 = 0; j0 < NP; j0 += 16) {
      float rv[16];
      tmem_ld16(tmem_addr(dR, warp, j0), rv);
      if (t < n) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (j0 + i < n) P.S[(long long)chunk * n + j0 + i + (long long)t * P.lds] = (j0 + i <= t) ? rv[i] : 0.f;
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tm);
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

void launch_tc_reduce(const TcParams& p, bool corrected, IoType in, IoType q, cudaStream_t s) {
  if (in != q) throw std::runtime_error("tsqr_b200: tensor-core reduce pass needs matching A and Q element types");
  if (in == IoType::f32) {
    if (corrected) launch_reduce_np<true, float, float>(p, s);
    else launch_reduce_np<false, float, float>(p, s);
  } else {
    if (corrected) launch_reduce_np<true, __half, __half>(p, s);
    else launch_reduce_np<false, __half, __half>(p, s);
  }
}

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
