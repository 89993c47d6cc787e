// Preconditioned-Cholesky TSQR for sm_100a (see pc_tsqr.h for the algorithm).
//
// Kernels (a tile = 128 rows, thread t of a 128-thread tile pipeline = tile row t = TMEM lane t; a CTA = three
// independent tile pipelines, one CTA per SM):
//   pc_sample_kernel : h = 8 NP rows of A -> R_pre (Householder sweep in the (column, row-segment) register layout
//                      of cs_reduce.cu, column norms / reflector dots by shuffles); writes the constant-memory image
//   pc_gram_kernel   : pass 1. TMA (128-byte swizzle) lands the tile in the Gram-operand layout; Z = A R_pre^-1 panel
//                      by panel: 16 x 16 diagonal solves on CUDA cores with packed FFMA2 (no cross-thread traffic:
//                      a row of Z depends on its own row of A only), off-diagonal updates U += Z_p R[p, trail] on
//                      tcgen05 with the A operand READ FROM TENSOR MEMORY (tcgen05.st by the row's thread, hi/lo
//                      split fused into the store) and the R blocks as shared-memory B operands; the thread
//                      overwrites its own elements of the landed tile with Z_hi / Z_lo; G_tile = [Z_hi | Z_lo]^T Z_hi
//                      as one M = 128 product; fp32 -> fp64 accumulation per pipeline
//   pc_finish_kernel : sum of the partials, chol, certificate, R = R_z R_pre, C = R_z^-1 (fp64)
//   pc_apply_kernel  : pass 2. Z again (same code, same bits), Q += Z_p C[16p:16p+16, :] on tcgen05 panel by panel,
//                      Q leaves through shared memory and one TMA store per tile
// Replaces, for the tensor-core modes, reference src/tcqr32x16.cu:1373-1469 (leaf QR), src/tsqr.cu:1102-1172
// (R tree) and src/tsqr.cu:143-1061 (backward products).
#include <cuda.h>
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include "common.cuh"
#include "tc05.cuh"
#include "pc_tsqr.h"
#include "driver.h"

namespace tsqr_b200 {
namespace {
using namespace tc;

constexpr unsigned kFull = 0xffffffffu;

struct PcParams {
  const void* A; long long lda; long long M; int n;
  void* Q; long long ldq;
  const float* Rpre;     // NP x NP, ld NP, upper, identity-padded beyond n
  const float* C;        // pass 2
  double* Gpart;         // pass 1
  int* status;
  int tiles, tiles_per_cta, use_tma, use_tma_q;
};

// Image of R_pre's diagonal blocks for the forward substitution, in CONSTANT memory: [k][16] = -R[k][16 (k/16) + j]
// (zero up to the diagonal), then 1 / R[k][k] at kImgInv + k. Every thread of a tile uses the same values: as
// constant-bank operands of the FFMAs they cost no load instruction and no L1TEX bandwidth (as 160 broadcast
// LDS.128 per thread and tile they kept the shared-memory pipe 70 % busy and bounded both passes). The image is
// written per call by a stream-ordered device-to-device copy; kImgSlots slots, guarded by events, keep calls on
// different streams apart; the kernels are instantiated per slot so that the operands are immediate addresses.
constexpr int kApTmemPerWg = 160;   // TMEM columns per tile pipeline, 144 used: U [0, 48), D or G [48, 112), Z hi [112, 128), Z lo [128, 144)
constexpr int kImgInv = 64 * 16;
constexpr int kImgFloats = kImgInv + 64;
constexpr int kImgSlots = 3;
__constant__ float c_img0[kImgFloats];
__constant__ float c_img1[kImgFloats];
__constant__ float c_img2[kImgFloats];
template <int SLOT>
__device__ __forceinline__ const float* cimg() {
  if constexpr (SLOT == 0) return c_img0;
  else if constexpr (SLOT == 1) return c_img1;
  else return c_img2;
}
// global staging copy of the image (source of the cudaMemcpyToSymbolAsync)
__device__ __forceinline__ void cimg_store(float* img, int NP, int k, int c, float v) {
  if ((k >> 4) == (c >> 4)) {
    img[k * 16 + (c & 15)] = (c > k) ? -v : 0.f;
    if (c == k) img[kImgInv + k] = __fdiv_rn(1.f, v);
  }
}
__global__ void pc_cimg_kernel(const float* __restrict__ Rpre, float* __restrict__ img, int NP) {
  for (int idx = threadIdx.x; idx < NP * NP; idx += blockDim.x) cimg_store(img, NP, idx % NP, idx / NP, Rpre[idx]);
}

__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// ======================================================================================= sample
// Householder QR of the h x NP sample with structured reflectors [e_j ; v_j] on top of a zero R (only R is
// wanted: it is a preconditioner, the Cholesky step absorbs whatever it misses). Thread (cl = t % 16,
// s = t / 16) owns rows [16 s, 16 s + 16) of the columns cl + 16 k.
struct HhScal { float beta, tau, rden; };
__device__ __forceinline__ HhScal hh_scalars(float alpha, float g) {
  HhScal h;
  const float nx2 = __fmaf_rn(alpha, alpha, g);
  const bool ok = nx2 > 1e-36f;
  const float nx2s = ok ? nx2 : 1.f;
  const float y = rsqrt_approx(nx2s);
  const float n0 = __fmul_rn(nx2s, y);
  const float norm = __fmaf_rn(__fmul_rn(0.5f, y), __fmaf_rn(-n0, n0, nx2s), n0);
  const float beta = -copysignf(norm, alpha);
  h.beta = ok ? beta : 0.f;
  h.rden = ok ? rcp_approx(__fsub_rn(alpha, beta)) : 0.f;
  h.tau = ok ? __fmaf_rn(fabsf(alpha), rcp_approx(norm), 1.f) : 2.f;
  return h;
}
template <int NP>
struct SmpCfg {
  static constexpr int H = 8 * NP, T = 8 * NP, NW = T / 32, PS = H + 4, LD = NP + 1, KP = NP / 16;
  static constexpr int PLD = NP + 4;     // row stride of the per-warp partials: stage-2 reads are at most 2-way conflicted
  static constexpr int GW = NW < 8 ? NW : 8;   // threads that share one column in stage 2, each sums NW / GW partials
  static constexpr uint32_t oTile = 0;
  static constexpr uint32_t oR = oTile + NP * PS * 4;
  static constexpr uint32_t oPart = oR + NP * LD * 4;            // [NW][PLD] partials
  static constexpr uint32_t oTot = oPart + NW * PLD * 4;         // 2 x [NP] column totals (double buffered)
  static constexpr uint32_t kBytes = oTot + 2 * NP * 4 + 16;
};

// One Householder step needs, for every live column c, g_c = x_j . a_c over all H rows. Each warp owns 32 rows:
// stage 1 = per-thread partial (16 rows) + one shuffle, written as part[warp][c] (consecutive lanes, consecutive
// words); stage 2 = thread (c = t / 8, g = t % 8) adds NW / 8 of the warp partials of column c, three shuffles
// finish the sum, tot[c] is published for everybody. Two block barriers per step, no bank conflict worth the name.
template <int PJ, int NP>
__device__ __forceinline__ void smp_sweep_panel(f32x2 (&a)[NP / 16][8], float* tileS, float* Rs, float* part, float* tot, int t) {
  using Cf = SmpCfg<NP>;
  constexpr int KP = Cf::KP, NW = Cf::NW, PS = Cf::PS, LD = Cf::LD, PLD = Cf::PLD, GW = Cf::GW, K0 = PJ;
  const int cl = t & 15, s = t >> 4, lane = t & 31, warp = t >> 5;
  const int c2 = t / GW, g2 = t % GW;
#pragma unroll 1
  for (int jl = 0; jl < 16; ++jl) {
    const int j = 16 * PJ + jl;
    __syncwarp();  // column j of this row segment was published by a lane of the same half warp
    f32x2 vj[8];
    {
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + j * PS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const ulonglong2 v = src[i]; vj[2 * i] = v.x; vj[2 * i + 1] = v.y; }
    }
    const float alpha = Rs[j * LD + j];
    float rjc[KP];
#pragma unroll
    for (int k = K0; k < KP; ++k) rjc[k] = Rs[j * LD + cl + 16 * k];
#pragma unroll
    for (int k = K0; k < KP; ++k) {
      f32x2 q0 = 0ull, q1 = 0ull;
#pragma unroll
      for (int i = 0; i < 8; i += 2) { q0 = fma2(vj[i], a[k][i], q0); q1 = fma2(vj[i + 1], a[k][i + 1], q1); }
      float p0, p1, p2, p3;
      unpack2(q0, p0, p1);
      unpack2(q1, p2, p3);
      float p = __fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3));
      p = __fadd_rn(p, __shfl_xor_sync(kFull, p, 16));
      if (lane < 16) part[warp * PLD + cl + 16 * k] = p;
    }
    __syncthreads();
    float* tb = tot + (jl & 1) * NP;
    {
      float q = 0.f;
      if (c2 < NP && c2 >= 16 * K0) {
#pragma unroll
        for (int w = 0; w < NW / GW; ++w) q = __fadd_rn(q, part[(g2 * (NW / GW) + w) * PLD + c2]);
      }
#pragma unroll
      for (int o = GW / 2; o > 0; o >>= 1) q = __fadd_rn(q, __shfl_xor_sync(kFull, q, o));
      if (g2 == 0 && c2 < NP) tb[c2] = q;
    }
    __syncthreads();
    const HhScal h = hh_scalars(alpha, tb[j]);
#pragma unroll
    for (int k = K0; k < KP; ++k) {
      const int c = cl + 16 * k;
      const float g = tb[c];
      const float coef = __fmul_rn(h.tau, __fmaf_rn(g, h.rden, rjc[k]));
      const bool active = (k > K0) || (cl > jl);
      const float ncv = active ? -__fmul_rn(coef, h.rden) : 0.f;
      const f32x2 ncv2 = pack2(ncv, ncv);
#pragma unroll
      for (int i = 0; i < 8; ++i) a[k][i] = fma2(ncv2, vj[i], a[k][i]);
      if (s == 0 && active) Rs[j * LD + c] = __fsub_rn(rjc[k], coef);
    }
    if (jl < 15 && cl == jl + 1) {  // publish the next pivot column of this panel
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + (j + 1) * PS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_ulonglong2(a[K0][2 * i], a[K0][2 * i + 1]);
    }
    if (s == 0 && cl == jl) Rs[j * LD + j] = h.beta;
  }
  if (PJ + 1 < KP) {  // first pivot column of the next panel
    constexpr int K1 = PJ + 1 < KP ? PJ + 1 : 0;
    if (cl == 0) {
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(tileS + (16 * K1) * PS + 16 * s);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_ulonglong2(a[K1][2 * i], a[K1][2 * i + 1]);
    }
  }
}
template <int PJ, int NP>
struct SmpSweepAll {
  static __device__ __forceinline__ void run(f32x2 (&a)[NP / 16][8], float* tileS, float* Rs, float* part, float* tot, int t) {
    smp_sweep_panel<PJ, NP>(a, tileS, Rs, part, tot, t);
    SmpSweepAll<PJ + 1, NP>::run(a, tileS, Rs, part, tot, t);
  }
};
template <int NP>
struct SmpSweepAll<NP / 16, NP> {
  static __device__ __forceinline__ void run(f32x2 (&)[NP / 16][8], float*, float*, float*, float*, int) {}
};

template <int NP, typename TIo>
__global__ void __launch_bounds__(8 * NP, 1) pc_sample_kernel(const TIo* __restrict__ A, long long lda, long long M, int n,
                                                             float* __restrict__ Rpre, float* __restrict__ img,
                                                             int* __restrict__ status) {
  using Cf = SmpCfg<NP>;
  constexpr int H = Cf::H, T = Cf::T, PS = Cf::PS, LD = Cf::LD, KP = Cf::KP;
  extern __shared__ __align__(1024) char sm[];
  float* tileS = reinterpret_cast<float*>(sm + Cf::oTile);
  float* Rs = reinterpret_cast<float*>(sm + Cf::oR);
  float* part = reinterpret_cast<float*>(sm + Cf::oPart);
  float* tot = reinterpret_cast<float*>(sm + Cf::oTot);
  const int t = threadIdx.x, cl = t & 15, s = t >> 4;
  const long long clk0 = clock64();
  if (t == 0) { status[0] = 0; status[1] = 0; }
  {
    // sample row t = row (t % 16) of cluster t / 16; the H / 16 clusters of 16 consecutive rows (one 64-byte piece
    // per column) are spread evenly over the matrix. All NP loads of a thread are issued before the first store.
    const long long cl0 = (((long long)(t >> 4) * (M - 16)) / (H / 16 - 1 > 0 ? H / 16 - 1 : 1));
    const long long row = cl0 + (t & 15);
    const TIo* ap = A + row;
    float v[NP];
#pragma unroll
    for (int c = 0; c < NP; ++c) v[c] = (c < n) ? io_load(ap + (long long)c * lda) : 0.f;
#pragma unroll
    for (int c = 0; c < NP; ++c) tileS[c * PS + t] = v[c];
  }
  for (int i = t; i < NP * LD; i += T) Rs[i] = 0.f;
  __syncthreads();
  const long long clk1 = clock64();
  f32x2 a[KP][8];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(tileS + (cl + 16 * k) * PS + 16 * s);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const ulonglong2 v = src[i]; a[k][2 * i] = v.x; a[k][2 * i + 1] = v.y; }
  }
  SmpSweepAll<0, NP>::run(a, tileS, Rs, part, tot, t);
  __syncthreads();
  if (t == 0) { status[8] = (int)(clk1 - clk0); status[9] = (int)(clock64() - clk1); }   // phase clocks (cycles): gather, sweep
  for (int i = t; i < NP * NP; i += T) {
    const int k = i % NP, c = i / NP;
    float v = 0.f;
    if (k < n && c < n) v = (k <= c) ? Rs[k * LD + c] : 0.f;
    else if (k == c) v = 1.f;
    Rpre[k + c * NP] = v;
    cimg_store(img, NP, k, c, v);
  }
}

// z R_pp = y for one 16-column panel of one matrix row (forward substitution along the row), shared by both passes so
// that they produce bit-identical Z. ci = constant-memory image of R_pre (-R_pp strictly above the diagonal, zero
// elsewhere; 1 / diag at kImgInv). Adjacent columns (y_2k, y_2k+1) are updated by ONE packed fma.rn.f32x2 whose
// operands are z (register, broadcast form) and the pair (c_2k, c_2k+1) (uniform register pair loaded by LDCU): 64
// FFMA2 + 16 FMUL per panel instead of 120 FFMA + 16 FMUL, no load instruction, IEEE per element (same bits as the
// scalar form). For an even pivot i the pair (i, i+1) is included: its first constant is zero, fma(z, 0, z) = z.
__device__ __forceinline__ void solve_panel16(float (&y)[16], const float* __restrict__ ci, int p) {
  f32x2 yp[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) yp[k] = pack2(y[2 * k], y[2 * k + 1]);
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float lo, hi;
    unpack2(yp[i / 2], lo, hi);
    const float z = __fmul_rn((i & 1) ? hi : lo, ci[kImgInv + 16 * p + i]);
    yp[i / 2] = (i & 1) ? pack2(lo, z) : pack2(z, hi);
    const f32x2 zz = pack2(z, z);
#pragma unroll
    for (int j2 = (i + 1) / 2; j2 < 8; ++j2) {
      if (2 * j2 + 1 <= i) continue;
      const f32x2 cc = pack2(ci[(16 * p + i) * 16 + 2 * j2], ci[(16 * p + i) * 16 + 2 * j2 + 1]);
      yp[j2] = fma2(zz, cc, yp[j2]);
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) unpack2(yp[k], y[2 * k], y[2 * k + 1]);
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 addp2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// ======================================================================================= passes
template <int NP>
__host__ __device__ constexpr uint32_t rblk_off(int p) {   // byte offset of R block p (rows [16p, 16p+16), cols >= 16(p+1))
  uint32_t o = 0;
  for (int q = 0; q < p; ++q) o += kb::bytes(16, NP - 16 * (q + 1));
  return o;
}
// ======================================================================================= pass 1
// G = Z^T Z, Z = A R_pre^-1, with THREE tiles in flight per SM (one CTA of 3 warp groups per SM, each an independent
// 128-row tile pipeline that works panel by panel like pc_apply_kernel). The Gram operand of a tile -- Z_hi and Z_lo,
// K-major with K = tile rows, 128-byte hardware swizzle: 64 KB -- is also the LANDING buffer of the tile: the TMA
// (tensor map with SWIZZLE_128B, one box of 32 rows x NP columns per K block) drops the raw rows of A exactly where
// the operand layout wants column c of rows [32 kb, 32 kb + 32), every thread reads its own row from there, solves it
// panel by panel and overwrites ITS OWN elements with Z_hi (Z_lo goes to line NP + c). The next tile may land as
// soon as the Gram product of this one has been read by the tensor core. U += Z_p R[p, trail] takes Z_p from tensor
// memory as in pass 2. Per tile: 144 TMEM columns (U 48, G 64, Z_p 32), 64 KB of shared memory.
constexpr int kGrWg = 3;
template <int NP>
struct GramSmem {
  static constexpr int NPAN = NP / 16;
  static constexpr uint32_t kRblk = rblk_off<NP>(NPAN - 1);
  static constexpr uint32_t oRblkHi = 0;
  static constexpr uint32_t oRblkLo = oRblkHi + kRblk;
  static constexpr uint32_t oBar = oRblkLo + kRblk;          // per warp group: full, mma, big; then the tmem slot
  static constexpr uint32_t kBarBytes = 128;
  static constexpr uint32_t kOp = sw::bytes(128);             // 65536
  static constexpr uint32_t oWg = (oBar + kBarBytes + 1023) / 1024 * 1024;
  static constexpr uint32_t kBytes = oWg + kGrWg * kOp;
};

template <int NP, bool COR, typename TIo, int SLOT>
__global__ void __launch_bounds__(128 * kGrWg, 1) pc_gram_kernel(const PcParams P, const __grid_constant__ CUtensorMap tmapG) {
  using Sm = GramSmem<NP>;
  constexpr int NPAN = NP / 16;
  extern __shared__ __align__(1024) char sm[];
  const int tid = threadIdx.x, wg = tid >> 7, t = tid & 127, warp = (tid >> 5) & 3;
  char* RbHi = sm + Sm::oRblkHi;
  char* RbLo = sm + Sm::oRblkLo;
  const float* ci = cimg<SLOT>();
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + Sm::oBar) + wg * 3;
  uint64_t* bar_full = bars;
  uint64_t* bar_mma = bars + 1;
  uint64_t* bar_big = bars + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + Sm::oBar + Sm::kBarBytes - 8);
  char* op = sm + Sm::oWg + wg * Sm::kOp;
  const int n = P.n;
  const TIo* A = reinterpret_cast<const TIo*>(P.A);
  // Corrected modes: the raw tile lands IN the operand (swizzled, fp32 only) once the previous Gram product has been
  // read. Uncorrected modes: the lo lines of the operand are unused -- 8 KB per K block -- and serve as a separate
  // landing buffer (plain 128 x 16 panel boxes, fp32 or fp16: panel p behind K block p), so the next tile is
  // requested as soon as this tile's raw values have been read and lands while this tile is being worked on.
  constexpr bool kLandLo = !COR;
  const bool use_tma = (P.use_tma != 0) && (kLandLo || sizeof(TIo) == 4);

  if ((smem_u32(sm) & 1023u) != 0) __trap();   // the swizzled operand needs 1024-byte aligned atoms
  if (tid < 32) tmem_alloc<512>(tmem_slot);
  if (t == 0) {
    mbar_init(bar_full, 1);
    mbar_init(bar_mma, 1);
    mbar_init(bar_big, 1);
    fence_mbar_init();
    if (use_tma && wg == 0) tma_prefetch_desc(&tmapG);
  }
  for (int idx = tid; idx < NP * NP; idx += 128 * kGrWg) {   // operand image of R_pre's off-diagonal blocks
    const int k = idx % NP, c = idx / NP;
    const int pk = k >> 4, pc = c >> 4;
    if (pc > pk) tile_store<COR>(RbHi, RbLo, rblk_off<NP>(pk) + kb::off(k & 15, c - 16 * (pk + 1), 16), P.Rpre[idx]);
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot + (uint32_t)(kApTmemPerWg * wg);
  const uint32_t tmU = tm, tmG = tm + 48, tmZh = tm + 112, tmZl = tm + 128;
  // this warp group's tiles: every kGrWg-th tile of the CTA's range
  const int tile_beg = blockIdx.x * P.tiles_per_cta;
  const int tile_end = min(tile_beg + P.tiles_per_cta, P.tiles);
  const int ntl_cta = max(tile_end - tile_beg, 0);
  const int ntl = ntl_cta > wg ? (ntl_cta - wg + kGrWg - 1) / kGrWg : 0;
  auto tile_of = [&](int i) { return tile_beg + wg + kGrWg * i; };
  auto issue_tile = [&](int i) {
    mbar_arrive_expect_tx(bar_full, 128u * NP * (uint32_t)sizeof(TIo));
    if (kLandLo) {   // NPAN boxes of 128 rows x 16 columns behind the hi lines of K block p
#pragma unroll
      for (int p = 0; p < NPAN; ++p) tma_load_2d(op + p * sw::kBlock + 8192, &tmapG, bar_full, tile_of(i) * kPcTileRows, 16 * p);
    } else {         // 4 boxes of 32 rows x NP columns: K block kb, lines [0, NP)
#pragma unroll
      for (int kb4 = 0; kb4 < 4; ++kb4) tma_load_2d(op + kb4 * sw::kBlock, &tmapG, bar_full, tile_of(i) * kPcTileRows + 32 * kb4, 0);
    }
  };
  if (use_tma && t == 0 && ntl > 0) issue_tile(0);
  uint32_t ph_full = 0, ph_mma = 0, ph_big = 0;
  const uint64_t dRh = smem_desc(smem_u32(RbHi), 128, 512), dRl = smem_desc(smem_u32(RbLo), 128, 512);
  float acc[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) acc[j] = 0.f;
  bool first_flush = true;
  const int glanes = (COR ? 2 : 1) * NP;
  double* gp = P.Gpart + ((size_t)blockIdx.x * kGrWg + wg) * NP * 128 + t;
  auto flush = [&]() {   // fp32 sums of at most 16 tiles -> the warp group's fp64 partial
    if (t < glanes) {
#pragma unroll
      for (int j = 0; j < NP; ++j) {
        const double prev = first_flush ? 0.0 : gp[(size_t)j * 128];
        gp[(size_t)j * 128] = prev + (double)acc[j];
      }
    }
#pragma unroll
    for (int j = 0; j < NP; ++j) acc[j] = 0.f;
    first_flush = false;
  };
  auto drain = [&]() {   // accumulator (lane = operand line: Z_hi columns, then Z_lo columns) -> registers
#pragma unroll
    for (int c0 = 0; c0 < NP; c0 += 16) {
      float g[16];
      tmem_ld16(tmem_addr(tmG, warp, c0), g);
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        const f32x2 d = addp2(pack2(acc[c0 + j], acc[c0 + j + 1]), pack2(g[j], g[j + 1]));
        unpack2(d, acc[c0 + j], acc[c0 + j + 1]);
      }
    }
  };
  // this thread's row inside the operand: K block t / 32, 16-byte chunk (t % 32) / 4 (XOR-swizzled with line % 8)
  char* myrow = op + (uint32_t)(t >> 5) * sw::kBlock + (uint32_t)((t & 3) << 2);
  const uint32_t q = (uint32_t)((t & 31) >> 2);
  auto stage = [&](int p, const float (&zh)[16], const float (&zl)[16]) {   // Z_hi at line c, Z_lo at line NP + c
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int c = 16 * p + i;
      const uint32_t o = (uint32_t)c * 128u + ((q ^ (uint32_t)(c & 7)) << 4);
      *reinterpret_cast<float*>(myrow + o) = zh[i];
      if (COR) *reinterpret_cast<float*>(myrow + o + (uint32_t)NP * 128u) = zl[i];
    }
  };
  auto issue_u = [&](int p) {   // thread 0: U[:, 16(p+1):] (+)= Z_p R[p, 16(p+1):], Z_p in tensor memory
    fence_after_sync();
    const int N = NP - 16 * (p + 1);
    const uint32_t id = idesc_tf32(128, N, false, false, false);
    const uint32_t d = tmU + 16 * p;
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      const uint64_t dh = dRh + (uint64_t)((rblk_off<NP>(p) + kk * 256) >> 4), dl = dRl + (uint64_t)((rblk_off<NP>(p) + kk * 256) >> 4);
      mma_tf32_ts(d, tmZh + 8 * kk, dh, id, p > 0 || kk > 0);
      if (COR) {
        mma_tf32_ts(d, tmZh + 8 * kk, dl, id, true);
        mma_tf32_ts(d, tmZl + 8 * kk, dh, id, true);
      }
    }
    mma_commit(bar_mma);
  };

  for (int it = 0; it < ntl; ++it) {
    const int tile = tile_of(it);
    const long long row = (long long)tile * kPcTileRows + t;
    const bool valid = row < P.M;
    if (use_tma) {
      mbar_wait(bar_full, ph_full);
      ph_full ^= 1;
    }
#pragma unroll
    for (int p = 0; p < NPAN; ++p) {
      float y[16];
      if (use_tma && kLandLo) {
        const TIo* land = reinterpret_cast<const TIo*>(op + p * sw::kBlock + 8192);
#pragma unroll
        for (int i = 0; i < 16; ++i) y[i] = io_plain(land + i * 128 + t);
      } else if (use_tma) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int c = 16 * p + i;
          y[i] = *reinterpret_cast<const float*>(myrow + (uint32_t)c * 128u + ((q ^ (uint32_t)(c & 7)) << 4));
        }
      } else {
        const TIo* ap = A + row + (long long)(16 * p) * P.lda;
#pragma unroll
        for (int i = 0; i < 16; ++i) y[i] = (valid && 16 * p + i < n) ? io_load(ap + (long long)i * P.lda) : 0.f;
      }
      if (p > 0) {   // y_p -= sum_{q<p} Z_q R[q, p]
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
        float u[16];
        tmem_ld16(tmem_addr(tmU, warp, 16 * (p - 1)), u);
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const f32x2 d = sub2(pack2(y[i], y[i + 1]), pack2(u[i], u[i + 1]));
          unpack2(d, y[i], y[i + 1]);
        }
      }
      solve_panel16(y, ci, p);
      float zh[16], zl[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) { zh[i] = to_tf32_rna(y[i]); zl[i] = __fsub_rn(y[i], zh[i]); }
      stage(p, zh, zl);
      if (p < NPAN - 1) {
        tmem_st16(tmem_addr(tmZh, warp, 0), zh);
        if (COR) tmem_st16(tmem_addr(tmZl, warp, 0), zl);
        tmem_wait_st();
        fence_before_sync();
        named_bar_sync(1 + wg, 128);
        if (t == 0) issue_u(p);
      }
    }
    // G_tile = [Z_hi | Z_lo]^T Z_hi: one M = 128 product over the 128 rows
    fence_async_smem();
    fence_before_sync();
    named_bar_sync(1 + wg, 128);
    if (t == 0) {
      fence_after_sync();
      sw::gram_issue(tmG, smem_u32(op), NP, 128);
      mma_commit(bar_big);
      // separate landing buffer: every thread has read its raw values (barrier above). The M = 128 product also
      // reads the lo lines, as rows of the accumulator nobody looks at: what it finds there does not matter.
      if (kLandLo && use_tma && it + 1 < ntl) issue_tile(it + 1);
    }
    mbar_wait(bar_big, ph_big);   // the tensor core has read the operand: the next tile may land on it
    ph_big ^= 1;
    fence_after_sync();
    if (!kLandLo && use_tma && t == 0 && it + 1 < ntl) issue_tile(it + 1);
    drain();
    if ((it & 15) == 15) flush();
  }
  flush();   // warp groups without tiles write zeros: the finish kernel sums every slot
  fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc<512>(*tmem_slot);
}

// ======================================================================================= pass 2
// Q = (A R_pre^-1) C with THREE tiles in flight per SM: one CTA of 3 warp groups per SM, each warp group an
// independent 128-row tile pipeline that works panel by panel (16 columns):
//   * the tile lands panel by panel (TMA boxes of 128 x 16 into a ring of kApSlots slots per warp group);
//   * the solved panel Z_p goes to TENSOR MEMORY (tcgen05.st, hi | lo: 32 columns, reused panel after panel) and is
//     the A operand of both products of the panel: U += Z_p R[p, trail] (substitution update) and
//     D += Z_p C[16p : 16p+16, :] (Q accumulates panel by panel, nothing of Z outlives its panel). With Z_p staged
//     in shared memory instead (measured) every corrected product re-reads a 4 KB operand slice: 400 KB of
//     shared-memory traffic per tile, which bounded the kernel; from tensor memory the A operand costs nothing;
//   * a tile needs 144 TMEM columns (U 48, D 64, Z_p 32): three tiles per SM;
//   * Q leaves through shared memory and ONE TMA store per tile (no per-column address arithmetic, no 64 strided
//     stores per thread); when ldq does not allow a tensor map the threads store their rows themselves.
constexpr int kApSlots = 3;
constexpr int kApWg = 3;
template <int NP, typename TIo>
struct ApplySmem {
  static constexpr int NPAN = NP / 16;
  static constexpr uint32_t kC = 2 * kb::bytes(NP, NP);
  static constexpr uint32_t kRblk = rblk_off<NP>(NPAN - 1);
  static constexpr uint32_t oC = 0;
  static constexpr uint32_t oRblkHi = oC + kC;
  static constexpr uint32_t oRblkLo = oRblkHi + kRblk;
  static constexpr uint32_t oBar = oRblkLo + kRblk;                      // per warp group: full[S], empty[S], mma; then the tmem slot
  static constexpr uint32_t kBarBytes = 256;
  static constexpr uint32_t kSlot = 128u * 16u * sizeof(TIo);           // one landed panel
  static constexpr uint32_t kQst = 128u * NP * sizeof(TIo);             // Q tile staging for the TMA store, [column][row]
  static constexpr uint32_t kWg = kApSlots * kSlot + kQst;
  static constexpr uint32_t oWg = (oBar + kBarBytes + 1023) / 1024 * 1024;
  static constexpr uint32_t kBytes = oWg + kApWg * kWg;
};

template <int NP, bool COR, typename TIo, int SLOT>
__global__ void __launch_bounds__(128 * kApWg, 1) pc_apply_kernel(const PcParams P, const __grid_constant__ CUtensorMap tmapP,
                                                                 const __grid_constant__ CUtensorMap tmapQ) {
  using Sm = ApplySmem<NP, TIo>;
  constexpr int NPAN = NP / 16, S = kApSlots;
  extern __shared__ __align__(1024) char sm[];
  if (*P.status != kPcOk) return;   // the Householder passes behind this launch do the work
  const int tid = threadIdx.x, wg = tid >> 7, t = tid & 127, warp = (tid >> 5) & 3;
  char* CHi = sm + Sm::oC;
  char* CLo = CHi + kb::bytes(NP, NP);
  char* RbHi = sm + Sm::oRblkHi;
  char* RbLo = sm + Sm::oRblkLo;
  const float* ci = cimg<SLOT>();
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + Sm::oBar) + wg * (2 * S + 1);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + S;
  uint64_t* bar_mma = bars + 2 * S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + Sm::oBar + Sm::kBarBytes - 8);
  char* wgbase = sm + Sm::oWg + wg * Sm::kWg;
  TIo* qst = reinterpret_cast<TIo*>(wgbase + S * Sm::kSlot);
  const int n = P.n;
  const TIo* A = reinterpret_cast<const TIo*>(P.A);
  TIo* Q = reinterpret_cast<TIo*>(P.Q);
  const bool use_tma = P.use_tma != 0, tma_q = P.use_tma_q != 0;

  if (tid < 32) tmem_alloc<512>(tmem_slot);
  if (t == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, 128); }
    mbar_init(bar_mma, 1);
    fence_mbar_init();
    if (wg == 0) {
      if (use_tma) tma_prefetch_desc(&tmapP);
      if (tma_q) tma_prefetch_desc(&tmapQ);
    }
  }
  for (int idx = tid; idx < NP * NP; idx += 128 * kApWg) {   // operand images of R_pre (off-diagonal blocks) and C
    const int k = idx % NP, c = idx / NP;
    const float v = P.Rpre[idx];
    const int pk = k >> 4, pc = c >> 4;
    if (pc > pk) tile_store<COR>(RbHi, RbLo, rblk_off<NP>(pk) + kb::off(k & 15, c - 16 * (pk + 1), 16), v);
    tile_store<COR>(CHi, CLo, kb::off(k, c, NP), P.C[idx]);
  }
  fence_async_smem();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = *tmem_slot + (uint32_t)(kApTmemPerWg * wg);
  const uint32_t tmU = tm, tmD = tm + 48, tmZh = tm + 112, tmZl = tm + 128;   // U column of tile column c (>= 16): tmU + c - 16
  // this warp group's tiles: every kApWg-th tile of the CTA's range, walked backwards (the rows this CTA read last
  // in pass 1 are the likeliest to be in L2)
  const int tile_beg = blockIdx.x * P.tiles_per_cta;
  const int tile_end = min(tile_beg + P.tiles_per_cta, P.tiles);
  const int ntl_cta = max(tile_end - tile_beg, 0);
  const int ntl = ntl_cta > wg ? (ntl_cta - wg + kApWg - 1) / kApWg : 0;
  auto tile_of = [&](int i) { return tile_end - 1 - (wg + kApWg * i); };
  const int npanels = ntl * NPAN;
  auto issue_load = [&](int g) {   // panel g of this warp group's stream -> slot g % S
    const int slot = g % S;
    mbar_arrive_expect_tx(bar_full + slot, Sm::kSlot);
    tma_load_2d(wgbase + slot * Sm::kSlot, &tmapP, bar_full + slot, tile_of(g / NPAN) * kPcTileRows, 16 * (g % NPAN));
  };
  if (use_tma && t == 0)
    for (int g = 0; g < S && g < npanels; ++g) issue_load(g);
  uint32_t ph_mma = 0;
  // B operand descriptors: fixed per CTA, everything else is a compile-time delta of the address field
  const uint64_t dRh = smem_desc(smem_u32(RbHi), 128, 512), dRl = smem_desc(smem_u32(RbLo), 128, 512);
  const uint64_t dCh = smem_desc(smem_u32(CHi), 128, kb::sbo(NP)), dCl = smem_desc(smem_u32(CLo), 128, kb::sbo(NP));

  for (int it = 0; it < ntl; ++it) {
    const int tile = tile_of(it);
    const long long row = (long long)tile * kPcTileRows + t;
    const bool valid = row < P.M;
#pragma unroll
    for (int p = 0; p < NPAN; ++p) {
      const int g = it * NPAN + p, slot = g % S;
      float y[16];
      if (use_tma) {
        mbar_wait(bar_full + slot, (uint32_t)(g / S) & 1u);
        const TIo* land = reinterpret_cast<const TIo*>(wgbase + slot * Sm::kSlot);
#pragma unroll
        for (int i = 0; i < 16; ++i) y[i] = io_plain(land + i * 128 + t);
      } else {
        const TIo* ap = A + row + (long long)(16 * p) * P.lda;
#pragma unroll
        for (int i = 0; i < 16; ++i) y[i] = (valid && 16 * p + i < n) ? io_plain(ap + (long long)i * P.lda) : 0.f;
      }
      if (p > 0) {   // y_p -= sum_{q<p} Z_q R[q, p]; all products of panel p-1 are complete after this wait
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        fence_after_sync();
        float u[16];
        tmem_ld16(tmem_addr(tmU, warp, 16 * (p - 1)), u);
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const f32x2 d = sub2(pack2(y[i], y[i + 1]), pack2(u[i], u[i + 1]));
          unpack2(d, y[i], y[i + 1]);
        }
      }
      solve_panel16(y, ci, p);
      {   // Z_p -> tensor memory, rounded exactly as pass 1 rounds it (both passes must compute the same Z)
        float zh[16], zl[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { zh[i] = to_tf32_rna(y[i]); zl[i] = __fsub_rn(y[i], zh[i]); }
        tmem_st16(tmem_addr(tmZh, warp, 0), zh);
        if (COR) tmem_st16(tmem_addr(tmZl, warp, 0), zl);
        tmem_wait_st();
      }
      if (use_tma) {   // hand the slot back: every loaded value has been consumed by the solve above; the fence
        fence_async_smem();   // orders the generic-proxy reads before the async-proxy write that reuses the slot
        mbar_arrive(bar_empty + slot);
        if (t == 0 && g + S < npanels) {
          mbar_wait(bar_empty + slot, (uint32_t)(g / S) & 1u);
          issue_load(g + S);
        }
      }
      fence_before_sync();
      named_bar_sync(1 + wg, 128);
      if (t == 0) {   // up to 12 products per panel
        fence_after_sync();
        if (p < NPAN - 1) {   // U[:, 16(p+1):] (+)= Z_p R[p, 16(p+1):]
          const int N = NP - 16 * (p + 1);
          const uint32_t id = idesc_tf32(128, N, false, false, false);
          const uint32_t d = tmU + 16 * p;
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const uint64_t dh = dRh + (uint64_t)((rblk_off<NP>(p) + kk * 256) >> 4), dl = dRl + (uint64_t)((rblk_off<NP>(p) + kk * 256) >> 4);
            mma_tf32_ts(d, tmZh + 8 * kk, dh, id, p > 0 || kk > 0);
            if (COR) {
              mma_tf32_ts(d, tmZh + 8 * kk, dl, id, true);
              mma_tf32_ts(d, tmZl + 8 * kk, dh, id, true);
            }
          }
        }
        {   // Q (+)= Z_p C[16p : 16p+16, :]
          const uint32_t id = idesc_tf32(128, NP, false, false, false);
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) {
            const uint64_t dh = dCh + (uint64_t)((p * 512 + kk * 256) >> 4), dl = dCl + (uint64_t)((p * 512 + kk * 256) >> 4);
            mma_tf32_ts(tmD, tmZh + 8 * kk, dh, id, p > 0 || kk > 0);
            if (COR) {
              mma_tf32_ts(tmD, tmZh + 8 * kk, dl, id, true);
              mma_tf32_ts(tmD, tmZl + 8 * kk, dh, id, true);
            }
          }
        }
        mma_commit(bar_mma);
      }
    }
    // ---- Q tile: accumulator row -> (shared memory -> one TMA store) or straight to global memory
    if (tma_q && it > 0) {   // the staging buffer is free once the previous tile's store has read it
      if (t == 0) tma_store_wait_read();
      named_bar_sync(1 + wg, 128);
    }
    mbar_wait(bar_mma, ph_mma);
    ph_mma ^= 1;
    fence_after_sync();
    uint32_t qr[NP];   // all loads of the accumulator row in flight behind one wait
#pragma unroll
    for (int c0 = 0; c0 < NP; c0 += 16) tmem_ld16_nowait(tmem_addr(tmD, warp, c0), qr + c0);
    tmem_wait_ld();
    if (tma_q) {
#pragma unroll
      for (int c = 0; c < NP; ++c) io_store(qst + c * 128 + t, __uint_as_float(qr[c]));
      fence_async_smem();
      named_bar_sync(1 + wg, 128);
      if (t == 0) {   // rows beyond M and columns beyond n are clipped by the tensor map
        tma_store_2d(&tmapQ, qst, tile * kPcTileRows, 0);
        tma_store_commit();
      }
    } else {
      TIo* qp = Q + row;
#pragma unroll
      for (int c = 0; c < NP; ++c) {
        if (valid && c < n) io_store(qp, __uint_as_float(qr[c]));
        qp += P.ldq;
      }
    }
  }
  if (tma_q && t == 0) tma_store_wait_all();
  fence_before_sync();
  __syncthreads();
  if (tid < 32) tmem_dealloc<512>(*tmem_slot);
}
// ======================================================================================= CUDA-core passes (*_notc modes)
// The same two passes in plain fp32 on the CUDA cores, one row per thread and no cross-thread traffic in the substitution:
//   z = a R_pre^-1 panel by panel: y_p = a_p - sum_{q<p} z_q R[q, p] (R from shared memory, broadcast), then the 16 x 16
//   diagonal solve with the constant-memory image (solve_panel16, the code the tensor-core passes use);
//   pass 1: the tile's Z goes through shared memory for G += Z^T Z as a register-blocked rank-128 update (thread (ty, tx)
//           owns a (NP/16) x (NP/8) block of G in fp32 over at most 16 tiles, then adds it to its fp64 partial);
//   pass 2: q = z C with C broadcast from shared memory.
// Both passes run solve_row() on the same inputs, so Z is the same to the bit. Grid = the partial slots of the layout
// (3 per SM), 128 threads, ~50 KB of shared memory: three CTAs per SM.
template <int NP>
__device__ __forceinline__ void solve_row(float (&y)[NP], const float* __restrict__ ci, const float* __restrict__ Rs) {
  constexpr int NPAN = NP / 16;
#pragma unroll
  for (int p = 0; p < NPAN; ++p) {
#pragma unroll
    for (int q = 0; q < p; ++q) {   // y_p -= z_q R[16q : 16q+16, 16p : 16p+16]; Rs[k][c] row-major, ld NP
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const f32x2 nz = pack2(-y[16 * q + i], -y[16 * q + i]);
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {   // packed FMAs: two columns each
          const float4 r = *reinterpret_cast<const float4*>(Rs + (16 * q + i) * NP + 16 * p + 4 * j4);
          const f32x2 d0 = fma2(nz, pack2(r.x, r.y), pack2(y[16 * p + 4 * j4 + 0], y[16 * p + 4 * j4 + 1]));
          const f32x2 d1 = fma2(nz, pack2(r.z, r.w), pack2(y[16 * p + 4 * j4 + 2], y[16 * p + 4 * j4 + 3]));
          unpack2(d0, y[16 * p + 4 * j4 + 0], y[16 * p + 4 * j4 + 1]);
          unpack2(d1, y[16 * p + 4 * j4 + 2], y[16 * p + 4 * j4 + 3]);
        }
      }
    }
    float yp[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) yp[j] = y[16 * p + j];
    solve_panel16(yp, ci, p);
#pragma unroll
    for (int j = 0; j < 16; ++j) y[16 * p + j] = yp[j];
  }
}

// N consecutive floats (N = 1, 2, 4, 8) from an address aligned to min(N, 4) floats
template <int N>
__device__ __forceinline__ void ld_vec(const float* __restrict__ p, float (&v)[N]) {
  if constexpr (N == 1) {
    v[0] = p[0];
  } else if constexpr (N == 2) {
    const float2 x = *reinterpret_cast<const float2*>(p);
    v[0] = x.x; v[1] = x.y;
  } else {
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
      const float4 x = *reinterpret_cast<const float4*>(p + 4 * i);
      v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
    }
  }
}

template <int NP>
struct NotcSmem {
  static constexpr int kZld = NP + 4;                              // row stride of the Z tile: conflict-free 128-bit row stores
  static constexpr uint32_t oR = 0;                                // R_pre, row-major [NP][NP]
  static constexpr uint32_t oZ = oR + NP * NP * 4;                 // pass 1: Z tile [128][kZld]; pass 2: C, row-major [NP][NP]
  static constexpr uint32_t kBytes1 = oZ + 128 * kZld * 4;
  static constexpr uint32_t kBytes2 = oZ + NP * NP * 4;
};

template <int NP, typename TIo, int SLOT>
__global__ void __launch_bounds__(128, 3) pc_gram_notc_kernel(const PcParams P) {
  using Sm = NotcSmem<NP>;
  constexpr int RW = NP / 16, CW = NP / 8;   // this thread's block of G: rows [RW ty, +RW), columns [CW tx, +CW)
  extern __shared__ __align__(16) char sm[];
  float* Rs = reinterpret_cast<float*>(sm + Sm::oR);
  float* Zs = reinterpret_cast<float*>(sm + Sm::oZ);
  const float* ci = cimg<SLOT>();
  const int t = threadIdx.x, ty = t >> 3, tx = t & 7;
  const TIo* A = reinterpret_cast<const TIo*>(P.A);
  for (int idx = t; idx < NP * NP; idx += 128) Rs[(idx % NP) * NP + idx / NP] = P.Rpre[idx];   // Rpre[k + c NP] -> Rs[k][c]
  __syncthreads();
  f32x2 acc[RW][CW / 2];   // fp32 pairs (columns 2j, 2j + 1)
#pragma unroll
  for (int i = 0; i < RW; ++i)
#pragma unroll
    for (int j = 0; j < CW / 2; ++j) acc[i][j] = 0ull;
  double* gp = P.Gpart + (size_t)blockIdx.x * NP * 128;   // slot of this CTA: entry (line i, column j) at j * 128 + i
  bool first_flush = true;
  auto flush = [&]() {
#pragma unroll
    for (int i = 0; i < RW; ++i)
#pragma unroll
      for (int j = 0; j < CW / 2; ++j) {
        float a0, a1;
        unpack2(acc[i][j], a0, a1);
        const int col = CW == 8 ? (2 * j < 4 ? 4 * tx + 2 * j : 32 + 4 * tx + 2 * j - 4) : CW * tx + 2 * j;   // see the b loads
        double* g = gp + (size_t)col * 128 + (RW * ty + i);
        g[0] = (first_flush ? 0.0 : g[0]) + (double)a0;
        g[128] = (first_flush ? 0.0 : g[128]) + (double)a1;
        acc[i][j] = 0ull;
      }
    first_flush = false;
  };
  int since = 0;
  for (int tile = blockIdx.x; tile < P.tiles; tile += gridDim.x) {
    const long long row = (long long)tile * kPcTileRows + t;
    const bool valid = row < P.M;
    float y[NP];
#pragma unroll
    for (int c = 0; c < NP; ++c) y[c] = (valid && c < P.n) ? io_plain(A + row + (long long)c * P.lda) : 0.f;
    solve_row<NP>(y, ci, Rs);
    __syncthreads();   // the previous tile's update has read Zs
#pragma unroll
    for (int c4 = 0; c4 < NP / 4; ++c4)
      *reinterpret_cast<float4*>(Zs + t * Sm::kZld + 4 * c4) = make_float4(y[4 * c4], y[4 * c4 + 1], y[4 * c4 + 2], y[4 * c4 + 3]);
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < 128; ++k) {
      float a[RW], b[CW];
      ld_vec<RW>(Zs + k * Sm::kZld + RW * ty, a);   // 128-/64-bit shared loads: the update is bound by them otherwise
      if constexpr (CW == 8) {   // two groups of 4 columns, 32 apart: the 8 lanes of a quarter-warp read 128 contiguous bytes
        float b0[4], b1[4];
        ld_vec<4>(Zs + k * Sm::kZld + 4 * tx, b0);
        ld_vec<4>(Zs + k * Sm::kZld + 32 + 4 * tx, b1);
#pragma unroll
        for (int j = 0; j < 4; ++j) { b[j] = b0[j]; b[4 + j] = b1[j]; }
      } else {
        ld_vec<CW>(Zs + k * Sm::kZld + CW * tx, b);
      }
#pragma unroll
      for (int i = 0; i < RW; ++i) {
        const f32x2 aa = pack2(a[i], a[i]);
#pragma unroll
        for (int j = 0; j < CW / 2; ++j) acc[i][j] = fma2(aa, pack2(b[2 * j], b[2 * j + 1]), acc[i][j]);
      }
    }
    if (++since == 16) { flush(); since = 0; }
  }
  flush();   // CTAs without tiles write zeros: the finish kernel sums every slot
}

template <int NP, typename TIo, int SLOT>
__global__ void __launch_bounds__(128, 3) pc_apply_notc_kernel(const PcParams P) {
  using Sm = NotcSmem<NP>;
  extern __shared__ __align__(16) char sm[];
  if (*P.status != kPcOk) return;   // the Householder passes behind this launch do the work
  float* Rs = reinterpret_cast<float*>(sm + Sm::oR);
  float* Cs = reinterpret_cast<float*>(sm + Sm::oZ);
  const float* ci = cimg<SLOT>();
  const int t = threadIdx.x;
  const TIo* A = reinterpret_cast<const TIo*>(P.A);
  TIo* Q = reinterpret_cast<TIo*>(P.Q);
  for (int idx = t; idx < NP * NP; idx += 128) {
    Rs[(idx % NP) * NP + idx / NP] = P.Rpre[idx];
    Cs[(idx % NP) * NP + idx / NP] = P.C[idx];   // C[k + c NP] -> Cs[k][c]
  }
  __syncthreads();
  for (int tile = blockIdx.x; tile < P.tiles; tile += gridDim.x) {
    const long long row = (long long)tile * kPcTileRows + t;
    const bool valid = row < P.M;
    float y[NP];
#pragma unroll
    for (int c = 0; c < NP; ++c) y[c] = (valid && c < P.n) ? io_plain(A + row + (long long)c * P.lda) : 0.f;
    solve_row<NP>(y, ci, Rs);
    f32x2 q2[NP / 2];   // q[2c], q[2c + 1]
#pragma unroll
    for (int c = 0; c < NP / 2; ++c) q2[c] = 0ull;
#pragma unroll 4
    for (int k = 0; k < NP; ++k) {   // (C = R_z^-1 B is a full matrix when a root factor B is applied); packed FMAs: two columns each
      const f32x2 zz = pack2(y[k], y[k]);
#pragma unroll
      for (int c4 = 0; c4 < NP / 4; ++c4) {
        const float4 cc = *reinterpret_cast<const float4*>(Cs + k * NP + 4 * c4);
        q2[2 * c4] = fma2(zz, pack2(cc.x, cc.y), q2[2 * c4]);
        q2[2 * c4 + 1] = fma2(zz, pack2(cc.z, cc.w), q2[2 * c4 + 1]);
      }
    }
    if (valid) {
      TIo* qp = Q + row;
#pragma unroll
      for (int c2 = 0; c2 < NP / 2; ++c2) {
        float q0, q1;
        unpack2(q2[c2], q0, q1);
        if (2 * c2 < P.n) io_store(qp + (long long)(2 * c2) * P.ldq, q0);
        if (2 * c2 + 1 < P.n) io_store(qp + (long long)(2 * c2 + 1) * P.ldq, q1);
      }
    }
  }
}

// Pass 2 of fp32_notc at NP = 64 with the Q product register-blocked like the Gram update: the tile's Z goes through
// shared memory (k-major), thread (ty, tx) computes rows [8 ty, +8) x columns {4 tx .. +3, 32 + 4 tx .. +3} of Q = Z C:
// 4 LDS.128 per 64 FMAs instead of 16 broadcast LDS.128 -- the thread-per-row form is bound by exactly those loads.
// Needs 16-byte aligned Q columns (the launcher checks); rows beyond M and columns beyond n are not stored.
struct Notc2Smem {
  static constexpr int kZld = 132;                       // k-major Z tile [64][kZld]
  static constexpr uint32_t oR = 0;                      // R_pre row-major [64][64]
  static constexpr uint32_t oC = oR + 64 * 64 * 4;       // C row-major [64][64]
  static constexpr uint32_t oZ = oC + 64 * 64 * 4;
  static constexpr uint32_t kBytes = oZ + 64 * kZld * 4;
};
template <int SLOT>
__global__ void __launch_bounds__(128, 3) pc_apply_notc2_kernel(const PcParams P) {
  constexpr int NP = 64;
  using Sm = Notc2Smem;
  extern __shared__ __align__(16) char sm[];
  if (*P.status != kPcOk) return;
  float* Rs = reinterpret_cast<float*>(sm + Sm::oR);
  float* Cs = reinterpret_cast<float*>(sm + Sm::oC);
  float* Zt = reinterpret_cast<float*>(sm + Sm::oZ);
  const float* ci = cimg<SLOT>();
  const int t = threadIdx.x, ty = t >> 3, tx = t & 7;
  const float* A = reinterpret_cast<const float*>(P.A);
  float* Q = reinterpret_cast<float*>(P.Q);
  for (int idx = t; idx < NP * NP; idx += 128) {
    Rs[(idx % NP) * NP + idx / NP] = P.Rpre[idx];
    Cs[(idx % NP) * NP + idx / NP] = P.C[idx];
  }
  __syncthreads();
  for (int tile = blockIdx.x; tile < P.tiles; tile += gridDim.x) {
    const long long row = (long long)tile * kPcTileRows + t;
    {
      const bool valid = row < P.M;
      float y[NP];
#pragma unroll
      for (int c = 0; c < NP; ++c) y[c] = (valid && c < P.n) ? A[row + (long long)c * P.lda] : 0.f;
      solve_row<NP>(y, ci, Rs);
      __syncthreads();   // the previous tile's product has read Zt
#pragma unroll
      for (int k = 0; k < NP; ++k) Zt[k * Sm::kZld + t] = y[k];
    }
    __syncthreads();
    f32x2 acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0ull;
#pragma unroll 2
    for (int k = 0; k < NP; ++k) {
      float a[8], b0[4], b1[4];
      ld_vec<8>(Zt + k * Sm::kZld + 8 * ty, a);
      ld_vec<4>(Cs + k * NP + 4 * tx, b0);
      ld_vec<4>(Cs + k * NP + 32 + 4 * tx, b1);
      const f32x2 bb[4] = {pack2(b0[0], b0[1]), pack2(b0[2], b0[3]), pack2(b1[0], b1[1]), pack2(b1[2], b1[3])};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const f32x2 aa = pack2(a[i], a[i]);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma2(aa, bb[j], acc[i][j]);
      }
    }
    // column of pair j, element e: j < 2: 4 tx + 2 j + e; else 32 + 4 tx + 2 (j - 2) + e; rows r0 .. r0 + 7
    const long long r0 = (long long)tile * kPcTileRows + 8 * ty;
    const bool full = r0 + 8 <= P.M;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int col = (j < 2 ? 4 * tx + 2 * j : 32 + 4 * tx + 2 * (j - 2)) + e;
        if (col >= P.n) continue;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float lo, hi;
          unpack2(acc[i][j], lo, hi);
          v[i] = e == 0 ? lo : hi;
        }
        float* qp = Q + r0 + (long long)col * P.ldq;
        if (full) {
          *reinterpret_cast<float4*>(qp) = make_float4(v[0], v[1], v[2], v[3]);
          *reinterpret_cast<float4*>(qp + 4) = make_float4(v[4], v[5], v[6], v[7]);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (r0 + i < P.M) qp[i] = v[i];
        }
      }
  }
}

// ======================================================================================= finish
// Stage 1, every CTA: fp64 sum of the per-CTA partials for 32 consecutive accumulator lanes of one Gram column
// (warp w adds partials w, w + 8, ...: coalesced 256-byte reads, then the 8 warp sums are added in fixed order).
// Stage 2, the CTA that finishes last (ticket): chol(G), certificate, R = R_z R_pre, C = R_z^-1, all in fp64 on
// 64 x 64 matrices in shared memory.
constexpr int kFinThreads = 512;
constexpr int kFinWarps = kFinThreads / 32;
__device__ __forceinline__ double rcp_f64(double x) {   // 1/x to fp64 accuracy from the fp32 approximation (x > 0, normal)
  double y = (double)rcp_approx((float)x);
  y = fma(y, fma(-x, y, 1.0), y);
  y = fma(y, fma(-x, y, 1.0), y);
  y = fma(y, fma(-x, y, 1.0), y);
  return y;
}
template <int NP>
__global__ void __launch_bounds__(kFinThreads) pc_finish_kernel(const double* __restrict__ Gpart, int nparts, double* __restrict__ Gsum,
                                                               const float* __restrict__ Rpre, int n, int cor, float max_est,
                                                               float* __restrict__ Rout, float* __restrict__ Cmat,
                                                               float* __restrict__ Rzinv, int* __restrict__ status, int stage) {
  // stage 0: sum the partials, the last CTA factors (single GPU); stage 1: sum only (the row-sharded driver all-reduces
  // Gsum across the ranks next); stage 2: factor Gsum only (one CTA)
  extern __shared__ __align__(16) char smraw[];
  double* G = reinterpret_cast<double*>(smraw);         // [NP][NP] row-major: trailing matrix of the factorisation, then G^-1
  double* Rz = G + NP * NP;                              // R_z (upper)
  constexpr int LX = NP + 1;                              // row stride of X: its transposed reads are conflict free
  double* X = Rz + NP * NP;                              // R_z^-1 (upper), [NP][LX]
  double* Rp = X + NP * LX;                              // Rp[k][j] = R_pre[k][j], upper triangle, fp64
  float* Xf = reinterpret_cast<float*>(Rp + NP * NP);    // fp32 copy of X for the certificate (this GPU runs ~4 DFMA / clock / SM)
  __shared__ double red[kFinThreads / 32][32];
  __shared__ double rdiag[NP];
  __shared__ int s_last, s_bad;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int lanes = (cor ? 2 : 1) * NP;
  if (stage != 2) {
    const int nlb = (lanes + 31) / 32;
    const int lb = blockIdx.x % nlb, j = blockIdx.x / nlb;
    const int l = 32 * lb + lane;
    double s = 0.0;
    const double* p = Gpart + (size_t)j * 128 + l;
    if (l < lanes) {
#pragma unroll 4
      for (int c = warp; c < nparts; c += kFinThreads / 32) s += __ldcg(p + (size_t)c * NP * 128);
    }
    red[warp][lane] = s;
    __syncthreads();
    if (warp == 0 && l < lanes) {
      double r = 0.0;
#pragma unroll
      for (int w = 0; w < kFinThreads / 32; ++w) r += red[w][lane];
      Gsum[l * NP + j] = r;
    }
  }
  if (stage == 1) return;
  if (stage == 0) {
    __threadfence();
    __syncthreads();
    if (t == 0) s_last = (atomicAdd(&status[1], 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
  }
  const long long fclk0 = clock64();
  if (t == 0) s_bad = 0;
  // R_pre has long left the L2: its loads are issued here and hide behind the factorisation (first use: the R product)
  for (int e = t; e < NP * NP; e += kFinThreads) {
    const int k = e % NP, j = e / NP;
    Rp[k * NP + j] = k <= j ? (double)Rpre[e] : 0.0;
  }
  // G = HH + LH + LH^T, symmetrised; padded columns get a unit diagonal
  double ssq = 0.0;
  for (int e = t; e < NP * NP; e += kFinThreads) {
    const int i = e / NP, j = e % NP;
    double g = 0.5 * (__ldcg(Gsum + i * NP + j) + __ldcg(Gsum + j * NP + i));
    if (cor) g += __ldcg(Gsum + (NP + i) * NP + j) + __ldcg(Gsum + (NP + j) * NP + i);
    if (i == j && i >= n) g = 1.0;
    if (i != j && (i >= n || j >= n)) g = 0.0;
    G[e] = g;
    if (i < n && j < n) ssq += g * g;
  }
  auto block_sum = [&](double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    __syncthreads();
    if (lane == 0) red[0][warp] = v;
    __syncthreads();
    double r = 0.0;
    for (int w = 0; w < kFinThreads / 32; ++w) r += red[0][w];
    return r;
  };
  const double nG = sqrt(block_sum(ssq));
  // right-looking Cholesky G = R_z^T R_z with ONE barrier per step: row k of R_z goes to its own array, the
  // trailing update uses the unscaled row k of G and 1 / G_kk. Thread (warp, lane) owns rows warp + 16a and columns
  // lane + 32b of the trailing matrix: row k is read with consecutive lanes, G[k][i] is a broadcast.
  for (int k = 0; k < NP; ++k) {
    const double piv = G[k * NP + k];
    const bool okp = piv > 0.0 && piv < 1e300;
    if (!okp && t == 0) s_bad = 1;
    const double inv = okp ? rcp_f64(piv) : 1.0;
    if (t < NP) {
      const double rs = sqrt(inv);                     // 1 / sqrt(G_kk)
      Rz[k * NP + t] = (t >= k) ? G[k * NP + t] * rs : 0.0;
      if (t == k) rdiag[k] = rs;                       // 1 / R_z[k][k]
    }
#pragma unroll
    for (int b2 = 0; b2 < (NP + 31) / 32; ++b2) {
      const int j = lane + 32 * b2;
      if (j > k && j < NP) {
        const double gkj = G[k * NP + j] * inv;
#pragma unroll
        for (int a2 = 0; a2 < (NP + kFinWarps - 1) / kFinWarps; ++a2) {
          const int i = warp + kFinWarps * a2;
          if (i > k && i <= j && i < NP) G[i * NP + j] -= G[k * NP + i] * gkj;
        }
      }
    }
    __syncthreads();
  }
  const long long fclk1 = clock64();
  // X = R_z^-1 (upper): column c by back substitution, 4 adjacent lanes share a column and split the dot products
  {
    const int c = t >> 2, g4 = t & 3;
    for (int k = NP - 1; k >= 0; --k) {
      double s = 0.0;
      if (k < c && c < NP)
        for (int j = k + 1 + g4; j <= c; j += 4) s -= Rz[k * NP + j] * X[j * LX + c];
      s += __shfl_xor_sync(kFull, s, 1);
      s += __shfl_xor_sync(kFull, s, 2);
      if (g4 == 0 && c < NP) {
        const double x = (k > c) ? 0.0 : ((k == c ? 1.0 : 0.0) + s) * rdiag[k];
        X[k * LX + c] = x;
        Xf[k * LX + c] = (float)x;
      }
      __syncwarp();
    }
  }
  __syncthreads();
  const long long fclk2 = clock64();
  const long long fclk3 = clock64();
  // |G^-1|_F for the certificate, G^-1 = X X^T, in fp32 (the value is compared with a threshold of 6): entries i <= j,
  // the off-diagonal ones counted twice; i = warp + 16a, j = lane + 32b. X is upper triangular with explicit zeros
  double ssi = 0.0;
  {
    float ssf = 0.f;
#pragma unroll 1
    for (int a2 = 0; a2 < (NP + kFinWarps - 1) / kFinWarps; ++a2) {
      const int i = warp + kFinWarps * a2;
#pragma unroll
      for (int b2 = 0; b2 < (NP + 31) / 32; ++b2) {
        const int j = lane + 32 * b2;
        if (i < n && j < n && i <= j) {
          float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 4
          for (int k = 0; k < NP; k += 4) {
            s0 = __fmaf_rn(Xf[i * LX + k], Xf[j * LX + k], s0);
            s1 = __fmaf_rn(Xf[i * LX + k + 1], Xf[j * LX + k + 1], s1);
            s2 = __fmaf_rn(Xf[i * LX + k + 2], Xf[j * LX + k + 2], s2);
            s3 = __fmaf_rn(Xf[i * LX + k + 3], Xf[j * LX + k + 3], s3);
          }
          const float v = (s0 + s1) + (s2 + s3);
          ssf = __fmaf_rn(i == j ? 1.f : 2.f, v * v, ssf);
        }
      }
    }
    ssi = (double)ssf;
  }
  const double nGi = sqrt(block_sum(ssi));
  const double est = nG * nGi / (double)n;
  const long long fclk4 = clock64();
  // R = R_z R_pre (n x n, ld n), C = R_z^-1
#pragma unroll 1
  for (int a2 = 0; a2 < (NP + kFinWarps - 1) / kFinWarps; ++a2) {
    const int i = warp + kFinWarps * a2;
#pragma unroll
    for (int b2 = 0; b2 < (NP + 31) / 32; ++b2) {
      const int j = lane + 32 * b2;
      if (i < NP && j < NP) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        if (i <= j) {   // k in [i, j]: R_z row i is zero left of the diagonal, column j of Rp below it
          int k = i;
          for (; k + 3 <= j; k += 4) {
            s0 = fma(Rz[i * NP + k], Rp[k * NP + j], s0);
            s1 = fma(Rz[i * NP + k + 1], Rp[(k + 1) * NP + j], s1);
            s2 = fma(Rz[i * NP + k + 2], Rp[(k + 2) * NP + j], s2);
            s3 = fma(Rz[i * NP + k + 3], Rp[(k + 3) * NP + j], s3);
          }
          for (; k <= j; ++k) s0 = fma(Rz[i * NP + k], Rp[k * NP + j], s0);
        }
        if (i < n && j < n) Rout[i + j * n] = (float)((s0 + s1) + (s2 + s3));
        const float x = (float)X[j * LX + i];              // element (j, i) of the column-major outputs
        Cmat[j + i * NP] = x;
        Rzinv[j + i * NP] = x;
      }
    }
  }
  __syncthreads();
  if (t == 0) {
    const bool ok = !s_bad && (est <= (double)max_est);   // a NaN est compares false
    status[0] = ok ? kPcOk : 1;
    status[1] = 0;
    status[2] = __float_as_int((float)est);
    status[4] = (int)(fclk1 - fclk0);          // phase clocks (cycles): G build + Cholesky, inverse, the rest
    status[5] = (int)(fclk2 - fclk1);
    status[6] = (int)(clock64() - fclk2);
    status[10] = (int)(fclk3 - fclk2);         // of which: (nothing since the transposed copies went), certificate product
    status[11] = (int)(fclk4 - fclk3);
  }
}

// C = R_z^-1 B for the row-sharded driver (B = this slab's block of the root factor)
template <int NP>
__global__ void __launch_bounds__(256) pc_makec_kernel(const float* __restrict__ Rzinv, const float* __restrict__ B, long long ldb,
                                                      int n, float* __restrict__ Cmat, const int* __restrict__ status) {
  if (*status != kPcOk) return;
  for (int e = threadIdx.x; e < NP * NP; e += 256) {
    const int i = e % NP, j = e / NP;
    float s = 0.f;
    if (i < n && j < n) {
      for (int k = i; k < n; ++k) s = __fmaf_rn(Rzinv[i + k * NP], B[k + (long long)j * ldb], s);
    } else if (i == j) {
      s = 1.f;
    }
    Cmat[i + j * NP] = s;
  }
}

// ======================================================================================= host side
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeFn encode_fn() {
  static EncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<EncodeFn>(p);
  }();
  return fn;
}
// 2-D map of a column-major M x n matrix: dimension 0 = rows (contiguous), dimension 1 = columns (stride lda);
// box = 128 rows x NP columns, out-of-range rows and columns read as zero
bool make_tile_map(CUtensorMap* map, const void* A, long long lda, long long M, int n, int np, IoType io, int box_cols = 0,
                   int box_rows = 128, bool swizzle128 = false) {
  EncodeFn fn = encode_fn();
  const size_t es = io == IoType::f16 ? 2 : 4;
  if (!fn) return false;
  if ((reinterpret_cast<uintptr_t>(A) & 15u) != 0 || ((size_t)lda * es) % 16 != 0) return false;
  if (M >= (1ll << 31)) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)M, (cuuint64_t)n};
  cuuint64_t gstr[1] = {(cuuint64_t)lda * es};
  cuuint32_t box[2] = {(cuuint32_t)box_rows, (cuuint32_t)(box_cols > 0 ? box_cols : np)};
  cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = fn(map, io == IoType::f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                        const_cast<void*>(A), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// TSQR_B200_TMA=0 makes the passes load their tiles with per-thread global loads (the path taken anyway when
// lda * sizeof(element) is not a multiple of 16 bytes)
bool tma_enabled() {
  static const bool on = [] { const char* e = std::getenv("TSQR_B200_TMA"); return !(e && e[0] == '0'); }();
  return on;
}

template <typename K>
void set_smem(K kern, int bytes) {
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
}
// the attribute belongs to the device's context: once per kernel instantiation AND device (`done` = bit per device)
template <typename K>
void set_smem_once(K kern, int bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  TSQR_CUDA_CHECK(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return;
  set_smem(kern, bytes);
  done.fetch_or(bit, std::memory_order_release);
}

// ---- constant-image slots: a slot is (re)written by a stream-ordered copy only after the event of its previous
// user has completed on the new user's stream
struct ImgSlots {
  std::mutex mu;
  cudaEvent_t ev[64][kImgSlots] = {};
  unsigned next[64] = {};
};
ImgSlots& img_slots() { static ImgSlots s; return s; }
int img_acquire(const float* img_global, cudaStream_t s) {
  ImgSlots& st = img_slots();
  int dev = 0;
  TSQR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) dev = 0;
  int slot;
  {
    std::lock_guard<std::mutex> lk(st.mu);
    slot = (int)(st.next[dev]++ % kImgSlots);
    if (!st.ev[dev][slot]) TSQR_CUDA_CHECK(cudaEventCreateWithFlags(&st.ev[dev][slot], cudaEventDisableTiming));
    else TSQR_CUDA_CHECK(cudaStreamWaitEvent(s, st.ev[dev][slot], 0));
  }
  const size_t bytes = kImgFloats * sizeof(float);
  switch (slot) {
    case 0: TSQR_CUDA_CHECK(cudaMemcpyToSymbolAsync(c_img0, img_global, bytes, 0, cudaMemcpyDeviceToDevice, s)); break;
    case 1: TSQR_CUDA_CHECK(cudaMemcpyToSymbolAsync(c_img1, img_global, bytes, 0, cudaMemcpyDeviceToDevice, s)); break;
    default: TSQR_CUDA_CHECK(cudaMemcpyToSymbolAsync(c_img2, img_global, bytes, 0, cudaMemcpyDeviceToDevice, s)); break;
  }
  return slot;
}
void img_release(int slot, cudaStream_t s) {
  ImgSlots& st = img_slots();
  int dev = 0;
  TSQR_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) dev = 0;
  std::lock_guard<std::mutex> lk(st.mu);
  TSQR_CUDA_CHECK(cudaEventRecord(st.ev[dev][slot], s));
}

template <int NP, bool COR, typename TIo, int SLOT>
void launch_pass1_t(const PcParams& p, const CUtensorMap& map, const CUtensorMap&, int grid, cudaStream_t s) {
  auto kern = pc_gram_kernel<NP, COR, TIo, SLOT>;
  static std::atomic<unsigned long long> smem_set{0};
  set_smem_once(kern, (int)GramSmem<NP>::kBytes, smem_set);
  kern<<<grid, 128 * kGrWg, GramSmem<NP>::kBytes, s>>>(p, map);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <int NP, bool COR, typename TIo, int SLOT>
void launch_apply_t(const PcParams& p, const CUtensorMap& map, const CUtensorMap& mapq, int grid, cudaStream_t s) {
  auto kern = pc_apply_kernel<NP, COR, TIo, SLOT>;
  static std::atomic<unsigned long long> smem_set{0};
  set_smem_once(kern, (int)ApplySmem<NP, TIo>::kBytes, smem_set);
  kern<<<grid, 128 * kApWg, ApplySmem<NP, TIo>::kBytes, s>>>(p, map, mapq);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <int PASS, int NP, bool COR, typename TIo>
void launch_slot(int slot, const PcParams& p, const CUtensorMap& map, const CUtensorMap& mapq, int grid, cudaStream_t s) {
  if (PASS == 1) {
    switch (slot) {
      case 0: launch_pass1_t<NP, COR, TIo, 0>(p, map, mapq, grid, s); break;
      case 1: launch_pass1_t<NP, COR, TIo, 1>(p, map, mapq, grid, s); break;
      default: launch_pass1_t<NP, COR, TIo, 2>(p, map, mapq, grid, s); break;
    }
  } else {
    switch (slot) {
      case 0: launch_apply_t<NP, COR, TIo, 0>(p, map, mapq, grid, s); break;
      case 1: launch_apply_t<NP, COR, TIo, 1>(p, map, mapq, grid, s); break;
      default: launch_apply_t<NP, COR, TIo, 2>(p, map, mapq, grid, s); break;
    }
  }
}
template <int PASS, int NP>
void launch_np(int slot, const PcParams& p, const CUtensorMap& map, const CUtensorMap& mapq, int grid, bool cor, IoType io, cudaStream_t s) {
  if (io == IoType::f16) launch_slot<PASS, NP, false, __half>(slot, p, map, mapq, grid, s);
  else if (cor) launch_slot<PASS, NP, true, float>(slot, p, map, mapq, grid, s);
  else launch_slot<PASS, NP, false, float>(slot, p, map, mapq, grid, s);
}
// CUDA-core passes of the *_notc modes: one CTA per partial slot of the layout
template <int PASS, int NP, typename TIo, int SLOT>
void launch_notc_t(const PcParams& p, int grid, cudaStream_t s) {
  static std::atomic<unsigned long long> smem_set{0};
  if (PASS == 1) {
    auto kern = pc_gram_notc_kernel<NP, TIo, SLOT>;
    set_smem_once(kern, (int)NotcSmem<NP>::kBytes1, smem_set);
    kern<<<grid, 128, NotcSmem<NP>::kBytes1, s>>>(p);
  } else {
    auto kern = pc_apply_notc_kernel<NP, TIo, SLOT>;
    set_smem_once(kern, (int)NotcSmem<NP>::kBytes2, smem_set);
    kern<<<grid, 128, NotcSmem<NP>::kBytes2, s>>>(p);
  }
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <int PASS, int NP, typename TIo>
void launch_notc_slot(int slot, const PcParams& p, int grid, cudaStream_t s) {
  switch (slot) {
    case 0: launch_notc_t<PASS, NP, TIo, 0>(p, grid, s); break;
    case 1: launch_notc_t<PASS, NP, TIo, 1>(p, grid, s); break;
    default: launch_notc_t<PASS, NP, TIo, 2>(p, grid, s); break;
  }
}
template <int SLOT>
void launch_notc2_t(const PcParams& p, int grid, cudaStream_t s) {
  auto kern = pc_apply_notc2_kernel<SLOT>;
  static std::atomic<unsigned long long> smem_set{0};
  set_smem_once(kern, (int)Notc2Smem::kBytes, smem_set);
  kern<<<grid, 128, Notc2Smem::kBytes, s>>>(p);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <int PASS>
void launch_notc(const PcCall& c, const PcParams& p, cudaStream_t s) {
  const int grid = (int)c.L.parts;
  const int slot = img_acquire(c.ws + c.L.cimg, s);
  // pass 2, fp32, 64 padded columns, Q columns 16-byte aligned: the register-blocked form of the Q product
  static const bool blocked = [] { const char* e = std::getenv("TSQR_B200_NOTC_BLOCKED"); return !(e && e[0] == '0'); }();
  if (PASS == 2 && blocked && c.io == IoType::f32 && c.L.np == 64 && (reinterpret_cast<uintptr_t>(c.Q) & 15u) == 0 && (c.ldq % 4) == 0) {
    switch (slot) {
      case 0: launch_notc2_t<0>(p, grid, s); break;
      case 1: launch_notc2_t<1>(p, grid, s); break;
      default: launch_notc2_t<2>(p, grid, s); break;
    }
    img_release(slot, s);
    return;
  }
  if (c.io == IoType::f16) {
    switch (c.L.np) {
      case 16: launch_notc_slot<PASS, 16, __half>(slot, p, grid, s); break;
      case 32: launch_notc_slot<PASS, 32, __half>(slot, p, grid, s); break;
      default: launch_notc_slot<PASS, 64, __half>(slot, p, grid, s); break;
    }
  } else {
    switch (c.L.np) {
      case 16: launch_notc_slot<PASS, 16, float>(slot, p, grid, s); break;
      case 32: launch_notc_slot<PASS, 32, float>(slot, p, grid, s); break;
      default: launch_notc_slot<PASS, 64, float>(slot, p, grid, s); break;
    }
  }
  img_release(slot, s);
}
// pass 1 (PASS = 1: pc_gram_kernel, swizzled 32-row boxes) or pass 2 (PASS = 2: pc_apply_kernel, 128 x 16 boxes)
template <int PASS>
void launch_pass(const PcCall& c, PcParams p, bool want_tma, cudaStream_t s) {
  if (c.notc) { launch_notc<PASS>(c, p, s); return; }
  const int np = c.L.np;
  CUtensorMap map, mapq;
  memset(&map, 0, sizeof(map));
  memset(&mapq, 0, sizeof(mapq));
  // pass 1, corrected modes: boxes of 32 rows x NP columns in the 128-byte-swizzled operand layout (fp32 only);
  // pass 1, uncorrected modes and pass 2: plain 128 x 16 panels
  if (PASS == 1 && c.cor) p.use_tma = (want_tma && tma_enabled() && c.io == IoType::f32 && make_tile_map(&map, c.A, c.lda, c.M, c.n, np, c.io, np, 32, true)) ? 1 : 0;
  else p.use_tma = (want_tma && tma_enabled() && make_tile_map(&map, c.A, c.lda, c.M, c.n, np, c.io, 16)) ? 1 : 0;
  p.use_tma_q = (PASS == 2 && want_tma && tma_enabled() && make_tile_map(&mapq, c.Q, c.ldq, c.M, c.n, np, c.io)) ? 1 : 0;
  // both passes: one CTA of three tile pipelines per SM
  const int grid = c.L.grid;
  p.tiles_per_cta = c.L.tiles_per_cta;
  const int slot = img_acquire(c.ws + c.L.cimg, s);
  switch (np) {
    case 16: launch_np<PASS, 16>(slot, p, map, mapq, grid, c.cor, c.io, s); break;
    case 32: launch_np<PASS, 32>(slot, p, map, mapq, grid, c.cor, c.io, s); break;
    default: launch_np<PASS, 64>(slot, p, map, mapq, grid, c.cor, c.io, s); break;
  }
  img_release(slot, s);
}

template <int NP, typename TIo>
void launch_sample_t(const PcCall& c, cudaStream_t s) {
  auto kern = pc_sample_kernel<NP, TIo>;
  static std::atomic<unsigned long long> smem_set{0};
  set_smem_once(kern, (int)SmpCfg<NP>::kBytes, smem_set);
  kern<<<1, 8 * NP, SmpCfg<NP>::kBytes, s>>>(reinterpret_cast<const TIo*>(c.A), c.lda, c.M, c.n, c.ws + c.L.rpre,
                                              c.ws + c.L.cimg, reinterpret_cast<int*>(c.ws + c.L.status));
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
template <int NP>
void launch_finish_t(const PcCall& c, cudaStream_t s, int stage = 0) {
  auto kern = pc_finish_kernel<NP>;
  const int bytes = (4 * NP * NP + NP) * (int)sizeof(double) + NP * (NP + 1) * (int)sizeof(float);   // G, R_z, X (row stride NP + 1), Rp, Xf
  static std::atomic<unsigned long long> smem_set{0};
  set_smem_once(kern, bytes, smem_set);
  const int lanes = (c.cor ? 2 : 1) * NP;
  const int grid = stage == 2 ? 1 : ((lanes + 31) / 32) * NP;
  kern<<<grid, kFinThreads, bytes, s>>>(reinterpret_cast<const double*>(c.ws + c.L.gpart), c.L.parts,
                                         reinterpret_cast<double*>(c.ws + c.L.gsum), c.ws + c.L.rpre, c.n, c.cor ? 1 : 0,
                                         kPcMaxEst, c.ws + c.L.rout, c.ws + c.L.cmat, c.ws + c.L.rzinv,
                                         reinterpret_cast<int*>(c.ws + c.L.status), stage);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

PcParams pass_params(const PcCall& c) {
  PcParams p{};
  p.A = c.A; p.lda = c.lda; p.M = c.M; p.n = c.n;
  p.Q = c.Q; p.ldq = c.ldq;
  p.Rpre = c.ws + c.L.rpre;
  p.C = c.ws + c.L.cmat;
  p.Gpart = reinterpret_cast<double*>(c.ws + c.L.gpart);
  p.status = reinterpret_cast<int*>(c.ws + c.L.status);
  p.tiles = (int)c.L.tiles;
  p.tiles_per_cta = c.L.tiles_per_cta;
  return p;
}

size_t align64(size_t x) { return (x + 63) / 64 * 64; }


}  // namespace

PcLayout pc_layout(long long m, int n) {
  PcLayout L;
  L.np = tc_np(n);
  L.tiles = (m + kPcTileRows - 1) / kPcTileRows;
  const long long want = device_sm_count();
  L.grid = (int)std::min<long long>(want, L.tiles);
  L.tiles_per_cta = (int)((L.tiles + L.grid - 1) / L.grid);
  L.grid = (int)((L.tiles + L.tiles_per_cta - 1) / L.tiles_per_cta);
  L.parts = 3 * L.grid;                                                   // one fp64 partial per tile pipeline (kGrWg per CTA)
  const size_t np2 = (size_t)L.np * L.np;
  size_t off = 0;
  L.rpre = off; off = align64(off + np2);
  L.status = off; off = align64(off + 16);
  L.rout = off; off = align64(off + np2);
  L.cmat = off; off = align64(off + np2);
  L.rzinv = off; off = align64(off + np2);
  L.cimg = off; off = align64(off + kImgFloats);
  L.gsum = off; off = align64(off + 2 * (2 * np2));                       // doubles
  // sized for the largest grid any device of this process can ask for (the query may run before a device is chosen)
  const size_t gmax = (size_t)std::max<long long>(L.parts, 3 * std::min<long long>(160, L.tiles));
  L.gpart = off; off = align64(off + 2 * gmax * (size_t)L.np * 128);      // doubles
  L.total = off;
  return L;
}

void pc_reduce(const PcCall& c, cudaStream_t s) {
  const int np = c.L.np;
  {
    ProfScope prof(s, 100);
    if (c.io == IoType::f16) {
      switch (np) { case 16: launch_sample_t<16, __half>(c, s); break; case 32: launch_sample_t<32, __half>(c, s); break; default: launch_sample_t<64, __half>(c, s); }
    } else {
      switch (np) { case 16: launch_sample_t<16, float>(c, s); break; case 32: launch_sample_t<32, float>(c, s); break; default: launch_sample_t<64, float>(c, s); }
    }
  }
  {
    ProfScope prof(s, 101);
    launch_pass<1>(c, pass_params(c), true, s);
  }
  {
    ProfScope prof(s, 102);
    switch (np) { case 16: launch_finish_t<16>(c, s); break; case 32: launch_finish_t<32>(c, s); break; default: launch_finish_t<64>(c, s); }
  }
}

namespace {
__global__ void pc_set_status_kernel(int* status, int v) { status[0] = v; }
void finish_stage(const PcCall& c, cudaStream_t s, int stage) {
  switch (c.L.np) {
    case 16: launch_finish_t<16>(c, s, stage); break;
    case 32: launch_finish_t<32>(c, s, stage); break;
    default: launch_finish_t<64>(c, s, stage); break;
  }
}
}  // namespace

// ---- the stages one by one, for the row-sharded driver (sharded.cu): sample on one rank, R_pre broadcast, pass 1,
// partial sums, all-reduce of Gsum, factorisation, pass 2
void pc_stage_sample(const PcCall& c, cudaStream_t s) {
  const int np = c.L.np;
  ProfScope prof(s, 100);
  if (c.io == IoType::f16) {
    switch (np) { case 16: launch_sample_t<16, __half>(c, s); break; case 32: launch_sample_t<32, __half>(c, s); break; default: launch_sample_t<64, __half>(c, s); }
  } else {
    switch (np) { case 16: launch_sample_t<16, float>(c, s); break; case 32: launch_sample_t<32, float>(c, s); break; default: launch_sample_t<64, float>(c, s); }
  }
}
void pc_stage_adopt_rpre(const PcCall& c, cudaStream_t s) {   // R_pre (in the workspace) came from elsewhere: image + status reset
  TSQR_CUDA_CHECK(cudaMemsetAsync(c.ws + c.L.status, 0, 16 * sizeof(int), s));
  pc_cimg_kernel<<<1, 256, 0, s>>>(c.ws + c.L.rpre, c.ws + c.L.cimg, c.L.np);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}
void pc_stage_pass1(const PcCall& c, cudaStream_t s) {
  ProfScope prof(s, 101);
  launch_pass<1>(c, pass_params(c), true, s);
}
void pc_stage_partial_sums(const PcCall& c, cudaStream_t s) {
  ProfScope prof(s, 102);
  finish_stage(c, s, 1);
}
void pc_stage_factor(const PcCall& c, cudaStream_t s) {
  ProfScope prof(s, 102);
  finish_stage(c, s, 2);
}

// pass 1 with a caller-supplied R_pre, then pass 2 with a caller-supplied C (tests of the operand layouts)
void pc_probe(const PcCall& c, const float* rpre, const float* cmat, double* gsum_out, bool use_tma, cudaStream_t s) {
  const int np = c.L.np;
  const size_t np2 = (size_t)np * np;
  int* status = reinterpret_cast<int*>(c.ws + c.L.status);
  TSQR_CUDA_CHECK(cudaMemsetAsync(status, 0, 16 * sizeof(int), s));
  TSQR_CUDA_CHECK(cudaMemcpyAsync(c.ws + c.L.rpre, rpre, np2 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  pc_cimg_kernel<<<1, 256, 0, s>>>(c.ws + c.L.rpre, c.ws + c.L.cimg, np);
  if (use_tma) {
    CUtensorMap probe_map;
    if (!make_tile_map(&probe_map, c.A, c.lda, c.M, c.n, np, c.io)) throw std::runtime_error("tsqr_b200: no TMA map for this matrix (alignment)");
  }
  launch_pass<1>(c, pass_params(c), use_tma, s);
  switch (np) { case 16: launch_finish_t<16>(c, s); break; case 32: launch_finish_t<32>(c, s); break; default: launch_finish_t<64>(c, s); }
  TSQR_CUDA_CHECK(cudaMemcpyAsync(gsum_out, c.ws + c.L.gsum, 2 * np2 * sizeof(double), cudaMemcpyDeviceToDevice, s));
  TSQR_CUDA_CHECK(cudaMemcpyAsync(c.ws + c.L.cmat, cmat, np2 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  pc_set_status_kernel<<<1, 1, 0, s>>>(status, kPcOk);
  launch_pass<2>(c, pass_params(c), use_tma, s);
}

void pc_build(const PcCall& c, const float* B, long long ldb, cudaStream_t s) {
  const int np = c.L.np;
  if (B) {
    const int* st = reinterpret_cast<const int*>(c.ws + c.L.status);
    switch (np) {
      case 16: pc_makec_kernel<16><<<1, 256, 0, s>>>(c.ws + c.L.rzinv, B, ldb, c.n, c.ws + c.L.cmat, st); break;
      case 32: pc_makec_kernel<32><<<1, 256, 0, s>>>(c.ws + c.L.rzinv, B, ldb, c.n, c.ws + c.L.cmat, st); break;
      default: pc_makec_kernel<64><<<1, 256, 0, s>>>(c.ws + c.L.rzinv, B, ldb, c.n, c.ws + c.L.cmat, st); break;
    }
    note_launch();
    TSQR_CUDA_CHECK(cudaGetLastError());
  }
  launch_pass<2>(c, pass_params(c), true, s);
}

}  // namespace tsqr_b200
