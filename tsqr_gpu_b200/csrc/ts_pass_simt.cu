// CUDA-core (fp32 FMA) TSQR level pass: the path for the *_notc compute modes and the
// numerically exact baseline every tensor-core variant is checked against.
//
// Thread-per-row layout: each of the 128 threads of a CTA owns RPT rows of the tile and
// keeps the 16 columns of the current Householder panel in registers. Column norms and
// reflector dot products are per-thread partial sums finished with a 16-wide transposed
// warp-shuffle butterfly (16 shuffles for 16 sums) and a fixed-order 4-warp combine, so
// the arithmetic is bit-reproducible between the reduce pass and the build pass.
//
// Math per tile ("TS" step, top block = running R, reflector j = [e_j ; v_j]):
//   left-looking over 16-column panels; earlier panels are applied with their 16x16 T
//   (W = T_q^T (R[Jq,J] + Y_q^T A_J); R[Jq,J] -= W; A_J -= Y_q W); the panel itself is an
//   unblocked Householder sweep that also yields T_p. The build pass additionally
//   accumulates the full N x N T (larft recurrence) and finishes with
//   Q_tile = -Y (T B), B <- B - T B.
// Replaces reference src/tcqr32x16.cu:1373-1469 (qr32x16_core), src/tsqr.cu:143-204 and
// :591-656 (generic backward kernels) with a formulation that never forms H.
#include "common.cuh"
#include "ts_pass.h"

namespace tsqr_b200 {
namespace {

constexpr int kT = kPassThreads;

// 16 per-lane partial sums -> 16 warp totals. After the butterfly lane L holds the total
// of index 8*b4 + 4*b3 + 2*b2 + b1 (bK = bit K of L); even lanes write dst[idx*4 + warp].
__device__ __forceinline__ void reduce16_store(float (&p)[16], float* dst, int lane, int warp) {
  bool hi = lane & 16;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float send = hi ? p[i] : p[i + 8];
    float keep = hi ? p[i + 8] : p[i];
    p[i] = __fadd_rn(keep, __shfl_xor_sync(kFullMask, send, 16));
  }
  hi = lane & 8;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float send = hi ? p[i] : p[i + 4];
    float keep = hi ? p[i + 4] : p[i];
    p[i] = __fadd_rn(keep, __shfl_xor_sync(kFullMask, send, 8));
  }
  hi = lane & 4;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float send = hi ? p[i] : p[i + 2];
    float keep = hi ? p[i + 2] : p[i];
    p[i] = __fadd_rn(keep, __shfl_xor_sync(kFullMask, send, 4));
  }
  hi = lane & 2;
  {
    float send = hi ? p[0] : p[1];
    float keep = hi ? p[1] : p[0];
    p[0] = __fadd_rn(keep, __shfl_xor_sync(kFullMask, send, 2));
  }
  p[0] = __fadd_rn(p[0], __shfl_xor_sync(kFullMask, p[0], 1));
  if ((lane & 1) == 0) {
    int idx = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
    dst[idx * 4 + warp] = p[0];
  }
}

__device__ __forceinline__ float sum4(const float* part4) {
  float4 v = *reinterpret_cast<const float4*>(part4);
  return __fadd_rn(__fadd_rn(v.x, v.y), __fadd_rn(v.z, v.w));
}

// part[c*64 + k*4 + warp] <- per-warp totals of sum_rows Y[:, c] * a[:, k], c,k in [0,16)
template <int RPT>
__device__ __forceinline__ void gram16(const float (&a)[RPT][16], const float* ycols, int H, int t,
                                       float* part, int lane, int warp) {
#pragma unroll 2
  for (int c = 0; c < 16; ++c) {
    float y[RPT];
#pragma unroll
    for (int r = 0; r < RPT; ++r) y[r] = ycols[c * H + r * kT + t];
    float p[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      p[k] = __fmul_rn(y[0], a[0][k]);
#pragma unroll
      for (int r = 1; r < RPT; ++r) p[k] = __fmaf_rn(y[r], a[r][k], p[k]);
    }
    reduce16_store(p, part + c * 64, lane, warp);
  }
}

struct Smem {
  float* Rs;    // np*np   running R, row-major
  float* Ys;    // np*H    reflector tails, column-major (column c at Ys + c*H)
  float* Ts;    // np*np   T, row-major (upper triangular)
  float* Bs;    // np*np   (build) B, row-major
  float* part;  // 2*1024  double-buffered per-warp partial sums
  float* tot;   // 256
  float* Wsm;   // 256
  float* Ssm;   // 768     (build) Y_prev^T V_p
  float* Xsm;   // 768     (build) S * T_p
  float* rowb;  // 2*16    (first tile) broadcast of the pivot row of the panel
};

__host__ __device__ inline size_t smem_floats(int np, int H, bool build) {
  size_t f = (size_t)np * np * 2 + (size_t)np * H + 2048 + 256 + 256 + 32;
  if (build) f += (size_t)np * np + 768 + 768;
  return f;
}


// One Householder column step of the panel sweep (J = column inside the panel, compile-time
// so that every register-array index is static).
//  FIRST == false ("TS" step): the pivot row lives in the running R (Rs), the reflector is
//      [e_jj ; v] with v = a[:,J]/(alpha-beta) over all tile rows.
//  FIRST == true (first tile of a chunk, plain Householder): the pivot row is tile row jj
//      (thread jj, r = 0); rows <= jj are masked out of the sums and keep their R entries.
template <int J, int RPT, bool FIRST>
__device__ __forceinline__ void sweep_step(float (&a)[RPT][16], const Smem& sm, int np, int c0, int t, int lane,
                                           int warp, int& buf) {
  const int jj = c0 + J;
  float rrow[16];
  if (!FIRST) {
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      float4 v4 = *reinterpret_cast<const float4*>(sm.Rs + jj * np + c0 + k4 * 4);
      rrow[k4 * 4 + 0] = v4.x; rrow[k4 * 4 + 1] = v4.y; rrow[k4 * 4 + 2] = v4.z; rrow[k4 * 4 + 3] = v4.w;
    }
  } else if (t == jj) {
#pragma unroll
    for (int x = 0; x < 16; ++x) sm.rowb[buf * 16 + x] = a[0][x];
  }
  const bool masked0 = FIRST && (t <= jj);  // row r=0 of this thread is at or above the pivot
  float aj[RPT];
#pragma unroll
  for (int r = 0; r < RPT; ++r) aj[r] = (r == 0 && masked0) ? 0.f : a[r][J];
  float pr[16];
#pragma unroll
  for (int x = 0; x < 16; ++x) {
    pr[x] = __fmul_rn(aj[0], a[0][x]);
#pragma unroll
    for (int r = 1; r < RPT; ++r) pr[x] = __fmaf_rn(aj[r], a[r][x], pr[x]);
  }
  float* part = sm.part + buf * 1024;
  reduce16_store(pr, part, lane, warp);
  __syncthreads();
  if (FIRST) {
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      float4 v4 = *reinterpret_cast<const float4*>(sm.rowb + buf * 16 + k4 * 4);
      rrow[k4 * 4 + 0] = v4.x; rrow[k4 * 4 + 1] = v4.y; rrow[k4 * 4 + 2] = v4.z; rrow[k4 * 4 + 3] = v4.w;
    }
  }
  buf ^= 1;
  float g[16];
#pragma unroll
  for (int x = 0; x < 16; ++x) g[x] = sum4(part + x * 4);
  const float alpha = rrow[J];
  const float nx2 = __fmaf_rn(alpha, alpha, g[J]);
  float tau = 0.f, rden = 0.f, beta = alpha;
  if (nx2 > 0.f) {
    beta = -copysignf(__fsqrt_rn(nx2), alpha);
    const float den = __fsub_rn(alpha, beta);
    rden = __frcp_rn(den);
    tau = __fdiv_rn(__fsub_rn(beta, alpha), beta);
  }
#pragma unroll
  for (int k = J + 1; k < 16; ++k) {
    const float s = __fmaf_rn(g[k], rden, rrow[k]);
    const float coef = __fmul_rn(tau, s);
    const float cv = __fmul_rn(coef, rden);
#pragma unroll
    for (int r = 0; r < RPT; ++r) a[r][k] = __fmaf_rn(-cv, aj[r], a[r][k]);
    if (FIRST) {
      if (t == jj) a[0][k] = __fsub_rn(rrow[k], coef);
    } else {
      if (t == k) sm.Rs[jj * np + c0 + k] = __fsub_rn(rrow[k], coef);
    }
  }
#pragma unroll
  for (int r = 0; r < RPT; ++r) a[r][J] = (r == 0 && masked0) ? a[r][J] : __fmul_rn(a[r][J], rden);
  if (t == jj) {  // jj < np <= 64 < 128 threads
    if (FIRST) a[0][J] = beta;
    else sm.Rs[jj * np + jj] = beta;
    sm.Ts[jj * np + jj] = tau;
  }
  // T_p column J: T[i][J] = -tau * sum_{l=i}^{J-1} T[i][l] * (v_l . v_J), computed by thread i < J
  if (t < J) {
    float acc = 0.f;
#pragma unroll
    for (int l = 0; l < J; ++l) {
      float z = __fmul_rn(g[l], rden);
      if (FIRST) z = __fadd_rn(z, rrow[l]);  // v_l at the pivot row times the unit entry of v_J
      if (l >= t) acc = __fmaf_rn(sm.Ts[(c0 + t) * np + c0 + l], z, acc);
    }
    sm.Ts[(c0 + t) * np + jj] = __fmul_rn(-tau, acc);
  }
}

template <int J, int RPT, bool FIRST>
struct Sweep {
  static __device__ __forceinline__ void run(float (&a)[RPT][16], const Smem& sm, int np, int c0, int t, int lane,
                                             int warp, int& buf) {
    sweep_step<J, RPT, FIRST>(a, sm, np, c0, t, lane, warp, buf);
    Sweep<J + 1, RPT, FIRST>::run(a, sm, np, c0, t, lane, warp, buf);
  }
};
template <int RPT, bool FIRST>
struct Sweep<16, RPT, FIRST> {
  static __device__ __forceinline__ void run(float (&)[RPT][16], const Smem&, int, int, int, int, int, int&) {}
};

// Factor [R ; A_tile]. On exit Rs holds the new R, Ys the reflector tails, Ts the 16x16
// diagonal T blocks (and, when BUILD, the full upper-triangular T).
template <int RPT, bool BUILD, bool FIRST, typename TIn>
__device__ __forceinline__ void factor_tile(const PassParams& P, const Smem& sm, long long row0,
                                            long long row_end, int& buf) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int np = P.np, H = kT * RPT;
  const TIn* A = reinterpret_cast<const TIn*>(P.A);
  const int npanels = np / kPanel;

  for (int p = 0; p < npanels; ++p) {
    const int c0 = p * kPanel;
    float a[RPT][16];
#pragma unroll
    for (int r = 0; r < RPT; ++r) {
      const long long row = row0 + r * kT + t;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const int col = c0 + c;
        a[r][c] = (row < row_end && col < P.n) ? io_load(A + row + (long long)col * P.lda) : 0.f;
      }
    }
    // ---- apply the earlier panels of this tile (left-looking) ----
    for (int q = 0; q < p; ++q) {
      const int q0 = q * kPanel;
      float* part = sm.part + buf * 1024;
      gram16<RPT>(a, sm.Ys + (size_t)q0 * H, H, t, part, lane, warp);
      __syncthreads();
      buf ^= 1;
#pragma unroll
      for (int v = t; v < 256; v += kT) {
        const int c = v >> 4, k = v & 15;
        const float gsum = sum4(part + v * 4);
        sm.tot[v] = FIRST ? gsum : __fadd_rn(gsum, sm.Rs[(q0 + c) * np + c0 + k]);
      }
      __syncthreads();
#pragma unroll
      for (int v = t; v < 256; v += kT) {
        const int c = v >> 4, k = v & 15;
        float w = 0.f;
        for (int d = 0; d <= c; ++d) w = __fmaf_rn(sm.Ts[(q0 + d) * np + q0 + c], sm.tot[d * 16 + k], w);
        sm.Wsm[v] = w;
        if (!FIRST) sm.Rs[(q0 + c) * np + c0 + k] = __fsub_rn(sm.Rs[(q0 + c) * np + c0 + k], w);
      }
      __syncthreads();
#pragma unroll 4
      for (int c = 0; c < 16; ++c) {
        float w[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          float4 v4 = *reinterpret_cast<const float4*>(sm.Wsm + c * 16 + k4 * 4);
          w[k4 * 4 + 0] = v4.x; w[k4 * 4 + 1] = v4.y; w[k4 * 4 + 2] = v4.z; w[k4 * 4 + 3] = v4.w;
        }
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
          const float y = sm.Ys[(size_t)(q0 + c) * H + r * kT + t];
#pragma unroll
          for (int k = 0; k < 16; ++k) a[r][k] = __fmaf_rn(-y, w[k], a[r][k]);
        }
      }
      // next use of tot/Wsm is behind the next gram16's __syncthreads
    }
    // ---- unblocked Householder sweep over the 16 panel columns ----
    Sweep<0, RPT, FIRST>::run(a, sm, np, c0, t, lane, warp, buf);
    if (FIRST && t < np) {
      // rows 0..np-1 of the first tile hold R on and above the diagonal: move it to Rs and
      // leave the unit-lower-trapezoidal reflector block in the registers
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const int k = c0 + c;
        if (t < k) { sm.Rs[t * np + k] = a[0][c]; a[0][c] = 0.f; }
        else if (t == k) { sm.Rs[k * np + k] = a[0][c]; a[0][c] = 1.f; }
      }
    }
    // ---- publish the reflector tails ----
#pragma unroll
    for (int r = 0; r < RPT; ++r)
#pragma unroll
      for (int c = 0; c < 16; ++c) sm.Ys[(size_t)(c0 + c) * H + r * kT + t] = a[r][c];
    __syncthreads();

    if (BUILD && p > 0) {
      // T[0:c0, J] = -T[0:c0,0:c0] * (Y_prev^T V_p) * T_p
      for (int q = 0; q < p; ++q) {
        const int q0 = q * kPanel;
        float* part = sm.part + buf * 1024;
        gram16<RPT>(a, sm.Ys + (size_t)q0 * H, H, t, part, lane, warp);
        __syncthreads();
        buf ^= 1;
#pragma unroll
        for (int v = t; v < 256; v += kT) sm.Ssm[q0 * 16 + v] = sum4(part + v * 4);
      }
      __syncthreads();
      for (int idx = t; idx < c0 * 16; idx += kT) {
        const int c = idx >> 4, k = idx & 15;
        float x = 0.f;
        for (int d = 0; d <= k; ++d) x = __fmaf_rn(sm.Ssm[c * 16 + d], sm.Ts[(c0 + d) * np + c0 + k], x);
        sm.Xsm[idx] = x;
      }
      __syncthreads();
      for (int idx = t; idx < c0 * 16; idx += kT) {
        const int c = idx >> 4, k = idx & 15;
        float x = 0.f;
        for (int d = c; d < c0; ++d) x = __fmaf_rn(sm.Ts[c * np + d], sm.Xsm[d * 16 + k], x);
        sm.Ts[c * np + c0 + k] = -x;
      }
      __syncthreads();
    }
  }
}

template <int RPT, bool BUILD, typename TIn, typename TOut>
__global__ void __launch_bounds__(kPassThreads) ts_pass_kernel(const PassParams P) {
  extern __shared__ __align__(16) float smem_raw[];
  const int t = threadIdx.x;
  const int np = P.np, n = P.n, H = kT * RPT;
  Smem sm;
  sm.Rs = smem_raw;
  sm.Ys = sm.Rs + np * np;
  sm.Ts = sm.Ys + (size_t)np * H;
  sm.part = sm.Ts + np * np;
  sm.tot = sm.part + 2048;
  sm.Wsm = sm.tot + 256;
  sm.rowb = sm.Wsm + 256;
  sm.Bs = sm.rowb + 32;
  sm.Ssm = sm.Bs + np * np;
  sm.Xsm = sm.Ssm + 768;

  const int chunk = blockIdx.x;
  const long long cbeg = (long long)chunk * P.rows_per_chunk;
  long long cend = cbeg + P.rows_per_chunk;
  if (cend > P.M) cend = P.M;
  const int ntiles = cend > cbeg ? (int)((cend - cbeg + H - 1) / H) : 0;
  const size_t ck_stride = (size_t)np * (np + 1) / 2;
  float* ck_base = P.ckpt + (size_t)chunk * P.tiles_per_chunk * ck_stride;
  int buf = 0;

  for (int i = t; i < np * np; i += kT) { sm.Rs[i] = 0.f; sm.Ts[i] = 0.f; }

  if (!BUILD) {
    __syncthreads();
    for (int tile = 0; tile < ntiles; ++tile) {
      if (tile > 0) {
        float* ck = ck_base + (size_t)tile * ck_stride;
        for (int j = 0; j < np; ++j) {
          const int k = j + t;
          if (k < np) ck[(size_t)j * np - (size_t)j * (j - 1) / 2 + (k - j)] = sm.Rs[j * np + k];
        }
      }
      if (tile == 0) factor_tile<RPT, false, true, TIn>(P, sm, cbeg, cend, buf);
      else factor_tile<RPT, false, false, TIn>(P, sm, cbeg + (long long)tile * H, cend, buf);
    }
    // stacked output block: rows [chunk*n, chunk*n+n)
    for (int i = t; i < n * n; i += kT) {
      const int j = i % n, k = i / n;
      P.S[(long long)chunk * n + j + (long long)k * P.lds] = (k >= j) ? sm.Rs[j * np + k] : 0.f;
    }
    return;
  } else {
    if (ntiles == 0) return;
    TOut* Q = reinterpret_cast<TOut*>(P.Q);
    for (int i = t; i < np * np; i += kT) {
      const int c = i / np, k = i % np;
      float b = 0.f;
      if (c < n && k < n) b = P.Bsrc ? P.Bsrc[(long long)chunk * n + c + (long long)k * P.ldb] : (c == k ? 1.f : 0.f);
      sm.Bs[i] = b;
    }
    for (int tile = ntiles - 1; tile >= 0; --tile) {
      __syncthreads();
      {
        const float* ck = ck_base + (size_t)tile * ck_stride;
        for (int i = t; i < np * np; i += kT) {
          const int j = i / np, k = i % np;
          float r = 0.f;
          if (tile > 0 && k >= j) r = ck[(size_t)j * np - (size_t)j * (j - 1) / 2 + (k - j)];
          sm.Rs[i] = r;
          sm.Ts[i] = 0.f;
        }
      }
      __syncthreads();
      const long long row0 = cbeg + (long long)tile * H;
      if (tile == 0) factor_tile<RPT, true, true, TIn>(P, sm, row0, cend, buf);
      else factor_tile<RPT, true, false, TIn>(P, sm, row0, cend, buf);
      const int lane = t & 31, warp = t >> 5;
      const float* Bin = sm.Bs;
      if (tile == 0) {
        // first tile: Q = [B;0] - V (T (V1^T B)); V1 = rows 0..np-1 of V. B' = V1^T B goes to Rs
        // (R of this tile is not needed any more in the build pass).
        for (int c = warp; c < np; c += 4) {
          float acc0 = 0.f, acc1 = 0.f;
          for (int i = c; i < np; ++i) {
            const float v = sm.Ys[(size_t)c * H + i];
            if (lane < np) acc0 = __fmaf_rn(v, sm.Bs[i * np + lane], acc0);
            if (lane + 32 < np) acc1 = __fmaf_rn(v, sm.Bs[i * np + lane + 32], acc1);
          }
          if (lane < np) sm.Rs[c * np + lane] = acc0;
          if (lane + 32 < np) sm.Rs[c * np + lane + 32] = acc1;
        }
        __syncthreads();
        Bin = sm.Rs;
      }
      // W2 = T * B, written over T row by row (a warp owns whole rows)
      {
        for (int c = warp; c < np; c += 4) {
          float acc0 = 0.f, acc1 = 0.f;
          for (int d = c; d < np; ++d) {
            const float tcd = sm.Ts[c * np + d];
            if (lane < np) acc0 = __fmaf_rn(tcd, Bin[d * np + lane], acc0);
            if (lane + 32 < np) acc1 = __fmaf_rn(tcd, Bin[d * np + lane + 32], acc1);
          }
          __syncwarp();
          if (lane < np) sm.Ts[c * np + lane] = acc0;
          if (lane + 32 < np) sm.Ts[c * np + lane + 32] = acc1;
        }
      }
      __syncthreads();
      // Q_tile = -Y * W2 (thread per row, 16 output columns at a time)
      for (int k0 = 0; k0 < np; k0 += 16) {
        float z[RPT][16];
#pragma unroll
        for (int r = 0; r < RPT; ++r)
#pragma unroll
          for (int k = 0; k < 16; ++k) z[r][k] = (tile == 0 && r == 0 && t < np) ? sm.Bs[t * np + k0 + k] : 0.f;
        for (int c = 0; c < np; ++c) {
          float w[16];
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            float4 v4 = *reinterpret_cast<const float4*>(sm.Ts + c * np + k0 + k4 * 4);
            w[k4 * 4 + 0] = v4.x; w[k4 * 4 + 1] = v4.y; w[k4 * 4 + 2] = v4.z; w[k4 * 4 + 3] = v4.w;
          }
#pragma unroll
          for (int r = 0; r < RPT; ++r) {
            const float y = sm.Ys[(size_t)c * H + r * kT + t];
#pragma unroll
            for (int k = 0; k < 16; ++k) z[r][k] = __fmaf_rn(-y, w[k], z[r][k]);
          }
        }
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
          const long long row = row0 + r * kT + t;
          if (row < cend) {
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (k0 + k < n) io_store(Q + row + (long long)(k0 + k) * P.ldq, z[r][k]);
          }
        }
      }
      __syncthreads();
      for (int i = t; i < np * np; i += kT) sm.Bs[i] = __fsub_rn(sm.Bs[i], sm.Ts[i]);
    }
  }
}

template <int RPT, bool BUILD, typename TIn, typename TOut>
void launch_one(const PassParams& p, cudaStream_t s) {
  const int H = kT * RPT;
  const size_t bytes = smem_floats(p.np, H, BUILD) * sizeof(float);
  auto kern = ts_pass_kernel<RPT, BUILD, TIn, TOut>;
  TSQR_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  kern<<<p.nchunks, kPassThreads, bytes, s>>>(p);
  note_launch();
  TSQR_CUDA_CHECK(cudaGetLastError());
}

template <bool BUILD, typename TIn, typename TOut>
void launch_rpt(const PassParams& p, cudaStream_t s) {
  switch (pass_rpt(p.np)) {
    case 4: launch_one<4, BUILD, TIn, TOut>(p, s); break;
    case 2: launch_one<2, BUILD, TIn, TOut>(p, s); break;
    default: launch_one<1, BUILD, TIn, TOut>(p, s); break;
  }
}

}  // namespace

void launch_pass_simt(const PassParams& p, bool build, IoType in, IoType out, cudaStream_t s) {
  if (!build) {
    if (in == IoType::f32) launch_rpt<false, float, float>(p, s);
    else launch_rpt<false, __half, float>(p, s);
  } else {
    if (in == IoType::f32 && out == IoType::f32) launch_rpt<true, float, float>(p, s);
    else if (in == IoType::f16 && out == IoType::f16) launch_rpt<true, __half, __half>(p, s);
    else if (in == IoType::f32 && out == IoType::f16) launch_rpt<true, float, __half>(p, s);
    else launch_rpt<true, __half, float>(p, s);
  }
}

int pass_simt_max_ctas_per_sm(int np, bool build) {
  const int H = kT * pass_rpt(np);
  const size_t bytes = smem_floats(np, H, build) * sizeof(float);
  int per_sm = (int)((227 * 1024) / (bytes + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  return per_sm;
}

}  // namespace tsqr_b200
