// Shared device/host helpers for the tsqr_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

namespace tsqr_b200 {

// CUDA failures surface as std::runtime_error, like the reference's CUTF_CHECK_ERROR
// (reference src/cutf/cutf/cuda.hpp:11-21); the C ABI maps them to negative codes.
inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
  if (e != cudaSuccess) {
    throw std::runtime_error(std::string("tsqr_b200 CUDA error: ") + cudaGetErrorString(e) +
                             " [" + what + "] at " + file + ":" + std::to_string(line));
  }
}
#define TSQR_CUDA_CHECK(x) ::tsqr_b200::cuda_check((x), #x, __FILE__, __LINE__)

__device__ __forceinline__ float io_load(const float* p) { return __ldg(p); }
__device__ __forceinline__ float io_load(const __half* p) { return __half2float(__ldg(p)); }
__device__ __forceinline__ float io_plain(const float* p) { return *p; }
__device__ __forceinline__ float io_plain(const __half* p) { return __half2float(*p); }
__device__ __forceinline__ void io_store(float* p, float v) { *p = v; }
__device__ __forceinline__ void io_store(__half* p, float v) { *p = __float2half_rn(v); }

constexpr unsigned kFullMask = 0xffffffffu;

// message behind tsqr_b200_last_error() of the calling thread (api.cu)
void set_error_message(const char* msg);

// every kernel launch of the library is counted (reported by tsqr_b200_launch_count)
void note_launch(unsigned n = 1);

// packed fp32 pairs (Blackwell FFMA2): two fused multiply-adds per instruction, same rounding
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

}  // namespace tsqr_b200
