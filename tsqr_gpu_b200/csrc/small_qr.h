#pragma once
#include <cuda_runtime.h>
#include <cstddef>
namespace tsqr_b200 {
// starts == nullptr: a single block made of rows [0, m)
template <typename QT, typename RT, typename AT>
void small_qr_launch(QT* q, size_t ldq, RT* r, size_t ldr, const AT* a, size_t lda, unsigned m, unsigned n,
                     size_t batch, const unsigned* starts, cudaStream_t s);
}  // namespace tsqr_b200
