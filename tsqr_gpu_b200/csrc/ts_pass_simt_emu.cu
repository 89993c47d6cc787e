// CUDA-core TSQR level pass with the operand rounding of the three _emu compute modes (reference src/matmul.hpp:26-209;
// see ts_pass_simt.cu / ts_pass_simt.inc).
#include "ts_pass_simt.inc"
}  // namespace

void launch_simt_emu(const PassParams& p, bool build, cudaStream_t s) {
  if (build) launch_rpt<true, true, float, float>(p, s);
  else launch_rpt<false, true, float, float>(p, s);
}

}  // namespace tsqr_b200
