// CUDA-core TSQR level pass with the operand rounding of the three _emu compute modes (reference src/matmul.hpp:26-209;
// see ts_pass_simt.cu / ts_pass_simt.inc): the R-reduce instantiations, and the dispatch to the Q-build ones
// (ts_pass_simt_emu_build.cu).
#include "ts_pass_simt.inc"
}  // namespace

void launch_simt_emu_build(const PassParams& p, cudaStream_t s);

void launch_simt_emu(const PassParams& p, bool build, cudaStream_t s) {
  if (build) launch_simt_emu_build(p, s);
  else launch_rpt<false, true, float, float>(p, s);
}

}  // namespace tsqr_b200
