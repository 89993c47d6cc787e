// CUDA-core TSQR level pass, Q-build instantiations of the _emu compute modes (see ts_pass_simt_emu.cu).
#include "ts_pass_simt.inc"
}  // namespace

void launch_simt_emu_build(const PassParams& p, cudaStream_t s) { launch_rpt<true, true, float, float>(p, s); }

}  // namespace tsqr_b200
