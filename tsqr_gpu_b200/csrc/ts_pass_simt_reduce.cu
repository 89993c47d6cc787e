// CUDA-core TSQR level pass, R-reduce instantiations (see ts_pass_simt.cu / ts_pass_simt.inc).
#include "ts_pass_simt.inc"
}  // namespace

void launch_simt_reduce(const PassParams& p, IoType in, cudaStream_t s) {
  if (in == IoType::f32) launch_rpt<false, false, float, float>(p, s);
  else launch_rpt<false, false, __half, float>(p, s);
}

}  // namespace tsqr_b200
