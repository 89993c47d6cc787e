// CUDA-core TSQR level pass, Q-build instantiations (see ts_pass_simt.cu / ts_pass_simt.inc).
#include "ts_pass_simt.inc"
}  // namespace

void launch_simt_build(const PassParams& p, IoType in, IoType out, cudaStream_t s) {
  if (in == IoType::f32 && out == IoType::f32) launch_rpt<true, false, float, float>(p, s);
  else if (in == IoType::f16 && out == IoType::f16) launch_rpt<true, false, __half, __half>(p, s);
  else if (in == IoType::f32 && out == IoType::f16) launch_rpt<true, false, float, __half>(p, s);
  else launch_rpt<true, false, __half, float>(p, s);
}

}  // namespace tsqr_b200
