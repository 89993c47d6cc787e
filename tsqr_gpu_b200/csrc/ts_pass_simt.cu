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
#include "ts_pass_simt.inc"
}  // namespace

// instantiated in ts_pass_simt_reduce.cu / ts_pass_simt_build.cu / ts_pass_simt_emu.cu
void launch_simt_reduce(const PassParams& p, IoType in, cudaStream_t s);
void launch_simt_build(const PassParams& p, IoType in, IoType out, cudaStream_t s);
void launch_simt_emu(const PassParams& p, bool build, cudaStream_t s);

void launch_pass_simt(const PassParams& p, bool build, IoType in, IoType out, cudaStream_t s) {
  if (p.emu != 0) {
    if (in != IoType::f32 || out != IoType::f32) throw std::runtime_error("tsqr_b200: the _emu modes take float matrices");
    launch_simt_emu(p, build, s);
  } else if (!build) {
    launch_simt_reduce(p, in, s);
  } else {
    launch_simt_build(p, in, out, s);
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
