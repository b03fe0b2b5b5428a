// placeholder until the fused DMMA kernel lands (next commit): reports "unsupported" so every handle
// runs the generic two-kernel path.
#include "sw2d_common.cuh"

int fused_supported(const bdg_sw2d *) { return 0; }
cudaError_t fused_prepare(bdg_sw2d *) { return cudaSuccess; }
cudaError_t launch_fused_stage(bdg_sw2d *, const double *, double *, double *, int, int, double, double, int, double *) {
  return cudaErrorNotSupported;
}
void fused_release(bdg_sw2d *) {}
