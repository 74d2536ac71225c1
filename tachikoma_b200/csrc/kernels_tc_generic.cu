// Generic- and raw-epilogue instantiations of the persistent tcgen05 kernel (tc_kernel.cuh): every
// rounding mode, compute dtype and the oneDNN-style float post-ops go through epilogue_one(); plain
// qnn.conv2d / qnn.dense / qnn.batch_matmul (int32 accumulators out) use the vectorised raw path.
#include "tc_kernel.cuh"

namespace tkb {

int launch_tc_other(int bn, int mode, const CUtensorMap& ta, const CUtensorMap& tb, TcKernelParams& p, cudaStream_t s) {
  if (mode == kModeRaw) {
    if (bn == 64) return launch_t<64, kModeRaw>(ta, tb, p, s);
    if (bn == 128) return launch_t<128, kModeRaw>(ta, tb, p, s);
    return launch_t<256, kModeRaw>(ta, tb, p, s);
  }
  if (bn == 64) return launch_t<64, kModeGeneric>(ta, tb, p, s);
  if (bn == 128) return launch_t<128, kModeGeneric>(ta, tb, p, s);
  return launch_t<256, kModeGeneric>(ta, tb, p, s);
}

}  // namespace tkb
