// Generic-epilogue instantiations of the persistent tcgen05 kernel (tc_kernel.cuh): every rounding mode,
// compute dtype, the oneDNN-style float post-ops and raw int32 output, through epilogue_one().
#include "tc_kernel.cuh"

namespace tkb {

int launch_tc_generic(int bn, const CUtensorMap& ta, const CUtensorMap& tb, TcKernelParams& p, cudaStream_t s) {
  if (bn == 64) return launch_t<64, false>(ta, tb, p, s);
  if (bn == 128) return launch_t<128, false>(ta, tb, p, s);
  return launch_t<256, false>(ta, tb, p, s);
}

}  // namespace tkb
