// tcgen05 tap GEMM -- placeholder until the UMMA kernel lands (next commit).
#include "tapgemm.cuh"
namespace spanon {
bool tapgemm_tc_available() { return false; }
int launch_tapgemm_tc(const TapGemmParams&, int, cudaStream_t) { return SPANON_ERR_UNSUPPORTED; }
}  // namespace spanon
