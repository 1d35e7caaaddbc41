// gemm_tc.cu — TF32 tcgen05/TMEM contractions (placeholder until the tensor-core path lands).
#include "kernels.hpp"
namespace sn {
bool tc_available() { return false; }
#define SN_TC_STUB fatal("TF32 tcgen05 kernels are not built into this library yet")
void tc_conv_fwd(cudaStream_t, const ConvGeom&, const float*, const float*, const float*, float*, bool) { SN_TC_STUB; }
void tc_conv_dgrad(cudaStream_t, const ConvGeom&, const float*, const float*, float*) { SN_TC_STUB; }
void tc_conv_wgrad(cudaStream_t, const ConvGeom&, const float*, const float*, float*, bool) { SN_TC_STUB; }
void tc_ip_fwd(cudaStream_t, int, int, int, const float*, const float*, const float*, float*, bool) { SN_TC_STUB; }
void tc_ip_dgrad(cudaStream_t, int, int, int, const float*, const float*, float*) { SN_TC_STUB; }
void tc_ip_wgrad(cudaStream_t, int, int, int, const float*, const float*, float*, bool) { SN_TC_STUB; }
}  // namespace sn
