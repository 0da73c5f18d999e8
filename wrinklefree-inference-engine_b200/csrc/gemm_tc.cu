// placeholder until the tcgen05 kernel lands: route to the mma.sync tile kernel
#include "linear_common.cuh"
namespace wf {
int launch_gemm_tc(const wf_linear_args* a, const EpiParams& ep, cudaStream_t s) { return launch_gemm_mma(a, ep, s); }
}
