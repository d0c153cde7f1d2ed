// placeholder until the sub-warp kernel lands
#include "admm_kernels.hpp"
namespace tmpc {
bool subwarp_kernel_supports(int, int, int, const AdmmSettings&) { return false; }
cudaError_t launch_admm_subwarp(const BatchView&, const AdmmSettings&, const HostMats&, int, int*, int, cudaStream_t) { return cudaErrorNotSupported; }
}
