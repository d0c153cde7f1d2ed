// Conversion between the two HBM formats of the carried ADMM state (admm_device.cuh):
//   element-major  g, v, q, y, z, r, pN as [element][ld] arrays  (coalesced for one thread per instance)
//   instance-major one contiguous block per instance, ImLayout     (whole sectors for one sub-warp per instance)
// Runs only when consecutive warm-started solves of a batch use kernels of different families, which the library
// never does on its own (tests do, to prove the formats carry the same state).
#include <cuda_runtime.h>

#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace tmpc {
namespace {

// T = double, or float for the float build (both formats then hold floats in the first half of their allocations)
template <class T>
__global__ void state_convert_kernel(const __grid_constant__ BatchView bv, int to_im) {
    const int nx = bv.nx, nu = bv.nu, N = bv.N;
    const int nxN = nx * N, nuN = nu * (N - 1);
    const int per = 3 * nxN + 3 * nuN + nx;              // scalars per instance
    const long long total = static_cast<long long>(per) * bv.B;
    const size_t ld = static_cast<size_t>(bv.ld);
    const ImLayout L(nx, nu, N, sizeof(T) == 4 ? 2 : 1);
    for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        // idx enumerates (field element, instance) with the instance fastest: coalesced on the element-major side
        const int b = static_cast<int>(idx % bv.B);
        int e = static_cast<int>(idx / bv.B);
        T* em;           // element-major address
        int im;          // offset in the instance block
        if (e < 3 * nxN) {
            const int f = e / nxN, k = e % nxN, i = k / nx, row = k % nx;       // f: 0 g, 1 v, 2 q
            em = reinterpret_cast<T*>(f == 0 ? bv.g : f == 1 ? bv.v[0] : bv.q) + static_cast<size_t>(k) * ld + b;
            im = L.x_off(f, i, row);
        } else if (e < 3 * nxN + 3 * nuN) {
            e -= 3 * nxN;
            const int f = e / nuN, k = e % nuN, i = k / nu, row = k % nu;       // f: 0 y, 1 z, 2 r
            em = reinterpret_cast<T*>(f == 0 ? bv.y : f == 1 ? bv.z[0] : bv.r) + static_cast<size_t>(k) * ld + b;
            im = L.u_off(f, i, row);
        } else {
            e -= 3 * nxN + 3 * nuN;
            em = reinterpret_cast<T*>(bv.pN) + static_cast<size_t>(e) * ld + b;
            im = L.p_off(e);
        }
        T* blk = reinterpret_cast<T*>(bv.state_im) + static_cast<size_t>(b) * bv.state_im_stride + im;
        if (to_im) *blk = *em; else *em = *blk;
    }
}

}  // namespace

cudaError_t launch_state_convert(const BatchView& bv, bool to_instance_major, cudaStream_t stream, int precision) {
    if (bv.B <= 0) return cudaSuccess;
    if (!bv.state_im) return cudaErrorInvalidValue;
    if (precision == 32) state_convert_kernel<float><<<148 * 8, 256, 0, stream>>>(bv, to_instance_major ? 1 : 0);
    else state_convert_kernel<double><<<148 * 8, 256, 0, stream>>>(bv, to_instance_major ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace tmpc
