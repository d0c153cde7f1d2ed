// Closed-loop helpers: the plant step and per-step bookkeeping of a batched MPC rollout.
//
// Reference pattern being batched (examples/cartpole_example_mpc.jl:35-51): after every solve the
// first control of the returned solution (the clamped slack znew(:,0)) drives the plant,
// x <- A x + B u + f, and the result becomes the next x0; nothing else in the solver is reset.
#include <cuda_runtime.h>

#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace tmpc {
namespace {

__global__ void plant_step_kernel(const __grid_constant__ BatchView bv, double* __restrict__ x0) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bv.B) return;
    const int nx = bv.nx, nu = bv.nu;
    double* x = x0 + static_cast<size_t>(b) * nx;
    const double* u = bv.Usol + static_cast<size_t>(b) * nu * (bv.N - 1);
    double xn[kThreadKernelMaxNx];
    for (int j = 0; j < nx; ++j) {
        double s1 = 0.0, s2 = 0.0;
        for (int k = 0; k < nx; ++k) s1 = fma(__ldg(bv.A + k * nx + j), x[k], s1);
        for (int k = 0; k < nu; ++k) s2 = fma(__ldg(bv.Bm + k * nx + j), u[k], s2);
        xn[j] = dadd(dadd(s1, s2), __ldg(bv.f + j));
    }
    for (int j = 0; j < nx; ++j) x[j] = xn[j];
}

// the same with each instance's own A, B, f (element-major [e][ld], tmpc_batch_set_instance_dynamics)
__global__ void plant_step_pi_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ PerInstanceMats pm,
                                     long long ld, double* __restrict__ x0) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bv.B) return;
    const int nx = bv.nx, nu = bv.nu;
    double* x = x0 + static_cast<size_t>(b) * nx;
    const double* u = bv.Usol + static_cast<size_t>(b) * nu * (bv.N - 1);
    auto at = [&](const double* a, int e) { return __ldg(a + static_cast<size_t>(e) * ld + b); };
    double xn[kThreadKernelMaxNx];
    for (int j = 0; j < nx; ++j) {
        double s1 = 0.0, s2 = 0.0;
        for (int k = 0; k < nx; ++k) s1 = fma(at(pm.A, k * nx + j), x[k], s1);
        for (int k = 0; k < nu; ++k) s2 = fma(at(pm.B, k * nx + j), u[k], s2);
        xn[j] = dadd(dadd(s1, s2), at(pm.f, j));
    }
    for (int j = 0; j < nx; ++j) x[j] = xn[j];
}

__global__ void record_step_kernel(const __grid_constant__ BatchView bv, int* __restrict__ status_out,
                                   int* __restrict__ iters_out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bv.B) return;
    if (status_out) status_out[b] = bv.status[b];
    if (iters_out) iters_out[b] = bv.iters[b];
}

// first control of every instance, packed nu x B: what a closed loop consumes from a solve (cartpole_example_mpc.jl:41-42)
// (optionally also the status / iteration words: with `out`, `st_out`, `it_out` in mapped pinned host memory this one
// kernel is the whole device-to-host leg of a closed-loop step)
__global__ void gather_u0_kernel(const double* __restrict__ Usol, long long stride, int nu, int B, double* __restrict__ out,
                                 const int* __restrict__ st_in, const int* __restrict__ it_in, int* __restrict__ st_out,
                                 int* __restrict__ it_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B) {
        if (st_out) st_out[i] = st_in[i];
        if (it_out) it_out[i] = it_in[i];
    }
    if (i >= B * nu) return;
    const int b = i / nu, j = i - b * nu;
    out[i] = Usol[static_cast<size_t>(b) * stride + j];
}

}  // namespace

cudaError_t launch_gather_u0(const double* Usol, long long stride, int nu, int B, double* out, cudaStream_t stream,
                             const int* st_in, const int* it_in, int* st_out, int* it_out) {
    if (B <= 0) return cudaSuccess;
    const int block = 256, grid = (B * nu + block - 1) / block;
    gather_u0_kernel<<<grid, block, 0, stream>>>(Usol, stride, nu, B, out, st_in, it_in, st_out, it_out);
    return cudaGetLastError();
}

cudaError_t launch_plant_step(const BatchView& bv, double* x0_rw, cudaStream_t stream) {
    const int block = 128, grid = (bv.B + block - 1) / block;
    plant_step_kernel<<<grid, block, 0, stream>>>(bv, x0_rw);
    return cudaGetLastError();
}

cudaError_t launch_plant_step_per_instance(const BatchView& bv, const PerInstanceMats& pm, long long ld, double* x0_rw,
                                           cudaStream_t stream) {
    const int block = 128, grid = (bv.B + block - 1) / block;
    plant_step_pi_kernel<<<grid, block, 0, stream>>>(bv, pm, ld, x0_rw);
    return cudaGetLastError();
}

cudaError_t launch_record_step(const BatchView& bv, int* status_out, int* iters_out, cudaStream_t stream) {
    const int block = 256, grid = (bv.B + block - 1) / block;
    record_step_kernel<<<grid, block, 0, stream>>>(bv, status_out, iters_out);
    return cudaGetLastError();
}

}  // namespace tmpc
