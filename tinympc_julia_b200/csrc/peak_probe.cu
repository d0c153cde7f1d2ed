// Measurement aid: the FP64 / FP32 FMA issue peak of the device the batch runs on.
//
// The ADMM solve loop is bound by the CUDA-core FMA pipe at the reference's `tinytype`
// (double, reference types.hpp:15), and MEASURED_PEAKS.json carries no CUDA-core figure, so the
// roofline denominator is measured live, on the same box and under the same clocks as the solve:
// register-resident, dependency-free FMA chains (8 independent accumulators per thread, enough
// warps per SM to cover the pipe latency), timed with CUDA events.  Nothing in the solve path
// calls this.
#include <cuda_runtime.h>

#include "../../include/tinympc_b200.h"

namespace {

template <typename T> struct Fma;
template <> struct Fma<double> {
    static __device__ __forceinline__ void op(double& a, double b, double c) { asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a) : "d"(b), "d"(c)); }
};
template <> struct Fma<float> {
    static __device__ __forceinline__ void op(float& a, float b, float c) { asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a) : "f"(b), "f"(c)); }
};

constexpr int kIlp = 8, kInner = 16;

template <typename T>
__global__ void __launch_bounds__(256) fma_peak_kernel(T* out, int iters, T b, T c) {
    T a[kIlp];
#pragma unroll
    for (int j = 0; j < kIlp; ++j) a[j] = static_cast<T>(threadIdx.x) * static_cast<T>(1e-6) + static_cast<T>(j);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < kInner; ++r) {
#pragma unroll
            for (int j = 0; j < kIlp; ++j) Fma<T>::op(a[j], b, c);
        }
    }
    T s = 0;
#pragma unroll
    for (int j = 0; j < kIlp; ++j) s += a[j];
    out[static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x] = s;
}

template <typename T>
int measure(int device, int reps, double* tflops) {
    if (!tflops) return TMPC_ERR_DIM;
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return TMPC_ERR_CUDA;
    cudaDeviceProp prop{};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return TMPC_ERR_CUDA;
    const int block = 256, grid = prop.multiProcessorCount * 8, iters = 4096;
    T* out = nullptr;
    if (cudaMalloc(reinterpret_cast<void**>(&out), sizeof(T) * static_cast<size_t>(grid) * block) != cudaSuccess) return TMPC_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0.0;
    if (reps < 1) reps = 1;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0, 0);
        fma_peak_kernel<T><<<grid, block>>>(out, iters, static_cast<T>(0.999999), static_cast<T>(1e-9));
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return TMPC_ERR_CUDA; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flop = 2.0 * static_cast<double>(grid) * block * kIlp * kInner * iters;
        const double tf = flop / (ms * 1e-3) / 1e12;
        if (r > 0 && tf > best) best = tf;      // first launch is the warm-up
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return cudaGetLastError() == cudaSuccess ? TMPC_OK : TMPC_ERR_CUDA;
}

}  // namespace

extern "C" int tmpc_measure_fma_peak(int device, int fp64, int reps, double* tflops) {
    return fp64 ? measure<double>(device, reps, tflops) : measure<float>(device, reps, tflops);
}
