// Longest-first instance queue of the on-chip solve kernels.
//
// The TMEM kernels are persistent: sub-warps pull instance ids from a global ticket counter.  ADMM iteration counts
// are data-dependent (reference admm.cpp:312-381: 5 ... max_iter), so an instance that needs 25 iterations and is
// pulled in the last wave holds one sub-warp for three average solves while 147 SMs idle (measured: the whole
// 0.90 weak-scaling efficiency of round 1, and ~10 % of single-GPU time on typical seeds).  Here the ticket indexes a
// permutation that lists the instances expected to run longest first:
//
//   * cold start: key = saturation of the unconstrained LQR input against the input box,
//         p_b = max_j |(Kinf x0_b)_j| / min(|u_min_j|, |u_max_j|)       (knot 0 of the bound table)
//     Checked against the compiled reference on seeds 3..10 of config 3: under descending p every instance above 12
//     iterations sits in queue positions 0..17 of 65,536 (easy); correlation with the iteration count 0.81 (hard).
//   * warm start: key = iteration count of the instance's previous solve.
//
// Keys are quantised to 256 buckets and the permutation is a counting sort (histogram kernel + scatter kernel, a few
// microseconds for 65,536 instances).  Results do not depend on the order (instances are independent), only the
// makespan does.
#include <cuda_runtime.h>

#include <cmath>

#include "admm_kernels.hpp"

namespace tmpc {
namespace {

constexpr int kBlock = 256;

__global__ void __launch_bounds__(kBlock)
order_key_kernel(const OrderArgs a) {
    __shared__ int hist[kOrderBuckets];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int b = blockIdx.x * kBlock + threadIdx.x;
    if (b < a.B) {
        int key = 0;
        if (a.prev_iters != nullptr) {
            key = a.prev_iters[b];
        } else {
            // max_j |Kinf(j,:) x0| / bound_j ; Kinf column-major nu x nx: shared (stride 1 between entries, kinf_ld = 0 marks
            // "one matrix") or per instance, element-major [e][ld]
            double p = 0.0;
            const double* x0 = a.x0 + static_cast<size_t>(b) * a.nx;
            for (int j = 0; j < a.nu; ++j) {
                const double lo = a.u_min ? fabs(a.u_min[(a.ub_stride ? static_cast<size_t>(b) * a.ub_stride : 0) + j]) : INFINITY;
                const double hi = a.u_max ? fabs(a.u_max[(a.ub_stride ? static_cast<size_t>(b) * a.ub_stride : 0) + j]) : INFINITY;
                const double bound = fmin(lo, hi);
                if (!(bound > 0.0) || isinf(bound)) continue;
                double s = 0.0;
                for (int k = 0; k < a.nx; ++k) {
                    const double kv = a.kinf_ld ? a.Kinf[static_cast<size_t>(k * a.nu + j) * a.kinf_ld + b] : a.Kinf[k * a.nu + j];
                    s = fma(kv, x0[k], s);
                }
                p = fmax(p, fabs(s) / bound);
            }
            p *= 64.0;                       // 1/64 of the bound per bucket
            key = p < 255.0 ? static_cast<int>(p) : 255;
        }
        key = key < 0 ? 0 : (key > kOrderBuckets - 1 ? kOrderBuckets - 1 : key);
        a.key[b] = static_cast<unsigned char>(key);
        atomicAdd(&hist[key], 1);
    }
    __syncthreads();
    if (hist[threadIdx.x]) atomicAdd(&a.hist[threadIdx.x], hist[threadIdx.x]);
}

__global__ void __launch_bounds__(kBlock)
order_scatter_kernel(const OrderArgs a) {
    __shared__ int base[kOrderBuckets];      // instances with a LARGER key (they go first)
    __shared__ int cnt[kOrderBuckets];       // of this block, per key
    __shared__ int start[kOrderBuckets];     // of this block's run inside the bucket
    const int t = threadIdx.x;
    // suffix sum over the 256 buckets: Hillis-Steele in shared memory
    base[t] = a.hist[t];
    cnt[t] = 0;
    __syncthreads();
    for (int off = 1; off < kOrderBuckets; off <<= 1) {
        const int v = (t + off < kOrderBuckets) ? base[t + off] : 0;
        __syncthreads();
        base[t] += v;
        __syncthreads();
    }
    const int b = blockIdx.x * kBlock + t;
    int key = 0, mine = 0;
    if (b < a.B) {
        key = a.key[b];
        mine = atomicAdd(&cnt[key], 1);
    }
    __syncthreads();
    if (cnt[t]) start[t] = atomicAdd(&a.cursor[t], cnt[t]);
    __syncthreads();
    if (b < a.B) a.order[base[key] - a.hist[key] + start[key] + mine] = b;   // base is inclusive: subtract the own bucket
}

}  // namespace

cudaError_t launch_build_order(const OrderArgs& a, cudaStream_t stream) {
    if (a.B <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(a.hist, 0, sizeof(int) * 2 * kOrderBuckets, stream);   // hist + cursor are adjacent
    if (e != cudaSuccess) return e;
    const int grid = (a.B + kBlock - 1) / kBlock;
    order_key_kernel<<<grid, kBlock, 0, stream>>>(a);
    order_scatter_kernel<<<grid, kBlock, 0, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace tmpc
