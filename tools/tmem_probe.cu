// Probe: can Tensor Memory (TMEM, 256 KB/SM) serve as per-lane private scratch for a
// non-MMA kernel?  The ADMM kernel's per-instance state (duals, slacks) is the capacity
// limit on resident instances per SM (DESIGN.md "Data layout"); TMEM would double the
// on-chip bytes available.  This probe (a) checks a tcgen05.st -> tcgen05.ld round trip
// with dynamically computed column addresses, (b) times back-to-back 32x32b.x4 loads and
// stores per SM.  Run under `timeout`; it is not part of the product.
//
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tmem_probe tmem_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ void tmem_st4(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" :: "r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

template <int NCOLS>
__global__ void __launch_bounds__(256) k_probe(unsigned* errors, long long* cyc, int iters, int nwarps_use) {
  __shared__ uint32_t s_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "l"((uint64_t)__cvta_generic_to_shared(&s_base)), "n"(NCOLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = s_base;
  // warp w may touch TMEM lanes 32*(w%4) .. +31; warps w and w+4 split the columns
  const int nwarps = blockDim.x >> 5;
  const int cols_per_warp = NCOLS / ((nwarps + 3) / 4);
  const uint32_t my = base + ((uint32_t)(warp & 3) * 32u << 16) + (uint32_t)((warp >> 2) * cols_per_warp);
  unsigned err = 0;
  // (a) round trip with runtime column index
  for (int c = 0; c < cols_per_warp; c += 4) {
    uint32_t v = (uint32_t)(blockIdx.x * 1000003u + threadIdx.x * 4099u + c * 17u);
    tmem_st4(my + c, v, v + 1, v + 2, v + 3);
  }
  tmem_wait_st();
  for (int c = cols_per_warp - 4; c >= 0; c -= 4) {
    uint32_t v = (uint32_t)(blockIdx.x * 1000003u + threadIdx.x * 4099u + c * 17u);
    uint32_t a, b, cc, d;
    tmem_ld4(my + c, a, b, cc, d);
    tmem_wait_ld();
    err += (a != v) + (b != v + 1) + (cc != v + 2) + (d != v + 3);
  }
  // (b) throughput: loads
  __syncthreads();
  uint32_t acc = 0;
  long long t0 = clock64();
  if (warp < nwarps_use) {
    for (int it = 0; it < iters; ++it) {
      uint32_t r[8][4];
#pragma unroll
      for (int k = 0; k < 8; ++k) tmem_ld4(my + ((it * 32 + k * 4) & (cols_per_warp - 1)), r[k][0], r[k][1], r[k][2], r[k][3]);
      tmem_wait_ld();
#pragma unroll
      for (int k = 0; k < 8; ++k) acc ^= r[k][0] ^ r[k][3];
    }
  }
  long long t1 = clock64();
  __syncthreads();
  long long t2 = clock64();
  if (warp < nwarps_use) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 8; ++k) tmem_st4(my + ((it * 32 + k * 4) & (cols_per_warp - 1)), acc, it, k, lane);
      tmem_wait_st();
    }
  }
  long long t3 = clock64();
  // (c) latency: dependent st -> wait -> ld -> wait chain
  long long t4 = clock64();
  if (warp == 0) {
    uint32_t a = lane, b = 1, c2 = 2, d = 3;
    for (int it = 0; it < 256; ++it) {
      tmem_st4(my + ((it * 4) & (cols_per_warp - 1)), a, b, c2, d);
      tmem_wait_st();
      tmem_ld4(my + ((it * 4) & (cols_per_warp - 1)), a, b, c2, d);
      tmem_wait_ld();
      a += 1;
    }
    acc ^= a;
  }
  long long t5 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) {
    cyc[blockIdx.x * 3 + 0] = t1 - t0;
    cyc[blockIdx.x * 3 + 1] = t3 - t2;
    cyc[blockIdx.x * 3 + 2] = t5 - t4;
  }
  atomicAdd(errors, err + (acc == 0xdeadbeefu));
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(base), "n"(NCOLS));
  }
}

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int sms = prop.multiProcessorCount;
  unsigned* d_err; long long* d_cyc;
  CK(cudaMalloc(&d_err, 4)); CK(cudaMalloc(&d_cyc, sizeof(long long) * sms * 3));
  const int iters = 2000;
  for (int nw : {1, 4, 8}) {
    CK(cudaMemset(d_err, 0, 4));
    k_probe<512><<<sms, 256>>>(d_err, d_cyc, iters, nw);
    CK(cudaDeviceSynchronize());
    unsigned herr; CK(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
    std::vector<long long> h(sms * 3);
    CK(cudaMemcpy(h.data(), d_cyc, sizeof(long long) * sms * 3, cudaMemcpyDeviceToHost));
    double ld = 0, st = 0, rt = 0;
    for (int i = 0; i < sms; ++i) { ld += h[i * 3]; st += h[i * 3 + 1]; rt += h[i * 3 + 2]; }
    ld /= sms; st /= sms; rt /= sms;
    double bytes = (double)nw * iters * 8 * 512.0;  // per SM: warps * iters * 8 instr * (32 lanes*16 B)
    printf("{\"test\": \"tmem_probe\", \"warps_active\": %d, \"roundtrip_errors\": %u, \"ld_bytes_per_clk_per_sm\": %.1f, "
           "\"st_bytes_per_clk_per_sm\": %.1f, \"ld_cycles_per_x4_instr\": %.2f, \"st_cycles_per_x4_instr\": %.2f, "
           "\"st_wait_ld_wait_chain_cycles\": %.1f}\n",
           nw, herr, bytes / ld, bytes / st, ld / (iters * 8.0), st / (iters * 8.0), rt / 256.0);
  }
  CK(cudaFree(d_err)); CK(cudaFree(d_cyc));
  return 0;
}
