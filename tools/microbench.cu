// Pipe microbenchmarks for the B200 (sm_100a) CUDA-core path the ADMM kernel lives on.
//
// The ADMM solve loop is bound by the FP64 FMA pipe and by how the
// broadcast operand of every FMA is delivered (shared-memory broadcast loads, warp
// shuffles, constant bank).  MEASURED_PEAKS.json has no CUDA-core FMA peak, so this
// program measures the denominators the roofline uses, plus the throughputs of the
// operand-delivery paths that decide the lane mapping:
//   dfma / ffma / dadd / dmul / dmnmx  throughput (per SM per clock) and DFMA latency
//   DFMA with a constant-bank operand
//   SHFL.IDX throughput
//   LDS.64 / LDS.128 with sub-warp broadcast patterns (groups of 4 / 8 lanes), private
//   16 B-per-lane pattern, and a mixed DFMA+LDS.128 stream.
// Output: one JSON object per line on stdout.
//
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o microbench microbench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <string>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

__constant__ double c_coef[64];

// ---------------------------------------------------------------- arithmetic pipes
template <int ILP>
__global__ void k_dfma_reg(double* out, int iters, double b, double c, long long* cyc) {
  double a[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) a[j] = threadIdx.x * 1e-9 + j;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[j]) : "d"(b), "d"(c));
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
__global__ void k_dfma_const(double* out, int iters, double c, long long* cyc) {
  double a[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) a[j] = threadIdx.x * 1e-9 + j;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) a[j] = fma(a[j], c_coef[(r * ILP + j) & 63], c);
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int ILP>
__global__ void k_ffma_reg(float* out, int iters, float b, float c, long long* cyc) {
  float a[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) a[j] = threadIdx.x * 1e-6f + j;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[j]) : "f"(b), "f"(c));
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// op: 0 add, 1 mul, 2 min/max pair
template <int ILP, int OP>
__global__ void k_dop(double* out, int iters, double b, long long* cyc) {
  double a[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) a[j] = threadIdx.x * 1e-9 + j;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        if (OP == 0) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(a[j]) : "d"(b));
        if (OP == 1) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(a[j]) : "d"(b));
        if (OP == 2) asm volatile("max.f64 %0, %0, %1;" : "+d"(a[j]) : "d"(b));
        if (OP == 3) asm volatile("abs.f64 %0, %0;" : "+d"(a[j]));
        if (OP == 4) asm volatile("{ .reg .pred p; setp.lt.f64 p, %0, %1; selp.f64 %0, %1, %0, p; }" : "+d"(a[j]) : "d"(b));
        if (OP == 5) asm volatile("{ .reg .s64 t, u; mov.b64 t, %0; mov.b64 u, %1; max.s64 t, t, u; mov.b64 %0, t; }" : "+d"(a[j]) : "d"(b));
        if (OP == 6) asm volatile("{ max.f64 %0, %0, %1; min.f64 %0, %0, %2; }" : "+d"(a[j]) : "d"(b), "d"(b + 1.0));
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// ---------------------------------------------------------------- shuffle
template <int ILP>
__global__ void k_shfl(int* out, int iters, long long* cyc) {
  int a[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) a[j] = threadIdx.x + j;
  int src = (threadIdx.x & 7) ^ 1;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) a[j] = __shfl_sync(0xffffffffu, a[j], src, 8);
    }
  }
  long long t1 = clock64();
  int s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// ---------------------------------------------------------------- shared-memory loads
// pattern: 0 all lanes one address; 1 groups of 8 lanes -> 4 distinct consecutive vectors;
//          2 groups of 4 lanes -> 8 distinct consecutive vectors; 3 private (lane*VEC bytes)
//          4 groups of 8 lanes, all groups the SAME address (matrix shared across instances)
template <int VEC /*8 or 16 bytes*/, int PATTERN>
__global__ void k_lds(double* out, int iters, long long* cyc) {
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x; i < 8192 / 8; i += blockDim.x) ((double*)smem)[i] = i * 0.5;
  __syncthreads();
  int lane = threadIdx.x & 31;
  int idx;
  if (PATTERN == 0) idx = 0;
  else if (PATTERN == 1) idx = lane >> 3;
  else if (PATTERN == 2) idx = lane >> 2;
  else if (PATTERN == 3) idx = lane;
  else idx = 0;
  unsigned base = (unsigned)__cvta_generic_to_shared(smem) + idx * VEC;
  double acc0 = 0, acc1 = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      // offsets stride by 512 B so that successive loads touch different rows but the same banks pattern
      if (VEC == 16) {
        double x, y;
        asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(base + r * 512));
        acc0 += x; acc1 += y;
      } else {
        double x;
        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(base + r * 512));
        acc0 += x;
      }
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// LDS without the dependent DADDs (pure load issue): results xor-folded through integer regs
template <int VEC, int PATTERN>
__global__ void k_lds_pure(unsigned* out, int iters, long long* cyc) {
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x; i < 8192 / 4; i += blockDim.x) ((unsigned*)smem)[i] = i;
  __syncthreads();
  int lane = threadIdx.x & 31;
  int idx;
  if (PATTERN == 0) idx = 0;
  else if (PATTERN == 1) idx = lane >> 3;
  else if (PATTERN == 2) idx = lane >> 2;
  else if (PATTERN == 3) idx = lane;
  else idx = 0;
  unsigned base = (unsigned)__cvta_generic_to_shared(smem) + idx * VEC;
  unsigned acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    unsigned v[16][4];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      if (VEC == 16)
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[r][0]), "=r"(v[r][1]), "=r"(v[r][2]), "=r"(v[r][3]) : "r"(base + r * 512));
      else
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v[r][0]), "=r"(v[r][1]) : "r"(base + r * 512));
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) acc ^= (VEC == 16) ? (v[r][0] ^ v[r][1] ^ v[r][2] ^ v[r][3]) : (v[r][0] ^ v[r][1]);
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// mixed: per LDS.128 (pattern groups-of-8) do NF DFMAs that consume the loaded values
template <int NF>
__global__ void k_mixed(double* out, int iters, long long* cyc) {
  extern __shared__ __align__(16) unsigned char smem[];
  for (int i = threadIdx.x; i < 8192 / 8; i += blockDim.x) ((double*)smem)[i] = 1.0 + i * 1e-6;
  __syncthreads();
  int lane = threadIdx.x & 31;
  unsigned base = (unsigned)__cvta_generic_to_shared(smem) + (lane >> 3) * 16;
  double a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = lane * 1e-9 + j;
  double m0 = 1.0000001, m1 = 0.9999999;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      double x, y;
      asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(base + r * 512));
#pragma unroll
      for (int f = 0; f < NF; ++f) {
        a[(f * 2) & 7] = fma(m0, x, a[(f * 2) & 7]);
        a[(f * 2 + 1) & 7] = fma(m1, y, a[(f * 2 + 1) & 7]);
        m0 += 0.0; // keep distinct regs
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// ---------------------------------------------------------------- host driver
struct Result { double ms; double cycles; };

template <typename F>
static Result time_kernel(F launch, long long* d_cyc, int grid) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  launch(); launch();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  std::vector<long long> h(grid);
  CK(cudaMemcpy(h.data(), d_cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost));
  double mean = 0; for (auto v : h) mean += (double)v; mean /= grid;
  CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1));
  return {best, mean};
}

static int g_sms = 148;

static void report(const char* name, Result r, double ops_per_thread, int grid, int block, const char* note) {
  double total = ops_per_thread * (double)grid * block;
  double per_clk_sm = total / (r.cycles * (double)grid) * ((double)grid / g_sms);  // ops per clock per SM, CTA-resident cycles
  // A cleaner figure: ops per SM per clock assuming all CTAs of an SM ran concurrently.
  double ctas_per_sm = (double)grid / g_sms;
  double ops_per_sm = total / g_sms;
  double clk_ghz = r.cycles / (r.ms * 1e6);
  printf("{\"test\": \"%s\", \"grid\": %d, \"block\": %d, \"ms\": %.4f, \"cta_cycles\": %.0f, \"eff_clock_ghz\": %.3f, "
         "\"ops_per_clk_per_sm\": %.2f, \"gops_per_s\": %.1f, \"ctas_per_sm\": %.1f, \"note\": \"%s\"}\n",
         name, grid, block, r.ms, r.cycles, clk_ghz, ops_per_sm / r.cycles, total / (r.ms * 1e6), ctas_per_sm, note);
  (void)per_clk_sm;
  fflush(stdout);
}

int main(int argc, char** argv) {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  g_sms = prop.multiProcessorCount;
  printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\", \"clock_khz\": %d, \"smem_per_sm\": %zu, \"regs_per_sm\": %d}\n",
         prop.name, g_sms, prop.major, prop.minor, prop.clockRate, prop.sharedMemPerMultiprocessor, prop.regsPerMultiprocessor);
  std::vector<double> hc(64); for (int i = 0; i < 64; ++i) hc[i] = 1.0 + 1e-9 * i;
  CK(cudaMemcpyToSymbol(c_coef, hc.data(), 64 * sizeof(double)));

  const int block = 256;
  const int maxgrid = g_sms * 8;
  double* d_out; long long* d_cyc;
  CK(cudaMalloc(&d_out, sizeof(double) * maxgrid * 1024));
  CK(cudaMalloc(&d_cyc, sizeof(long long) * maxgrid));
  const int iters = 2000;

  // ---- arithmetic throughput: 4 CTAs x 256 threads per SM = 32 warps/SM
  {
    int grid = g_sms * 4;
    auto r = time_kernel([&] { k_dfma_reg<8><<<grid, block>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_reg_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 FMA, register operands, 32 warps/SM");
    r = time_kernel([&] { k_dfma_const<8><<<grid, block>>>(d_out, iters, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_const_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 FMA, constant-bank multiplicand");
    r = time_kernel([&] { k_ffma_reg<8><<<grid, block>>>((float*)d_out, iters, 1.0000001f, 1e-6f, d_cyc); }, d_cyc, grid);
    report("ffma_reg_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP32 FMA, register operands");
    r = time_kernel([&] { k_dop<8, 0><<<grid, block>>>(d_out, iters, 1e-9, d_cyc); }, d_cyc, grid);
    report("dadd_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 add");
    r = time_kernel([&] { k_dop<8, 1><<<grid, block>>>(d_out, iters, 1.0000001, d_cyc); }, d_cyc, grid);
    report("dmul_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 mul");
    r = time_kernel([&] { k_dop<8, 2><<<grid, block>>>(d_out, iters, 0.5, d_cyc); }, d_cyc, grid);
    report("dmax_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 max");
    r = time_kernel([&] { k_dop<8, 3><<<grid, block>>>(d_out, iters, 0.5, d_cyc); }, d_cyc, grid);
    report("dabs_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 abs");
    r = time_kernel([&] { k_dop<8, 4><<<grid, block>>>(d_out, iters, 0.5, d_cyc); }, d_cyc, grid);
    report("dsetp_selp_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 setp.lt + selp");
    r = time_kernel([&] { k_dop<8, 5><<<grid, block>>>(d_out, iters, 0.5, d_cyc); }, d_cyc, grid);
    report("max_s64_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "integer max.s64 on the bit pattern");
    r = time_kernel([&] { k_dop<8, 6><<<grid, block>>>(d_out, iters, 0.5, d_cyc); }, d_cyc, grid);
    report("dclamp_ilp8_32w", r, 8.0 * 8 * iters, grid, block, "FP64 max then min (one clamp = 1 op)");
    r = time_kernel([&] { k_dfma_reg<8><<<grid, block>>>(d_out, iters * 40, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_reg_ilp8_32w_long", r, 8.0 * 8 * iters * 40, grid, block, "FP64 FMA sustained (~45 ms kernel)");
    r = time_kernel([&] { k_ffma_reg<8><<<grid, block>>>((float*)d_out, iters * 40, 1.0000001f, 1e-6f, d_cyc); }, d_cyc, grid);
    report("ffma_reg_ilp8_32w_long", r, 8.0 * 8 * iters * 40, grid, block, "FP32 FMA sustained");
  }
  // ---- DFMA with few warps: 1 warp per SMSP (128 threads/SM), ILP 1..8 -> latency and the ILP needed
  {
    int grid = g_sms;
    auto r = time_kernel([&] { k_dfma_reg<1><<<grid, 32>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_latency_1warp_ilp1", r, 1.0 * 8 * iters, grid, 32, "cta_cycles/(8*iters) = DFMA dependent-issue latency");
    printf("{\"derived\": \"dfma_latency_cycles\", \"value\": %.2f}\n", r.cycles / (8.0 * iters));
    r = time_kernel([&] { k_ffma_reg<1><<<grid, 32>>>((float*)d_out, iters, 1.0000001f, 1e-6f, d_cyc); }, d_cyc, grid);
    printf("{\"derived\": \"ffma_latency_cycles\", \"value\": %.2f}\n", r.cycles / (8.0 * iters));
    r = time_kernel([&] { k_dfma_reg<2><<<grid, 128>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_4w_ilp2", r, 2.0 * 8 * iters, grid, 128, "1 warp/SMSP ILP2");
    r = time_kernel([&] { k_dfma_reg<4><<<grid, 128>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_4w_ilp4", r, 4.0 * 8 * iters, grid, 128, "1 warp/SMSP ILP4");
    r = time_kernel([&] { k_dfma_reg<8><<<grid, 128>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_4w_ilp8", r, 8.0 * 8 * iters, grid, 128, "1 warp/SMSP ILP8");
    r = time_kernel([&] { k_dfma_reg<4><<<grid, 256>>>(d_out, iters, 1.0000001, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_8w_ilp4", r, 4.0 * 8 * iters, grid, 256, "2 warps/SMSP ILP4");
    r = time_kernel([&] { k_dfma_const<4><<<grid, 256>>>(d_out, iters, 1e-9, d_cyc); }, d_cyc, grid);
    report("dfma_const_8w_ilp4", r, 4.0 * 8 * iters, grid, 256, "2 warps/SMSP ILP4 constant operand");
  }
  // ---- shuffle
  {
    int grid = g_sms * 4;
    auto r = time_kernel([&] { k_shfl<8><<<grid, block>>>((int*)d_out, iters, d_cyc); }, d_cyc, grid);
    report("shfl_idx_b32_32w", r, 8.0 * 8 * iters, grid, block, "SHFL.IDX width 8; thread-results per clk per SM");
  }
  // ---- shared memory loads, 8 warps per SM (one CTA of 256 threads per SM) and 32 warps per SM
  {
    int grid = g_sms;
    const int smem = 8192 + 512;
    const int it2 = 1000;
#define RUN_LDS(VEC, PAT, NAME, NOTE) { \
      auto r = time_kernel([&] { k_lds_pure<VEC, PAT><<<grid, 256, smem>>>((unsigned*)d_out, it2, d_cyc); }, d_cyc, grid); \
      report(NAME "_8w", r, 16.0 * it2, grid, 256, NOTE " (ops = lane-loads; warp-instr/clk/SM = value/32)"); \
      auto r4 = time_kernel([&] { k_lds_pure<VEC, PAT><<<grid * 4, 256, smem>>>((unsigned*)d_out, it2, d_cyc); }, d_cyc, grid * 4); \
      report(NAME "_32w", r4, 16.0 * it2, grid * 4, 256, NOTE); }
    RUN_LDS(8, 0, "lds64_bcast_all", "LDS.64 all lanes same address")
    RUN_LDS(8, 1, "lds64_grp8", "LDS.64 4 groups of 8 lanes")
    RUN_LDS(8, 2, "lds64_grp4", "LDS.64 8 groups of 4 lanes")
    RUN_LDS(8, 3, "lds64_private", "LDS.64 one address per lane, contiguous")
    RUN_LDS(16, 0, "lds128_bcast_all", "LDS.128 all lanes same address")
    RUN_LDS(16, 1, "lds128_grp8", "LDS.128 4 groups of 8 lanes")
    RUN_LDS(16, 2, "lds128_grp4", "LDS.128 8 groups of 4 lanes")
    RUN_LDS(16, 3, "lds128_private", "LDS.128 one address per lane, contiguous 512 B")
    {
      auto r = time_kernel([&] { k_lds<16, 1><<<grid, 256, smem>>>(d_out, it2, d_cyc); }, d_cyc, grid);
      report("lds128_grp8_dadd_8w", r, 16.0 * it2, grid, 256, "LDS.128 grp8 feeding 2 DADD each");
    }
    {
      auto r = time_kernel([&] { k_mixed<2><<<grid, 256, smem>>>(d_out, it2, d_cyc); }, d_cyc, grid);
      report("mixed_lds128_4dfma_8w", r, 16.0 * it2 * 4, grid, 256, "per LDS.128: 4 DFMA (ops = DFMA)");
      r = time_kernel([&] { k_mixed<4><<<grid, 256, smem>>>(d_out, it2, d_cyc); }, d_cyc, grid);
      report("mixed_lds128_8dfma_8w", r, 16.0 * it2 * 8, grid, 256, "per LDS.128: 8 DFMA (ops = DFMA)");
      r = time_kernel([&] { k_mixed<1><<<grid, 256, smem>>>(d_out, it2, d_cyc); }, d_cyc, grid);
      report("mixed_lds128_2dfma_8w", r, 16.0 * it2 * 2, grid, 256, "per LDS.128: 2 DFMA (ops = DFMA)");
    }
  }
  CK(cudaFree(d_out)); CK(cudaFree(d_cyc));
  return 0;
}
