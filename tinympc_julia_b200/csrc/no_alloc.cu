// tmpc_solve_noalloc: one batched cold solve that allocates nothing on the device.
//
// Added; it has no counterpart in the reference.  For callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The solve is the thread-per-instance
// kernel (admm_thread.cu) — the kernel of this library whose working set is scratch in global memory: the
// element-major carried state g, v, q, y, z, r, p_N and the feed-forward scratch d of every instance —
// with that scratch in the file-scope array below instead of buffers owned by a tmpc_batch handle.
//
// * The array is a fixed 32 MiB per GPU that loads this module, held until the process exits.
// * Its address is taken once per device with cudaGetSymbolAddress while that device is current (in host code the
//   array's name is only a stand-in for the symbol, not a device address) and handed to the kernel as pointers.
// * Every call shares the array and finds the previous call's contents in it: the part a solve uses is cleared on the
//   caller's stream in front of the kernel.  Two calls that overlap on the GPU (different streams, no ordering) race.
// * No cudaMalloc / cudaFree / stream or event creation here; cudaMemsetAsync, (for shapes that are not compiled into
//   the thread kernel, and only when the problem constants changed) one cudaMemcpyAsync, one kernel launch.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/tinympc_b200.h"
#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace {

constexpr size_t kScratchDoubles = static_cast<size_t>(TMPC_NOALLOC_SCRATCH_BYTES) / sizeof(double);
static __device__ __align__(256) double g_noalloc_scratch[kScratchDoubles];

// head of the array: problem constants for the run-time-dimension kernel (nx <= 32, nu <= 16), 256-byte aligned pieces
constexpr int kMaxNx = tmpc::kThreadKernelMaxNx, kMaxNu = tmpc::kThreadKernelMaxNu;
constexpr size_t kConstDoubles = 3 * kMaxNx * kMaxNx + 2 * kMaxNx * kMaxNu + kMaxNu * kMaxNu + 3 * kMaxNx + 2 * kMaxNu;   // 4480
constexpr size_t kConstRegion = (kConstDoubles + 31) / 32 * 32;
static_assert(kConstRegion < kScratchDoubles, "scratch array too small");

constexpr int kMaxDevices = 16;
std::mutex g_mu;
double* g_dev_ptr[kMaxDevices] = {};            // device address of the array, per device
double g_staged[kMaxDevices][kConstDoubles];    // constants last uploaded to that device
size_t g_staged_n[kMaxDevices] = {};
tmpc::HostMats g_hm;                            // reused between calls (vector capacity stays)

int fail(char* errbuf, size_t errlen, int code, const char* msg) {
    std::fprintf(stderr, "tinympc_b200: tmpc_solve_noalloc: %s\n", msg);
    if (errbuf && errlen) { std::snprintf(errbuf, errlen, "%s", msg); }
    return code;
}

bool finite_all(const double* p, size_t n) {
    for (size_t i = 0; i < n; ++i) if (!std::isfinite(p[i])) return false;
    return true;
}

}  // namespace

extern "C" int tmpc_solve_noalloc(int nx, int nu, int N, int B, double rho, const double* A, const double* Bdyn,
                                  const double* f, const double* Kinf, const double* Pinf, const double* Quu_inv,
                                  const double* AmBKt, const double* APf, const double* BPf, const double* Qd,
                                  const double* Rd, const double* d_x0, const double* d_x_ref, const double* d_u_ref,
                                  const double* d_x_min, const double* d_x_max, const double* d_u_min,
                                  const double* d_u_max, double abs_pri_tol, double abs_dua_tol, int max_iter,
                                  int check_termination, double* d_states, double* d_controls, int* d_status,
                                  int* d_iters, void* cuda_stream, char* errbuf, size_t errbuf_len) {
    if (errbuf && errbuf_len) errbuf[0] = '\0';
    if (!A || !Bdyn || !f || !Kinf || !Pinf || !Quu_inv || !AmBKt || !APf || !BPf || !Qd || !Rd)
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, "null problem matrix (host pointers: A, B, f and the outputs of tmpc_compute_cache)");
    if (!d_x0 || !d_states || !d_controls || !d_status || !d_iters)
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, "null device pointer (x0, states, controls, status, iters are required)");
    if (nx < 1 || nu < 1 || N < 2 || B < 1 || !tmpc::thread_kernel_supports(nx, nu))
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, "need 1 <= nx <= 32, 1 <= nu <= 16, N >= 2, B >= 1");
    if ((d_x_min == nullptr) != (d_x_max == nullptr) || (d_u_min == nullptr) != (d_u_max == nullptr))
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, "bounds come in (min, max) pairs");
    if (max_iter < 1 || check_termination < 1 || !(rho > 0.0) || !std::isfinite(rho))
        return fail(errbuf, errbuf_len, TMPC_ERR_STATE, "need max_iter >= 1, check_termination >= 1, rho > 0");
    const size_t n = nx, m = nu;
    if (!finite_all(A, n * n) || !finite_all(Bdyn, n * m) || !finite_all(f, n) || !finite_all(Kinf, m * n) ||
        !finite_all(Pinf, n * n) || !finite_all(Quu_inv, m * m) || !finite_all(AmBKt, n * n) || !finite_all(APf, n) ||
        !finite_all(BPf, m) || !finite_all(Qd, n) || !finite_all(Rd, m))
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, "non-finite problem data");

    // ---- the one limit of this entry point: the working set must fit the static array
    const size_t ld = (static_cast<size_t>(B) + 31) / 32 * 32;
    const size_t sx = n * N * ld, su = m * (N - 1) * ld;
    const size_t state = 4 * sx + 5 * su + n * ld;          // g v0 v1 q | y z0 z1 r d | pN
    if (kConstRegion + state > kScratchDoubles) {
        char msg[256];
        std::snprintf(msg, sizeof msg,
                      "batch too large for the static scratch array: (4 nx N + 5 nu (N-1) + nx) x %zu instances = %zu bytes, "
                      "limit %zu bytes (TMPC_NOALLOC_SCRATCH_BYTES minus %zu for constants); use tmpc_batch_create / tmpc_batch_solve",
                      ld, state * sizeof(double), (kScratchDoubles - kConstRegion) * sizeof(double), kConstRegion * sizeof(double));
        return fail(errbuf, errbuf_len, TMPC_ERR_DIM, msg);
    }

    std::lock_guard<std::mutex> lock(g_mu);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, "device ordinal beyond 15");
    if (!g_dev_ptr[dev]) {          // once per device, with that device current
        double* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "cudaGetSymbolAddress returned null");
        g_dev_ptr[dev] = p;
    }
    double* const base = g_dev_ptr[dev];
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);

    // ---- constants: by value in the kernel parameters for the compiled shapes; the run-time-dimension kernel reads them
    // from the head of the array, uploaded only when they differ from what this device already holds
    g_hm.nx = nx; g_hm.nu = nu;
    g_hm.A.assign(A, A + n * n); g_hm.B.assign(Bdyn, Bdyn + n * m); g_hm.f.assign(f, f + n);
    g_hm.Kinf.assign(Kinf, Kinf + m * n); g_hm.Pinf.assign(Pinf, Pinf + n * n); g_hm.Quu_inv.assign(Quu_inv, Quu_inv + m * m);
    g_hm.AmBKt.assign(AmBKt, AmBKt + n * n); g_hm.APf.assign(APf, APf + n); g_hm.BPf.assign(BPf, BPf + m);
    g_hm.Qd.assign(Qd, Qd + n); g_hm.Rd.assign(Rd, Rd + m);
    const size_t oA = 0, oB = oA + n * n, oK = oB + n * m, oP = oK + m * n, oQu = oP + n * n, oAm = oQu + m * m,
                 of = oAm + n * n, oAPf = of + n, oBPf = oAPf + n, oQd = oBPf + m, oRd = oQd + n, ncst = oRd + m;
    const bool compiled = (nx == 12 && nu == 4) || (nx == 4 && nu == 1) || (nx == 6 && nu == 3);
    if (!compiled) {
        static double tmp[kConstDoubles];
        std::memcpy(tmp + oA, A, sizeof(double) * n * n); std::memcpy(tmp + oB, Bdyn, sizeof(double) * n * m);
        std::memcpy(tmp + oK, Kinf, sizeof(double) * m * n); std::memcpy(tmp + oP, Pinf, sizeof(double) * n * n);
        std::memcpy(tmp + oQu, Quu_inv, sizeof(double) * m * m); std::memcpy(tmp + oAm, AmBKt, sizeof(double) * n * n);
        std::memcpy(tmp + of, f, sizeof(double) * n); std::memcpy(tmp + oAPf, APf, sizeof(double) * n);
        std::memcpy(tmp + oBPf, BPf, sizeof(double) * m); std::memcpy(tmp + oQd, Qd, sizeof(double) * n);
        std::memcpy(tmp + oRd, Rd, sizeof(double) * m);
        if (g_staged_n[dev] != ncst || std::memcmp(g_staged[dev], tmp, sizeof(double) * ncst) != 0) {
            e = cudaMemcpyAsync(base, tmp, sizeof(double) * ncst, cudaMemcpyHostToDevice, s);   // pageable source: staged before the call returns
            if (e != cudaSuccess) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, cudaGetErrorString(e));
            std::memcpy(g_staged[dev], tmp, sizeof(double) * ncst);
            g_staged_n[dev] = ncst;
        }
    }

    // ---- clear what must start at zero (the whole working set of this solve), then the kernel, all on the caller's stream
    double* st0 = base + kConstRegion;
    e = cudaMemsetAsync(st0, 0, sizeof(double) * state, s);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, cudaGetErrorString(e));

    tmpc::BatchView v{};
    v.B = B; v.ld = static_cast<int>(ld); v.nx = nx; v.nu = nu; v.N = N; v.rho = rho;
    v.x0 = d_x0;
    v.Xref = d_x_ref; v.xref_stride = 0; v.Uref = d_u_ref; v.uref_stride = 0;
    v.x_min = d_x_min; v.x_max = d_x_max; v.xb_stride = 0; v.u_min = d_u_min; v.u_max = d_u_max; v.ub_stride = 0;
    v.A = base + oA; v.Bm = base + oB; v.Kinf = base + oK; v.Pinf = base + oP; v.Quu_inv = base + oQu; v.AmBKt = base + oAm;
    v.f = base + of; v.APf = base + oAPf; v.BPf = base + oBPf; v.Qd = base + oQd; v.Rd = base + oRd;
    double* p = st0;
    v.g = p; p += sx; v.v[0] = p; p += sx; v.v[1] = p; p += sx; v.q = p; p += sx;
    v.y = p; p += su; v.z[0] = p; p += su; v.z[1] = p; p += su; v.r = p; p += su; v.d = p; p += su;
    v.pN = p;
    v.Xsol = d_states; v.Usol = d_controls; v.status = d_status; v.iters = d_iters;

    tmpc::AdmmSettings st{};
    st.abs_pri_tol = abs_pri_tol; st.abs_dua_tol = abs_dua_tol; st.max_iter = max_iter; st.check_termination = check_termination;
    st.en_state_bound = d_x_min != nullptr; st.en_input_bound = d_u_min != nullptr;

    e = tmpc::launch_admm_thread(v, st, g_hm, s, 64);
    if (e != cudaSuccess) return fail(errbuf, errbuf_len, TMPC_ERR_CUDA, cudaGetErrorString(e));
    return TMPC_OK;
}
