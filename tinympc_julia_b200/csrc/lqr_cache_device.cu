// Batched setup for per-instance dynamics: the infinite-horizon LQR cache of B DIFFERENT
// systems (A, B, f, Q, R per instance; rho shared), one thread per instance.
//
// It is the device twin of lqr_cache.cpp, i.e. of the reference's tiny_setup / tiny_precompute_and_set_cache
// (tinympc/TinyMPC/src/tinympc/tiny_api.cpp:107-108, 244-318): same loop order, every product and sum rounded
// separately (__dmul_rn / __dadd_rn: nvcc must not contract them, the host twin is built with -ffp-contract=off), same
// LU inverse, same stop rule (max |Kinf - Kinf_prev| < 1e-5 after <= 1000 sweeps from P = rho I).  The Riccati stop
// rule makes the cache only ~1e-5-converged, so the solve loop reproduces the reference's iteration
// counts only if the cache is reproduced operation by operation; tests compare this kernel with the host twin bit for bit.
//
// Inputs in the Julia layout (instance = slowest dimension), outputs element-major [e][ld] so that the solve kernel's
// one-thread-per-instance reads are coalesced.
#include <cuda_runtime.h>

#include <cmath>

#include "admm_kernels.hpp"

namespace tmpc {
namespace {

constexpr int PN = kPerInstanceMaxNx, PM = kPerInstanceMaxNu;

__device__ __forceinline__ double mul_(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub_(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double div_(double a, double b) { return __ddiv_rn(a, b); }

// Z (r x c) = X (r x k) * Y (k x c), column-major, column AXPYs with ascending inner index per entry (lqr_cache.cpp matmul)
__device__ void mm(double* Z, const double* X, const double* Y, int r, int k, int c) {
    for (int j = 0; j < c; ++j) {
        for (int i = 0; i < r; ++i) Z[j * r + i] = 0.0;
        for (int l = 0; l < k; ++l) {
            const double y = Y[j * k + l];
            for (int i = 0; i < r; ++i) Z[j * r + i] = add_(Z[j * r + i], mul_(X[l * r + i], y));
        }
    }
}
__device__ void tr(double* T, const double* X, int r, int c) {      // T (c x r) = X'
    for (int j = 0; j < c; ++j)
        for (int i = 0; i < r; ++i) T[i * c + j] = X[j * r + i];
}
// LU with partial pivoting + substitution on unit vectors (lqr_cache.cpp invert); n <= PM
__device__ bool inv(const double* X, double* Xinv, int n) {
    double lu[PM * PM];
    int perm[PM];
    for (int k = 0; k < n * n; ++k) lu[k] = X[k];
    for (int i = 0; i < n; ++i) perm[i] = i;
    for (int k = 0; k < n; ++k) {
        int piv = k;
        double best = fabs(lu[k * n + k]);
        for (int i = k + 1; i < n; ++i)
            if (fabs(lu[k * n + i]) > best) { best = fabs(lu[k * n + i]); piv = i; }
        if (!(best > 0.0) || !isfinite(best)) return false;
        if (piv != k) {
            for (int j = 0; j < n; ++j) { const double t = lu[j * n + k]; lu[j * n + k] = lu[j * n + piv]; lu[j * n + piv] = t; }
            const int t = perm[k]; perm[k] = perm[piv]; perm[piv] = t;
        }
        for (int i = k + 1; i < n; ++i) {
            lu[k * n + i] = div_(lu[k * n + i], lu[k * n + k]);
            const double l = lu[k * n + i];
            for (int j = k + 1; j < n; ++j) lu[j * n + i] = sub_(lu[j * n + i], mul_(l, lu[j * n + k]));
        }
    }
    for (int c = 0; c < n; ++c) {
        for (int i = 0; i < n; ++i) {
            double s = (perm[i] == c) ? 1.0 : 0.0;
            for (int j = 0; j < i; ++j) s = sub_(s, mul_(lu[j * n + i], Xinv[c * n + j]));
            Xinv[c * n + i] = s;
        }
        for (int i = n - 1; i >= 0; --i) {
            double s = Xinv[c * n + i];
            for (int j = i + 1; j < n; ++j) s = sub_(s, mul_(lu[j * n + i], Xinv[c * n + j]));
            Xinv[c * n + i] = div_(s, lu[i * n + i]);
        }
    }
    return true;
}

__global__ void __launch_bounds__(64)
riccati_cache_kernel(int n, int m, int B, long long ld, double rho, const double* __restrict__ Ain,
                     const double* __restrict__ Bin, const double* __restrict__ fin, const double* __restrict__ Qin,
                     long long q_stride, const double* __restrict__ Rin, long long r_stride, PerInstanceMats out,
                     int* __restrict__ fail) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double A[PN * PN], Bm[PN * PM], At[PN * PN], Bt[PM * PN], P[PN * PN], Pn[PN * PN], K[PM * PN], Kp[PM * PN];
    double T1[PN * PN], T2[PN * PN], G[PM * PM], Gi[PM * PM], Qd[PN], Rd[PM], fv[PN];
    for (int k = 0; k < n * n; ++k) A[k] = Ain[static_cast<size_t>(b) * n * n + k];
    for (int k = 0; k < n * m; ++k) Bm[k] = Bin[static_cast<size_t>(b) * n * m + k];
    for (int k = 0; k < n; ++k) fv[k] = fin ? fin[static_cast<size_t>(b) * n + k] : 0.0;
    for (int i = 0; i < n; ++i) Qd[i] = add_(Qin[static_cast<size_t>(b) * q_stride + i * n + i], rho);
    for (int i = 0; i < m; ++i) Rd[i] = add_(Rin[static_cast<size_t>(b) * r_stride + i * m + i], rho);
    tr(At, A, n, n);
    tr(Bt, Bm, n, m);
    for (int k = 0; k < n * n; ++k) P[k] = 0.0;
    for (int i = 0; i < n; ++i) P[i * n + i] = rho;
    for (int k = 0; k < m * n; ++k) { K[k] = 0.0; Kp[k] = 0.0; }
    for (int k = 0; k < n * n; ++k) Pn[k] = 0.0;
    bool ok = true;

    // G = R1 + (B' P) B with R1 = diag(Rd + rho); `add(X, Y)` of the host twin is X + 1.0 * Y
    auto form_G = [&](const double* Pm) {
        mm(T1, Bt, Pm, m, n, n);                       // B' P        (m x n)
        mm(G, T1, Bm, m, n, m);                        // (B' P) B    (m x m)
        for (int j = 0; j < m; ++j)
            for (int i = 0; i < m; ++i) G[j * m + i] = add_((i == j) ? add_(Rd[i], rho) : 0.0, mul_(1.0, G[j * m + i]));
    };
    for (int it = 0; it < 1000 && ok; ++it) {
        form_G(P);
        if (!inv(G, Gi, m)) { ok = false; break; }
        mm(T1, Gi, Bt, m, m, n);                       // G^-1 B'
        mm(T2, T1, P, m, n, n);                        // (G^-1 B') P
        mm(K, T2, A, m, n, n);                         // ((G^-1 B') P) A
        mm(T1, Bm, K, n, m, n);                        // B K
        for (int k = 0; k < n * n; ++k) T1[k] = add_(A[k], mul_(-1.0, T1[k]));      // A - B K
        mm(T2, At, P, n, n, n);                        // A' P
        mm(Pn, T2, T1, n, n, n);                       // (A' P)(A - B K)
        for (int j = 0; j < n; ++j)
            for (int i = 0; i < n; ++i) Pn[j * n + i] = add_((i == j) ? add_(Qd[i], rho) : 0.0, mul_(1.0, Pn[j * n + i]));
        double delta = 0.0;
        for (int k = 0; k < m * n; ++k) delta = fmax(delta, fabs(sub_(K[k], Kp[k])));
        if (delta < 1e-5) break;                       // Pn of this sweep is the one kept
        for (int k = 0; k < m * n; ++k) Kp[k] = K[k];
        for (int k = 0; k < n * n; ++k) P[k] = Pn[k];
    }
    if (ok) {
        form_G(Pn);
        ok = inv(G, Gi, m);                            // Quu_inv
    }
    if (!ok) { atomicExch(fail, 1); return; }
    // AmBKt = (A - B K)' ; APf = (AmBKt Pinf) f ; BPf = (B' Pinf) f
    mm(T1, Bm, K, n, m, n);
    for (int k = 0; k < n * n; ++k) T1[k] = add_(A[k], mul_(-1.0, T1[k]));
    tr(T2, T1, n, n);                                  // AmBKt
#define PUT(arr, e, val) (arr)[static_cast<size_t>(e) * ld + b] = (val)
    for (int k = 0; k < n * n; ++k) { PUT(out.A, k, A[k]); PUT(out.Pinf, k, Pn[k]); PUT(out.AmBKt, k, T2[k]); }
    for (int k = 0; k < n * m; ++k) { PUT(out.B, k, Bm[k]); PUT(out.Kinf, k, K[k]); }
    for (int k = 0; k < m * m; ++k) PUT(out.Quu_inv, k, Gi[k]);
    for (int i = 0; i < n; ++i) { PUT(out.f, i, fv[i]); PUT(out.Qd, i, Qd[i]); }
    for (int i = 0; i < m; ++i) PUT(out.Rd, i, Rd[i]);
    mm(T1, T2, Pn, n, n, n);                           // AmBKt Pinf
    mm(At, T1, fv, n, n, 1);                           // APf (At reused as scratch)
    for (int i = 0; i < n; ++i) PUT(out.APf, i, At[i]);
    mm(T1, Bt, Pn, m, n, n);                           // B' Pinf
    mm(G, T1, fv, m, n, 1);                            // BPf
    for (int i = 0; i < m; ++i) PUT(out.BPf, i, G[i]);
#undef PUT
}

}  // namespace

cudaError_t launch_riccati_cache(int nx, int nu, int B, long long ld, double rho, const double* A, const double* Bdyn,
                                 const double* f, const double* Q, long long q_stride, const double* R,
                                 long long r_stride, const PerInstanceMats& out, int* fail, cudaStream_t stream) {
    if (nx > PN || nu > PM || nx < 1 || nu < 1) return cudaErrorInvalidValue;
    const int block = 64, grid = (B + block - 1) / block;
    riccati_cache_kernel<<<grid, block, 0, stream>>>(nx, nu, B, ld, rho, A, Bdyn, f, Q, q_stride, R, r_stride, out, fail);
    return cudaGetLastError();
}

}  // namespace tmpc
