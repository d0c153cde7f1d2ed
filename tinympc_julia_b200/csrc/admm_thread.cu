// ADMM solve kernel, mapping "one thread per problem instance".
//
// This is the general-purpose kernel of the library: any nx, nu, N, every constraint family
// (box, second-order cone, half-space), warm start.  An instance's carried state streams through
// L2/HBM once per ADMM iteration in the element-major layout (fully coalesced across the warp),
// the per-column vectors live in registers, and for the compiled-in shapes the problem matrices
// are kernel parameters, i.e. constant-bank operands folded straight into the DFMAs (no operand
// delivery instructions at all).  Its ceiling is HBM bandwidth (~11-19 KB per instance-iteration);
// the shared-memory-resident sub-warp kernel (admm_subwarp.cu) removes that traffic for the
// box-constrained shapes.
//
// Algorithm: the ADMM iteration of the reference, restated in SURVEY.md §8a —
//   backward_pass_grad   reference admm.cpp:13-20
//   forward_pass         reference admm.cpp:25-32
//   update_slack         reference admm.cpp:81-175   (project_soc :39-60, project_hyperplane :70-73)
//   update_dual          reference admm.cpp:181-208
//   update_linear_cost   reference admm.cpp:214-247
//   termination_condition reference admm.cpp:253-271
//   solve                reference admm.cpp:274-389
// The slack/dual/cost/residual updates are elementwise per knot point, so they are fused into the
// forward sweep column by column; x, u and p never leave registers.
//
// Adaptive rho (reference admm.cpp:331-357 + rho_benchmark.cpp) runs here as well: the
// MatsAdaptive operand provider gives every instance its own rho, Kinf and Pinf (the reference updates them in place
// in the solver's cache, so in a batch they are per-instance carried state), and after every fifth iteration the
// thread re-rolls its trajectory, forms the reference's residual ratio and applies the first-order update.
#include <cuda_runtime.h>

#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace tmpc {

namespace {

// ---------------------------------------------------------------- arithmetic at the reference's `tinytype`
// The reference's precision is one typedef (types.hpp:15, `typedef double tinytype`; float is the other supported
// value).  The kernel is instantiated for both: T = double is the default build of the reference, T = float its
// float build.  In the float instantiation every input is rounded to float where
// src/bindings.cpp casts (:50-55, :82, :105, :128) and results widen back to double (:172, :194).  Inputs, solutions
// and the C ABI stay double; the carried state, which only this kernel reads in float mode, is stored as T.
// (the traits class Ar<T> lives in admm_device.cuh)

// Blocks per SM the register allocation aims at (M::kMinBlocks).  The kernel streams its state through L2 / HBM and hides
// that latency with resident warps only: with the compiled-in shapes at one block per 242 registers (8 warps per SM) the
// quadrotor batch runs 80 ms on the hard workload, at four blocks (128 registers, ~0.7 KB of L1-resident spills per
// thread) 49 ms; five and more lose again.  The adaptive-rho provider keeps one: its run time is the tail of the few
// instances that never converge, i.e. the latency of single warps, which the spills lengthen.
#ifndef TMPC_THREAD_MINBLOCKS
#define TMPC_THREAD_MINBLOCKS 4
#endif
#ifndef TMPC_THREAD_MINBLOCKS_PI
#define TMPC_THREAD_MINBLOCKS_PI 0
#endif
// ---------------------------------------------------------------- matrix operand providers
template <class T, int NX, int NU>
struct MatsConst {           // by-value kernel parameter -> constant bank
    static constexpr int kMaxNx = NX, kMaxNu = NU;
    static constexpr bool kStatic = true;
    static constexpr bool kAdaptive = false;
    static constexpr int kMinBlocks = TMPC_THREAD_MINBLOCKS;
    __device__ __forceinline__ const MatsConst& bound(int) const { return *this; }
    T A_[NX * NX], B_[NX * NU], Kinf_[NU * NX], AmBKt_[NX * NX], Quu_inv_[NU * NU], Pinf_[NX * NX];
    T f_[NX], APf_[NX], BPf_[NU], Qd_[NX], Rd_[NU];
    __device__ __forceinline__ int nx() const { return NX; }
    __device__ __forceinline__ int nu() const { return NU; }
    __device__ __forceinline__ T A(int i, int k) const { return A_[k * NX + i]; }
    __device__ __forceinline__ T B(int i, int k) const { return B_[k * NX + i]; }
    __device__ __forceinline__ T Kinf(int i, int k) const { return Kinf_[k * NU + i]; }
    __device__ __forceinline__ T AmBKt(int i, int k) const { return AmBKt_[k * NX + i]; }
    __device__ __forceinline__ T Quu_inv(int i, int k) const { return Quu_inv_[k * NU + i]; }
    __device__ __forceinline__ T Pinf(int i, int k) const { return Pinf_[k * NX + i]; }
    __device__ __forceinline__ T f(int i) const { return f_[i]; }
    __device__ __forceinline__ T APf(int i) const { return APf_[i]; }
    __device__ __forceinline__ T BPf(int i) const { return BPf_[i]; }
    __device__ __forceinline__ T Qd(int i) const { return Qd_[i]; }
    __device__ __forceinline__ T Rd(int i) const { return Rd_[i]; }
};

struct MatsGlobal {          // runtime dimensions: uniform loads through the read-only path
    static constexpr int kMaxNx = kThreadKernelMaxNx, kMaxNu = kThreadKernelMaxNu;
    static constexpr bool kStatic = false;
    static constexpr bool kAdaptive = false;
    static constexpr int kMinBlocks = 0;        // unspecified (rolled loops: 56-64 registers)
    __device__ __forceinline__ const MatsGlobal& bound(int) const { return *this; }
    int nx_, nu_;
    const double *A_, *B_, *Kinf_, *AmBKt_, *Quu_inv_, *Pinf_, *f_, *APf_, *BPf_, *Qd_, *Rd_;
    __device__ __forceinline__ int nx() const { return nx_; }
    __device__ __forceinline__ int nu() const { return nu_; }
    __device__ __forceinline__ double A(int i, int k) const { return __ldg(A_ + k * nx_ + i); }
    __device__ __forceinline__ double B(int i, int k) const { return __ldg(B_ + k * nx_ + i); }
    __device__ __forceinline__ double Kinf(int i, int k) const { return __ldg(Kinf_ + k * nu_ + i); }
    __device__ __forceinline__ double AmBKt(int i, int k) const { return __ldg(AmBKt_ + k * nx_ + i); }
    __device__ __forceinline__ double Quu_inv(int i, int k) const { return __ldg(Quu_inv_ + k * nu_ + i); }
    __device__ __forceinline__ double Pinf(int i, int k) const { return __ldg(Pinf_ + k * nx_ + i); }
    __device__ __forceinline__ double f(int i) const { return __ldg(f_ + i); }
    __device__ __forceinline__ double APf(int i) const { return __ldg(APf_ + i); }
    __device__ __forceinline__ double BPf(int i) const { return __ldg(BPf_ + i); }
    __device__ __forceinline__ double Qd(int i) const { return __ldg(Qd_ + i); }
    __device__ __forceinline__ double Rd(int i) const { return __ldg(Rd_ + i); }
};

// per-instance dynamics: element-major arrays [e][ld], one coalesced load per operand (bound to the thread's instance)
template <int NX, int NU>            // NX = 0: run-time dimensions (rolled loops); else compiled-in shape, fully unrolled
struct MatsPerInstance {
    static constexpr int kMaxNx = NX ? NX : kPerInstanceMaxNx, kMaxNu = NX ? NU : kPerInstanceMaxNu;
    static constexpr bool kStatic = NX != 0;
    static constexpr bool kAdaptive = false;
    static constexpr int kMinBlocks = NX ? TMPC_THREAD_MINBLOCKS_PI : 0;
    int nx_, nu_;
    long long ld_;
    PerInstanceMats p_;
    long long b_;
    __device__ __forceinline__ MatsPerInstance bound(int b) const { MatsPerInstance m = *this; m.b_ = b; return m; }
    __device__ __forceinline__ int nx() const { return NX ? NX : nx_; }
    __device__ __forceinline__ int nu() const { return NX ? NU : nu_; }
    __device__ __forceinline__ double at(const double* a, int e) const { return __ldg(a + e * ld_ + b_); }
    __device__ __forceinline__ double A(int i, int k) const { return at(p_.A, k * nx() + i); }
    __device__ __forceinline__ double B(int i, int k) const { return at(p_.B, k * nx() + i); }
    __device__ __forceinline__ double Kinf(int i, int k) const { return at(p_.Kinf, k * nu() + i); }
    __device__ __forceinline__ double AmBKt(int i, int k) const { return at(p_.AmBKt, k * nx() + i); }
    __device__ __forceinline__ double Quu_inv(int i, int k) const { return at(p_.Quu_inv, k * nu() + i); }
    __device__ __forceinline__ double Pinf(int i, int k) const { return at(p_.Pinf, k * nx() + i); }
    __device__ __forceinline__ double f(int i) const { return at(p_.f, i); }
    __device__ __forceinline__ double APf(int i) const { return at(p_.APf, i); }
    __device__ __forceinline__ double BPf(int i) const { return at(p_.BPf, i); }
    __device__ __forceinline__ double Qd(int i) const { return at(p_.Qd, i); }
    __device__ __forceinline__ double Rd(int i) const { return at(p_.Rd, i); }
};

// adaptive rho: shared dynamics, but rho, Kinf and Pinf belong to the instance (element-major [e][ld], written by the
// owning thread during the solve: plain loads, not the read-only path)
template <int NX, int NU>            // NX = 0: run-time dimensions
struct MatsAdaptive {
    static constexpr int kMaxNx = NX ? NX : kThreadKernelMaxNx, kMaxNu = NX ? NU : kThreadKernelMaxNu;
    static constexpr bool kStatic = NX != 0;
    static constexpr bool kAdaptive = true;
    static constexpr int kMinBlocks = 0;
    int nx_, nu_;
    long long ld_, b_;
    const double *A_, *B_, *AmBKt_, *Quu_inv_, *f_, *APf_, *BPf_, *Qd_, *Rd_;
    AdaptiveRho ad_;
    __device__ __forceinline__ MatsAdaptive bound(int b) const { MatsAdaptive m = *this; m.b_ = b; return m; }
    __device__ __forceinline__ int nx() const { return NX ? NX : nx_; }
    __device__ __forceinline__ int nu() const { return NX ? NU : nu_; }
    __device__ __forceinline__ double A(int i, int k) const { return __ldg(A_ + k * nx() + i); }
    __device__ __forceinline__ double B(int i, int k) const { return __ldg(B_ + k * nx() + i); }
    __device__ __forceinline__ double Kinf(int i, int k) const { return ad_.Kinf[(k * nu() + i) * ld_ + b_]; }
    __device__ __forceinline__ double AmBKt(int i, int k) const { return __ldg(AmBKt_ + k * nx() + i); }
    __device__ __forceinline__ double Quu_inv(int i, int k) const { return __ldg(Quu_inv_ + k * nu() + i); }
    __device__ __forceinline__ double Pinf(int i, int k) const { return ad_.Pinf[(k * nx() + i) * ld_ + b_]; }
    __device__ __forceinline__ double f(int i) const { return __ldg(f_ + i); }
    __device__ __forceinline__ double APf(int i) const { return __ldg(APf_ + i); }
    __device__ __forceinline__ double BPf(int i) const { return __ldg(BPf_ + i); }
    __device__ __forceinline__ double Qd(int i) const { return __ldg(Qd_ + i); }
    __device__ __forceinline__ double Rd(int i) const { return __ldg(Rd_ + i); }
};

// EXTRA = false: box constraints only - the cone / half-space code and the registers of its slack columns are compiled out
// (the launchers pick the instantiation from the settings).
template <class T, class M, bool EXTRA>
__global__ void __launch_bounds__(kThreadKernelBlock, M::kMinBlocks)
admm_thread_kernel(const __grid_constant__ M mt_param, const __grid_constant__ BatchView bv,
                   const __grid_constant__ AdmmSettings st) {
    using R = Ar<T>;
    constexpr int MX = M::kMaxNx, MU = M::kMaxNu;
    constexpr int UNR = M::kStatic ? 64 : 1;   // full unroll for compiled-in shapes, rolled loops otherwise
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bv.B) return;
    const auto& mt = mt_param.bound(b);        // per-instance dynamics: operands of THIS instance; else the parameter itself
    const int nx = mt.nx(), nu = mt.nu(), N = bv.N;
    const size_t ld = static_cast<size_t>(bv.ld);
    T rho = static_cast<T>(bv.rho);                             // bindings.cpp:55
    if constexpr (M::kAdaptive) rho = static_cast<T>(mt.ad_.rho[b]);   // cache->rho of this instance (adapted by earlier solves)
    const T tol_pri = static_cast<T>(st.abs_pri_tol), tol_dua = static_cast<T>(st.abs_dua_tol);   // types.hpp:64-65
    // Carried state and scratch are the kernel's private format: element-major arrays of T.  The float instantiation
    // uses the first half of each (double-sized) allocation, which halves the bytes an iteration streams through L2 / HBM.
#define ST(arr, e) reinterpret_cast<T*>(arr)[static_cast<size_t>(e) * ld + b]

    if (st.max_iter <= 0) {   // the reference's loop does not run: solution keeps the previous slacks
        bv.status[b] = 1; bv.iters[b] = 0;
        return;
    }

    const bool box_x = st.en_state_bound && bv.x_min != nullptr;
    const bool box_u = st.en_input_bound && bv.u_min != nullptr;
    const bool soc_x = EXTRA && st.en_state_soc && bv.ncx > 0;
    const bool soc_u = EXTRA && st.en_input_soc && bv.ncu > 0;
    const bool lin_x = EXTRA && st.en_state_linear != 0;
    const bool lin_u = EXTRA && st.en_input_linear != 0;

    const double* xref = bv.Xref ? bv.Xref + static_cast<size_t>(b) * bv.xref_stride : nullptr;
    const double* uref = bv.Uref ? bv.Uref + static_cast<size_t>(b) * bv.uref_stride : nullptr;
    const double* xmin = box_x ? bv.x_min + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* xmax = box_x ? bv.x_max + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* umin = box_u ? bv.u_min + static_cast<size_t>(b) * bv.ub_stride : nullptr;
    const double* umax = box_u ? bv.u_max + static_cast<size_t>(b) * bv.ub_stride : nullptr;

    T p[MX], x[MX], tmp[MX], acc[MX], u[MU], t[MU], rr[MU];
    T vc[MX];   // cone / half-space slack column (state or input)

#pragma unroll UNR
    for (int j = 0; j < MX; ++j) if (j < nx) p[j] = ST(bv.pN, j);

    int cur = 0;          // bv.v[cur] / bv.z[cur] hold the previous slacks, the new ones go to [cur ^ 1]
    int iter = 0;
    bool converged = false;

    for (int it = 0; it < st.max_iter; ++it) {
        double* __restrict__ v_old = bv.v[cur];
        double* __restrict__ v_new = bv.v[cur ^ 1];
        double* __restrict__ z_old = bv.z[cur];
        double* __restrict__ z_new = bv.z[cur ^ 1];

        // ---------------- backward pass: d_i, p_i for i = N-2 .. 0
        // r_i and q_i do not depend on the recursion: they are requested one knot point ahead, BEFORE this knot's store of
        // d_i (the compiler must assume the store may alias them and would otherwise start every load after it - a lone
        // warp then pays one HBM round trip per group of loads, ~10 per knot point)
        T qn[MX], rn[MU];
#pragma unroll UNR
        for (int j = 0; j < MU; ++j) if (j < nu) rn[j] = ST(bv.r, (N - 2) * nu + j);
#pragma unroll UNR
        for (int j = 0; j < MX; ++j) if (j < nx) qn[j] = ST(bv.q, (N - 2) * nx + j);
        for (int i = N - 2; i >= 0; --i) {
            T qi[MX];
#pragma unroll UNR
            for (int j = 0; j < MU; ++j) if (j < nu) rr[j] = rn[j];
#pragma unroll UNR
            for (int j = 0; j < MX; ++j) if (j < nx) qi[j] = qn[j];
            if (i > 0) {
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) rn[j] = ST(bv.r, (i - 1) * nu + j);
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) qn[j] = ST(bv.q, (i - 1) * nx + j);
            }
#pragma unroll UNR
            for (int j = 0; j < MU; ++j) if (j < nu) {
                T s = 0;
#pragma unroll UNR
                for (int k = 0; k < MX; ++k) if (k < nx) s = R::fma(mt.B(k, j), p[k], s);
                t[j] = R::add(R::add(s, rr[j]), mt.BPf(j));
            }
#pragma unroll UNR
            for (int j = 0; j < MU; ++j) if (j < nu) {
                T s = 0;
#pragma unroll UNR
                for (int k = 0; k < MU; ++k) if (k < nu) s = R::fma(mt.Quu_inv(j, k), t[k], s);
                ST(bv.d, i * nu + j) = s;
            }
#pragma unroll UNR
            for (int j = 0; j < MX; ++j) if (j < nx) {
                T s1 = 0.0, s2 = 0.0;
#pragma unroll UNR
                for (int k = 0; k < MX; ++k) if (k < nx) s1 = R::fma(mt.AmBKt(j, k), p[k], s1);
#pragma unroll UNR
                for (int k = 0; k < MU; ++k) if (k < nu) s2 = R::fma(mt.Kinf(k, j), rr[k], s2);
                tmp[j] = R::add(R::sub(R::add(qi[j], s1), s2), mt.APf(j));
            }
#pragma unroll UNR
            for (int j = 0; j < MX; ++j) if (j < nx) p[j] = tmp[j];
        }

        // ---------------- forward pass fused with slack / dual / linear-cost / residual updates
#pragma unroll UNR
        for (int j = 0; j < MX; ++j) if (j < nx) x[j] = bv.x0[static_cast<size_t>(b) * nx + j];
        T pri_x = 0, dua_x = 0, pri_u = 0, dua_u = 0;

        for (int i = 0; i < N; ++i) {
            const bool terminal = (i == N - 1);
            // everything this knot point reads of the box-constrained state is requested here, before its first store
            // (see the backward pass): d_i, y_i, z_i (previous), g_i, v_i (previous)
            T dk[MU], yk[MU], zo[MU], gk[MX], vo[MX];
            if (!terminal) {
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) {
                    dk[j] = ST(bv.d, i * nu + j); yk[j] = ST(bv.y, i * nu + j); zo[j] = ST(z_old, i * nu + j);
                }
            }
#pragma unroll UNR
            for (int j = 0; j < MX; ++j) if (j < nx) { gk[j] = ST(bv.g, i * nx + j); vo[j] = ST(v_old, i * nx + j); }
            T gck[MX], glk[MX], yck[MU], ylk[MU];          // cone / half-space duals of the knot point, likewise
            if (soc_u && !terminal) {
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) yck[j] = ST(bv.yc, i * nu + j);
            }
            if (lin_u && !terminal) {
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) ylk[j] = ST(bv.yl, i * nu + j);
            }
            if (soc_x) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) gck[j] = ST(bv.gc, i * nx + j);
            }
            if (lin_x) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) glk[j] = ST(bv.gl, i * nx + j);
            }
            if (!terminal) {
                // u_i = -Kinf x_i - d_i
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) {
                    T s = 0;
#pragma unroll UNR
                    for (int k = 0; k < MX; ++k) if (k < nx) s = R::fma(mt.Kinf(j, k), x[k], s);
                    u[j] = R::sub(-s, dk[j]);
                }
                // ---- input column i
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) {
                    const int e = i * nu + j;
                    const T yj = yk[j];
                    T znew = R::add(u[j], yj);
                    if (box_u) znew = R::clamp(znew, __ldg(umin + e), __ldg(umax + e));
                    const T ynew = R::sub(R::add(yj, u[j]), znew);
                    pri_u = R::max(pri_u, R::abs(R::sub(u[j], znew)));
                    dua_u = R::max(dua_u, R::abs(R::sub(zo[j], znew)));
                    T racc = -R::mul(uref ? __ldg(uref + e) : 0.0, mt.Rd(j));
                    racc = R::sub(racc, R::mul(rho, R::sub(znew, ynew)));
                    ST(bv.y, e) = ynew;
                    ST(z_new, e) = znew;
                    t[j] = racc;
                }
                if (soc_u) {
#pragma unroll UNR
                    for (int j = 0; j < MU; ++j) if (j < nu) vc[j] = R::add(u[j], yck[j]);
                    if (st.en_input_soc)
                        for (int k = 0; k < bv.ncu; ++k) project_soc3(vc + bv.Acu[k], bv.cu[k]);
#pragma unroll UNR
                    for (int j = 0; j < MU; ++j) if (j < nu) {
                        const int e = i * nu + j;
                        const T ycn = R::sub(R::add(yck[j], u[j]), vc[j]);
                        t[j] = R::sub(t[j], R::mul(rho, R::sub(vc[j], ycn)));
                        ST(bv.yc, e) = ycn;
                    }
                }
                if (lin_u) {
#pragma unroll UNR
                    for (int j = 0; j < MU; ++j) if (j < nu) vc[j] = R::add(u[j], ylk[j]);
                    project_halfspaces(vc, nu, bv.Alin_u, bv.blin_u, bv.nlu);
#pragma unroll UNR
                    for (int j = 0; j < MU; ++j) if (j < nu) {
                        const int e = i * nu + j;
                        const T yln = R::sub(R::add(ylk[j], u[j]), vc[j]);
                        t[j] = R::sub(t[j], R::mul(rho, R::sub(vc[j], yln)));
                        ST(bv.yl, e) = yln;
                    }
                }
#pragma unroll UNR
                for (int j = 0; j < MU; ++j) if (j < nu) ST(bv.r, i * nu + j) = t[j];
            }

            // ---- state column i
#pragma unroll UNR
            for (int j = 0; j < MX; ++j) if (j < nx) {
                const int e = i * nx + j;
                const T gj = gk[j];
                T vnew = R::add(x[j], gj);
                if (box_x) vnew = R::clamp(vnew, __ldg(xmin + e), __ldg(xmax + e));
                const T gnew = R::sub(R::add(gj, x[j]), vnew);
                pri_x = R::max(pri_x, R::abs(R::sub(x[j], vnew)));
                dua_x = R::max(dua_x, R::abs(R::sub(vo[j], vnew)));
                ST(bv.g, e) = gnew;
                ST(v_new, e) = vnew;
                acc[j] = R::mul(rho, R::sub(vnew, gnew));       // rho * (vnew - g)
            }
            // linear cost: q_i = -(Xref_i .* Qd) - rho (vnew - g) [- cone] [- half-space];
            // terminal column: p_{N-1} = -(Xref_{N-1}' Pinf)' - ...   (reference admm.cpp:238-246)
            if (!terminal) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx)
                    acc[j] = R::sub(-R::mul(xref ? __ldg(xref + i * nx + j) : 0.0, mt.Qd(j)), acc[j]);
            } else {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) {
                    T s = 0;
                    if (xref) {
#pragma unroll UNR
                        for (int k = 0; k < MX; ++k) if (k < nx) s = R::fma(__ldg(xref + i * nx + k), mt.Pinf(k, j), s);
                    }
                    acc[j] = R::sub(-s, acc[j]);
                }
            }
            if (soc_x) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) vc[j] = R::add(x[j], gck[j]);
                if (st.en_state_soc)
                    for (int k = 0; k < bv.ncx; ++k) project_soc3(vc + bv.Acx[k], bv.cx[k]);
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) {
                    const int e = i * nx + j;
                    const T gcn = R::sub(R::add(gck[j], x[j]), vc[j]);
                    acc[j] = R::sub(acc[j], R::mul(rho, R::sub(vc[j], gcn)));
                    ST(bv.gc, e) = gcn;
                }
            }
            if (lin_x) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) vc[j] = R::add(x[j], glk[j]);
                project_halfspaces(vc, nx, bv.Alin_x, bv.blin_x, bv.nlx);
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) {
                    const int e = i * nx + j;
                    const T gln = R::sub(R::add(glk[j], x[j]), vc[j]);
                    acc[j] = R::sub(acc[j], R::mul(rho, R::sub(vc[j], gln)));
                    ST(bv.gl, e) = gln;
                }
            }
            if (!terminal) {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) ST(bv.q, i * nx + j) = acc[j];
                // x_{i+1} = A x_i + B u_i + f
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) {
                    T s1 = 0.0, s2 = 0.0;
#pragma unroll UNR
                    for (int k = 0; k < MX; ++k) if (k < nx) s1 = R::fma(mt.A(j, k), x[k], s1);
#pragma unroll UNR
                    for (int k = 0; k < MU; ++k) if (k < nu) s2 = R::fma(mt.B(j, k), u[k], s2);
                    tmp[j] = R::add(R::add(s1, s2), mt.f(j));
                }
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) x[j] = tmp[j];
            } else {
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) p[j] = acc[j];
            }
        }

        iter += 1;

        // ---------------- adaptive rho (reference admm.cpp:331-357): after iterations 6, 11, 16, ... and BEFORE the
        // termination test, which then multiplies the dual residuals by the new rho.  benchmark_rho_adaptation
        // (rho_benchmark.cpp:213-249) writes the horizon as one dense problem [x_0 u_0 x_1 ... x_{N-1}] with an identity
        // row block per input and a row block [A B -I] per step; only the non-zero blocks are evaluated here, on a second
        // roll-out of the trajectory (same operations as the forward sweep, so the same x and u).  Quirks kept: the
        // dynamics rows carry no f and are compared with the SLACK vnew_{i+1} (:97-98, :158), the duals are the updated
        // ones, q = Qd .* x at the terminal knot too (:135), Pinf enters before it is updated (:113).
        if constexpr (M::kAdaptive) {
            if (mt.ad_.enabled && it > 0 && it % 5 == 0) {
                T pri_res = 0, max_Ax = 0, max_z = 0, dual_res = 0, max_Px = 0, max_ATy = 0, max_q = 0;
#pragma unroll UNR
                for (int j = 0; j < MX; ++j) if (j < nx) x[j] = bv.x0[static_cast<size_t>(b) * nx + j];
                for (int i = 0; i < N; ++i) {
                    const bool terminal = (i == N - 1);
                    if (!terminal) {
#pragma unroll UNR
                        for (int j = 0; j < MU; ++j) if (j < nu) {
                            T s = 0;
#pragma unroll UNR
                            for (int k = 0; k < MX; ++k) if (k < nx) s = R::fma(mt.Kinf(j, k), x[k], s);
                            u[j] = R::sub(-s, ST(bv.d, i * nu + j));
                            const T zj = ST(z_new, i * nu + j);               // input rows: (A x) = u_i, z = znew_i
                            pri_res = R::max(pri_res, R::abs(R::sub(u[j], zj)));
                            max_Ax = R::max(max_Ax, R::abs(u[j]));
                            max_z = R::max(max_z, R::abs(zj));
                        }
#pragma unroll UNR
                        for (int j = 0; j < MX; ++j) if (j < nx) acc[j] = ST(bv.g, (i + 1) * nx + j);     // g_{i+1}
#pragma unroll UNR
                        for (int j = 0; j < MX; ++j) if (j < nx) {
                            T s1 = 0.0, s2 = 0.0;
#pragma unroll UNR
                            for (int k = 0; k < MX; ++k) if (k < nx) s1 = R::fma(mt.A(j, k), x[k], s1);
#pragma unroll UNR
                            for (int k = 0; k < MU; ++k) if (k < nu) s2 = R::fma(mt.B(j, k), u[k], s2);
                            const T ab = R::add(s1, s2);
                            tmp[j] = R::add(ab, mt.f(j));                       // x_{i+1} as the forward sweep computed it
                            const T dyn = R::sub(ab, tmp[j]);                   // dynamics row: A x_i + B u_i - x_{i+1}
                            const T vn = ST(v_new, (i + 1) * nx + j);           // ... against z = vnew_{i+1}
                            pri_res = R::max(pri_res, R::abs(R::sub(dyn, vn)));
                            max_Ax = R::max(max_Ax, R::abs(dyn));
                            max_z = R::max(max_z, R::abs(vn));
                        }
#pragma unroll UNR
                        for (int j = 0; j < MU; ++j) if (j < nu) {              // u_i block of P x + q + A' y
                            const T Px = R::mul(mt.Rd(j), u[j]);                // = q (zero reference, :141-143)
                            T aty = ST(bv.y, i * nu + j);
#pragma unroll UNR
                            for (int k = 0; k < MX; ++k) if (k < nx) aty = R::fma(mt.B(k, j), acc[k], aty);
                            dual_res = R::max(dual_res, R::abs(R::add(R::add(Px, Px), aty)));
                            max_Px = R::max(max_Px, R::abs(Px));
                            max_q = R::max(max_q, R::abs(Px));
                            max_ATy = R::max(max_ATy, R::abs(aty));
                        }
                    }
#pragma unroll UNR
                    for (int j = 0; j < MX; ++j) if (j < nx) {                  // x_i block
                        const T qv = R::mul(mt.Qd(j), x[j]);
                        T Px = qv;
                        if (terminal) {
                            Px = 0;
#pragma unroll UNR
                            for (int k = 0; k < MX; ++k) if (k < nx) Px = R::fma(mt.Pinf(j, k), x[k], Px);
                        }
                        T aty = 0;
                        if (i >= 1) aty = -ST(bv.g, i * nx + j);                // -I of the previous dynamics block (y = g_i)
                        if (!terminal) {
#pragma unroll UNR
                            for (int k = 0; k < MX; ++k) if (k < nx) aty = R::fma(mt.A(k, j), acc[k], aty);   // A' g_{i+1}
                        }
                        dual_res = R::max(dual_res, R::abs(R::add(R::add(Px, qv), aty)));
                        max_Px = R::max(max_Px, R::abs(Px));
                        max_q = R::max(max_q, R::abs(qv));
                        max_ATy = R::max(max_ATy, R::abs(aty));
                    }
                    if (!terminal) {
#pragma unroll UNR
                        for (int j = 0; j < MX; ++j) if (j < nx) x[j] = tmp[j];
                    }
                }
                // predict_rho (rho_benchmark.cpp:179-202)
                const T eps = static_cast<T>(1e-10);
                const T pri_norm = R::max(max_Ax, max_z);
                const T dual_norm = R::max(R::max(max_Px, max_ATy), max_q);
                const T n_pri = R::div(pri_res, R::add(pri_norm, eps));
                const T n_dua = R::div(dual_res, R::add(dual_norm, eps));
                T new_rho = R::mul(rho, R::sqrt(R::div(n_pri, R::add(n_dua, eps))));
                if (mt.ad_.clip) {                                              // std::min(std::max(new_rho, lo), hi)
                    const T lo = static_cast<T>(mt.ad_.rho_min), hi = static_cast<T>(mt.ad_.rho_max);
                    new_rho = (new_rho < lo) ? lo : new_rho;
                    new_rho = (hi < new_rho) ? hi : new_rho;
                }
                // update_matrices_with_derivatives (:204-218): first-order update of this instance's Kinf and Pinf.
                // (C1 / C2 are updated there as well; nothing in the solve path reads them - the loop uses Quu_inv / AmBKt.)
                const T delta = R::sub(new_rho, rho);
                for (int e = 0; e < nu * nx; ++e) {
                    double* p_ = mt.ad_.Kinf + static_cast<size_t>(e) * ld + b;
                    *p_ = R::add(static_cast<T>(*p_), R::mul(delta, static_cast<T>(__ldg(mt.ad_.dK + e))));
                }
                for (int e = 0; e < nx * nx; ++e) {
                    double* p_ = mt.ad_.Pinf + static_cast<size_t>(e) * ld + b;
                    *p_ = R::add(static_cast<T>(*p_), R::mul(delta, static_cast<T>(__ldg(mt.ad_.dP + e))));
                }
                rho = new_rho;
            }
        }

        if (iter % st.check_termination == 0) {
            converged = pri_x < tol_pri && pri_u < tol_pri &&
                        R::mul(dua_x, rho) < tol_dua && R::mul(dua_u, rho) < tol_dua;
        }
        if (converged) break;     // v, z deliberately NOT advanced (reference admm.cpp:375 precedes :379-380)
        cur ^= 1;
    }

    // ---------------- results and carried state
    const int sol = converged ? (cur ^ 1) : cur;     // buffer holding vnew / znew of the last iteration
    const int nxN = nx * N, nuN = nu * (N - 1);
    double* Xs = bv.Xsol + static_cast<size_t>(b) * nxN;
    double* Us = bv.Usol + static_cast<size_t>(b) * nuN;
    // copies in groups of eight loads, then eight stores (a load behind a store that may alias it would wait for nothing less
    // than the previous round trip)
    auto copy_out = [&](double* dst, const double* src_arr, int n) {            // dst[e] = element e of the thread's column
        int e = 0;
        for (; e + 8 <= n; e += 8) {
            T w[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) w[k] = reinterpret_cast<const T*>(src_arr)[static_cast<size_t>(e + k) * ld + b];
#pragma unroll
            for (int k = 0; k < 8; ++k) dst[e + k] = w[k];
        }
        for (; e < n; ++e) dst[e] = reinterpret_cast<const T*>(src_arr)[static_cast<size_t>(e) * ld + b];
    };
    auto copy_col = [&](double* dst_arr, const double* src_arr, int n) {        // element-major to element-major
        int e = 0;
        for (; e + 8 <= n; e += 8) {
            T w[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) w[k] = reinterpret_cast<const T*>(src_arr)[static_cast<size_t>(e + k) * ld + b];
#pragma unroll
            for (int k = 0; k < 8; ++k) reinterpret_cast<T*>(dst_arr)[static_cast<size_t>(e + k) * ld + b] = w[k];
        }
        for (; e < n; ++e) reinterpret_cast<T*>(dst_arr)[static_cast<size_t>(e) * ld + b] = reinterpret_cast<const T*>(src_arr)[static_cast<size_t>(e) * ld + b];
    };
    copy_out(Xs, bv.v[sol], nxN);
    copy_out(Us, bv.z[sol], nuN);
    if (cur == 1) {   // the carried previous-slack pair must end up in buffer 0
        copy_col(bv.v[0], bv.v[1], nxN);
        copy_col(bv.z[0], bv.z[1], nuN);
    }
#pragma unroll UNR
    for (int j = 0; j < MX; ++j) if (j < nx) ST(bv.pN, j) = p[j];
    bv.status[b] = converged ? 0 : 1;
    bv.iters[b] = iter;
    if constexpr (M::kAdaptive) mt.ad_.rho[b] = rho;
#undef ST
}

// any cone or half-space active? (same conditions as in the kernel)
inline bool needs_extra(const BatchView& bv, const AdmmSettings& st) {
    return (st.en_state_soc && bv.ncx > 0) || (st.en_input_soc && bv.ncu > 0) || st.en_state_linear != 0 || st.en_input_linear != 0;
}

template <class T, int NX, int NU>
void fill_const(MatsConst<T, NX, NU>& m, const HostMats& h) {
    auto cp = [](T* dst, const std::vector<double>& src) { for (size_t i = 0; i < src.size(); ++i) dst[i] = static_cast<T>(src[i]); };
    cp(m.A_, h.A); cp(m.B_, h.B); cp(m.Kinf_, h.Kinf); cp(m.AmBKt_, h.AmBKt); cp(m.Quu_inv_, h.Quu_inv);
    cp(m.Pinf_, h.Pinf); cp(m.f_, h.f); cp(m.APf_, h.APf); cp(m.BPf_, h.BPf); cp(m.Qd_, h.Qd); cp(m.Rd_, h.Rd);
}

template <class T, int NX, int NU>
cudaError_t launch_const(const BatchView& bv, const AdmmSettings& st, const HostMats& h, cudaStream_t stream) {
    MatsConst<T, NX, NU> m;
    fill_const(m, h);
    const int grid = (bv.B + kThreadKernelBlock - 1) / kThreadKernelBlock;
    if (needs_extra(bv, st)) admm_thread_kernel<T, MatsConst<T, NX, NU>, true><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
    else admm_thread_kernel<T, MatsConst<T, NX, NU>, false><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
    return cudaGetLastError();
}

template <class T>
cudaError_t launch_shared(const BatchView& bv, const AdmmSettings& st, const HostMats& h, cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (bv.nx == 12 && bv.nu == 4) return launch_const<T, 12, 4>(bv, st, h, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch_const<T, 4, 1>(bv, st, h, stream);
    if (bv.nx == 6 && bv.nu == 3) return launch_const<T, 6, 3>(bv, st, h, stream);
    MatsGlobal m;
    m.nx_ = bv.nx; m.nu_ = bv.nu;
    m.A_ = bv.A; m.B_ = bv.Bm; m.Kinf_ = bv.Kinf; m.AmBKt_ = bv.AmBKt; m.Quu_inv_ = bv.Quu_inv; m.Pinf_ = bv.Pinf;
    m.f_ = bv.f; m.APf_ = bv.APf; m.BPf_ = bv.BPf; m.Qd_ = bv.Qd; m.Rd_ = bv.Rd;
    const int grid = (bv.B + kThreadKernelBlock - 1) / kThreadKernelBlock;
    if (needs_extra(bv, st)) admm_thread_kernel<T, MatsGlobal, true><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
    else admm_thread_kernel<T, MatsGlobal, false><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
    return cudaGetLastError();
}

template <class T>
cudaError_t launch_per_instance(const BatchView& bv, const AdmmSettings& st, const PerInstanceMats& pm,
                                cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (bv.nx > kPerInstanceMaxNx || bv.nu > kPerInstanceMaxNu) return cudaErrorInvalidValue;
    const int grid = (bv.B + kThreadKernelBlock - 1) / kThreadKernelBlock;
    auto go = [&](auto m) {
        m.nx_ = bv.nx; m.nu_ = bv.nu; m.ld_ = bv.ld; m.p_ = pm; m.b_ = 0;
        if (needs_extra(bv, st)) admm_thread_kernel<T, decltype(m), true><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
        else admm_thread_kernel<T, decltype(m), false><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
        return cudaGetLastError();
    };
    if (bv.nx == 12 && bv.nu == 4) return go(MatsPerInstance<12, 4>{});
    if (bv.nx == 4 && bv.nu == 1) return go(MatsPerInstance<4, 1>{});
    return go(MatsPerInstance<0, 0>{});
}

// broadcast of the shared cache into the per-instance adaptive state (fresh / reset batch, new cache terms)
__global__ void adaptive_init_kernel(AdaptiveRho ad, const double* __restrict__ Kinf, const double* __restrict__ Pinf,
                                     double rho, int nkinf, int npinf, int B, long long ld, bool keep_rho) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    for (int e = 0; e < nkinf; ++e) ad.Kinf[e * ld + b] = __ldg(Kinf + e);
    for (int e = 0; e < npinf; ++e) ad.Pinf[e * ld + b] = __ldg(Pinf + e);
    if (!keep_rho) ad.rho[b] = rho;
}

// instance-major copies of the adapted cache for the getter: Kinf nu x nx x B, Pinf nx x nx x B
__global__ void adaptive_gather_kernel(AdaptiveRho ad, double* __restrict__ Kinf, double* __restrict__ Pinf, int nkinf,
                                       int npinf, int B, long long ld) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    if (Kinf) for (int e = 0; e < nkinf; ++e) Kinf[static_cast<size_t>(b) * nkinf + e] = ad.Kinf[e * ld + b];
    if (Pinf) for (int e = 0; e < npinf; ++e) Pinf[static_cast<size_t>(b) * npinf + e] = ad.Pinf[e * ld + b];
}

}  // namespace

cudaError_t launch_adaptive_init(const BatchView& bv, const AdaptiveRho& ad, cudaStream_t stream, bool keep_rho) {
    if (bv.B <= 0) return cudaSuccess;
    adaptive_init_kernel<<<(bv.B + 127) / 128, 128, 0, stream>>>(ad, bv.Kinf, bv.Pinf, bv.rho, bv.nu * bv.nx, bv.nx * bv.nx,
                                                                 bv.B, bv.ld, keep_rho);
    return cudaGetLastError();
}

cudaError_t launch_adaptive_gather(const BatchView& bv, const AdaptiveRho& ad, double* Kinf, double* Pinf,
                                   cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    adaptive_gather_kernel<<<(bv.B + 127) / 128, 128, 0, stream>>>(ad, Kinf, Pinf, bv.nu * bv.nx, bv.nx * bv.nx, bv.B, bv.ld);
    return cudaGetLastError();
}

cudaError_t launch_admm_thread_adaptive(const BatchView& bv, const AdmmSettings& st, const AdaptiveRho& ad,
                                        cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (!thread_kernel_supports(bv.nx, bv.nu)) return cudaErrorInvalidValue;
    const int grid = (bv.B + kThreadKernelBlock - 1) / kThreadKernelBlock;
    auto go = [&](auto m) {
        m.nx_ = bv.nx; m.nu_ = bv.nu; m.ld_ = bv.ld; m.b_ = 0; m.ad_ = ad;
        m.A_ = bv.A; m.B_ = bv.Bm; m.AmBKt_ = bv.AmBKt; m.Quu_inv_ = bv.Quu_inv; m.f_ = bv.f; m.APf_ = bv.APf;
        m.BPf_ = bv.BPf; m.Qd_ = bv.Qd; m.Rd_ = bv.Rd;
        if (needs_extra(bv, st)) admm_thread_kernel<double, decltype(m), true><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
        else admm_thread_kernel<double, decltype(m), false><<<grid, kThreadKernelBlock, 0, stream>>>(m, bv, st);
        return cudaGetLastError();
    };
    if (bv.nx == 12 && bv.nu == 4) return go(MatsAdaptive<12, 4>{});
    if (bv.nx == 4 && bv.nu == 1) return go(MatsAdaptive<4, 1>{});
    return go(MatsAdaptive<0, 0>{});
}

bool thread_kernel_supports(int nx, int nu) {
    return nx >= 1 && nu >= 1 && nx <= kThreadKernelMaxNx && nu <= kThreadKernelMaxNu;
}

cudaError_t launch_admm_thread(const BatchView& bv, const AdmmSettings& st, const HostMats& h, cudaStream_t stream,
                               int precision) {
    return precision == 32 ? launch_shared<float>(bv, st, h, stream) : launch_shared<double>(bv, st, h, stream);
}

cudaError_t launch_admm_thread_per_instance(const BatchView& bv, const AdmmSettings& st, const PerInstanceMats& pm,
                                            cudaStream_t stream, int precision) {
    return precision == 32 ? launch_per_instance<float>(bv, st, pm, stream)
                           : launch_per_instance<double>(bv, st, pm, stream);
}

}  // namespace tmpc
