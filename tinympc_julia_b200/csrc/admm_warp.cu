// The LATENCY kernel: a block of one or four warps per problem instance.  It is what a single solve runs on (the B = 1
// singleton behind the reference's sixteen symbols: set_x0 -> solve_mpc -> get_controls) and what small batches run on, where
// the throughput kernels leave the GPU empty and a solve costs what one 4-lane sub-warp needs for it.
//
// Reference: tinympc/TinyMPC/src/tinympc/admm.cpp:13-389 (solve, backward_pass_grad, forward_pass, update_slack box part,
// update_dual, update_linear_cost, termination_condition).  Box constraints only (like the sub-warp and TMEM kernels).
//
// What is sequential in an ADMM iteration is two recursions over the horizon; everything else is elementwise over all knot
// points at once.  The kernel separates the two:
//   * recursions on warp 0, one lane per matrix row, the vector of the previous knot point broadcast through shared memory:
//       backward   p_i = AmBKt p_{i+1} + c_i               c_i = (q_i - Kinf' r_i) + APf does not depend on the recursion
//       forward    x_{i+1} = (A - B Kinf) x_i + e_i        e_i = f - B d_i likewise: the closed-loop form of
//                                                           x_{i+1} = A x_i + B u_i + f with u_i = -Kinf x_i - d_i, which
//                                                           needs no second broadcast (of u) per knot point
//     measured 125-135 cycles per knot point (B200, 1.9 GHz): one shared-memory round trip, NX FMAs in four interleaved
//     chains and the issue of ~35 instructions by a lone warp;
//   * phases over all (knot, row) pairs on every warp of the block: d_i = Quu_inv (B' p_{i+1} + r_i + BPf) after the backward
//     recursion; u_i, slack, dual, linear cost and the residual tests after the forward one; convergence by block vote
//     (__syncthreads_and).
// Measured per iteration of the quadrotor shape with four warps (clock64 around the sections, -DTMPC_WARP_PROFILE): backward
// 2,600 cycles, d phase 370, forward 2,570, slack phases 970, barrier 260 = 6.8 k cycles = 3.5 us; one warp alone 8.4 k.
// The elementwise formulas are the thread kernel's (admm_thread.cu), operation for operation; the dot products are summed in
// another order than the reference's ascending chains (split chains, closed-loop forward form), so results agree with the
// other kernels to rounding (1e-13 relative on the tests' problems; identical status and iteration counts on all of them),
// not bit for bit.
// Carried state: the element-major arrays of the thread kernel (admm_device.cuh), read once at the start of a warm-started
// solve and written once at the end.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>

#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace tmpc {
namespace {

// shared-memory carve-up, in doubles (nx-sized arrays first: their rows are read with 128-bit loads)
template <int NX, int NU>
struct WarpSmem {
    static constexpr int kNxArrays = 9;     // G V VN Q PS XS XLO XHI CQ
    static constexpr int kNuArrays = 8;     // Y Z ZN R D ULO UHI CR
    static __host__ __device__ size_t doubles(int N) {
        return static_cast<size_t>(kNxArrays) * NX * N + static_cast<size_t>(kNuArrays) * NU * (N - 1) + NX /*TQ*/ +
               32 /*scratch*/ + 2;
    }
};

// chains of the dot products: measured on the quadrotor shape, 2, 3, 4 or 6 chains and either assignment of k to chains
// all give the same step time (the step is bound by the in-order issue of its ~35 dependent-latency instructions, not by
// the depth of a chain)
template <int NX> struct Chains { static constexpr int value = (NX % 4 == 0 && NX >= 8) ? 4 : ((NX % 3 == 0) ? 3 : 2); };

// s + sum_k m[k] v[k] in Chains<NX> interleaved FMA chains (the first one starts from `s`), summed pairwise
template <int NX>
__device__ __forceinline__ double dot_chains(const double (&m)[NX], const double (&v)[NX], double s) {
    constexpr int CH = Chains<NX>::value;
    double a[CH];
    a[0] = s;
#pragma unroll
    for (int c = 1; c < CH; ++c) a[c] = 0.0;
#pragma unroll
    for (int k = 0; k < NX; ++k) a[k % CH] = fma(m[k], v[k], a[k % CH]);
#pragma unroll
    for (int w = 1; w < CH; w *= 2) {
#pragma unroll
        for (int c = 0; c + w < CH; c += 2 * w) a[c] = dadd(a[c], a[c + w]);
    }
    return a[0];
}

// row i of an [N][NX] shared array into registers (every lane the same addresses: broadcast)
template <int NX>
__device__ __forceinline__ void load_row(const double* row, double (&v)[NX]) {
    static_assert(NX % 2 == 0, "rows are read as double2");
#pragma unroll
    for (int k = 0; k < NX; k += 2) {
        const double2 t = *reinterpret_cast<const double2*>(row + k);
        v[k] = t.x; v[k + 1] = t.y;
    }
}

// NT threads per instance: warp 0 runs the recursions, all NT / 32 warps the phases over the knot points.  Four warps when
// every instance of the batch is resident at once (lowest latency: the phases are a third of a one-warp iteration), one warp
// beyond that (four times the instances per SM).
template <int NX, int NU, int NT>
__global__ void __launch_bounds__(NT)
admm_warp_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st, const int cold) {
    static_assert(32 % NU == 0, "the lanes of a knot point's input rows sit in one aligned group");
    extern __shared__ __align__(16) double sm[];
    const int tid = threadIdx.x, lane = tid & 31, b = blockIdx.x, N = bv.N;
    const bool w0 = tid < 32;
    const int nxN = NX * N, nuN = NU * (N - 1);
    const size_t ld = static_cast<size_t>(bv.ld);
    const double rho = bv.rho;
    const double tol_pri = st.abs_pri_tol, tol_dua = st.abs_dua_tol;
    const double inf = INFINITY;

    double* G = sm;            double* V = G + nxN;     double* VN = V + nxN;   double* Q = VN + nxN;
    double* PS = Q + nxN;      double* XS = PS + nxN;   double* XLO = XS + nxN; double* XHI = XLO + nxN;
    double* CQ = XHI + nxN;
    double* Y = CQ + nxN;      double* Z = Y + nuN;     double* ZN = Z + nuN;   double* R = ZN + nuN;
    double* D = R + nuN;       double* ULO = D + nuN;   double* UHI = ULO + nuN; double* CR = UHI + nuN;
    double* TQ = CR + nuN;     double* DUM = TQ + NX;   // 32 scratch slots: where the lanes without a matrix row store in the recursions

    const bool box_x = st.en_state_bound && bv.x_min != nullptr;
    const bool box_u = st.en_input_bound && bv.u_min != nullptr;
    const double* xref = bv.Xref ? bv.Xref + static_cast<size_t>(b) * bv.xref_stride : nullptr;
    const double* uref = bv.Uref ? bv.Uref + static_cast<size_t>(b) * bv.uref_stride : nullptr;
    const double* xmin = box_x ? bv.x_min + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* xmax = box_x ? bv.x_max + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* umin = box_u ? bv.u_min + static_cast<size_t>(b) * bv.ub_stride : nullptr;
    const double* umax = box_u ? bv.u_max + static_cast<size_t>(b) * bv.ub_stride : nullptr;
    const bool warm = (cold == 0);

    // ---------------- the thread's matrix rows (global loads first: they are in flight while the state is staged)
    const int jr = lane < NX ? lane : 0;             // recursions (warp 0): row of the state
    const int ju = tid % NU;                         // phases over (knot, input row): the thread's input row never changes
    double Mb[NX], Mf[NX], Bc[NX], Kr[NX], Br[NU], Qi[NU], Kc[NU];
#pragma unroll
    for (int k = 0; k < NX; ++k) {
        Mb[k] = w0 ? __ldg(bv.AmBKt + k * NX + jr) : 0.0;      // AmBKt(jr, k)
        Mf[k] = w0 ? __ldg(bv.A + k * NX + jr) : 0.0;          // A(jr, k); - B(jr, :) Kinf(:, k) below
        Bc[k] = __ldg(bv.Bm + ju * NX + k);                    // B(k, ju)
        Kr[k] = __ldg(bv.Kinf + k * NU + ju);                  // Kinf(ju, k)
    }
#pragma unroll
    for (int k = 0; k < NU; ++k) {
        Br[k] = w0 ? __ldg(bv.Bm + k * NX + jr) : 0.0;         // B(jr, k)
        Kc[k] = w0 ? __ldg(bv.Kinf + jr * NU + k) : 0.0;       // Kinf(k, jr)
        Qi[k] = __ldg(bv.Quu_inv + k * NU + ju);               // Quu_inv(ju, k)
    }
    const double f_j = __ldg(bv.f + jr), APf_j = __ldg(bv.APf + jr), BPf_j = __ldg(bv.BPf + ju);
    // closed-loop matrix of the forward recursion from A, B and Kinf themselves, not from the cached AmBKt: a cache replaced
    // through set_cache_terms need not be consistent, and the reference's forward pass reads A, B, Kinf
    if (w0) {
#pragma unroll
        for (int k = 0; k < NX; ++k) {
#pragma unroll
            for (int m = 0; m < NU; ++m) Mf[k] = fma(-Br[m], __ldg(bv.Kinf + k * NU + m), Mf[k]);
        }
    }

    // ---------------- stage: carried state, bounds, reference terms.  Two rounds of 128 elements at a time, every global
    // load of the two issued before the first shared-memory store
    for (int tb = 0; tb < nxN; tb += 2 * NT) {
        double g2[2], v2[2], q2[2], lo2[2], hi2[2], xr2[2], qd2[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int t = tb + tid + NT * u;
            const int tc = t < nxN ? t : 0;
            const size_t e = static_cast<size_t>(tc) * ld + b;
            g2[u] = warm ? bv.g[e] : 0.0;
            v2[u] = warm ? bv.v[0][e] : 0.0;
            q2[u] = (warm && tc < nxN - NX) ? bv.q[e] : 0.0;
            lo2[u] = box_x ? __ldg(xmin + tc) : -inf;
            hi2[u] = box_x ? __ldg(xmax + tc) : inf;
            xr2[u] = xref ? __ldg(xref + tc) : 0.0;
            qd2[u] = __ldg(bv.Qd + tc % NX);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int t = tb + tid + NT * u;
            if (t < nxN) {
                G[t] = g2[u]; V[t] = v2[u]; Q[t] = q2[u]; XLO[t] = lo2[u]; XHI[t] = hi2[u];
                CQ[t] = -dmul(xr2[u], qd2[u]);                                  // admm.cpp:218
            }
        }
    }
    for (int tb = 0; tb < nuN; tb += NT) {
        const int t = tb + tid;
        if (t < nuN) {
            const size_t e = static_cast<size_t>(t) * ld + b;
            const double y1 = warm ? bv.y[e] : 0.0, z1 = warm ? bv.z[0][e] : 0.0, r1 = warm ? bv.r[e] : 0.0;
            const double lo1 = box_u ? __ldg(umin + t) : -inf, hi1 = box_u ? __ldg(umax + t) : inf;
            const double ur1 = uref ? __ldg(uref + t) : 0.0, rd1 = __ldg(bv.Rd + t % NU);
            Y[t] = y1; Z[t] = z1; R[t] = r1; ULO[t] = lo1; UHI[t] = hi1;
            CR[t] = -dmul(ur1, rd1);                                            // admm.cpp:228
        }
    }
    if (tid < NX) {
        double s = 0.0;                                                          // -(Xref_{N-1}' Pinf)'   (admm.cpp:238)
        if (xref) {
#pragma unroll
            for (int k = 0; k < NX; ++k) s = fma(__ldg(xref + (N - 1) * NX + k), __ldg(bv.Pinf + tid * NX + k), s);
        }
        TQ[tid] = -s;
        XS[tid] = bv.x0[static_cast<size_t>(b) * NX + tid];                      // x_0, the same in every iteration
        PS[(N - 1) * NX + tid] = warm ? bv.pN[static_cast<size_t>(tid) * ld + b] : 0.0;
    }
    __syncthreads();

    int iter = 0;
    bool conv = false;
#ifdef TMPC_WARP_PROFILE
    long long tk[6] = {0, 0, 0, 0, 0, 0}, t_prev = clock64();
#define TICK(n) do { const long long t_now = clock64(); tk[n] += t_now - t_prev; t_prev = t_now; } while (0)
#else
#define TICK(n) do { } while (0)
#endif
    TICK(0);
    // where a lane of warp 0 stores in the recursions: its row of the knot point, or (lanes without a row) a scratch slot:
    // no divergent region around the step
    double* const pw = lane < NX ? PS + lane : DUM + lane;
    double* const xw = lane < NX ? XS + lane : DUM + lane;
    const int sw = lane < NX ? NX : 0;

    for (int it = 0; it < st.max_iter; ++it) {
        // ---------------- backward recursion (admm.cpp:13-20) on warp 0: p_i = c_i + AmBKt p_{i+1}, i = N-2 .. 0, with
        // c_i = (q_i - Kinf' r_i) + APf formed one knot point ahead, in the shadow of the shared-memory round trip
        if (w0) {
            for (int i = N - 2; i >= 0; --i) {
                double pv[NX];
                load_row<NX>(PS + (i + 1) * NX, pv);
                // c_i = (q_i - Kinf' r_i) + APf: independent of the recursion, joins after the dot product
                double s2 = 0.0;
#pragma unroll
                for (int k = 0; k < NU; ++k) s2 = fma(Kc[k], R[i * NU + k], s2);
                const double c = dadd(dsub(Q[i * NX + jr], s2), APf_j);
                pw[i * sw] = dadd(dot_chains<NX>(Mb, pv, 0.0), c);
                __syncwarp();
            }
        }
        TICK(1);
        __syncthreads();
        // ---------------- d_i = Quu_inv (B' p_{i+1} + r_i + BPf) for every knot point
        for (int tb = 0; tb < nuN; tb += NT) {       // block-uniform trip count: the shuffles below need every lane of a warp
            const int t = tb + tid;
            const int tc = t < nuN ? t : 0, i = tc / NU;
            double pv[NX];
            load_row<NX>(PS + (i + 1) * NX, pv);
            const double tj = dadd(dadd(dot_chains<NX>(Bc, pv, 0.0), R[tc]), BPf_j);
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < NU; ++k) s = fma(Qi[k], __shfl_sync(0xffffffffu, tj, (lane & ~(NU - 1)) + k), s);
            if (t < nuN) D[t] = s;
        }
        TICK(2);
        __syncthreads();
        // ---------------- forward recursion (admm.cpp:25-32) on warp 0, closed-loop form: x_{i+1} = AmBK x_i + (f - B d_i),
        // the second term formed one knot point ahead
        if (w0) {
            for (int i = 0; i < N - 1; ++i) {
                double xv[NX];
                load_row<NX>(XS + i * NX, xv);
                // e_i = f - B d_i: independent of the recursion, joins after the dot product
                double s2 = 0.0;
#pragma unroll
                for (int k = 0; k < NU; ++k) s2 = fma(Br[k], D[i * NU + k], s2);
                xw[(i + 1) * sw] = dadd(dot_chains<NX>(Mf, xv, 0.0), dsub(f_j, s2));
                __syncwarp();
            }
        }
        TICK(3);
        __syncthreads();
        // ---------------- input rows of every knot point: u_i, slack, dual, r_i, residual tests
        // (admm.cpp:84-90, :184-185, :228-233, :257-265)
        bool ok = true;
        for (int tb = 0; tb < nuN; tb += NT) {       // block-uniform trip count (shuffles)
            const int t = tb + tid;
            const int tc = t < nuN ? t : 0, i = tc / NU;
            double xv[NX];
            load_row<NX>(XS + i * NX, xv);
            const double u = dsub(-dot_chains<NX>(Kr, xv, 0.0), D[tc]);
            const double yj = Y[tc];
            const double znew = clampd(dadd(u, yj), ULO[tc], UHI[tc]);
            const double ynew = dsub(dadd(yj, u), znew);
            const bool pass = (fabs(dsub(u, znew)) < tol_pri) && (dmul(fabs(dsub(Z[tc], znew)), rho) < tol_dua);
            if (t < nuN) {
                ok = ok && pass;
                R[t] = dsub(CR[tc], dmul(rho, dsub(znew, ynew)));
                Y[t] = ynew;
                ZN[t] = znew;
            }
        }
        // ---------------- state rows: slack, dual, q_i / terminal p, residual tests
        // (admm.cpp:92-98, :186-187, :218-219, :238-246); two rounds in flight
        for (int tb = 0; tb < nxN; tb += 2 * NT) {
            double x_[2], g_[2], lo_[2], hi_[2], vo_[2], cq_[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int t = tb + tid + NT * u;
                const int tc = t < nxN ? t : 0;
                x_[u] = XS[tc]; g_[u] = G[tc]; lo_[u] = XLO[tc]; hi_[u] = XHI[tc]; vo_[u] = V[tc];
                cq_[u] = tc >= nxN - NX ? TQ[tc - (nxN - NX)] : CQ[tc];
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int t = tb + tid + NT * u;
                const double vnew = clampd(dadd(x_[u], g_[u]), lo_[u], hi_[u]);
                const double gnew = dsub(dadd(g_[u], x_[u]), vnew);
                const bool pass = (fabs(dsub(x_[u], vnew)) < tol_pri) && (dmul(fabs(dsub(vo_[u], vnew)), rho) < tol_dua);
                const double q = dsub(cq_[u], dmul(rho, dsub(vnew, gnew)));      // q_i, or the terminal p
                if (t < nxN) {
                    ok = ok && pass;
                    G[t] = gnew;
                    VN[t] = vnew;
                    if (t < nxN - NX) Q[t] = q; else PS[t] = q;
                }
            }
        }
        iter += 1;
        TICK(4);
        if (iter % st.check_termination == 0) conv = __syncthreads_and(ok) != 0;
        else __syncthreads();
        TICK(5);
        if (conv) break;             // v, z deliberately NOT advanced (admm.cpp:375 precedes :379-380)
        double* tp = V; V = VN; VN = tp;
        tp = Z; Z = ZN; ZN = tp;
    }

    // ---------------- results and carried state
    const double* Xn = conv ? VN : V;        // vnew / znew of the last iteration
    const double* Zn = conv ? ZN : Z;
    double* Xs = bv.Xsol + static_cast<size_t>(b) * nxN;
    double* Us = bv.Usol + static_cast<size_t>(b) * nuN;
    for (int t = tid; t < nxN; t += NT) {
        const size_t e = static_cast<size_t>(t) * ld + b;
        Xs[t] = Xn[t];
        bv.g[e] = G[t];
        bv.v[0][e] = V[t];
        if (t < nxN - NX) bv.q[e] = Q[t];
    }
    for (int t = tid; t < nuN; t += NT) {
        const size_t e = static_cast<size_t>(t) * ld + b;
        Us[t] = Zn[t];
        bv.y[e] = Y[t];
        bv.z[0][e] = Z[t];
        bv.r[e] = R[t];
    }
    if (tid < NX) bv.pN[static_cast<size_t>(tid) * ld + b] = PS[(N - 1) * NX + tid];
    if (tid == 0) { bv.status[b] = conv ? 0 : 1; bv.iters[b] = iter; }
#ifdef TMPC_WARP_PROFILE
    TICK(0);
    if (tid == 0 && b == 0)
        printf("warp kernel cycles: stage+writeout %lld, per iteration: backward %lld, d phase (incl. barrier) %lld, forward (incl. barrier) %lld, "
               "slack phases (incl. barrier) %lld, final barrier %lld  (%d iterations)\n", tk[0], tk[1] / iter, tk[2] / iter, tk[3] / iter,
               tk[4] / iter, tk[5] / iter, iter);
#endif
#undef TICK
}

template <int NX, int NU>
bool fits(int N) { return N >= 2 && WarpSmem<NX, NU>::doubles(N) * sizeof(double) <= 200 * 1024; }

template <int NX, int NU, int NT>
cudaError_t launch_nt(const BatchView& bv, const AdmmSettings& st, int cold, cudaStream_t stream) {
    const size_t smem = WarpSmem<NX, NU>::doubles(bv.N) * sizeof(double);
    static size_t opted = 0;      // largest dynamic shared-memory size the kernel has been opted into (per process)
    if (smem > 48 * 1024 && smem > opted) {
        const cudaError_t e = cudaFuncSetAttribute(admm_warp_kernel<NX, NU, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        opted = smem;
    }
    admm_warp_kernel<NX, NU, NT><<<bv.B, NT, smem, stream>>>(bv, st, cold);
    return cudaGetLastError();
}

template <int NX, int NU>
cudaError_t launch(const BatchView& bv, const AdmmSettings& st, int cold, int num_sms, cudaStream_t stream) {
    // four warps per instance while two such blocks per SM hold the whole batch (registers: 2 x 128 x ~240)
    if (bv.B <= 2 * num_sms) return launch_nt<NX, NU, 128>(bv, st, cold, stream);
    return launch_nt<NX, NU, 32>(bv, st, cold, stream);
}

}  // namespace

bool warp_kernel_supports(const BatchView& bv, const AdmmSettings& st) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (bv.nx == 12 && bv.nu == 4) return fits<12, 4>(bv.N);
    if (bv.nx == 4 && bv.nu == 1) return fits<4, 1>(bv.N);
    return false;
}

cudaError_t launch_admm_warp(const BatchView& bv, const AdmmSettings& st, int cold, int num_sms, cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return cudaErrorNotSupported;     // the caller routes these to the general kernel
    if (bv.nx == 12 && bv.nu == 4) return launch<12, 4>(bv, st, cold, num_sms, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch<4, 1>(bv, st, cold, num_sms, stream);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
