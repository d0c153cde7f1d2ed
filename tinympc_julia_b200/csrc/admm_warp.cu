// One WARP per problem instance, built for latency: the kernel behind a single solve (the B = 1 singleton of the reference's
// sixteen symbols: set_x0 -> solve_mpc -> get_controls) and behind small batches, where the throughput kernels leave the GPU
// empty and a solve costs what one 4-lane sub-warp needs for it.
//
// Reference: tinympc/TinyMPC/src/tinympc/admm.cpp:13-389 (solve, backward_pass_grad, forward_pass, update_slack box part,
// update_dual, update_linear_cost, termination_condition).  Box constraints only (like the sub-warp and TMEM kernels).
//
// What is sequential in an ADMM iteration is two recursions over the horizon; everything else is elementwise over all knot
// points at once.  The kernel separates the two:
//   * recursions, one lane per matrix row, the vector of the previous knot point broadcast through shared memory:
//       backward   p_i = c_i + AmBKt p_{i+1}               c_i = (q_i - Kinf' r_i) + APf formed beforehand for every i
//       forward    x_{i+1} = (A - B Kinf) x_i - B d_i + f  (the closed-loop form of x_{i+1} = A x_i + B u_i + f with
//                                                           u_i = -Kinf x_i - d_i: no second broadcast of u per knot point)
//     a knot point costs one shared-memory round trip plus NX FMAs in three interleaved chains (~110 cycles);
//   * phases over all (knot, row) pairs, 32 at a time: d_i = Quu_inv (B' p_{i+1} + r_i + BPf) after the backward recursion;
//     u_i, slack, dual, linear cost, c_i and the residual tests after the forward one; convergence by warp vote.
// The elementwise formulas are the thread kernel's (admm_thread.cu), operation for operation; the dot products are summed in
// another order than the reference's ascending chains (split chains, closed-loop forward form), so results agree with the
// other kernels to rounding (1e-13 relative on the tests' problems), not bit for bit.
// Carried state: the element-major arrays of the thread kernel (admm_device.cuh), read once at the start of a warm-started
// solve and written once at the end.
#include <cuda_runtime.h>

#include <cmath>

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
               NU * NX /*Kinf*/ + NX /*APf*/ + 32 /*scratch*/ + 2;
    }
};

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
    if constexpr (CH == 4) return dadd(dadd(a[0], a[1]), dadd(a[2], a[3]));
    else if constexpr (CH == 3) return dadd(dadd(a[0], a[1]), a[2]);
    else return dadd(a[0], a[1]);
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

template <int NX, int NU>
__global__ void __launch_bounds__(32)
admm_warp_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st, const int cold) {
    static_assert(32 % NU == 0, "the lanes of a knot point's input rows sit in one aligned group");
    extern __shared__ __align__(16) double sm[];
    const int lane = threadIdx.x, b = blockIdx.x, N = bv.N;
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
    double* TQ = CR + nuN;     double* KS = TQ + NX;    double* APF = KS + NU * NX;
    double* DUM = APF + NX;    // 32 scratch slots: where the lanes without a matrix row store in the recursions

    const bool box_x = st.en_state_bound && bv.x_min != nullptr;
    const bool box_u = st.en_input_bound && bv.u_min != nullptr;
    const double* xref = bv.Xref ? bv.Xref + static_cast<size_t>(b) * bv.xref_stride : nullptr;
    const double* uref = bv.Uref ? bv.Uref + static_cast<size_t>(b) * bv.uref_stride : nullptr;
    const double* xmin = box_x ? bv.x_min + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* xmax = box_x ? bv.x_max + static_cast<size_t>(b) * bv.xb_stride : nullptr;
    const double* umin = box_u ? bv.u_min + static_cast<size_t>(b) * bv.ub_stride : nullptr;
    const double* umax = box_u ? bv.u_max + static_cast<size_t>(b) * bv.ub_stride : nullptr;
    const bool warm = (cold == 0);

    // ---------------- stage: carried state, bounds, reference terms, the small constants the phases index by row.
    // Four rounds of 32 elements at a time, every global load of the four issued before the first shared-memory store
    // (the stores may alias the next loads as far as the compiler knows: round by round each one would wait for an L2
    // round trip)
    for (int tb = 0; tb < nxN; tb += 128) {
        const int t0 = tb + lane;
        double g4[4], v4[4], q4[4], lo4[4], hi4[4], xr4[4], qd4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + 32 * u;
            const bool on = t < nxN;
            const int tc = on ? t : 0;
            const size_t e = static_cast<size_t>(tc) * ld + b;
            g4[u] = warm ? bv.g[e] : 0.0;
            v4[u] = warm ? bv.v[0][e] : 0.0;
            q4[u] = (warm && tc < nxN - NX) ? bv.q[e] : 0.0;
            lo4[u] = box_x ? __ldg(xmin + tc) : -inf;
            hi4[u] = box_x ? __ldg(xmax + tc) : inf;
            xr4[u] = xref ? __ldg(xref + tc) : 0.0;
            qd4[u] = __ldg(bv.Qd + tc % NX);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + 32 * u;
            if (t < nxN) {
                G[t] = g4[u]; V[t] = v4[u]; Q[t] = q4[u]; XLO[t] = lo4[u]; XHI[t] = hi4[u];
                CQ[t] = -dmul(xr4[u], qd4[u]);                                  // admm.cpp:218
            }
        }
    }
    for (int tb = 0; tb < nuN; tb += 128) {
        const int t0 = tb + lane;
        double y4[4], z4[4], r4[4], lo4[4], hi4[4], ur4[4], rd4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + 32 * u;
            const bool on = t < nuN;
            const int tc = on ? t : 0;
            const size_t e = static_cast<size_t>(tc) * ld + b;
            y4[u] = warm ? bv.y[e] : 0.0;
            z4[u] = warm ? bv.z[0][e] : 0.0;
            r4[u] = warm ? bv.r[e] : 0.0;
            lo4[u] = box_u ? __ldg(umin + tc) : -inf;
            hi4[u] = box_u ? __ldg(umax + tc) : inf;
            ur4[u] = uref ? __ldg(uref + tc) : 0.0;
            rd4[u] = __ldg(bv.Rd + tc % NU);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + 32 * u;
            if (t < nuN) {
                Y[t] = y4[u]; Z[t] = z4[u]; R[t] = r4[u]; ULO[t] = lo4[u]; UHI[t] = hi4[u];
                CR[t] = -dmul(ur4[u], rd4[u]);                                  // admm.cpp:228
            }
        }
    }
    for (int t = lane; t < NU * NX; t += 32) KS[t] = __ldg(bv.Kinf + t);        // Kinf(k, j) at j NU + k
    if (lane < NX) {
        APF[lane] = __ldg(bv.APf + lane);
        double s = 0.0;                                                          // -(Xref_{N-1}' Pinf)'   (admm.cpp:238)
        if (xref) {
#pragma unroll
            for (int k = 0; k < NX; ++k) s = fma(__ldg(xref + (N - 1) * NX + k), __ldg(bv.Pinf + lane * NX + k), s);
        }
        TQ[lane] = -s;
        XS[lane] = bv.x0[static_cast<size_t>(b) * NX + lane];                    // x_0, the same in every iteration
        PS[(N - 1) * NX + lane] = warm ? bv.pN[static_cast<size_t>(lane) * ld + b] : 0.0;
    }

    // ---------------- the lane's matrix rows
    const int jr = lane < NX ? lane : 0;             // recursions: row of the state
    const int ju = lane % NU;                        // phases over (knot, input row): the lane's input row never changes
    double Mb[NX], Mf[NX], Bc[NX], Kr[NX], Br[NU], Qi[NU];
#pragma unroll
    for (int k = 0; k < NX; ++k) {
        Mb[k] = __ldg(bv.AmBKt + k * NX + jr);       // AmBKt(jr, k)
        Mf[k] = __ldg(bv.A + k * NX + jr);           // A(jr, k); - B(jr, :) Kinf(:, k) below
        Bc[k] = __ldg(bv.Bm + ju * NX + k);          // B(k, ju)
        Kr[k] = __ldg(bv.Kinf + k * NU + ju);        // Kinf(ju, k)
    }
#pragma unroll
    for (int k = 0; k < NU; ++k) {
        Br[k] = __ldg(bv.Bm + k * NX + jr);          // B(jr, k)
        Qi[k] = __ldg(bv.Quu_inv + k * NU + ju);     // Quu_inv(ju, k)
    }
    // closed-loop matrix of the forward recursion from A, B and Kinf themselves, not from the cached AmBKt: a cache replaced
    // through set_cache_terms need not be consistent, and the reference's forward pass reads A, B, Kinf
#pragma unroll
    for (int k = 0; k < NX; ++k) {
#pragma unroll
        for (int m = 0; m < NU; ++m) Mf[k] = fma(-Br[m], __ldg(bv.Kinf + k * NU + m), Mf[k]);
    }
    const double f_j = __ldg(bv.f + jr), BPf_j = __ldg(bv.BPf + ju);
    __syncwarp();

    // c_i = (q_i - Kinf' r_i) + APf for the first backward recursion, from the carried q and r
    for (int t = lane; t < nxN - NX; t += 32) {
        const int i = t / NX, j = t % NX;
        double s2 = 0.0;
#pragma unroll
        for (int k = 0; k < NU; ++k) s2 = fma(KS[j * NU + k], R[i * NU + k], s2);
        PS[t] = dadd(dsub(Q[t], s2), APF[j]);
    }
    __syncwarp();

    int iter = 0;
    bool conv = false;

    // where a lane stores in the recursions: its row of the knot point, or (lanes without a row) a scratch slot: no
    // divergent region around the step
    double* const pw = lane < NX ? PS + lane : DUM + lane;
    double* const xw = lane < NX ? XS + lane : DUM + lane;
    const int sw = lane < NX ? NX : 0;
    constexpr int UI = 3;        // rounds of the input-row phases in flight together (19 knot points x 4 rows = 3 rounds)
    constexpr int UX = 4;        // rounds of the state-row phase in flight together

    for (int it = 0; it < st.max_iter; ++it) {
        // ---------------- backward recursion (admm.cpp:13-20): p_i over c_i in place, i = N-2 .. 0
        for (int i = N - 2; i >= 0; --i) {
            double pv[NX];
            load_row<NX>(PS + (i + 1) * NX, pv);
            const double c = PS[i * NX + jr];
            pw[i * sw] = dot_chains<NX>(Mb, pv, c);
            __syncwarp();
        }
        // ---------------- d_i = Quu_inv (B' p_{i+1} + r_i + BPf) for every knot point
        for (int tb = 0; tb < nuN; tb += 32 * UI) {      // warp-uniform trip count: the shuffles below need every lane
            const int t0 = tb + lane;
            double tj[UI];
#pragma unroll
            for (int u = 0; u < UI; ++u) {
                const int t = t0 + 32 * u;
                const int tc = t < nuN ? t : 0, i = tc / NU;
                double pv[NX];
                load_row<NX>(PS + (i + 1) * NX, pv);
                tj[u] = dadd(dadd(dot_chains<NX>(Bc, pv, 0.0), R[tc]), BPf_j);
            }
#pragma unroll
            for (int u = 0; u < UI; ++u) {
                const int t = t0 + 32 * u;
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < NU; ++k) s = fma(Qi[k], __shfl_sync(0xffffffffu, tj[u], (lane & ~(NU - 1)) + k), s);
                if (t < nuN) D[t] = s;
            }
        }
        __syncwarp();
        // ---------------- forward recursion (admm.cpp:25-32), closed-loop form: x_{i+1} = (AmBK x_i - B d_i) + f
        for (int i = 0; i < N - 1; ++i) {
            double xv[NX];
            load_row<NX>(XS + i * NX, xv);
            double s2 = 0.0;
#pragma unroll
            for (int k = 0; k < NU; ++k) s2 = fma(Br[k], D[i * NU + k], s2);
            xw[(i + 1) * sw] = dadd(dot_chains<NX>(Mf, xv, -s2), f_j);
            __syncwarp();
        }
        // ---------------- input rows of every knot point: u_i, slack, dual, r_i, residual tests
        // (admm.cpp:84-90, :184-185, :228-233, :257-265)
        bool ok = true;
        for (int tb = 0; tb < nuN; tb += 32 * UI) {
            const int t0 = tb + lane;
            double u_[UI], y_[UI], lo_[UI], hi_[UI], zo_[UI], cr_[UI];
#pragma unroll
            for (int u = 0; u < UI; ++u) {
                const int t = t0 + 32 * u;
                const int tc = t < nuN ? t : 0, i = tc / NU;
                double xv[NX];
                load_row<NX>(XS + i * NX, xv);
                u_[u] = dsub(-dot_chains<NX>(Kr, xv, 0.0), D[tc]);
                y_[u] = Y[tc]; lo_[u] = ULO[tc]; hi_[u] = UHI[tc]; zo_[u] = Z[tc]; cr_[u] = CR[tc];
            }
#pragma unroll
            for (int u = 0; u < UI; ++u) {
                const int t = t0 + 32 * u;
                const double znew = clampd(dadd(u_[u], y_[u]), lo_[u], hi_[u]);
                const double ynew = dsub(dadd(y_[u], u_[u]), znew);
                const bool pass = (fabs(dsub(u_[u], znew)) < tol_pri) && (dmul(fabs(dsub(zo_[u], znew)), rho) < tol_dua);
                if (t < nuN) {
                    ok = ok && pass;
                    R[t] = dsub(cr_[u], dmul(rho, dsub(znew, ynew)));
                    Y[t] = ynew;
                    ZN[t] = znew;
                }
            }
        }
        __syncwarp();
        // ---------------- state rows: slack, dual, q_i, the next c_i / terminal p, residual tests
        // (admm.cpp:92-98, :186-187, :218-219, :238-246)
        for (int tb = 0; tb < nxN; tb += 32 * UX) {
            const int t0 = tb + lane;
            double x_[UX], g_[UX], lo_[UX], hi_[UX], vo_[UX], cq_[UX], s2_[UX], af_[UX];
#pragma unroll
            for (int u = 0; u < UX; ++u) {
                const int t = t0 + 32 * u;
                const int tc = t < nxN ? t : 0, i = tc / NX, j = tc % NX;
                x_[u] = XS[tc]; g_[u] = G[tc]; lo_[u] = XLO[tc]; hi_[u] = XHI[tc]; vo_[u] = V[tc];
                const bool term = i >= N - 1;
                cq_[u] = term ? TQ[j] : CQ[tc];
                af_[u] = APF[j];
                const int ir = term ? 0 : i;
                double s2 = 0.0;
#pragma unroll
                for (int k = 0; k < NU; ++k) s2 = fma(KS[j * NU + k], R[ir * NU + k], s2);
                s2_[u] = s2;
            }
#pragma unroll
            for (int u = 0; u < UX; ++u) {
                const int t = t0 + 32 * u;
                const double vnew = clampd(dadd(x_[u], g_[u]), lo_[u], hi_[u]);
                const double gnew = dsub(dadd(g_[u], x_[u]), vnew);
                const bool pass = (fabs(dsub(x_[u], vnew)) < tol_pri) && (dmul(fabs(dsub(vo_[u], vnew)), rho) < tol_dua);
                const double q = dsub(cq_[u], dmul(rho, dsub(vnew, gnew)));      // q_i, or the terminal p
                if (t < nxN) {
                    ok = ok && pass;
                    G[t] = gnew;
                    VN[t] = vnew;
                    if (t < nxN - NX) { Q[t] = q; PS[t] = dadd(dsub(q, s2_[u]), af_[u]); }
                    else PS[t] = q;
                }
            }
        }
        __syncwarp();
        iter += 1;
        if (iter % st.check_termination == 0) conv = __all_sync(0xffffffffu, ok);
        if (conv) break;             // v, z deliberately NOT advanced (admm.cpp:375 precedes :379-380)
        double* tp = V; V = VN; VN = tp;
        tp = Z; Z = ZN; ZN = tp;
    }

    // ---------------- results and carried state
    const double* Xn = conv ? VN : V;        // vnew / znew of the last iteration
    const double* Zn = conv ? ZN : Z;
    double* Xs = bv.Xsol + static_cast<size_t>(b) * nxN;
    double* Us = bv.Usol + static_cast<size_t>(b) * nuN;
    for (int t = lane; t < nxN; t += 32) {
        const size_t e = static_cast<size_t>(t) * ld + b;
        Xs[t] = Xn[t];
        bv.g[e] = G[t];
        bv.v[0][e] = V[t];
        if (t < nxN - NX) bv.q[e] = Q[t];
    }
    for (int t = lane; t < nuN; t += 32) {
        const size_t e = static_cast<size_t>(t) * ld + b;
        Us[t] = Zn[t];
        bv.y[e] = Y[t];
        bv.z[0][e] = Z[t];
        bv.r[e] = R[t];
    }
    if (lane < NX) bv.pN[static_cast<size_t>(lane) * ld + b] = PS[(N - 1) * NX + lane];
    if (lane == 0) { bv.status[b] = conv ? 0 : 1; bv.iters[b] = iter; }
}

template <int NX, int NU>
bool fits(int N) { return N >= 2 && WarpSmem<NX, NU>::doubles(N) * sizeof(double) <= 200 * 1024; }

template <int NX, int NU>
cudaError_t launch(const BatchView& bv, const AdmmSettings& st, int cold, cudaStream_t stream) {
    const size_t smem = WarpSmem<NX, NU>::doubles(bv.N) * sizeof(double);
    static size_t opted = 0;      // largest dynamic shared-memory size the kernel has been opted into (per process)
    if (smem > 48 * 1024 && smem > opted) {
        const cudaError_t e = cudaFuncSetAttribute(admm_warp_kernel<NX, NU>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        opted = smem;
    }
    admm_warp_kernel<NX, NU><<<bv.B, 32, smem, stream>>>(bv, st, cold);
    return cudaGetLastError();
}

}  // namespace

bool warp_kernel_supports(const BatchView& bv, const AdmmSettings& st) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (bv.nx == 12 && bv.nu == 4) return fits<12, 4>(bv.N);
    if (bv.nx == 4 && bv.nu == 1) return fits<4, 1>(bv.N);
    return false;
}

cudaError_t launch_admm_warp(const BatchView& bv, const AdmmSettings& st, int cold, cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return cudaErrorNotSupported;     // the caller routes these to the general kernel
    if (bv.nx == 12 && bv.nu == 4) return launch<12, 4>(bv, st, cold, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch<4, 1>(bv, st, cold, stream);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
