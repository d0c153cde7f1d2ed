// ADMM solve kernel, mapping "one problem instance per 4-lane sub-warp, state resident on chip".
//
// This is the flagship kernel for the box-constrained shapes (quadrotor 12/4, cartpole 4/1).
//
//  * A warp owns a TILE of 8 instances.  lane = sub * 8 + inst: the four lanes {inst, inst+8, inst+16,
//    inst+24} cooperate on instance `inst`; lane `sub` owns state rows {r*4 + sub} and input row `sub`.
//    All four lanes are busy in every mat-vec of the Riccati recursions (100 % lane use for nx % 4 == 0).
//  * The whole per-instance ADMM state (duals g, y; box slacks v, z; feed-forward d) lives in SHARED
//    MEMORY for the entire solve, laid out [element][8 instances] so that a warp-wide access is 32
//    consecutive doubles (conflict free) and an 8-instance tile of the element-major HBM arrays maps
//    onto it 1:1.  HBM is touched once per solve: carried state in, solution + carried state out.
//  * The lane's rows of the cached matrices (AmBKt, Kinf', B', Quu_inv for the backward pass; A, B, Kinf
//    for the forward pass) are REGISTER resident for the duration of a pass; the broadcast operand of
//    every FMA (p_{i+1} or x_i, 12 values) is exchanged through a padded, double-buffered shared-memory
//    slot read back with 128-bit loads; the 4-vectors (r_i, B'p+r, u_i) are exchanged with warp shuffles.
//  * update_slack / update_dual / update_linear_cost and the residual tests are fused into the forward
//    sweep; the linear cost q_i, r_i is never stored — it is recomputed from v, g (z, y) when the
//    backward sweep needs it, which is bit-identical to storing it.
//  * Convergence is decided per instance with warp votes; converged instances stop mutating their
//    state, the warp leaves when its 8 instances are done, and warps pull tiles from a global queue so
//    that bimodal iteration counts do not idle SMs.
//
// Reference being reproduced: tinympc/TinyMPC/src/tinympc/admm.cpp:13-32 (Riccati sweeps), :81-98 (box
// slack), :181-187 (dual), :214-247 (linear cost), :253-271 (termination), :274-389 (solve).
// Residual tests: the reference compares max|.| (times rho for the dual residuals) against the
// tolerances; max(a_i) < tol  <=>  all(a_i < tol), and fl(rho * a) is monotone in a, so the per-element
// tests against the precomputed thresholds below decide exactly the same predicate without FP64 max.
#include <cuda_runtime.h>

#include <cmath>

#include "admm_device.cuh"
#include "admm_kernels.hpp"

namespace tmpc {
namespace {

constexpr int kTile = 8;            // instances per warp
constexpr unsigned kFull = 0xffffffffu;

struct SubwarpParams {
    double pri_thresh;      // primal residuals pass  <=>  |x - vnew| <  pri_thresh   (= abs_pri_tol)
    double dua_thresh;      // dual residuals pass    <=>  |v - vnew| <= dua_thresh   (largest a with fl(rho*a) < abs_dua_tol)
    int cold;               // carried state known to be zero: skip its loads
    int num_tiles;
    int* counter;           // tile queue
    int warps_per_block;
    // shared-memory carve-up (in doubles)
    int off_consts, off_bounds, off_refs, off_warp0, warp_stride;
    int shared_bounds, shared_refs;   // staged in shared memory (stride 0) vs read per instance from global
};

template <int NX>
struct Pad {   // row pitch of the broadcast slot: an odd number of 16-byte chunks => 8 instances hit 8 bank groups
    static constexpr int value = ((NX / 2) % 2 == 1) ? NX : NX + 2;
};

template <int NX, int NU>
struct ConstLayout {      // shared staging of the problem constants (doubles)
    static constexpr int A_rm = 0;                       // A(row, k) row-major
    static constexpr int AmBKt_rm = A_rm + NX * NX;      // AmBKt(row, k) row-major
    static constexpr int B_rm = AmBKt_rm + NX * NX;      // B(row, k) row-major  (NX x NU)
    static constexpr int B_cm = B_rm + NX * NU;          // B(k, j) column-major: column j = row j of B'
    static constexpr int Kinf_rm = B_cm + NX * NU;       // Kinf(j, k) row-major (NU x NX)
    static constexpr int Kinf_cm = Kinf_rm + NX * NU;    // Kinf(k, row): column `row` = row of Kinf'
    static constexpr int Quu_rm = Kinf_cm + NX * NU;     // Quu_inv(j, k) row-major
    static constexpr int Pinf_cm = Quu_rm + NU * NU;     // Pinf(k, row): column `row`
    static constexpr int f = Pinf_cm + NX * NX;
    static constexpr int APf = f + NX;
    static constexpr int BPf = APf + NX;
    static constexpr int Qd = BPf + NU;
    static constexpr int Rd = Qd + NX;
    static constexpr int size = ((Rd + NU) + 1) / 2 * 2;
};

__device__ __forceinline__ double shfl_d(double v, int src_lane) { return __shfl_sync(kFull, v, src_lane); }

__device__ __forceinline__ void lds128(const double* p, double& a, double& b) {
    const double2 t = *reinterpret_cast<const double2*>(p);
    a = t.x; b = t.y;
}

template <int NX, int NU, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1)
admm_subwarp_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st,
                    const __grid_constant__ SubwarpParams sp) {
    static_assert(NX % 4 == 0 && NX >= 4, "state rows are dealt 4 lanes x RX rows");
    static_assert(NU >= 1 && NU <= 4, "one input row per lane");
    constexpr int RX = NX / 4;
    constexpr int XP = Pad<NX>::value;
    using CL = ConstLayout<NX, NU>;

    extern __shared__ __align__(16) double smem[];
    const int N = bv.N;
    const int nxN = NX * N, nuN = NU * (N - 1);
    const size_t ld = static_cast<size_t>(bv.ld);
    const double rho = bv.rho;

    // ---------------- stage the shared constants once per block
    double* sc = smem + sp.off_consts;
    for (int idx = threadIdx.x; idx < NX * NX; idx += blockDim.x) {
        const int row = idx / NX, k = idx % NX;
        sc[CL::A_rm + idx] = bv.A[k * NX + row];
        sc[CL::AmBKt_rm + idx] = bv.AmBKt[k * NX + row];
        sc[CL::Pinf_cm + idx] = bv.Pinf[idx];
    }
    for (int idx = threadIdx.x; idx < NX * NU; idx += blockDim.x) {
        const int row = idx / NU, k = idx % NU;          // B(row, k)
        sc[CL::B_rm + idx] = bv.Bm[k * NX + row];
        sc[CL::B_cm + idx] = bv.Bm[idx];
        const int j = idx / NX, kk = idx % NX;           // Kinf(j, kk)
        sc[CL::Kinf_rm + idx] = bv.Kinf[kk * NU + j];
        sc[CL::Kinf_cm + idx] = bv.Kinf[idx];
    }
    for (int idx = threadIdx.x; idx < NU * NU; idx += blockDim.x)
        sc[CL::Quu_rm + idx] = bv.Quu_inv[(idx % NU) * NU + idx / NU];
    for (int idx = threadIdx.x; idx < NX; idx += blockDim.x) {
        sc[CL::f + idx] = bv.f[idx]; sc[CL::APf + idx] = bv.APf[idx]; sc[CL::Qd + idx] = bv.Qd[idx];
    }
    for (int idx = threadIdx.x; idx < NU; idx += blockDim.x) { sc[CL::BPf + idx] = bv.BPf[idx]; sc[CL::Rd + idx] = bv.Rd[idx]; }

    const bool box_x = st.en_state_bound && bv.x_min != nullptr;
    const bool box_u = st.en_input_bound && bv.u_min != nullptr;
    const bool has_xref = bv.Xref != nullptr, has_uref = bv.Uref != nullptr;
    double* s_xmin = smem + sp.off_bounds;            // shared (stride-0) bounds staged once
    double* s_xmax = s_xmin + nxN;
    double* s_umin = s_xmax + nxN;
    double* s_umax = s_umin + nuN;
    if (sp.shared_bounds) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) { s_xmin[e] = bv.x_min[e]; s_xmax[e] = bv.x_max[e]; }
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) { s_umin[e] = bv.u_min[e]; s_umax[e] = bv.u_max[e]; }
    }
    double* s_xref = smem + sp.off_refs;
    double* s_uref = s_xref + nxN;
    if (sp.shared_refs) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) s_xref[e] = has_xref ? bv.Xref[e] : 0.0;
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) s_uref[e] = has_uref ? bv.Uref[e] : 0.0;
    }
    __syncthreads();

    // ---------------- per-warp carve-up
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inst = lane & 7, sub = lane >> 3;
    const bool urow = sub < NU;                       // this lane owns an input row
    double* wbase = smem + sp.off_warp0 + static_cast<size_t>(warp) * sp.warp_stride;
    double* s_g = wbase;
    double* s_v = s_g + nxN * kTile;
    double* s_y = s_v + nxN * kTile;
    double* s_z = s_y + nuN * kTile;
    double* s_d = s_z + nuN * kTile;
    double* s_bc = s_d + nuN * kTile;                 // 2 x kTile x XP broadcast slots
    // state entry (column i, owned row r):  [(i*NX + r*4)*8 + lane];  input entry (column i): [i*NU*8 + lane]
#define SX(i, r) (((i) * NX + (r) * 4) * kTile + lane)
#define SU(i) ((i) * NU * kTile + lane)

    const double fr_BPf = urow ? sc[CL::BPf + sub] : 0.0;
    const double fr_Rd = urow ? sc[CL::Rd + sub] : 0.0;
    double fr_f[RX], fr_APf[RX], fr_Qd[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { fr_f[r] = sc[CL::f + r * 4 + sub]; fr_APf[r] = sc[CL::APf + r * 4 + sub]; fr_Qd[r] = sc[CL::Qd + r * 4 + sub]; }

    int bcast_phase = 0;

    for (;;) {
        // ---------------- next tile from the queue
        int tile = 0;
        if (lane == 0) tile = atomicAdd(sp.counter, 1);
        tile = __shfl_sync(kFull, tile, 0);
        if (tile >= sp.num_tiles) break;
        const int b = tile * kTile + inst;
        const bool valid = b < bv.B;
        const size_t bb = valid ? static_cast<size_t>(b) : 0;

        const double* xref_g = bv.Xref + bb * bv.xref_stride;   // only dereferenced when !shared_refs && has_xref
        const double* uref_g = bv.Uref + bb * bv.uref_stride;
        const double* xmin_g = bv.x_min + bb * bv.xb_stride;    // only dereferenced when !shared_bounds && box_x
        const double* xmax_g = bv.x_max + bb * bv.xb_stride;
        const double* umin_g = bv.u_min + bb * bv.ub_stride;
        const double* umax_g = bv.u_max + bb * bv.ub_stride;

        // ---------------- carried state -> shared memory (each lane moves exactly the entries it owns)
        double x0r[RX], p_term[RX];
#pragma unroll
        for (int r = 0; r < RX; ++r) {
            x0r[r] = valid ? bv.x0[bb * NX + r * 4 + sub] : 0.0;
            p_term[r] = (valid && !sp.cold) ? bv.pN[(r * 4 + sub) * ld + bb] : 0.0;
        }
        if (sp.cold || !valid) {
            for (int i = 0; i < N; ++i) {
#pragma unroll
                for (int r = 0; r < RX; ++r) { s_g[SX(i, r)] = 0.0; s_v[SX(i, r)] = 0.0; }
            }
            if (urow) for (int i = 0; i < N - 1; ++i) { s_y[SU(i)] = 0.0; s_z[SU(i)] = 0.0; s_d[SU(i)] = 0.0; }
        } else {
            for (int i = 0; i < N; ++i) {
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    const size_t ge = static_cast<size_t>(i * NX + r * 4 + sub) * ld + bb;
                    s_g[SX(i, r)] = bv.g[ge]; s_v[SX(i, r)] = bv.v[0][ge];
                }
            }
            if (urow) for (int i = 0; i < N - 1; ++i) {
                const size_t ge = static_cast<size_t>(i * NU + sub) * ld + bb;
                s_y[SU(i)] = bv.y[ge]; s_z[SU(i)] = bv.z[0][ge]; s_d[SU(i)] = 0.0;
            }
        }

        bool done = !valid;       // a finished (or padding) instance stops mutating its state
        int my_iters = 0, my_status = 1;
        bool first = true;
        int iter = 0;

        for (int it = 0; it < st.max_iter; ++it) {
            const bool may_converge = ((iter + 1) % st.check_termination) == 0;

            // ======================= backward sweep: d_i, p_i for i = N-2 .. 0  (reference admm.cpp:13-20)
            {
                double Mab[RX][NX], Mkt[RX][NU], Mbt[NX], Mq[NU];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
#pragma unroll
                    for (int k = 0; k < NX; k += 2) lds128(sc + CL::AmBKt_rm + (r * 4 + sub) * NX + k, Mab[r][k], Mab[r][k + 1]);
#pragma unroll
                    for (int k = 0; k < NU; ++k) Mkt[r][k] = sc[CL::Kinf_cm + (r * 4 + sub) * NU + k];
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) Mbt[k] = urow ? sc[CL::B_cm + sub * NX + k] : 0.0;
#pragma unroll
                for (int k = 0; k < NU; ++k) Mq[k] = urow ? sc[CL::Quu_rm + sub * NU + k] : 0.0;

                double p[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) p[r] = p_term[r];

                for (int i = N - 2; i >= 0; --i) {
                    // broadcast p_{i+1} to the 4 lanes of the instance
                    double* slot = s_bc + (bcast_phase & 1) * (kTile * XP) + inst * XP;
                    ++bcast_phase;
#pragma unroll
                    for (int r = 0; r < RX; ++r) slot[r * 4 + sub] = p[r];
                    __syncwarp();
                    double pb[NX];
#pragma unroll
                    for (int k = 0; k < NX; k += 2) lds128(slot + k, pb[k], pb[k + 1]);

                    // r_i (own row) and q_i (own rows): carried values on the first sweep of a solve,
                    // afterwards recomputed from the slacks and duals (reference admm.cpp:218-233)
                    double r_own = 0.0, q_own[RX];
                    if (first) {
                        if (sp.cold || !valid) {
#pragma unroll
                            for (int r = 0; r < RX; ++r) q_own[r] = 0.0;
                        } else {
                            if (urow) r_own = bv.r[static_cast<size_t>(i * NU + sub) * ld + bb];
#pragma unroll
                            for (int r = 0; r < RX; ++r) q_own[r] = bv.q[static_cast<size_t>(i * NX + r * 4 + sub) * ld + bb];
                        }
                    } else {
                        if (urow) {
                            const int e = i * NU + sub;
                            const double ur = sp.shared_refs ? s_uref[e] : (has_uref ? __ldg(uref_g + e) : 0.0);
                            r_own = dsub(-dmul(ur, fr_Rd), dmul(rho, dsub(s_z[SU(i)], s_y[SU(i)])));
                        }
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            const int e = i * NX + r * 4 + sub;
                            const double xr = sp.shared_refs ? s_xref[e] : (has_xref ? __ldg(xref_g + e) : 0.0);
                            q_own[r] = dsub(-dmul(xr, fr_Qd[r]), dmul(rho, dsub(s_v[SX(i, r)], s_g[SX(i, r)])));
                        }
                    }

                    // d_i = Quu_inv (B' p_{i+1} + r_i + BPf)
                    double s = 0.0;
#pragma unroll
                    for (int k = 0; k < NX; ++k) s = fma(Mbt[k], pb[k], s);
                    const double t_own = dadd(dadd(s, r_own), fr_BPf);
                    double dsum = 0.0, rr[NU];
#pragma unroll
                    for (int k = 0; k < NU; ++k) {
                        dsum = fma(Mq[k], shfl_d(t_own, k * kTile + inst), dsum);
                        rr[k] = shfl_d(r_own, k * kTile + inst);
                    }
                    if (urow) s_d[SU(i)] = dsum;

                    // p_i = q_i + AmBKt p_{i+1} - Kinf' r_i + APf
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        double s1 = 0.0, s2 = 0.0;
#pragma unroll
                        for (int k = 0; k < NX; ++k) s1 = fma(Mab[r][k], pb[k], s1);
#pragma unroll
                        for (int k = 0; k < NU; ++k) s2 = fma(Mkt[r][k], rr[k], s2);
                        p[r] = dadd(dsub(dadd(q_own[r], s1), s2), fr_APf[r]);
                    }
                }
            }

            // ======================= forward sweep fused with slack / dual / residual updates
            bool ok = true;          // every residual this lane has seen is under its tolerance
            {
                double Ma[RX][NX], Mb[RX][NU], Mk[NX];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
#pragma unroll
                    for (int k = 0; k < NX; k += 2) lds128(sc + CL::A_rm + (r * 4 + sub) * NX + k, Ma[r][k], Ma[r][k + 1]);
#pragma unroll
                    for (int k = 0; k < NU; ++k) Mb[r][k] = sc[CL::B_rm + (r * 4 + sub) * NU + k];
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) Mk[k] = urow ? sc[CL::Kinf_rm + sub * NX + k] : 0.0;

                double x[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) x[r] = x0r[r];

                for (int i = 0; i < N; ++i) {
                    const bool terminal = (i == N - 1);
                    double xb[NX], ub[NU];
                    if (!terminal) {
                        double* slot = s_bc + (bcast_phase & 1) * (kTile * XP) + inst * XP;
                        ++bcast_phase;
#pragma unroll
                        for (int r = 0; r < RX; ++r) slot[r * 4 + sub] = x[r];
                        __syncwarp();
#pragma unroll
                        for (int k = 0; k < NX; k += 2) lds128(slot + k, xb[k], xb[k + 1]);

                        // u_i = -Kinf x_i - d_i ; input slack, dual, residuals (admm.cpp:28, :87, :97, :187, :259-260)
                        double s = 0.0;
#pragma unroll
                        for (int k = 0; k < NX; ++k) s = fma(Mk[k], xb[k], s);
                        double u_own = 0.0;
                        if (urow) {
                            const int e = i * NU + sub;
                            u_own = dsub(-s, s_d[SU(i)]);
                            const double yj = s_y[SU(i)];
                            const double t = dadd(u_own, yj);
                            double znew = t;
                            if (box_u) {
                                const double lo = sp.shared_bounds ? s_umin[e] : __ldg(umin_g + e);
                                const double hi = sp.shared_bounds ? s_umax[e] : __ldg(umax_g + e);
                                znew = clampd(t, lo, hi);
                            }
                            const double zold = s_z[SU(i)];
                            ok = ok && (fabs(dsub(u_own, znew)) < sp.pri_thresh) && (fabs(dsub(zold, znew)) <= sp.dua_thresh);
                            if (!done) {
                                s_y[SU(i)] = dsub(t, znew);          // y + u - znew  (y + u == u + y bit for bit)
                                s_z[SU(i)] = znew;
                                if (may_converge) bv.z[0][static_cast<size_t>(e) * ld + bb] = zold;   // carried z lags one iteration
                            }
                        }
#pragma unroll
                        for (int k = 0; k < NU; ++k) ub[k] = shfl_d(u_own, k * kTile + inst);
                    }

                    // state column i: slack, dual, residuals (admm.cpp:84, :92, :184, :257-258)
                    double vg[RX];     // vnew - g(new), feeds the terminal cost
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const int e = i * NX + r * 4 + sub;
                        const double gj = s_g[SX(i, r)];
                        const double t = dadd(x[r], gj);
                        double vnew = t;
                        if (box_x) {
                            const double lo = sp.shared_bounds ? s_xmin[e] : __ldg(xmin_g + e);
                            const double hi = sp.shared_bounds ? s_xmax[e] : __ldg(xmax_g + e);
                            vnew = clampd(t, lo, hi);
                        }
                        const double vold = s_v[SX(i, r)];
                        const double gnew = dsub(t, vnew);
                        ok = ok && (fabs(dsub(x[r], vnew)) < sp.pri_thresh) && (fabs(dsub(vold, vnew)) <= sp.dua_thresh);
                        if (!done) {
                            s_g[SX(i, r)] = gnew;
                            s_v[SX(i, r)] = vnew;
                            if (may_converge) bv.v[0][static_cast<size_t>(e) * ld + bb] = vold;       // carried v lags one iteration
                        }
                        vg[r] = dsub(vnew, gnew);
                    }

                    if (!terminal) {
                        // x_{i+1} = A x_i + B u_i + f
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            double s1 = 0.0, s2 = 0.0;
#pragma unroll
                            for (int k = 0; k < NX; ++k) s1 = fma(Ma[r][k], xb[k], s1);
#pragma unroll
                            for (int k = 0; k < NU; ++k) s2 = fma(Mb[r][k], ub[k], s2);
                            x[r] = dadd(dadd(s1, s2), fr_f[r]);
                        }
                    } else if (!done) {
                        // p_{N-1} = -(Xref_{N-1}' Pinf)' - rho (vnew - g)   (admm.cpp:238-239)
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            double s = 0.0;
                            if (has_xref) {
                                for (int k = 0; k < NX; ++k) {
                                    const double xr = sp.shared_refs ? s_xref[i * NX + k] : __ldg(xref_g + i * NX + k);
                                    s = fma(xr, sc[CL::Pinf_cm + (r * 4 + sub) * NX + k], s);
                                }
                            }
                            p_term[r] = dsub(-s, dmul(rho, vg[r]));
                        }
                    }
                }
            }

            // ======================= termination (admm.cpp:253-271, :364-380)
            iter += 1;
            const unsigned votes = __ballot_sync(kFull, ok);
            const unsigned mine = 0x01010101u << inst;
            const bool inst_ok = ((iter % st.check_termination) == 0) && ((votes & mine) == mine);
            if (!done) {
                my_iters = iter;
                if (inst_ok) { done = true; my_status = 0; }
            }
            first = false;
            if (__all_sync(kFull, done)) break;
        }

        // ---------------- results and carried state back to HBM (each lane writes the entries it owns)
        if (valid) {
            double* Xs = bv.Xsol + bb * nxN;
            double* Us = bv.Usol + bb * nuN;
            const bool unconverged = (my_status != 0);
            for (int i = 0; i < N; ++i) {
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    const int e = i * NX + r * 4 + sub;
                    const size_t ge = static_cast<size_t>(e) * ld + bb;
                    const double vv = s_v[SX(i, r)], gg = s_g[SX(i, r)];
                    Xs[e] = vv;                                  // solution = vnew of the last iteration (admm.cpp:370,386)
                    bv.g[ge] = gg;
                    if (unconverged) bv.v[0][ge] = vv;           // v = vnew happened (admm.cpp:379); else the lagging copy is already there
                    if (i < N - 1) {                             // q as update_linear_cost left it (admm.cpp:218-219)
                        const double xr = sp.shared_refs ? s_xref[e] : (has_xref ? __ldg(xref_g + e) : 0.0);
                        bv.q[ge] = dsub(-dmul(xr, fr_Qd[r]), dmul(rho, dsub(vv, gg)));
                    }
                }
            }
            if (urow) for (int i = 0; i < N - 1; ++i) {
                const int e = i * NU + sub;
                const size_t ge = static_cast<size_t>(e) * ld + bb;
                const double zz = s_z[SU(i)], yy = s_y[SU(i)];
                Us[e] = zz;
                bv.y[ge] = yy;
                if (unconverged) bv.z[0][ge] = zz;
                const double ur = sp.shared_refs ? s_uref[e] : (has_uref ? __ldg(uref_g + e) : 0.0);
                bv.r[ge] = dsub(-dmul(ur, fr_Rd), dmul(rho, dsub(zz, yy)));
            }
#pragma unroll
            for (int r = 0; r < RX; ++r) bv.pN[(r * 4 + sub) * ld + bb] = p_term[r];
            if (sub == 0) { bv.status[b] = my_status; bv.iters[b] = my_iters; }
        }
    }
#undef SX
#undef SU
}

// Largest a >= 0 with fl(rho * a) < tol, so that  rho * max|.| < tol  <=>  every |.| <= a.
double dual_threshold(double rho, double tol) {
    if (!(tol > 0.0) || !(rho > 0.0)) return -1.0;
    double a = tol / rho;
    while (rho * a < tol) a = std::nextafter(a, INFINITY);
    while (!(rho * a < tol)) a = std::nextafter(a, -INFINITY);
    return a;
}

struct Plan { int warps; size_t smem_bytes; SubwarpParams sp; };

template <int NX, int NU, int MAXW>
bool make_plan(const BatchView& bv, Plan& pl) {
    constexpr int XP = Pad<NX>::value;
    const int N = bv.N, nxN = NX * N, nuN = NU * (N - 1);
    SubwarpParams sp{};
    int off = 0;
    sp.off_consts = off; off += ConstLayout<NX, NU>::size;
    sp.shared_bounds = (bv.x_min != nullptr && bv.xb_stride == 0 && bv.ub_stride == 0) ? 1 : 0;
    sp.off_bounds = off; if (sp.shared_bounds) off += 2 * nxN + 2 * nuN;
    sp.shared_refs = ((bv.Xref == nullptr || bv.xref_stride == 0) && (bv.Uref == nullptr || bv.uref_stride == 0)) ? 1 : 0;
    sp.off_refs = off; if (sp.shared_refs) off += nxN + nuN;
    off = (off + 1) / 2 * 2;
    sp.off_warp0 = off;
    int per_warp = (2 * nxN + 3 * nuN) * kTile + 2 * kTile * XP;
    per_warp = (per_warp + 1) / 2 * 2;
    sp.warp_stride = per_warp;
    const size_t limit = 227 * 1024;
    const size_t fixed = static_cast<size_t>(off) * sizeof(double);
    if (fixed + static_cast<size_t>(per_warp) * sizeof(double) > limit) return false;
    int warps = static_cast<int>((limit - fixed) / (static_cast<size_t>(per_warp) * sizeof(double)));
    if (warps > MAXW) warps = MAXW;
    pl.warps = warps;
    pl.smem_bytes = fixed + static_cast<size_t>(warps) * per_warp * sizeof(double);
    sp.warps_per_block = warps;
    pl.sp = sp;
    return true;
}

template <int NX, int NU, int MAXW>
cudaError_t launch(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms, cudaStream_t stream) {
    Plan pl;
    if (!make_plan<NX, NU, MAXW>(bv, pl)) return cudaErrorInvalidConfiguration;
    pl.sp.cold = cold;
    pl.sp.num_tiles = (bv.B + kTile - 1) / kTile;
    pl.sp.counter = counter;
    pl.sp.pri_thresh = st.abs_pri_tol;
    pl.sp.dua_thresh = dual_threshold(bv.rho, st.abs_dua_tol);
    auto kern = admm_subwarp_kernel<NX, NU, MAXW>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.smem_bytes));
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    int grid = (pl.sp.num_tiles + pl.warps - 1) / pl.warps;
    if (grid > num_sms) grid = num_sms;            // persistent: one CTA per SM, warps pull tiles from the queue
    kern<<<grid, pl.warps * 32, pl.smem_bytes, stream>>>(bv, st, pl.sp);
    return cudaGetLastError();
}

template <int NX, int NU, int MAXW>
bool fits(int N) {
    BatchView v{};
    v.N = N;
    Plan pl;
    return make_plan<NX, NU, MAXW>(v, pl);
}

}  // namespace

bool subwarp_kernel_supports(int nx, int nu, int N, const AdmmSettings& st) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (nx == 12 && nu == 4) return fits<12, 4, 8>(N);
    if (nx == 4 && nu == 1) return fits<4, 1, 16>(N);
    return false;
}

cudaError_t launch_admm_subwarp(const BatchView& bv, const AdmmSettings& st, const HostMats& h, int cold, int* counter,
                                int num_sms, cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return launch_admm_thread(bv, st, h, stream);   // degenerate settings: general kernel
    if (bv.nx == 12 && bv.nu == 4) return launch<12, 4, 8>(bv, st, cold, counter, num_sms, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch<4, 1, 16>(bv, st, cold, counter, num_sms, stream);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
