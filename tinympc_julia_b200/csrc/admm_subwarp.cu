// ADMM solve kernel, mapping "one problem instance per 4-lane sub-warp, state resident on chip".
//
// This is the flagship kernel for the box-constrained shapes (quadrotor 12/4, cartpole 4/1).
//
//  * A warp holds 8 instances.  lane = sub * 8 + inst: the four lanes {inst, inst+8, inst+16,
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
//  * Convergence is decided per instance with warp votes; a finished instance is written out and its
//    sub-warp pulls the next instance id from a global queue at once (per-instance refill: only the 4
//    lanes of the finished instance run the pass), so bimodal iteration counts idle neither lanes nor SMs.
//  * Templated on the scalar: T = double (the reference's default tinytype) and T = float (its float
//    build, DESIGN.md 3.8: half the shared memory per instance, 8 resident warps for the quadrotor shape).
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

template <class T>
struct SubwarpParams {
    T pri_thresh;      // primal residuals pass  <=>  |x - vnew| <  pri_thresh   (= abs_pri_tol)
    T dua_thresh;           // dual residuals pass    <=>  |v - vnew| <= dua_thresh   (largest a with fl(rho*a) < abs_dua_tol)
    int cold;               // carried state known to be zero: skip its loads
    int num_tiles;
    int* counter;           // tile queue
    int warps_per_block;
    // shared-memory carve-up (in doubles)
    int off_consts, off_bounds, off_refs, off_warp0, warp_stride;
    int shared_bounds, shared_refs;   // staged in shared memory (stride 0) vs read per instance from global
};

template <class T, int NX>
struct Pad {   // row pitch of the broadcast slot: an odd number of 16-byte chunks => the 8 instances hit 8 different bank groups
    static constexpr int kChunk = 16 / sizeof(T);                       // values per 128-bit load: 2 doubles, 4 floats
    static constexpr int value = ((NX / kChunk) % 2 == 1) ? NX : NX + kChunk;
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

template <class T>
__device__ __forceinline__ T shfl_d(T v, int src_lane) { return __shfl_sync(kFull, v, src_lane); }

// two consecutive values with one load (128-bit for double, 64-bit for float)
__device__ __forceinline__ void lds128(const double* p, double& a, double& b) {
    const double2 t = *reinterpret_cast<const double2*>(p);
    a = t.x; b = t.y;
}
__device__ __forceinline__ void lds128(const float* p, float& a, float& b) {
    const float2 t = *reinterpret_cast<const float2*>(p);
    a = t.x; b = t.y;
}
// two consecutive values with one store
__device__ __forceinline__ void sts_pair(double* p, double a, double b) { *reinterpret_cast<double2*>(p) = make_double2(a, b); }
__device__ __forceinline__ void sts_pair(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
// NX consecutive values from a 16-byte aligned row, 128 bits per load
template <int NX>
__device__ __forceinline__ void ldrow(const double* p, double* out) {
#pragma unroll
    for (int k = 0; k < NX; k += 2) lds128(p + k, out[k], out[k + 1]);
}
template <int NX>
__device__ __forceinline__ void ldrow(const float* p, float* out) {
    static_assert(NX % 4 == 0, "float rows are read four values at a time");
#pragma unroll
    for (int k = 0; k < NX; k += 4) {
        const float4 t = *reinterpret_cast<const float4*>(p + k);
        out[k] = t.x; out[k + 1] = t.y; out[k + 2] = t.z; out[k + 3] = t.w;
    }
}

template <class T, int NX, int NU, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1)
admm_subwarp_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st,
                    const __grid_constant__ SubwarpParams<T> sp) {
    static_assert(NX % 4 == 0 && NX >= 4, "state rows are dealt 4 lanes x RX rows");
    static_assert(NU >= 1 && NU <= 4, "one input row per lane");
    constexpr int RX = NX / 4;
    constexpr int XP = Pad<T, NX>::value;
    using R = Ar<T>;
    using CL = ConstLayout<NX, NU>;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* const smem = reinterpret_cast<T*>(smem_raw);
    const int N = bv.N;
    const int nxN = NX * N, nuN = NU * (N - 1);
    const size_t ld = static_cast<size_t>(bv.ld);
    const T rho = static_cast<T>(bv.rho);

    // ---------------- stage the shared constants once per block
    T* sc = smem + sp.off_consts;
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
    T* s_xmin = smem + sp.off_bounds;            // shared (stride-0) bounds staged once
    T* s_xmax = s_xmin + nxN;
    T* s_umin = s_xmax + nxN;
    T* s_umax = s_umin + nuN;
    if (sp.shared_bounds) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) { s_xmin[e] = bv.x_min[e]; s_xmax[e] = bv.x_max[e]; }
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) { s_umin[e] = bv.u_min[e]; s_umax[e] = bv.u_max[e]; }
    }
    T* s_xref = smem + sp.off_refs;
    T* s_uref = s_xref + nxN;
    if (sp.shared_refs) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) s_xref[e] = has_xref ? bv.Xref[e] : T(0);
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) s_uref[e] = has_uref ? bv.Uref[e] : T(0);
    }
    __syncthreads();

    // ---------------- per-warp carve-up
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inst = lane & 7, sub = lane >> 3;
    const bool urow = sub < NU;                       // this lane owns an input row
    T* wbase = smem + sp.off_warp0 + static_cast<size_t>(warp) * sp.warp_stride;
    T* s_gv = wbase;                             // (g, v) pairs: one load / store moves both
    T* s_yz = s_gv + 2 * nxN * kTile;            // (y, z) pairs
    T* s_d = s_yz + 2 * nuN * kTile;
    T* s_bc = s_d + nuN * kTile;                 // 2 x kTile x XP broadcast slots
    // state entry (column i, owned row r):  [(i*NX + r*4)*8 + lane];  input entry (column i): [i*NU*8 + lane]
#define SX(i, r) (2 * (((i) * NX + (r) * 4) * kTile + lane))
#define SU(i) (2 * ((i) * NU * kTile + lane))
#define SD(i) ((i) * NU * kTile + lane)

    const T fr_BPf = urow ? sc[CL::BPf + sub] : T(0);
    const T fr_Rd = urow ? sc[CL::Rd + sub] : T(0);
    T fr_f[RX], fr_APf[RX], fr_Qd[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { fr_f[r] = sc[CL::f + r * 4 + sub]; fr_APf[r] = sc[CL::APf + r * 4 + sub]; fr_Qd[r] = sc[CL::Qd + r * 4 + sub]; }

    int bcast_phase = 0;

    // ---------------- per-instance refill: a sub-warp whose instance has finished writes it out and pulls the next
    // instance id from the global queue at once, so a warp never idles on the slowest of its 8 instances (iteration
    // counts are bimodal on hard workloads).  Only the 4 lanes of a finished instance run the pass (shared memory needs
    // no warp-collective access); the sweeps carry no "finished" predication, parked lanes (queue empty) just own no
    // instance (valid = false) and store nothing.
    // Refill policy.  TILE = false: per instance, as described.  TILE = true: the warp waits until all its 8 instances have
    // finished and takes 8 CONSECUTIVE instances: stores to the element-major carried-state arrays then cover 8
    // adjacent entries per row instead of one (whole sectors), which matters more than the idle lanes when the
    // write-out is a large part of a short solve.  Measured on B200, config 3, B = 65,536: double easy 3.67 ms (tile) vs
    // 4.10 ms (per instance), hard 45.0 vs 38.0 ms; float easy 1.337 vs 1.308 ms, hard 15.8 vs 11.5 ms.
    constexpr bool TILE = sizeof(T) == 8;
    int b = 0, iter = 0, my_status = 1;
    bool valid = false, fin = true, first = true;
    size_t bb = 0;
    T x0r[RX], p_term[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { x0r[r] = T(0); p_term[r] = T(0); }
    // per-instance references / bounds (only dereferenced when !shared_refs && has_xref, !shared_bounds && box_x):
    // addresses are formed where they are used, from the instance index and the constant-bank strides
#define xref_g (bv.Xref + bb * bv.xref_stride)
#define uref_g (bv.Uref + bb * bv.uref_stride)
#define xmin_g (bv.x_min + bb * bv.xb_stride)
#define xmax_g (bv.x_max + bb * bv.xb_stride)
#define umin_g (bv.u_min + bb * bv.ub_stride)
#define umax_g (bv.u_max + bb * bv.ub_stride)

    for (;;) {
        const unsigned finmask = __ballot_sync(kFull, fin);
        const bool pass = TILE ? (__ballot_sync(kFull, fin || !valid) == kFull) : fin;   // TILE: every instance of the warp finished (or parked)
        if (pass) {
            // ---------------- results and carried state of the finished instance back to HBM (each lane writes the entries it owns)
            if (valid) {
                double* Xs = bv.Xsol + bb * nxN;
                double* Us = bv.Usol + bb * nuN;
                const bool unconverged = (my_status != 0);
                for (int i = 0; i < N; ++i) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const int e = i * NX + r * 4 + sub;
                        const size_t ge = static_cast<size_t>(e) * ld + bb;
                        T vv, gg;
                        lds128(s_gv + SX(i, r), gg, vv);
                        Xs[e] = vv;                                  // solution = vnew of the last iteration (admm.cpp:370,386)
                        reinterpret_cast<T*>(bv.g)[ge] = gg;
                        if (unconverged) reinterpret_cast<T*>(bv.v[0])[ge] = vv;   // v = vnew happened (admm.cpp:379); else the lagging copy is already there
                        if (i < N - 1) {                             // q as update_linear_cost left it (admm.cpp:218-219)
                            const T xr = sp.shared_refs ? s_xref[e] : (has_xref ? __ldg(xref_g + e) : T(0));
                            reinterpret_cast<T*>(bv.q)[ge] = R::sub(-R::mul(xr, fr_Qd[r]), R::mul(rho, R::sub(vv, gg)));
                        }
                    }
                }
                if (urow) for (int i = 0; i < N - 1; ++i) {
                    const int e = i * NU + sub;
                    const size_t ge = static_cast<size_t>(e) * ld + bb;
                    T zz, yy;
                    lds128(s_yz + SU(i), yy, zz);
                    Us[e] = zz;
                    reinterpret_cast<T*>(bv.y)[ge] = yy;
                    if (unconverged) reinterpret_cast<T*>(bv.z[0])[ge] = zz;
                    const T ur = sp.shared_refs ? s_uref[e] : (has_uref ? __ldg(uref_g + e) : T(0));
                    reinterpret_cast<T*>(bv.r)[ge] = R::sub(-R::mul(ur, fr_Rd), R::mul(rho, R::sub(zz, yy)));
                }
#pragma unroll
                for (int r = 0; r < RX; ++r) reinterpret_cast<T*>(bv.pN)[(r * 4 + sub) * ld + bb] = p_term[r];
                if (sub == 0) { bv.status[b] = my_status; bv.iters[b] = iter; }
            }
            // ---------------- next instance from the queue
            int nb = 0;
            if (TILE) {
                if (lane == 0) nb = atomicAdd(sp.counter, kTile);
                nb = __shfl_sync(kFull, nb, 0) + inst;
            } else {
                if (sub == 0) nb = atomicAdd(sp.counter, 1);
                nb = __shfl_sync(finmask, nb, inst);
            }
            valid = nb < bv.B;
            b = valid ? nb : 0;
            bb = static_cast<size_t>(b);

            // ---------------- carried state -> shared memory (each lane moves exactly the entries it owns)
#pragma unroll
            for (int r = 0; r < RX; ++r) {
                x0r[r] = valid ? static_cast<T>(bv.x0[bb * NX + r * 4 + sub]) : T(0);
                p_term[r] = (valid && !sp.cold) ? reinterpret_cast<T*>(bv.pN)[(r * 4 + sub) * ld + bb] : T(0);
            }
            if (sp.cold || !valid) {
                for (int i = 0; i < N; ++i) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) sts_pair(s_gv + SX(i, r), T(0), T(0));
                }
                if (urow) for (int i = 0; i < N - 1; ++i) { sts_pair(s_yz + SU(i), T(0), T(0)); s_d[SD(i)] = T(0); }
            } else {
                for (int i = 0; i < N; ++i) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const size_t ge = static_cast<size_t>(i * NX + r * 4 + sub) * ld + bb;
                        sts_pair(s_gv + SX(i, r), reinterpret_cast<T*>(bv.g)[ge], reinterpret_cast<T*>(bv.v[0])[ge]);
                    }
                }
                if (urow) for (int i = 0; i < N - 1; ++i) {
                    const size_t ge = static_cast<size_t>(i * NU + sub) * ld + bb;
                    sts_pair(s_yz + SU(i), reinterpret_cast<T*>(bv.y)[ge], reinterpret_cast<T*>(bv.z[0])[ge]); s_d[SD(i)] = T(0);
                }
            }
            iter = 0; my_status = 1; first = true; fin = false;             // a parked lane (queue empty) owns no instance: valid = false
        }
        __syncwarp();
        if (!__any_sync(kFull, valid)) break;
        const bool live = valid && !fin;     // owns an unfinished instance (TILE: finished ones wait here without mutating their state)

        {
            const bool may_converge = ((iter + 1) % st.check_termination) == 0;

            // ======================= backward sweep: d_i, p_i for i = N-2 .. 0  (reference admm.cpp:13-20)
            {
                T Mab[RX][NX], Mkt[RX][NU], Mbt[NX], Mq[NU];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    ldrow<NX>(sc + CL::AmBKt_rm + (r * 4 + sub) * NX, Mab[r]);
#pragma unroll
                    for (int k = 0; k < NU; ++k) Mkt[r][k] = sc[CL::Kinf_cm + (r * 4 + sub) * NU + k];
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) Mbt[k] = urow ? sc[CL::B_cm + sub * NX + k] : T(0);
#pragma unroll
                for (int k = 0; k < NU; ++k) Mq[k] = urow ? sc[CL::Quu_rm + sub * NU + k] : T(0);

                T p[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) p[r] = p_term[r];

                for (int i = N - 2; i >= 0; --i) {
                    // broadcast p_{i+1} to the 4 lanes of the instance
                    T* slot = s_bc + (bcast_phase & 1) * (kTile * XP) + inst * XP;
                    ++bcast_phase;
#pragma unroll
                    for (int r = 0; r < RX; ++r) slot[r * 4 + sub] = p[r];
                    __syncwarp();
                    T pb[NX];
                    ldrow<NX>(slot, pb);

                    // r_i (own row) and q_i (own rows): carried values on the first sweep of a solve,
                    // afterwards recomputed from the slacks and duals (reference admm.cpp:218-233)
                    T r_own = T(0), q_own[RX];
                    {
                        const bool carried = first && valid && !sp.cold;      // first sweep of a warm-started solve: the stored q, r
                        T r_car = T(0), q_car[RX];
#pragma unroll
                        for (int r = 0; r < RX; ++r) q_car[r] = T(0);
                        if (carried) {
                            if (urow) r_car = reinterpret_cast<T*>(bv.r)[static_cast<size_t>(i * NU + sub) * ld + bb];
#pragma unroll
                            for (int r = 0; r < RX; ++r) q_car[r] = reinterpret_cast<T*>(bv.q)[static_cast<size_t>(i * NX + r * 4 + sub) * ld + bb];
                        }
                        if (urow) {
                            const int e = i * NU + sub;
                            const T ur = sp.shared_refs ? s_uref[e] : (has_uref ? __ldg(uref_g + e) : T(0));
                            T yy, zz;
                            lds128(s_yz + SU(i), yy, zz);
                            r_own = R::sub(-R::mul(ur, fr_Rd), R::mul(rho, R::sub(zz, yy)));
                        }
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            const int e = i * NX + r * 4 + sub;
                            const T xr = sp.shared_refs ? s_xref[e] : (has_xref ? __ldg(xref_g + e) : T(0));
                            T gg, vv;
                            lds128(s_gv + SX(i, r), gg, vv);
                            q_own[r] = R::sub(-R::mul(xr, fr_Qd[r]), R::mul(rho, R::sub(vv, gg)));
                        }
                        // first sweep: the carried values (zeros after setup / reset: admm.cpp sees q = r = p = 0 whatever Xref is)
                        if (first) {
                            r_own = r_car;
#pragma unroll
                            for (int r = 0; r < RX; ++r) q_own[r] = q_car[r];
                        }
                    }

                    // d_i = Quu_inv (B' p_{i+1} + r_i + BPf)
                    T s = T(0);
#pragma unroll
                    for (int k = 0; k < NX; ++k) s = R::fma(Mbt[k], pb[k], s);
                    const T t_own = R::add(R::add(s, r_own), fr_BPf);
                    T dsum = T(0), rr[NU];
#pragma unroll
                    for (int k = 0; k < NU; ++k) {
                        dsum = R::fma(Mq[k], shfl_d(t_own, k * kTile + inst), dsum);
                        rr[k] = shfl_d(r_own, k * kTile + inst);
                    }
                    if (urow) s_d[SD(i)] = dsum;

                    // p_i = q_i + AmBKt p_{i+1} - Kinf' r_i + APf
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        T s1 = T(0), s2 = T(0);
#pragma unroll
                        for (int k = 0; k < NX; ++k) s1 = R::fma(Mab[r][k], pb[k], s1);
#pragma unroll
                        for (int k = 0; k < NU; ++k) s2 = R::fma(Mkt[r][k], rr[k], s2);
                        p[r] = R::add(R::sub(R::add(q_own[r], s1), s2), fr_APf[r]);
                    }
                }
            }

            // ======================= forward sweep fused with slack / dual / residual updates
            bool ok = true;          // every residual this lane has seen is under its tolerance
            {
                T Ma[RX][NX], Mb[RX][NU], Mk[NX];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    ldrow<NX>(sc + CL::A_rm + (r * 4 + sub) * NX, Ma[r]);
#pragma unroll
                    for (int k = 0; k < NU; ++k) Mb[r][k] = sc[CL::B_rm + (r * 4 + sub) * NU + k];
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) Mk[k] = urow ? sc[CL::Kinf_rm + sub * NX + k] : T(0);

                T x[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) x[r] = x0r[r];

                for (int i = 0; i < N; ++i) {
                    const bool terminal = (i == N - 1);
                    T xb[NX], ub[NU];
                    if (!terminal) {
                        T* slot = s_bc + (bcast_phase & 1) * (kTile * XP) + inst * XP;
                        ++bcast_phase;
#pragma unroll
                        for (int r = 0; r < RX; ++r) slot[r * 4 + sub] = x[r];
                        __syncwarp();
                        ldrow<NX>(slot, xb);

                        // u_i = -Kinf x_i - d_i ; input slack, dual, residuals (admm.cpp:28, :87, :97, :187, :259-260)
                        T s = T(0);
#pragma unroll
                        for (int k = 0; k < NX; ++k) s = R::fma(Mk[k], xb[k], s);
                        T u_own = T(0);
                        if (urow) {
                            const int e = i * NU + sub;
                            u_own = R::sub(-s, s_d[SD(i)]);
                            T yj, zold;
                            lds128(s_yz + SU(i), yj, zold);
                            const T t = R::add(u_own, yj);
                            T znew = t;
                            if (box_u) {
                                const T lo = sp.shared_bounds ? s_umin[e] : __ldg(umin_g + e);
                                const T hi = sp.shared_bounds ? s_umax[e] : __ldg(umax_g + e);
                                znew = R::clamp(t, lo, hi);
                            }
                            ok = ok && (R::abs(R::sub(u_own, znew)) < sp.pri_thresh) && (R::abs(R::sub(zold, znew)) <= sp.dua_thresh);
                            if (live) {
                                sts_pair(s_yz + SU(i), R::sub(t, znew), znew);   // y + u - znew  (y + u == u + y bit for bit)
                                if (may_converge) reinterpret_cast<T*>(bv.z[0])[static_cast<size_t>(e) * ld + bb] = zold;   // carried z lags one iteration
                            }
                        }
#pragma unroll
                        for (int k = 0; k < NU; ++k) ub[k] = shfl_d(u_own, k * kTile + inst);
                    }

                    // state column i: slack, dual, residuals (admm.cpp:84, :92, :184, :257-258)
                    T vg[RX];     // vnew - g(new), feeds the terminal cost
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const int e = i * NX + r * 4 + sub;
                        T gj, vold;
                        lds128(s_gv + SX(i, r), gj, vold);
                        const T t = R::add(x[r], gj);
                        T vnew = t;
                        if (box_x) {
                            const T lo = sp.shared_bounds ? s_xmin[e] : __ldg(xmin_g + e);
                            const T hi = sp.shared_bounds ? s_xmax[e] : __ldg(xmax_g + e);
                            vnew = R::clamp(t, lo, hi);
                        }
                        const T gnew = R::sub(t, vnew);
                        ok = ok && (R::abs(R::sub(x[r], vnew)) < sp.pri_thresh) && (R::abs(R::sub(vold, vnew)) <= sp.dua_thresh);
                        if (live) {
                            sts_pair(s_gv + SX(i, r), gnew, vnew);
                            if (may_converge) reinterpret_cast<T*>(bv.v[0])[static_cast<size_t>(e) * ld + bb] = vold;       // carried v lags one iteration
                        }
                        vg[r] = R::sub(vnew, gnew);
                    }

                    if (!terminal) {
                        // x_{i+1} = A x_i + B u_i + f
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            T s1 = T(0), s2 = T(0);
#pragma unroll
                            for (int k = 0; k < NX; ++k) s1 = R::fma(Ma[r][k], xb[k], s1);
#pragma unroll
                            for (int k = 0; k < NU; ++k) s2 = R::fma(Mb[r][k], ub[k], s2);
                            x[r] = R::add(R::add(s1, s2), fr_f[r]);
                        }
                    } else if (live) {
                        // p_{N-1} = -(Xref_{N-1}' Pinf)' - rho (vnew - g)   (admm.cpp:238-239)
#pragma unroll
                        for (int r = 0; r < RX; ++r) {
                            T s = T(0);
                            if (has_xref) {
                                for (int k = 0; k < NX; ++k) {
                                    const T xr = sp.shared_refs ? s_xref[i * NX + k] : __ldg(xref_g + i * NX + k);
                                    s = R::fma(xr, sc[CL::Pinf_cm + (r * 4 + sub) * NX + k], s);
                                }
                            }
                            p_term[r] = R::sub(-s, R::mul(rho, vg[r]));
                        }
                    }
                }
            }

            // ======================= termination (admm.cpp:253-271, :364-380), per instance
            if (live) iter += 1;
            const unsigned votes = __ballot_sync(kFull, ok);
            const unsigned mine = 0x01010101u << inst;
            const bool inst_ok = ((iter % st.check_termination) == 0) && ((votes & mine) == mine);
            if (live) {
                if (inst_ok) { fin = true; my_status = 0; }
                else if (iter >= st.max_iter) { fin = true; my_status = 1; }
            }
            if (live) first = false;
        }
    }
#undef xref_g
#undef uref_g
#undef xmin_g
#undef xmax_g
#undef umin_g
#undef umax_g
#undef SX
#undef SU
#undef SD
}

// Largest a >= 0 with fl(rho * a) < tol (in T), so that  rho * max|.| < tol  <=>  every |.| <= a.
template <class T>
T dual_threshold(T rho, T tol) {
    if (!(tol > T(0)) || !(rho > T(0))) return T(-1);
    T a = tol / rho;
    auto prod = [&](T x) { volatile T pr = rho * x; return static_cast<T>(pr); };   // one rounding to T, no excess precision
    while (prod(a) < tol) a = std::nextafter(a, static_cast<T>(INFINITY));
    while (!(prod(a) < tol)) a = std::nextafter(a, static_cast<T>(-INFINITY));
    return a;
}

template <class T> struct Plan { int warps; size_t smem_bytes; SubwarpParams<T> sp; };

template <class T, int NX, int NU, int MAXW>
bool make_plan(const BatchView& bv, Plan<T>& pl) {
    constexpr int XP = Pad<T, NX>::value;
    const int N = bv.N, nxN = NX * N, nuN = NU * (N - 1);
    SubwarpParams<T> sp{};
    int off = 0;
    sp.off_consts = off; off += (ConstLayout<NX, NU>::size + 3) / 4 * 4;
    sp.shared_bounds = (bv.x_min != nullptr && bv.xb_stride == 0 && bv.ub_stride == 0) ? 1 : 0;
    sp.off_bounds = off; if (sp.shared_bounds) off += 2 * nxN + 2 * nuN;
    sp.shared_refs = ((bv.Xref == nullptr || bv.xref_stride == 0) && (bv.Uref == nullptr || bv.uref_stride == 0)) ? 1 : 0;
    sp.off_refs = off; if (sp.shared_refs) off += nxN + nuN;
    off = (off + 3) / 4 * 4;
    sp.off_warp0 = off;
    int per_warp = (2 * nxN + 3 * nuN) * kTile + 2 * kTile * XP;
    per_warp = (per_warp + 3) / 4 * 4;
    sp.warp_stride = per_warp;
    const size_t limit = 227 * 1024;
    const size_t fixed = static_cast<size_t>(off) * sizeof(T);
    if (fixed + static_cast<size_t>(per_warp) * sizeof(T) > limit) return false;
    int warps = static_cast<int>((limit - fixed) / (static_cast<size_t>(per_warp) * sizeof(T)));
    if (warps > MAXW) warps = MAXW;
    pl.warps = warps;
    pl.smem_bytes = fixed + static_cast<size_t>(warps) * per_warp * sizeof(T);
    sp.warps_per_block = warps;
    pl.sp = sp;
    return true;
}

template <class T, int NX, int NU, int MAXW>
cudaError_t launch(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms, cudaStream_t stream) {
    Plan<T> pl;
    if (!make_plan<T, NX, NU, MAXW>(bv, pl)) return cudaErrorInvalidConfiguration;
    pl.sp.cold = cold;
    pl.sp.num_tiles = (bv.B + kTile - 1) / kTile;      // grid sizing only: the queue hands out instances
    pl.sp.counter = counter;
    pl.sp.pri_thresh = static_cast<T>(st.abs_pri_tol);
    pl.sp.dua_thresh = dual_threshold<T>(static_cast<T>(bv.rho), static_cast<T>(st.abs_dua_tol));
    auto kern = admm_subwarp_kernel<T, NX, NU, MAXW>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.smem_bytes));
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    int grid = (pl.sp.num_tiles + pl.warps - 1) / pl.warps;
    if (grid > num_sms) grid = num_sms;            // persistent: one CTA per SM, warps pull tiles from the queue
    kern<<<grid, pl.warps * 32, pl.smem_bytes, stream>>>(bv, st, pl.sp);
    return cudaGetLastError();
}

template <class T, int NX, int NU, int MAXW>
bool fits(int N) {
    BatchView v{};
    v.N = N;
    Plan<T> pl;
    return make_plan<T, NX, NU, MAXW>(v, pl);
}

}  // namespace

// float: shared memory holds twice the instances, so the resident warps double (8 -> 16 for the quadrotor shape)
bool subwarp_kernel_supports(int nx, int nu, int N, const AdmmSettings& st, int precision) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (precision == 32) {
        if (nx == 12 && nu == 4) return fits<float, 12, 4, 8>(N);
        if (nx == 4 && nu == 1) return fits<float, 4, 1, 16>(N);
        return false;
    }
    if (nx == 12 && nu == 4) return fits<double, 12, 4, 8>(N);
    if (nx == 4 && nu == 1) return fits<double, 4, 1, 16>(N);
    return false;
}

cudaError_t launch_admm_subwarp(const BatchView& bv, const AdmmSettings& st, const HostMats& h, int cold, int* counter,
                                int num_sms, cudaStream_t stream, int precision) {
    if (bv.B <= 0) return cudaSuccess;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return launch_admm_thread(bv, st, h, stream, precision);   // degenerate settings: general kernel
    if (precision == 32) {
        if (bv.nx == 12 && bv.nu == 4) return launch<float, 12, 4, 8>(bv, st, cold, counter, num_sms, stream);
        if (bv.nx == 4 && bv.nu == 1) return launch<float, 4, 1, 16>(bv, st, cold, counter, num_sms, stream);
        return cudaErrorNotSupported;
    }
    if (bv.nx == 12 && bv.nu == 4) return launch<double, 12, 4, 8>(bv, st, cold, counter, num_sms, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch<double, 4, 1, 16>(bv, st, cold, counter, num_sms, stream);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
