// ADMM solve kernel v2: "one instance per 4-lane sub-warp, duals and slacks split between Tensor Memory and
// shared memory, per-instance refill".
//
// Why it exists (measured, profiles/r01_a_*): the shared-memory-only sub-warp kernel (admm_subwarp.cu) holds 708
// doubles per instance in shared memory, which caps an SM at 32 resident instances = 4 warps = ONE warp per
// scheduler; every DFMA dependency chain is then exposed (issue active 22 %, FP64 pipe 11.6 %).  Blackwell's
// 256 KB of Tensor Memory per SM is idle in a kernel without MMAs, and tcgen05.ld/st give every lane private
// access to its own TMEM row (measured: 200-256 B/clk/SM loads, 177-212 B/clk/SM stores, st->ld round trip
// 34 cycles, profiles/r01_tmem_probe.jsonl).  This kernel keeps the state-row duals and slacks (g, v: 480 of the
// 708 doubles) in TMEM and only the input-row arrays (y, z, r/t) in shared memory, which doubles the residency to
// 8 warps per SM = two warps per scheduler, and it
//   * refills a sub-warp with the next instance the moment its instance finishes (global instance queue), so a
//     warp never idles on the slowest of its 8 instances (iteration counts are bimodal on the hard workload);
//   * moves the Quu_inv mat-vec into the forward sweep (the backward sweep leaves t = B'p + r + BPf in shared
//     memory, the forward sweep reads the 4-vector back and forms d), and lets the forward sweep leave the new
//     r in the same slot, so the backward sweep needs no warp shuffles at all;
//   * drops all "finished instance" predication from the sweeps: a finished instance is written out and replaced.
//
// Lane map: lane = sub * 8 + inst; lane `sub` of instance `inst` owns state rows {4 r + sub} and input row `sub`.
// TMEM map: warp w uses TMEM lanes 32 (w % 4) .. +31 (the only ones it may touch) and columns 256 (w / 4) .. +255;
// lane-private column 12 i + {0..5} = g(i, rows), {6..11} = v(i, rows) as 32-bit halves of the doubles.
//
// Arithmetic is the same sequence of correctly rounded operations as admm_subwarp.cu / admm_thread.cu, so all
// kernels agree bit for bit on the carried state (tests switch kernels in the middle of a closed loop).
//
// The kernel is generic in the scalar T = the reference's `tinytype` (types.hpp:15): double (default build) or float.
// A value is two Tensor-Memory words as a double and one as a float, shared-memory byte offsets are multiples of
// sizeof(T), arithmetic comes from Ar<T>; the float instantiation runs 12 warps per SM.
//
// Reference being reproduced: tinympc/TinyMPC/src/tinympc/admm.cpp:13-32 (Riccati sweeps), :81-98 (box slack),
// :181-187 (dual), :214-247 (linear cost), :253-271 (termination), :274-389 (solve).
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <type_traits>

#include "admm_device.cuh"
#include "admm_kernels.hpp"
#include "tmem_util.cuh"

namespace tmpc {
namespace {

using namespace tmu;

// backward sweep: all-gather p by warp shuffles instead of STS -> LDS.  Measured on B200: +1.7 % for the cartpole
// shape (one row per lane: 4 shuffles replace 1 store + 2 loads), -0.7 % for the quadrotor shape (24 SHFL vs 9
// shared-memory instructions per knot point), so it is on for NX = 4 only.
#ifndef TMPC_FENCE_A
#define TMPC_FENCE_A 0
#endif
#ifndef TMPC_FENCE_C
#define TMPC_FENCE_C 0
#endif
#ifndef TMPC_BWD_SHFL_MAX_NX
#define TMPC_BWD_SHFL_MAX_NX 4
#endif

#ifndef TMPC_QWARPS
#define TMPC_QWARPS 8
#endif
#ifndef TMPC_QWARPS_F32
#define TMPC_QWARPS_F32 12
#endif
// warps per CTA (one CTA per SM).  Quadrotor shape: two per scheduler (238 registers per thread, 240 of 256 Tensor-Memory
// columns and 16.4 KB of shared memory per warp).  Cartpole shape (one state row per lane: 139 registers, 80 TMEM columns,
// 15.4 KB per warp): three per scheduler - the kernel is latency-bound (ncu: 2.0 fixed-latency stall cycles per issue
// at 8 warps, profiles/r02_a_ncu_full_cartpole_summary.csv).
// Float build (tinytype = float): half the shared memory and Tensor-Memory columns per warp and one register per value:
// 12 warps (three per scheduler, <= 170 registers per thread; measured 3.5 % faster than 16 warps at 128 registers).
template <class T, int NX, bool PI = false> struct QWarps {
    static constexpr int value = sizeof(T) == 4 ? TMPC_QWARPS_F32 : ((NX == 4 && !PI) ? 12 : TMPC_QWARPS);   // per-instance operands: + 4.6 KB of shared memory per warp
};
constexpr int kQInst = 8;            // instances per warp
template <class T, int NX, bool PI = false> struct QTmemCols { static constexpr int value = (512 / ((QWarps<T, NX, PI>::value + 3) / 4)) & ~1; };   // columns per warp

struct QuadParams {
    double pri_thresh;      // primal residual passes  <=>  |x - vnew| <  pri_thresh
    double dua_thresh;      // dual residual passes    <=>  |v - vnew| <= dua_thresh  (largest a with fl(rho a) < tol)
    int* counter;           // instance queue: ticket counter ...
    const int* order;       // ... and the permutation the tickets index (longest-first, queue_order.cu); null = identity
    int off_consts, off_bounds, off_refs, off_warp0, warp_stride;   // shared-memory carve-up, in doubles
    // fused closed loop (ROLL): every instance runs `nsteps` MPC steps on the sub-warp that owns it
    int nsteps;             // MPC steps per instance
    int cold_install;       // the batch was reset: instances start from zeros instead of their carried block
    int slide;              // the reference window (per-instance Xref, REFS = 2) moves one column per step
    double* x0_rw;          // plant state, nx x B: read at install, written back after the last step
    int* status_out;        // [nsteps][B] per-step solve status / iteration count (may be null)
    int* iters_out;
    // per-instance dynamics (PI): one QConst-layout block of constants per instance in global memory
    const double* pi_blocks;
    int off_pi;             // per-warp offset (doubles) of the lanes' private B / Quu_inv row slots
};


// ptxas schedules a basic block by critical-path priority and lets the chain that feeds the shuffles run ahead
// serially (one DFMA per 8 cycles) instead of interleaving it with the independent chains.  A never-taken uniform
// branch ends the basic block, so inside it all chains have the same distance to the end and are issued round-robin.
#define SCHED_FENCE() do { if (st.max_iter == -12345) asm volatile("trap;"); } while (0)

// REFS: 0 = no references (Xref = Uref = 0), 1 = references shared by all instances (staged in shared memory as
// the products the linear cost needs), 2 = per-instance references read through the read-only path.
// WARPS != 0 overrides the warps per CTA of the shape (QWarps): the long-horizon form runs FOUR warps per SM, each with all
// 512 Tensor-Memory columns of its lane quarter, which takes the horizon of the quadrotor shape from N <= 21 to N <= 42 at
// one warp per scheduler.
template <class T, int NX, int NU, bool COLD, int REFS, bool ROLL = false, bool PI = false, int WARPS = 0>
__global__ void __launch_bounds__((WARPS ? WARPS : QWarps<T, NX, PI>::value) * 32, 1)
admm_tmem_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st,
                 const __grid_constant__ QuadParams qp) {
    // A global access in uniform control flow before anything else: ptxas then keeps the global-memory descriptor in a
    // uniform register for the whole kernel.  Without it the first global access sits in a divergent region (the
    // staging loop), the descriptor lands in a vector register pair and EVERY load/store pays two R2UR (7 % of all
    // instructions issued).  The queue counter never reaches the value tested.
    if (*reinterpret_cast<const volatile int*>(qp.counter) == 0x7ffffff0) return;
    static_assert(NX % 4 == 0 && NX >= 4, "state rows are dealt 4 lanes x RX rows");
    static_assert(NU >= 1 && NU <= 4, "one input row per lane");
    constexpr int RX = NX / 4;
    constexpr int XP = QPad<NX, sizeof(T)>::value;
    using R = Ar<T>;                                  // arithmetic at the reference's tinytype (types.hpp:15)
    constexpr int S = sizeof(T);                      // bytes per value: every shared-memory byte offset below is a multiple
    constexpr int W = 2 * RX * (S / 4);               // 32-bit TMEM words per knot point per lane: g(RX), v(RX)
    static_assert(W == 12 || W == 4 || W == 6 || W == 2, "tm_wait_ld is specialised for 3 or 1 state rows per lane");
    static_assert(S == 8 || (!ROLL && !PI), "the fused closed loop and per-instance dynamics are built for 64-bit only");
    using CL = QConst<NX, NU>;
    using PL = QLane<NX, NU>;
#ifndef TMPC_CREG
#define TMPC_CREG 1
#endif
    constexpr bool CREG = TMPC_CREG != 0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* const smem = reinterpret_cast<T*>(smem_raw);
    __shared__ uint32_t s_tmem_base;
    __shared__ uint32_t s_smem_base;
    const int N = bv.N;
    const int nxN = NX * N, nuN = NU * (N - 1);
    const size_t ld = static_cast<size_t>(bv.ld);
    const T rho = static_cast<T>(bv.rho), nrho = -rho;      // bindings.cpp:55 casts rho to tinytype
    const T pri_thresh = static_cast<T>(qp.pri_thresh), dua_thresh = static_cast<T>(qp.dua_thresh);

    // ---------------- Tensor Memory: the whole 512-column file of this SM (one CTA per SM)
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "l"(static_cast<uint64_t>(__cvta_generic_to_shared(&s_tmem_base))) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }

    // ---------------- stage the shared constants once per CTA
    T* sc = smem + qp.off_consts;
    static_assert(!PI || REFS != 1, "per-instance dynamics take references, shared or not, through the per-instance path");
    if (!PI)
    for (int idx = threadIdx.x; idx < NX * NX; idx += blockDim.x) {
        const int row = idx / NX, k = idx % NX;
        sc[CL::A_rm + idx] = bv.A[k * NX + row];
        sc[CL::AmBKt_rm + idx] = bv.AmBKt[k * NX + row];
        sc[CL::Pinf_cm + idx] = bv.Pinf[idx];
    }
    if (!PI)
    for (int idx = threadIdx.x; idx < NX * NU; idx += blockDim.x) {
        const int row = idx / NU, k = idx % NU;          // B(row, k)
        sc[CL::B_rm + idx] = bv.Bm[k * NX + row];
        sc[CL::B_cm + idx] = bv.Bm[idx];
        const int j = idx / NX, kk = idx % NX;           // Kinf(j, kk)
        sc[CL::Kinf_rm + idx] = bv.Kinf[kk * NU + j];
        sc[CL::Kinf_cm + idx] = bv.Kinf[idx];
    }
    if (!PI) {
    for (int idx = threadIdx.x; idx < NU * NU; idx += blockDim.x)
        sc[CL::Quu_rm + idx] = bv.Quu_inv[(idx % NU) * NU + idx / NU];
    for (int idx = threadIdx.x; idx < NX; idx += blockDim.x) {
        sc[CL::f + idx] = bv.f[idx]; sc[CL::APf + idx] = bv.APf[idx]; sc[CL::Qd + idx] = bv.Qd[idx];
    }
    for (int idx = threadIdx.x; idx < NU; idx += blockDim.x) { sc[CL::BPf + idx] = bv.BPf[idx]; sc[CL::Rd + idx] = bv.Rd[idx]; }
    }

    // box bounds as (lo, hi) pairs; a disabled side is (-inf, +inf), for which the clamp is the identity
    T* s_xb = smem + qp.off_bounds;              // [N][NX][2]
    T* s_ub = s_xb + 2 * nxN;                    // [N-1][NU][2]
    {
        const bool bx = st.en_state_bound && bv.x_min != nullptr, bu = st.en_input_bound && bv.u_min != nullptr;
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) {
            s_xb[2 * e] = static_cast<T>(bx ? bv.x_min[e] : -INFINITY); s_xb[2 * e + 1] = static_cast<T>(bx ? bv.x_max[e] : INFINITY);
        }
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) {
            s_ub[2 * e] = static_cast<T>(bu ? bv.u_min[e] : -INFINITY); s_ub[2 * e + 1] = static_cast<T>(bu ? bv.u_max[e] : INFINITY);
        }
    }
    // shared references, stored as the terms the linear cost uses: cq = -(Xref .* Qd), cr = -(Uref .* Rd),
    // tq = -(Xref_{N-1}' Pinf)'   (reference admm.cpp:218, :228, :238)
    T* s_cq = smem + qp.off_refs;                // [N][NX]
    T* s_cr = s_cq + nxN;                        // [N-1][NU]
    T* s_tq = s_cr + nuN;                        // [NX]
    if (REFS == 1) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) s_cq[e] = -R::mul(static_cast<T>(bv.Xref ? bv.Xref[e] : 0.0), static_cast<T>(bv.Qd[e % NX]));
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) s_cr[e] = -R::mul(static_cast<T>(bv.Uref ? bv.Uref[e] : 0.0), static_cast<T>(bv.Rd[e % NU]));
        for (int row = threadIdx.x; row < NX; row += blockDim.x) {
            T s = T(0);
            if (bv.Xref) for (int k = 0; k < NX; ++k) s = R::fma(static_cast<T>(bv.Xref[(N - 1) * NX + k]), static_cast<T>(bv.Pinf[row * NX + k]), s);
            s_tq[row] = -s;
        }
    }
    if (threadIdx.x == 0) s_smem_base = smem_addr(smem);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    // The shared-window address of smem[], read back through a volatile load: ptxas otherwise rematerialises it inside
    // the sweeps when registers are short (S2UR SR_CgaCtaId + ULEA per knot point, the S2UR latency exposed: 4 % of
    // the samples of the forward sweep).  A loaded value can only be kept.
    uint32_t sbase;
    asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(sbase) : "r"(smem_addr(&s_smem_base)) : "memory");
    auto sa = [&](const T* q) { return sbase + static_cast<uint32_t>((q - smem) * S); };

    // ---------------- per-warp carve-up
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inst = lane & 7, sub = lane >> 3;
    const bool urow = (NU == 4) || (sub < NU);        // this lane owns an input row
    const int usub = urow ? sub : 0;
    const uint32_t tbase = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) +
                           static_cast<uint32_t>((warp >> 2) * (WARPS ? ((512 / ((WARPS + 3) / 4)) & ~1) : QTmemCols<T, NX, PI>::value));
    T* wbase = smem + qp.off_warp0 + static_cast<size_t>(warp) * qp.warp_stride;
    T* s_yz = wbase;                             // [N-1][32 lanes][2] : (y, z) of the lane's input row
    T* s_rt = s_yz + (N - 1) * 64;               // [N-1][2 halves][8 inst][2] : r_i (after a forward sweep) / t_i (after a
                                                      // backward sweep); rows (0,1) and (2,3) of the 8 instances are each 128
                                                      // contiguous bytes, so the 4-vector comes back in two conflict-free LDS.128
    T* s_bc = s_rt + (N - 1) * 32;               // 2 x 8 x XP broadcast slots
    T* my_rt = s_rt + (sub >> 1) * 16 + inst * 2 + (sub & 1);   // + i * 32 : own row

    // 32-bit shared addresses of everything the sweeps touch (byte strides per knot point in the comments)
    const uint32_t a_yz = sa(s_yz) + lane * 2 * S;                            // + i * 64 S
    const uint32_t a_rt4 = sa(s_rt) + inst * 2 * S;                           // + i * 32 S ; rows 2,3: + 16 S
    const uint32_t a_myrt = sa(my_rt);                                        // + i * 32 S
    const uint32_t a_slotA = sa(s_bc) + inst * XP * S;
    const uint32_t a_slot_sum = 2 * a_slotA + kQInst * XP * S;                       // slot B = sum - slot A
    const uint32_t a_ub = sa(s_ub) + usub * 2 * S;                            // + i * NU * 2 S
    const uint32_t a_xb = sa(s_xb) + sub * 2 * S;                             // + i * NX * 2 S ; row r: + r * 8 S
    const uint32_t a_sc = sa(sc);
    // B rows / Quu_inv row read every knot point: the shared staging, or (PI) the lane's private slot of 18 doubles
    // (9 16-byte chunks: conflict-free LDS.128): [B row r = 0..RX-1 (NU each)] [Quu_inv row]
    constexpr int MBS = PI ? NU * S : 4 * NU * S;                                    // bytes between the lane's B rows
    T* s_pi = wbase + qp.off_pi + lane * 18;
    const uint32_t a_Mb = PI ? sa(s_pi) : a_sc + (CL::B_rm + sub * NU) * S;   // row r: + r * MBS
    const uint32_t a_Mq = PI ? sa(s_pi) + RX * NU * S : a_sc + (CL::Quu_rm + usub * NU) * S;
    const T* scl = sc;                           // constants of the lane's instance: shared staging, or (PI) its block + 2 sub
    // element of the constants: offset `so` in the shared staging (QConst), element `li` of the lane's array (QLane)
    auto cst = [&](int so, int li) -> const T* { return PI ? scl + PL::at(li, 0) : scl + so; };
    const uint32_t a_cq = sa(s_cq) + sub * S;                                 // + i * NX * S ; row r: + r * 4 S
    const uint32_t a_cr = sa(s_cr) + usub * S;                                // + i * NU * S
    uint32_t a_slot = a_slotA;

    T fr_BPf = T(0), fr_Rd = T(0);
    T fr_f[RX], fr_APf[RX], fr_Qd[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { fr_f[r] = T(0); fr_APf[r] = T(0); fr_Qd[r] = T(0); }
    if (!PI) {
        fr_BPf = urow ? sc[CL::BPf + usub] : T(0);
        fr_Rd = urow ? sc[CL::Rd + usub] : T(0);
#pragma unroll
        for (int r = 0; r < RX; ++r) { fr_f[r] = sc[CL::f + r * 4 + sub]; fr_APf[r] = sc[CL::APf + r * 4 + sub]; fr_Qd[r] = sc[CL::Qd + r * 4 + sub]; }
    }

    // ---------------- per-lane instance state
    int b = 0, iter = 0, status = 1;
    bool valid = false, first = true, fin = true;     // fin = true: the sub-warp wants a (new) instance
    T x0r[RX], p_term[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { x0r[r] = T(0); p_term[r] = T(0); }
    // carried state of an instance: one contiguous block (instance-major, ImLayout in admm_kernels.hpp).
    // A lane stores 8-byte chunks of VW values (one double, two floats), the chunks of the 4 lanes of an instance
    // interleaved so that a store of the sub-warp fills whole sectors.  Lane pointers carry + VW sub:
    //   knot i < N    : + XK i : chunk c at + CS c; chunks 0..GC-1 = g_0..g_{RX-1}, q_0..q_{RX-1}; chunks GC..GC+VC-1 = v
    //   knot i < N - 1: IB + UK i : y, r (+ NU), z (+ 2 NU) of the lane's input row as doubles; as floats the chunk (y, r)
    //                   and z at + 2 NU - sub                                    PB + 4 r - (VW - 1) sub: terminal p
    constexpr int VW = 8 / S, CS = 4 * VW;
    constexpr int GC = (2 * RX + VW - 1) / VW, VC = (RX + VW - 1) / VW, XK = (GC + VC) * CS;
    constexpr int UK = VW == 1 ? 3 * NU : ((3 * NU + 1) & ~1);
    const int IB = XK * N, PB = IB + UK * (N - 1);
    T* const state_im = reinterpret_cast<T*>(bv.state_im);
    T* sblk = state_im + VW * sub;               // block of the lane's instance, + VW sub
    const int zo = 2 * NU - (VW - 1) * sub;      // z of the lane's input row, from its (y, r)
    const int po = PB - (VW - 1) * sub;          // terminal p of the lane's row r: + 4 r
    const double* xref_g = nullptr;              // REFS == 2 only
    const double* uref_g = nullptr;
    // fused closed loop (ROLL): MPC step of the lane's instance; cont = the instance carries on from its previous
    // step with its state on chip; fix = that step converged, so the carried previous slacks are the lagging copies
    // parked in the instance's block, to be put in place after the first backward sweep (which needs vnew, via q)
    int step = 0;
    bool cont = false, fix = false;
    static_assert(!ROLL || !COLD, "the fused closed loop takes the cold start as a run-time flag");

    for (;;) {
        // =====================================================================================================
        // ROLL: step transition of finished solves (examples/cartpole_example_mpc.jl:35-51): record the step, drive the
        // plant with the first control of the solution, x <- (A x + B znew(:,0)) + f, and either start the next
        // solve on the state left on chip or fall through to the write-out below after the last step.
        // Warp-uniform because of the shuffles; lanes whose solve is not finished ignore the results.
        // =====================================================================================================
        if (ROLL && __any_sync(kAll, fin && valid)) {
            const bool tr = fin && valid;
            if (tr && sub == 0) {
                const size_t o = static_cast<size_t>(step) * bv.B + b;
                if (qp.status_out) qp.status_out[o] = status;
                if (qp.iters_out) qp.iters_out[o] = iter;
            }
            T zz0 = T(0);
            if (urow) zz0 = s_yz[lane * 2 + 1];           // znew of knot 0, own input row
            T xs[NX], u0[NU];
#pragma unroll
            for (int k = 0; k < NX; ++k) xs[k] = shfl_dq(x0r[k / 4], (k % 4) * kQInst + inst);
#pragma unroll
            for (int k = 0; k < NU; ++k) u0[k] = shfl_dq(zz0, k * kQInst + inst);
#pragma unroll
            for (int r = 0; r < RX; ++r) {
                T s1 = T(0), s2 = T(0);
#pragma unroll
                for (int k = 0; k < NX; ++k) s1 = R::fma(*cst(CL::A_rm + (r * 4 + sub) * NX + k, PL::A + r * NX + k), xs[k], s1);
#pragma unroll
                for (int k = 0; k < NU; ++k) s2 = R::fma(*cst(CL::B_rm + (r * 4 + sub) * NU + k, PL::B_rm + r * NU + k), u0[k], s2);
                const T xn = R::add(R::add(s1, s2), fr_f[r]);
                if (tr) x0r[r] = xn;
            }
            if (tr) {
                step += 1;
                if (step < qp.nsteps) {
                    fix = (status == 0);
                    cont = true; fin = false; iter = 0; first = true; status = 1;
                    if (REFS == 2 && qp.slide && xref_g) xref_g += NX;
                } else {
#pragma unroll
                    for (int r = 0; r < RX; ++r) qp.x0_rw[static_cast<size_t>(b) * NX + r * 4 + sub] = static_cast<double>(x0r[r]);
                }
            }
        }
        // =====================================================================================================
        // write-out of finished instances + refill from the queue.  Warp-uniform (tcgen05.ld/st are .aligned):
        // every lane walks the knot points, lanes whose instance is not finished store back what they loaded.
        // Global addresses advance by running offsets (one add per knot point instead of a multiply per store).
        // =====================================================================================================
        if (__any_sync(kAll, fin)) {
            int nb = 0x7fffffff;                          // queue ticket: issued first, consumed after the input rows are out
            if (fin && sub == 0) nb = atomicAdd(qp.counter, 1);
            const bool wr = fin && valid;                 // results to write
            const bool unconv = (status != 0);
            const size_t bb = wr ? static_cast<size_t>(b) : 0;
            T* const ob = sblk;                      // block of the finished instance (+ VW sub)
            if (wr && urow) {
                double* Us = bv.Usol + bb * nuN + sub;
                T* ou = ob + IB;                     // knot i: + UK i
                for (int i = 0; i < N - 1; ++i) {
                    T yy, zz;
                    ld2(s_yz + (i * 32 + lane) * 2, yy, zz);
                    *Us = static_cast<double>(zz);                // solution = znew of the last iteration (admm.cpp:371,387)
                    if constexpr (VW == 2) st2(ou, yy, my_rt[i * 32]);   // r as update_linear_cost left it (admm.cpp:228-233)
                    else { ou[0] = yy; ou[NU] = my_rt[i * 32]; }
                    if (unconv) ou[zo] = zz;                      // z = znew happened (admm.cpp:380); else the lagging copy is there
                    Us += NU; ou += UK;
                }
            }
            if (wr) {
#pragma unroll
                for (int r = 0; r < RX; ++r) ob[po + r * 4] = p_term[r];
                if (sub == 0) { bv.status[b] = status; bv.iters[b] = iter; }
            }
            // next instance for every finished sub-warp (the atomic was issued at the top of the pass); the ticket
            // indexes the longest-first permutation when there is one
            if (qp.order != nullptr && nb < bv.B) nb = __ldg(qp.order + nb);
            nb = __shfl_sync(kAll, nb, inst);
            const bool nvalid = fin && nb < bv.B;
            const size_t nbb = nvalid ? static_cast<size_t>(nb) : 0;
            const bool warm = !COLD && nvalid && !(ROLL && qp.cold_install);
            T* const nblk = state_im + nbb * static_cast<size_t>(bv.state_im_stride) + VW * sub;

            double* Xs = bv.Xsol + bb * nxN + sub;
            const double* xr_old = ROLL ? xref_g : ((REFS == 2 && bv.Xref) ? bv.Xref + bb * bv.xref_stride + sub : nullptr);
            T* op = ob;                              // knot i of the old / new instance: + XK i ; chunks (g, q) +0, v + CS GC
            const T* np = nblk;
            const bool wr_v = wr && unconv;
            auto walk = [&](auto anyv) {
            constexpr bool ANYV = decltype(anyv)::value;
            for (int i = 0; i < N; ++i) {
                uint32_t w[W];
                tm_load<W>(tbase + i * W, w);
                T gn[VW * VC], vn[VW * VC];              // the new instance's g (head of its (g, q) sequence) and v
#pragma unroll
                for (int r = 0; r < VW * VC; ++r) { gn[r] = T(0); vn[r] = T(0); }
                if (warm) {
#pragma unroll
                    for (int c = 0; c < VC; ++c) { ldc<VW>(np + CS * c, gn + VW * c); ldc<VW>(np + CS * (GC + c), vn + VW * c); }
                }
                tm_wait_ld<W>(w);
                T sq[VW * GC], sv[VW * VC];              // the finished instance's (g, q) sequence and v
                sq[VW * GC - 1] = T(0); sv[VW * VC - 1] = T(0);
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    // branch-free: every lane forms the values, the stores are predicated
                    const T gg = tm_get(w, r, T());
                    const T vv = tm_get(w, RX + r, T());
                    stg_if(Xs + r * 4, static_cast<double>(vv), wr);   // solution = vnew of the last iteration (admm.cpp:370,386)
                    T q;                                     // q as update_linear_cost left it (admm.cpp:218-219), all N columns
                    if (REFS == 0) q = R::mul(nrho, R::sub(vv, gg));
                    else {
                        const T cq = (REFS == 1) ? s_cq[i * NX + r * 4 + sub]
                                                      : -R::mul(static_cast<T>((xr_old && wr) ? __ldg(xr_old + i * NX + r * 4) : 0.0), fr_Qd[r]);
                        q = R::sub(cq, R::mul(rho, R::sub(vv, gg)));
                    }
                    sq[r] = gg; sq[RX + r] = q; sv[r] = vv;
                    tm_put_if(w, r, gn[r], fin);
                    tm_put_if(w, RX + r, vn[r], fin);
                }
#pragma unroll
                for (int c = 0; c < GC; ++c) stgc_if<VW>(op + CS * c, sq + VW * c, wr);
                if (ANYV) {                                  // v = vnew happened (admm.cpp:379); else the lagging copy is there
#pragma unroll
                    for (int c = 0; c < VC; ++c) stgc_if<VW>(op + CS * (GC + c), sv + VW * c, wr_v);
                }
                tm_store<W>(tbase + i * W, w);
                Xs += NX; op += XK; np += XK;
            }
            };
            // the v stores exist only in the variant that has an unconverged instance to write (none in a batch that converges:
            // a predicated-off store still costs its issue slot, 3 of the 12 per knot point)
            if (__any_sync(kAll, wr_v)) walk(std::true_type{}); else walk(std::false_type{});
            if (fin) {
                const T* nu_p = nblk + IB;
                for (int i = 0; i < N - 1; ++i) {
                    T yy = T(0), zz = T(0), rr = T(0);
                    if (warm && urow) {
                        if constexpr (VW == 2) ld2(nu_p, yy, rr); else { yy = nu_p[0]; rr = nu_p[NU]; }
                        zz = nu_p[zo];
                    }
                    st2(s_yz + (i * 32 + lane) * 2, yy, zz);
                    my_rt[i * 32] = rr;
                    nu_p += UK;
                }
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    x0r[r] = nvalid ? static_cast<T>(bv.x0[nbb * NX + r * 4 + sub]) : T(0);
                    p_term[r] = warm ? nblk[po + r * 4] : T(0);
                }
                b = nvalid ? nb : 0;
                sblk = nblk;
                valid = nvalid; iter = 0; first = true; status = 1; fin = false;
                step = 0; cont = false; fix = false;
                if (PI) {                                 // constants of the new instance: registers + the lane's B / Quu_inv slot
                    scl = reinterpret_cast<const T*>(qp.pi_blocks) + nbb * PL::size + 2 * sub;
                    fr_BPf = urow ? *cst(0, PL::BPf) : T(0);
                    fr_Rd = urow ? *cst(0, PL::Rd) : T(0);
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        fr_f[r] = *cst(0, PL::f + r); fr_APf[r] = *cst(0, PL::APf + r); fr_Qd[r] = *cst(0, PL::Qd + r);
#pragma unroll
                        for (int k = 0; k < NU; ++k) s_pi[r * NU + k] = *cst(0, PL::B_rm + r * NU + k);
                    }
#pragma unroll
                    for (int k = 0; k < NU; ++k) s_pi[RX * NU + k] = *cst(0, PL::Quu + k);
                }
                if (REFS == 2) {
                    xref_g = bv.Xref ? bv.Xref + nbb * bv.xref_stride + sub : nullptr;      // + i NX + 4 r
                    uref_g = (bv.Uref && urow) ? bv.Uref + nbb * bv.uref_stride + sub : nullptr;   // + i NU
                }
            }
            tm_wait_st();
            __syncwarp();
        }
        if (!__any_sync(kAll, valid)) break;

        const size_t bb = static_cast<size_t>(b);

        // =====================================================================================================
        // backward sweep: p_i for i = N-2 .. 0, leaving t_i = B' p_{i+1} + r_i + BPf for the forward sweep
        // (reference admm.cpp:13-20; d_i = Quu_inv t_i is formed where it is consumed)
        // =====================================================================================================
        {
            T Mab[RX][NX], Mkt[RX][NU], Mbt[NX];
#pragma unroll
            for (int r = 0; r < RX; ++r) {
#pragma unroll
                for (int k = 0; k < NX; k += 2) ld2(cst(CL::AmBKt_rm + (r * 4 + sub) * NX + k, PL::AmBKt + r * NX + k), Mab[r][k], Mab[r][k + 1]);
#pragma unroll
                for (int k = 0; k < NU; ++k) Mkt[r][k] = *cst(CL::Kinf_cm + (r * 4 + sub) * NU + k, PL::Kinf_cm + r * NU + k);
            }
#pragma unroll
            for (int k = 0; k < NX; ++k) Mbt[k] = urow ? *cst(CL::B_cm + usub * NX + k, PL::B_cm + k) : T(0);

            T p[RX];
#pragma unroll
            for (int r = 0; r < RX; ++r) p[r] = p_term[r];

            uint32_t w[W];
            tm_load<W>(tbase + (N - 2) * W, w);
            T qcar[RX];                              // carried q of the first sweep of a warm-started solve
#pragma unroll
            for (int r = 0; r < RX; ++r) qcar[r] = T(0);
            const bool ldq = !COLD && first && valid && !(ROLL && (cont || qp.cold_install));
            // ROLL, first sweep of a solve that continues on chip: q is what update_linear_cost left at the end of the
            // previous step, i.e. built from the slacks and duals in Tensor Memory and the PREVIOUS reference window
            const bool usecar = first && !(ROLL && cont);
            const double* xref_b = (ROLL && REFS == 2 && cont && first && qp.slide && xref_g) ? xref_g - NX : xref_g;
            const T* qp_g = sblk + (N - 2) * XK;                       // knot N-2; q = the second half of the (g, q) sequence
            auto load_q = [&]() {
                T sq[VW * GC];
#pragma unroll
                for (int c = RX / VW; c < GC; ++c) ldc<VW>(qp_g + CS * c, sq + VW * c);
#pragma unroll
                for (int r = 0; r < RX; ++r) qcar[r] = sq[RX + r];
            };
            if (ldq) load_q();
            T t_def = T(0);
            uint32_t o_rt = (N - 2) * 32 * S;                // byte offset of knot i in the r/t array
            uint32_t o_cq = (N - 2) * NX * S;

            for (int i = N - 2; i >= 0; --i) {
                T cq[RX];
                if (REFS == 2) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) cq[r] = -R::mul(static_cast<T>(xref_b ? __ldg(xref_b + i * NX + r * 4) : 0.0), fr_Qd[r]);
                }
                // all-gather p_{i+1} among the 4 lanes of the instance
                T pb[NX];
                if (NX <= TMPC_BWD_SHFL_MAX_NX) {
#pragma unroll
                for (int k = 0; k < NX; ++k) pb[k] = shfl_dq(p[k / 4], (k % 4) * kQInst + inst);
                __syncwarp();
                if (i < N - 2 && urow) sts1<32 * S>(a_myrt + o_rt, t_def);    // t_{i+1}: every lane has read r_{i+1} by now
                } else {
                a_slot = a_slot_sum - a_slot;
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (r == 0) sts1<0>(a_slot + sub * S, p[r]);
                    else if (r == 1) sts1<4 * S>(a_slot + sub * S, p[r]);
                    else sts1<8 * S>(a_slot + sub * S, p[r]);
                }
                __syncwarp();
                if (i < N - 2 && urow) sts1<32 * S>(a_myrt + o_rt, t_def);    // t_{i+1}: every lane has read r_{i+1} by now
                if constexpr (S == 4) {      // four floats per 128-bit load
                    lds4<0>(a_slot, pb[0], pb[1], pb[2], pb[3]);
                    if constexpr (NX > 4) lds4<16>(a_slot, pb[4], pb[5], pb[6], pb[7]);
                    if constexpr (NX > 8) lds4<32>(a_slot, pb[8], pb[9], pb[10], pb[11]);
                } else {
                lds2<0>(a_slot, pb[0], pb[1]);
                if (NX > 2) lds2<2 * S>(a_slot, pb[2], pb[3]);
                if (NX > 4) { lds2<4 * S>(a_slot, pb[4], pb[5]); lds2<6 * S>(a_slot, pb[6], pb[7]); }
                if (NX > 8) { lds2<8 * S>(a_slot, pb[8], pb[9]); lds2<10 * S>(a_slot, pb[10], pb[11]); }
                }
                }
                T rr[4];
                lds2<0>(a_rt4 + o_rt, rr[0], rr[1]);
                lds2<16 * S>(a_rt4 + o_rt, rr[2], rr[3]);
                const T r_own = lds1t<0, T>(a_myrt + o_rt);
                T cqs[RX];
                if (REFS == 1) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) cqs[r] = (r == 0) ? lds1t<0, T>(a_cq + o_cq) : (r == 1) ? lds1t<4 * S, T>(a_cq + o_cq) : lds1t<8 * S, T>(a_cq + o_cq);
                }

                // q_i (own rows): carried on the first sweep of a solve, otherwise rebuilt from slacks and duals
                tm_wait_ld<W>(w);
                T q_own[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    const T gg = tm_get(w, r, T());
                    const T vv = tm_get(w, RX + r, T());
                    T q;
                    if (REFS == 0) q = R::mul(nrho, R::sub(vv, gg));
                    else q = R::sub(REFS == 1 ? cqs[r] : cq[r], R::mul(rho, R::sub(vv, gg)));
                    q_own[r] = usecar ? qcar[r] : q;
                }
                if (i > 0) {
                    tm_load<W>(tbase + (i - 1) * W, w);
                    if (ldq) { qp_g -= XK; load_q(); }
                }

                // own row of B' p_{i+1} and the lane's rows of AmBKt p_{i+1}: RX + 1 independent chains, interleaved
                T s = T(0), s1[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) s1[r] = T(0);
#pragma unroll
                for (int k = 0; k < NX; ++k) {
                    s = R::fma(Mbt[k], pb[k], s);
#pragma unroll
                    for (int r = 0; r < RX; ++r) s1[r] = R::fma(Mab[r][k], pb[k], s1[r]);
                }
                t_def = R::add(R::add(s, r_own), fr_BPf);     // t_i = B' p_{i+1} + r_i + BPf

                // p_i = q_i + AmBKt p_{i+1} - Kinf' r_i + APf
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    T s2 = T(0);
#pragma unroll
                    for (int k = 0; k < NU; ++k) s2 = R::fma(Mkt[r][k], rr[k], s2);
                    p[r] = R::add(R::sub(R::add(q_own[r], s1[r]), s2), fr_APf[r]);
                }
                o_rt -= 32 * S; o_cq -= NX * S;
            }
            __syncwarp();
            if (urow) sts1<0>(a_myrt, t_def);
        }

        // ROLL: the previous step of some instance converged, so its carried previous slacks are the lagging copies
        // the last sweep parked in its block (admm.cpp:375 returns before :379-380).  The backward sweep above
        // was the last reader of vnew; put v_old / z_old in place for the dual residual of this solve's first iteration.
        if (ROLL && __any_sync(kAll, fix)) {
            const T* vl = sblk + CS * GC;            // v chunks of knot i: + XK i
            for (int i = 0; i < N; ++i) {
                uint32_t w[W];
                tm_load<W>(tbase + i * W, w);
                T vo[VW * VC];
#pragma unroll
                for (int r = 0; r < VW * VC; ++r) vo[r] = T(0);
#pragma unroll
                for (int c = 0; c < VC; ++c) { if (fix) ldc<VW>(vl + CS * c, vo + VW * c); }
                tm_wait_ld<W>(w);
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    tm_put_if(w, RX + r, vo[r], fix);
                }
                tm_store<W>(tbase + i * W, w);
                vl += XK;
            }
            if (fix && urow) {
                const T* zl = sblk + IB + zo;             // z of knot i: + UK i
                for (int i = 0; i < N - 1; ++i) { s_yz[(i * 32 + lane) * 2 + 1] = *zl; zl += UK; }
            }
            fix = false;
            tm_wait_st();
            __syncwarp();
        }

        // =====================================================================================================
        // forward sweep fused with slack / dual / linear-cost / residual updates
        // (reference admm.cpp:25-32, :84-98, :184-187, :218-239, :257-265)
        // =====================================================================================================
        bool ok = true;              // every residual this lane has seen in this sweep is under its tolerance
        {
            // register-resident: the lane's rows of A and Kinf (48 doubles).  B and Quu_inv rows are re-read from
            // shared memory every knot point (8 broadcast LDS.128 off the critical path): keeping them in registers
            // too pushed the sweep over the register file and the compiler into rematerialising addresses.
            T Ma[RX][NX], Mk[NX];
#pragma unroll
            for (int r = 0; r < RX; ++r) {
#pragma unroll
                for (int k = 0; k < NX; k += 2) ld2(cst(CL::A_rm + (r * 4 + sub) * NX + k, PL::A + r * NX + k), Ma[r][k], Ma[r][k + 1]);
            }
#pragma unroll
            for (int k = 0; k < NX; ++k) Mk[k] = urow ? *cst(CL::Kinf_rm + usub * NX + k, PL::Kinf_rm + k) : T(0);
            // rows of B and of Quu_inv: registers where the variant has room for them (CREG), else re-read from
            // shared memory at every knot point.  Shared-memory bandwidth (4 bytes per lane per clock) is what the
            // two warps of a scheduler and their six neighbours contend for: 128 of the 336 bytes a lane reads per
            // forward knot point are these constants.
            T Mb[CREG ? RX : 1][NU], Mq[NU];
            if (CREG) {
#pragma unroll
                for (int r = 0; r < RX; ++r)
#pragma unroll
                    for (int k = 0; k < NU; ++k) Mb[r][k] = *cst(CL::B_rm + (r * 4 + sub) * NU + k, PL::B_rm + r * NU + k);
#pragma unroll
                for (int k = 0; k < NU; ++k) Mq[k] = *cst(CL::Quu_rm + usub * NU + k, PL::Quu + k);
            }

            T x[RX];
#pragma unroll
            for (int r = 0; r < RX; ++r) x[r] = x0r[r];
            uint32_t w[W];
            tm_load<W>(tbase, w);
            // the carried "previous slacks" must lag the solution by one iteration when the solve converges
            // (admm.cpp:375 returns before :379-380): while this sweep can still end in convergence the lane parks
            // them in the instance's carried-state block in HBM ([pointer + immediate] stores, whole 32-byte sectors)
            const bool lag = valid && ((iter + 1) % st.check_termination == 0);
            T* vlag = sblk + CS * GC;                                    // v chunks of knot i: + XK i
            T* zlag = sblk + IB + zo;                                    // z of knot i: + UK i
            T r_def = T(0);
            T vg[RX];
            uint32_t o_xb = 0;                                                // byte offset of knot i in the state-bound table
            uint32_t tcol = tbase;

            // Residual tests are evaluated lazily: as soon as every instance of the warp has failed one test of
            // this sweep (warp vote after each knot point), or when this iteration cannot terminate the solve
            // (check_termination), the rest of the sweep runs the variant without tests and without lag stores.
            // The decision per instance is unchanged: a failed lane keeps ok = false.

            // state column: slack, dual, [residuals]; new (g, v) back to Tensor Memory
            auto state_column = [&](auto chk) {
                constexpr bool CHK = decltype(chk)::value;
                T lo[RX], hi[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (r == 0) lds2<0>(a_xb + o_xb, lo[r], hi[r]);
                    else if (r == 1) lds2<8 * S>(a_xb + o_xb, lo[r], hi[r]);
                    else lds2<16 * S>(a_xb + o_xb, lo[r], hi[r]);
                }
                tm_wait_ld<W>(w);
                uint32_t wn[W];
                T vol[VW * VC];                              // previous slacks of the lane's rows, parked chunk by chunk
                vol[VW * VC - 1] = T(0);
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    const T gold = tm_get(w, r, T());
                    const T t = R::add(x[r], gold);
                    const T vnew = R::clamp(t, lo[r], hi[r]);
                    const T gnew = R::sub(t, vnew);
                    if (CHK) {
                        const T vold = tm_get(w, RX + r, T());
                        const bool p1 = R::abs(R::sub(x[r], vnew)) < pri_thresh;
                        const bool p2 = R::abs(R::sub(vold, vnew)) <= dua_thresh;
                        ok = ok & p1 & p2;
                        vol[r] = vold;
                        if ((r + 1) % VW == 0 || r == RX - 1) { if (lag & ok) stc<VW>(vlag + CS * (r / VW), vol + (r / VW) * VW); }
                    }
                    tm_put(wn, r, gnew);
                    tm_put(wn, RX + r, vnew);
                    vg[r] = R::sub(vnew, gnew);
                }
                tm_store<W>(tcol, wn);
            };

            uint32_t o_rt = 0, o_yz = 0, o_ub = 0, o_cr = 0;
            int i = 0;
            // one non-terminal knot point
            auto knot = [&](auto chk) {
                constexpr bool CHK = decltype(chk)::value;
                T cr = T(0);
                if (REFS == 2) cr = -R::mul(static_cast<T>(uref_g ? __ldg(uref_g + i * NU) : 0.0), fr_Rd);
                a_slot = a_slot_sum - a_slot;
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (r == 0) sts1<0>(a_slot + sub * S, x[r]);
                    else if (r == 1) sts1<4 * S>(a_slot + sub * S, x[r]);
                    else sts1<8 * S>(a_slot + sub * S, x[r]);
                }
                __syncwarp();
                if (i > 0 && urow) sts1<-32 * S>(a_myrt + o_rt, r_def);         // r_{i-1}: every lane has read t_{i-1} by now
                T xb[NX];
                if constexpr (S == 4) {      // four floats per 128-bit load
                    lds4<0>(a_slot, xb[0], xb[1], xb[2], xb[3]);
                    if constexpr (NX > 4) lds4<16>(a_slot, xb[4], xb[5], xb[6], xb[7]);
                    if constexpr (NX > 8) lds4<32>(a_slot, xb[8], xb[9], xb[10], xb[11]);
                } else {
                lds2<0>(a_slot, xb[0], xb[1]);
                if (NX > 2) lds2<2 * S>(a_slot, xb[2], xb[3]);
                if (NX > 4) { lds2<4 * S>(a_slot, xb[4], xb[5]); lds2<6 * S>(a_slot, xb[6], xb[7]); }
                if (NX > 8) { lds2<8 * S>(a_slot, xb[8], xb[9]); lds2<10 * S>(a_slot, xb[10], xb[11]); }
                }
                T tt[4], mq[4];
                lds2<0>(a_rt4 + o_rt, tt[0], tt[1]);
                lds2<16 * S>(a_rt4 + o_rt, tt[2], tt[3]);
                if (CREG) {
#pragma unroll
                    for (int k = 0; k < NU; ++k) mq[k] = Mq[k];
                } else if (NU == 4) { lds2<0>(a_Mq, mq[0], mq[1]); lds2<2 * S>(a_Mq, mq[2], mq[3]); }
                else {
#pragma unroll
                    for (int k = 0; k < NU; ++k) mq[k] = *cst(CL::Quu_rm + usub * NU + k, PL::Quu + k);
                }
                T yj = T(0), zold = T(0), ulo = T(0), uhi = T(0), crs = T(0);
                if (urow) {
                    lds2<0>(a_yz + o_yz, yj, zold);
                    lds2<0>(a_ub + o_ub, ulo, uhi);
                    if (REFS == 1) crs = lds1t<0, T>(a_cr + o_cr);
                }

                state_column(chk);                            // needs only the lane's own x: covers the broadcast round trip
                tcol += W;
                tm_load<W>(tcol, w);                          // next knot point; consumed after the next broadcast
#if TMPC_FENCE_A
                SCHED_FENCE();
#endif

                // d_i = Quu_inv t_i ;  Kinf x_i (own row) ;  A x_i (own rows): RX + 2 independent chains, interleaved
                T d = T(0);
#pragma unroll
                for (int k = 0; k < NU; ++k) d = R::fma(mq[k], tt[k], d);
                T s = T(0), s1[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) s1[r] = T(0);
#pragma unroll
                for (int k = 0; k < NX; ++k) {
                    s = R::fma(Mk[k], xb[k], s);
#pragma unroll
                    for (int r = 0; r < RX; ++r) s1[r] = R::fma(Ma[r][k], xb[k], s1[r]);
                }

                SCHED_FENCE();
                // u_i = -Kinf x_i - d_i, broadcast at once; then input slack, dual, [residuals]
                // (admm.cpp:28, :87, :97, :187, :259-260)
                const T u_own = urow ? R::sub(-s, d) : T(0);
                T ub[NU];
#pragma unroll
                for (int k = 0; k < NU; ++k) ub[k] = shfl_dq(u_own, k * kQInst + inst);
                // rows of B for the tail of this knot point: issued now, consumed after the state column
                T mb[RX][4];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (CREG) {
#pragma unroll
                        for (int k = 0; k < NU; ++k) mb[r][k] = Mb[r][k];
                    } else if (NU == 4) {
                        if (r == 0) { lds2<0>(a_Mb, mb[r][0], mb[r][1]); lds2<2 * S>(a_Mb, mb[r][2], mb[r][3]); }
                        else if (r == 1) { lds2<MBS>(a_Mb, mb[r][0], mb[r][1]); lds2<MBS + 2 * S>(a_Mb, mb[r][2], mb[r][3]); }
                        else { lds2<2 * MBS>(a_Mb, mb[r][0], mb[r][1]); lds2<2 * MBS + 2 * S>(a_Mb, mb[r][2], mb[r][3]); }
                    } else {
#pragma unroll
                        for (int k = 0; k < NU; ++k) mb[r][k] = *cst(CL::B_rm + (r * 4 + sub) * NU + k, PL::B_rm + r * NU + k);
                    }
                }
                if (urow) {
                    const T t = R::add(u_own, yj);
                    const T znew = R::clamp(t, ulo, uhi);
                    const T ynew = R::sub(t, znew);                  // y + u - znew
                    if (CHK) {
                        const bool p1 = R::abs(R::sub(u_own, znew)) < pri_thresh;
                        const bool p2 = R::abs(R::sub(zold, znew)) <= dua_thresh;
                        ok = ok & p1 & p2;
                        if (lag & ok) *zlag = zold;
                    }
                    sts2<0>(a_yz + o_yz, ynew, znew);
                    // r_i = -(Uref .* Rd) - rho (znew - y)    (admm.cpp:228-229)
                    if (REFS == 0) r_def = R::mul(nrho, R::sub(znew, ynew));
                    else r_def = R::sub(REFS == 1 ? crs : cr, R::mul(rho, R::sub(znew, ynew)));
                }

#if TMPC_FENCE_C
                SCHED_FENCE();
#endif
                // x_{i+1} = A x_i + B u_i + f
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    T s2 = T(0);
#pragma unroll
                    for (int k = 0; k < NU; ++k) s2 = R::fma(mb[r][k], ub[k], s2);
                    x[r] = R::add(R::add(s1[r], s2), fr_f[r]);
                }
                vlag += XK; zlag += UK;
                o_rt += 32 * S; o_yz += 64 * S; o_ub += NU * 2 * S; o_cr += NU * S; o_xb += NX * 2 * S;
            };

            bool checking = __any_sync(kAll, lag);            // no instance of the warp can terminate: no tests at all
            if (!checking) ok = false;
            while (checking && i < N - 1) {
                knot(std::true_type{});
                ++i;
                const unsigned v = __ballot_sync(kAll, ok && lag);
                checking = (v & (v >> 8) & (v >> 16) & (v >> 24) & 0xffu) != 0;   // some instance has all 4 lanes ok
            }
            for (; i < N - 1; ++i) knot(std::false_type{});
            // terminal knot point: p_{N-1} = -(Xref_{N-1}' Pinf)' - rho (vnew - g)   (admm.cpp:238-239)
            if (checking) state_column(std::true_type{}); else state_column(std::false_type{});
#pragma unroll
            for (int r = 0; r < RX; ++r) {
                if (REFS == 0) p_term[r] = R::mul(nrho, vg[r]);
                else if (REFS == 1) p_term[r] = R::sub(s_tq[r * 4 + sub], R::mul(rho, vg[r]));
                else {
                    T s = T(0);
                    if (xref_g) for (int k = 0; k < NX; ++k) s = R::fma(static_cast<T>(__ldg(xref_g - sub + (N - 1) * NX + k)), *cst(CL::Pinf_cm + (r * 4 + sub) * NX + k, PL::Pinf + r * NX + k), s);
                    p_term[r] = R::sub(-s, R::mul(rho, vg[r]));
                }
            }
            tm_wait_st();
            __syncwarp();
            if (urow) sts1<-32 * S>(a_myrt + o_rt, r_def);       // r_{N-2}
        }

        // =====================================================================================================
        // termination (admm.cpp:253-271, :364-388), per instance
        // =====================================================================================================
        iter += 1;
        first = false;
        const unsigned votes = __ballot_sync(kAll, ok);
        const unsigned mine = 0x01010101u << inst;
        const bool conv = valid && (iter % st.check_termination == 0) && ((votes & mine) == mine);
        status = conv ? 0 : 1;
        fin = valid && (conv || iter >= st.max_iter);
    }

    __syncthreads();
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(s_tmem_base));
    }
}

// Largest a >= 0 with fl(rho * a) < tol, so that  rho * max|.| < tol  <=>  every |.| <= a.
template <class T>
T dual_threshold_q(T rho, T tol) {
    if (!(tol > T(0)) || !(rho > T(0))) return T(-1);
    auto prod = [&](T a) { volatile T p = rho * a; return static_cast<T>(p); };      // rounded to T, no excess precision
    T a = tol / rho;
    while (prod(a) < tol) a = std::nextafter(a, static_cast<T>(INFINITY));
    while (!(prod(a) < tol)) a = std::nextafter(a, static_cast<T>(-INFINITY));
    return a;
}

struct QPlan { size_t smem_bytes; QuadParams qp; };

constexpr int kQLongWarps = 4;       // the long-horizon form (kernel template parameter WARPS)

template <class T, int NX, int NU>
bool make_qplan(int N, int refs, QPlan& pl, bool pi = false, int warps_override = 0) {
    constexpr int XP = QPad<NX, sizeof(T)>::value;
    constexpr int W = 2 * (NX / 4) * (sizeof(T) / 4);
    const int warps = warps_override ? warps_override : (pi ? QWarps<T, NX, true>::value : QWarps<T, NX, false>::value);
    const int cols = warps_override ? ((512 / ((warps_override + 3) / 4)) & ~1) : (pi ? QTmemCols<T, NX, true>::value : QTmemCols<T, NX, false>::value);
    if (N < 2 || N * W > cols) return false;   // g, v of a lane must fit the warp's TMEM columns
    const int nxN = NX * N, nuN = NU * (N - 1);
    QuadParams qp{};
    int off = 0;
    qp.off_consts = off; off += QConst<NX, NU>::size;
    qp.off_bounds = off; off += 2 * (nxN + nuN);
    qp.off_refs = off; if (refs == 1) off += nxN + nuN + NX;
    off = (off + 3) / 4 * 4;                                       // 16-byte aligned at either scalar
    qp.off_warp0 = off;
    int per_warp = (N - 1) * 96 + 2 * kQInst * XP;
    per_warp = (per_warp + 3) / 4 * 4;
    qp.off_pi = per_warp;
    if (pi) per_warp += 32 * 18;                                   // the lanes' private B / Quu_inv row slots
    qp.warp_stride = per_warp;
    pl.smem_bytes = (static_cast<size_t>(off) + static_cast<size_t>(per_warp) * warps) * sizeof(T);
    pl.qp = qp;
    return pl.smem_bytes <= 227 * 1024;
}

int refs_mode(const BatchView& bv) {
    if (bv.Xref == nullptr && bv.Uref == nullptr) return 0;
    const bool xs = bv.Xref == nullptr || bv.xref_stride == 0, us = bv.Uref == nullptr || bv.uref_stride == 0;
    return (xs && us) ? 1 : 2;
}

template <class T, int NX, int NU, bool COLD, int REFS, bool ROLL = false, bool PI = false, int WARPS = 0>
cudaError_t launch_q(const BatchView& bv, const AdmmSettings& st, const QPlan& pl, int num_sms, cudaStream_t stream) {
    auto kern = admm_tmem_kernel<T, NX, NU, COLD, REFS, ROLL, PI, WARPS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.smem_bytes));
    if (e != cudaSuccess) return e;
    constexpr int NW = WARPS ? WARPS : QWarps<T, NX, PI>::value;
    const int per_cta = NW * kQInst;
    int grid = (bv.B + per_cta - 1) / per_cta;
    if (grid > num_sms) grid = num_sms;            // persistent: one CTA per SM, sub-warps pull instances from the queue
    kern<<<grid, NW * 32, pl.smem_bytes, stream>>>(bv, st, pl.qp);
    return cudaGetLastError();
}

// shapes that have the long-horizon form compiled: the quadrotor shape in double (N = 22 .. 42)
template <class T, int NX> constexpr bool long_horizon_shape() { return sizeof(T) == 8 && NX == 12; }

template <class T, int NX, int NU>
cudaError_t launch_shape(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                         cudaStream_t stream, const int* order) {
    QPlan pl;
    const int refs = refs_mode(bv);
    bool long_form = false;
    if (!make_qplan<T, NX, NU>(bv.N, refs, pl)) {
        if (!long_horizon_shape<T, NX>() || !make_qplan<T, NX, NU>(bv.N, refs, pl, false, kQLongWarps)) return cudaErrorInvalidConfiguration;
        long_form = true;
    }
    pl.qp.counter = counter;
    pl.qp.order = order;
    pl.qp.pri_thresh = static_cast<T>(st.abs_pri_tol);                                   // types.hpp:64-65: tolerances are tinytype
    pl.qp.dua_thresh = dual_threshold_q<T>(static_cast<T>(bv.rho), static_cast<T>(st.abs_dua_tol));
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    if constexpr (long_horizon_shape<T, NX>()) {
        if (long_form) {
            if (cold) {
                if (refs == 0) return launch_q<T, NX, NU, true, 0, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
                if (refs == 1) return launch_q<T, NX, NU, true, 1, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
                return launch_q<T, NX, NU, true, 2, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
            }
            if (refs == 0) return launch_q<T, NX, NU, false, 0, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
            if (refs == 1) return launch_q<T, NX, NU, false, 1, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
            return launch_q<T, NX, NU, false, 2, false, false, kQLongWarps>(bv, st, pl, num_sms, stream);
        }
    }
    if (cold) {
        if (refs == 0) return launch_q<T, NX, NU, true, 0>(bv, st, pl, num_sms, stream);
        if (refs == 1) return launch_q<T, NX, NU, true, 1>(bv, st, pl, num_sms, stream);
        return launch_q<T, NX, NU, true, 2>(bv, st, pl, num_sms, stream);
    }
    if (refs == 0) return launch_q<T, NX, NU, false, 0>(bv, st, pl, num_sms, stream);
    if (refs == 1) return launch_q<T, NX, NU, false, 1>(bv, st, pl, num_sms, stream);
    return launch_q<T, NX, NU, false, 2>(bv, st, pl, num_sms, stream);
}

template <int NX, int NU>
cudaError_t launch_shape_pi(const BatchView& bv, const AdmmSettings& st, int cold, const double* pi_blocks, int* counter,
                            int num_sms, cudaStream_t stream, const int* order) {
    QPlan pl;
    const int refs = refs_mode(bv) == 0 ? 0 : 2;       // references, shared or not, go through the per-instance path
    if (!make_qplan<double, NX, NU>(bv.N, refs, pl, true)) return cudaErrorInvalidConfiguration;
    pl.qp.counter = counter;
    pl.qp.pri_thresh = st.abs_pri_tol;
    pl.qp.dua_thresh = dual_threshold_q<double>(bv.rho, st.abs_dua_tol);
    pl.qp.pi_blocks = pi_blocks;
    pl.qp.order = order;
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    if (cold) {
        if (refs == 0) return launch_q<double, NX, NU, true, 0, false, true>(bv, st, pl, num_sms, stream);
        return launch_q<double, NX, NU, true, 2, false, true>(bv, st, pl, num_sms, stream);
    }
    if (refs == 0) return launch_q<double, NX, NU, false, 0, false, true>(bv, st, pl, num_sms, stream);
    return launch_q<double, NX, NU, false, 2, false, true>(bv, st, pl, num_sms, stream);
}

// element-major per-instance cache [e][ld]  ->  one lane-interleaved block per instance (QLane: what each lane of the
// sub-warp reads, in 16-byte chunks dealt round-robin to the four lanes)
template <int NX, int NU>
__global__ void pi_pack_kernel(PerInstanceMats pm, long long ld, int B, double* __restrict__ blocks) {
    using PL = QLane<NX, NU>;
    constexpr int RX = NX / 4;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double* o = blocks + static_cast<size_t>(b) * PL::size;
    auto at = [&](const double* a, int e) { return a[static_cast<size_t>(e) * ld + b]; };
    for (int sub = 0; sub < 4; ++sub) {
        auto put = [&](int idx, double v) { o[PL::at(idx, sub)] = v; };
        for (int idx = 0; idx < PL::per_lane; ++idx) put(idx, 0.0);
        for (int r = 0; r < RX; ++r) {
            const int row = r * 4 + sub;
            for (int k = 0; k < NX; ++k) {
                put(PL::A + r * NX + k, at(pm.A, k * NX + row));
                put(PL::AmBKt + r * NX + k, at(pm.AmBKt, k * NX + row));
                put(PL::Pinf + r * NX + k, at(pm.Pinf, row * NX + k));
            }
            for (int k = 0; k < NU; ++k) {
                put(PL::B_rm + r * NU + k, at(pm.B, k * NX + row));
                put(PL::Kinf_cm + r * NU + k, at(pm.Kinf, row * NU + k));
            }
            put(PL::f + r, at(pm.f, row)); put(PL::APf + r, at(pm.APf, row)); put(PL::Qd + r, at(pm.Qd, row));
        }
        if (sub < NU) {
            for (int k = 0; k < NX; ++k) {
                put(PL::B_cm + k, at(pm.B, sub * NX + k));
                put(PL::Kinf_rm + k, at(pm.Kinf, k * NU + sub));
            }
            for (int k = 0; k < NU; ++k) put(PL::Quu + k, at(pm.Quu_inv, k * NU + sub));
            put(PL::BPf, at(pm.BPf, sub)); put(PL::Rd, at(pm.Rd, sub));
        }
    }
}

template <int NX, int NU>
cudaError_t launch_shape_roll(const BatchView& bv, const AdmmSettings& st, const RolloutArgs& ra, int* counter,
                              int num_sms, cudaStream_t stream) {
    QPlan pl;
    const int refs = ra.pi_blocks ? (refs_mode(bv) == 0 ? 0 : 2) : refs_mode(bv);
    if (!make_qplan<double, NX, NU>(bv.N, refs, pl, ra.pi_blocks != nullptr)) return cudaErrorInvalidConfiguration;
    pl.qp.counter = counter;
    pl.qp.pri_thresh = st.abs_pri_tol;
    pl.qp.dua_thresh = dual_threshold_q<double>(bv.rho, st.abs_dua_tol);
    pl.qp.pi_blocks = ra.pi_blocks;
    pl.qp.order = ra.order;
    pl.qp.nsteps = ra.nsteps; pl.qp.cold_install = ra.cold; pl.qp.slide = ra.slide;
    pl.qp.x0_rw = ra.x0_rw; pl.qp.status_out = ra.status_out; pl.qp.iters_out = ra.iters_out;
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    if (ra.pi_blocks) {
        if (refs == 0) return launch_q<double, NX, NU, false, 0, true, true>(bv, st, pl, num_sms, stream);
        return launch_q<double, NX, NU, false, 2, true, true>(bv, st, pl, num_sms, stream);
    }
    if (refs == 0) return launch_q<double, NX, NU, false, 0, true>(bv, st, pl, num_sms, stream);
    if (refs == 1) return launch_q<double, NX, NU, false, 1, true>(bv, st, pl, num_sms, stream);
    return launch_q<double, NX, NU, false, 2, true>(bv, st, pl, num_sms, stream);
}

}  // namespace

cudaError_t launch_admm_tmem_rollout(const BatchView& bv, const AdmmSettings& st, const RolloutArgs& ra, int* counter,
                                     int num_sms, cudaStream_t stream) {
    if (bv.B <= 0 || ra.nsteps <= 0) return cudaSuccess;
    if (bv.nx == 12 && bv.nu == 4) return launch_shape_roll<12, 4>(bv, st, ra, counter, num_sms, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch_shape_roll<4, 1>(bv, st, ra, counter, num_sms, stream);
    return cudaErrorNotSupported;
}

long long tmem_pi_block_doubles(int nx, int nu) {
    if (nx == 12 && nu == 4) return QLane<12, 4>::size;
    if (nx == 4 && nu == 1) return QLane<4, 1>::size;
    return 0;
}

cudaError_t launch_tmem_pi_pack(int nx, int nu, const PerInstanceMats& pm, long long ld, int B, double* blocks,
                                cudaStream_t stream) {
    const int block = 128, grid = (B + block - 1) / block;
    if (nx == 12 && nu == 4) pi_pack_kernel<12, 4><<<grid, block, 0, stream>>>(pm, ld, B, blocks);
    else if (nx == 4 && nu == 1) pi_pack_kernel<4, 1><<<grid, block, 0, stream>>>(pm, ld, B, blocks);
    else return cudaErrorNotSupported;
    return cudaGetLastError();
}

cudaError_t launch_admm_tmem_pi(const BatchView& bv, const AdmmSettings& st, int cold, const double* pi_blocks,
                                int* counter, int num_sms, cudaStream_t stream, const int* order) {
    if (bv.B <= 0) return cudaSuccess;
    if (bv.nx == 12 && bv.nu == 4) return launch_shape_pi<12, 4>(bv, st, cold, pi_blocks, counter, num_sms, stream, order);
    if (bv.nx == 4 && bv.nu == 1) return launch_shape_pi<4, 1>(bv, st, cold, pi_blocks, counter, num_sms, stream, order);
    return cudaErrorNotSupported;
}

bool tmem_pi_kernel_supports(const BatchView& bv, const AdmmSettings& st) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return false;
    if (bv.x_min != nullptr && (bv.xb_stride != 0 || bv.ub_stride != 0)) return false;
    QPlan pl;
    const int refs = refs_mode(bv) == 0 ? 0 : 2;
    if (bv.nx == 12 && bv.nu == 4) return make_qplan<double, 12, 4>(bv.N, refs, pl, true);
    if (bv.nx == 4 && bv.nu == 1) return make_qplan<double, 4, 1>(bv.N, refs, pl, true);
    return false;
}

bool tmem_kernel_fused_supports(const BatchView& bv, const AdmmSettings& st, int precision) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return false;
    if (bv.x_min != nullptr && (bv.xb_stride != 0 || bv.ub_stride != 0)) return false;   // per-instance bounds: other kernels
    QPlan pl;
    if (precision == 32) {
        if (bv.nx == 12 && bv.nu == 4) return make_qplan<float, 12, 4>(bv.N, refs_mode(bv), pl);
        if (bv.nx == 4 && bv.nu == 1) return make_qplan<float, 4, 1>(bv.N, refs_mode(bv), pl);
        return false;
    }
    if (bv.nx == 12 && bv.nu == 4) return make_qplan<double, 12, 4>(bv.N, refs_mode(bv), pl);
    if (bv.nx == 4 && bv.nu == 1) return make_qplan<double, 4, 1>(bv.N, refs_mode(bv), pl);
    return false;
}

// ... or the long-horizon form (plain solves only: the fused closed loop and per-instance dynamics keep the default form)
bool tmem_kernel_supports(const BatchView& bv, const AdmmSettings& st, int precision) {
    if (tmem_kernel_fused_supports(bv, st, precision)) return true;
    if (precision != 64 || bv.nx != 12 || bv.nu != 4) return false;
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return false;
    if (bv.x_min != nullptr && (bv.xb_stride != 0 || bv.ub_stride != 0)) return false;
    QPlan pl;
    return make_qplan<double, 12, 4>(bv.N, refs_mode(bv), pl, false, kQLongWarps);
}

cudaError_t launch_admm_tmem(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                             cudaStream_t stream, const int* order, int precision) {
    if (bv.B <= 0) return cudaSuccess;
    if (precision == 32) {
        if (bv.nx == 12 && bv.nu == 4) return launch_shape<float, 12, 4>(bv, st, cold, counter, num_sms, stream, order);
        if (bv.nx == 4 && bv.nu == 1) return launch_shape<float, 4, 1>(bv, st, cold, counter, num_sms, stream, order);
        return cudaErrorNotSupported;
    }
    if (bv.nx == 12 && bv.nu == 4) return launch_shape<double, 12, 4>(bv, st, cold, counter, num_sms, stream, order);
    if (bv.nx == 4 && bv.nu == 1) return launch_shape<double, 4, 1>(bv, st, cold, counter, num_sms, stream, order);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
