// ADMM solve kernel v3: "two instances per lane" on top of the Tensor-Memory sub-warp design of admm_tmem.cu.
//
// Why it exists (measured, profiles/r01_d_*): admm_tmem.cu is capacity-bound at 8 warps per SM = two warps per
// scheduler, and a warp's knot point is one long dependency chain (broadcast -> 12-deep DFMA chain -> shuffle ->
// 4-deep chain -> adds): 460 + 660 cycles per knot-point pair for 2 x 328 cycles of FP64-pipe work per scheduler,
// i.e. the FP64 pipe idles 43 % of the time inside the sweeps.  More warps do not fit (registers: the lane's rows
// of the cached matrices are 96-120 registers; on-chip state: 5.6 KB per instance).  This kernel doubles the
// independent work per warp instead:
//   * every lane owns the same rows of TWO instances (group 0 / group 1 of the warp, 16 instances per warp,
//     128 per SM).  The matrix rows in registers and the B / Quu_inv / bound rows read from shared memory are
//     shared by the pair, so each costs half as much per instance, and every phase of a knot point has two
//     independent dependency chains to interleave;
//   * to make room, the resident state is ONE double per element instead of two: the pair (dual g, slack v) of an
//     element is a function of the pre-clamp value t = x + g_old alone (v = clamp(t), g = t - v, both exactly what
//     the reference computes, admm.cpp:92-97 and :184-187), likewise (y, z) of an input element.  A knot point of
//     a lane is then 6 words of Tensor Memory per instance, and the two instances of a lane sit side by side in
//     the same 12 columns, so one tcgen05.ld/st pair per knot point serves both.  The values a freshly installed
//     instance starts from (zeros, or the carried g / v / y / z of a warm start) are NOT such a pair in general
//     (e.g. bounds that exclude zero, bounds changed between solves); they enter through a per-instance "first
//     sweep" select and never through the compressed form;
//   * the write-out pass prefetches the next knot point's Tensor Memory row and, for cold starts, no longer
//     stores to Tensor Memory at all.
//
// Lane map: lane = sub * 8 + inst; lane `sub` owns state rows {4 r + sub} and input row `sub` of instance `inst` of
// BOTH groups.  TMEM map: warp w uses TMEM lanes 32 (w % 4) .. +31 and columns 256 (w / 4) .. +255; knot point i
// of a lane is columns 4 RX i .. : [group][row r] t as two 32-bit halves.
//
// Arithmetic is the same sequence of correctly rounded operations as the other kernels, so all kernels agree bit
// for bit on the carried state (tests switch kernels in the middle of a closed loop).
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

constexpr int kDWarps = 8;           // warps per CTA (one CTA per SM): two per scheduler
constexpr int kDInst = 8;            // instances per group
constexpr int kG = 2;                // groups per warp = instances per lane
constexpr int kDTmemColsPerWarp = 256;

struct DualParams {
    double pri_thresh;      // primal residual passes  <=>  |x - vnew| <  pri_thresh
    double dua_thresh;      // dual residual passes    <=>  |v - vnew| <= dua_thresh  (largest a with fl(rho a) < tol)
    int* counter;           // instance queue
    int off_consts, off_bounds, off_refs, off_warp0, warp_stride;   // shared-memory carve-up, in doubles
};

// shared memory by 32-bit address; constant offsets are added in C++ and folded into the instruction by ptxas
__device__ __forceinline__ void xlds2(uint32_t a, double& x, double& y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(a));
}
__device__ __forceinline__ double xlds1(uint32_t a) {
    double x;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(a));
    return x;
}
__device__ __forceinline__ void xsts1(uint32_t a, double x) { asm volatile("st.shared.f64 [%0], %1;" :: "r"(a), "d"(x) : "memory"); }
__device__ __forceinline__ void xsts2(uint32_t a, double x, double y) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" :: "r"(a), "d"(x), "d"(y) : "memory");
}

// ptxas schedules a basic block by critical-path priority and lets the chain that feeds the shuffles run ahead
// serially (one DFMA per 8 cycles) instead of interleaving it with the independent chains.  A never-taken uniform
// branch ends the basic block, so inside it all chains have the same distance to the end and are issued round-robin.
#define SCHED_FENCE() do { if (st.max_iter == -12345) asm volatile("trap;"); } while (0)

// REFS: 0 = no references (Xref = Uref = 0), 1 = references shared by all instances (staged in shared memory as
// the products the linear cost needs), 2 = per-instance references read through the read-only path.
template <int NX, int NU, bool COLD, int REFS>
__global__ void __launch_bounds__(kDWarps * 32, 1)
admm_tmem2_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st,
                  const __grid_constant__ DualParams dp) {
    // A global access in uniform control flow before anything else: ptxas then keeps the global-memory descriptor in a
    // uniform register for the whole kernel.  Without it the first global access sits in a divergent region (the
    // staging loop), the descriptor lands in a vector register pair and EVERY load/store pays two R2UR (7 % of all
    // instructions issued).  The queue counter never reaches the value tested.
    if (*reinterpret_cast<const volatile int*>(dp.counter) == 0x7ffffff0) return;
    static_assert(NX % 4 == 0 && NX >= 4, "state rows are dealt 4 lanes x RX rows");
    static_assert(NU >= 1 && NU <= 4, "one input row per lane");
    constexpr int RX = NX / 4;
    constexpr int XP = QPad<NX>::value;
    constexpr int WG = 2 * RX;                        // 32-bit TMEM words per knot point per lane per group: t(RX)
    constexpr int W = kG * WG;                        // per knot point per lane
    static_assert(W == 12 || W == 4, "tm_wait_ld is specialised for 12 or 4 words");
    constexpr int GS = kDInst * XP * 8;               // bytes between the groups' broadcast slots
    using CL = QConst<NX, NU>;

    extern __shared__ __align__(16) double smem[];
    __shared__ uint32_t s_tmem_base;
    const int N = bv.N;
    const int nxN = NX * N, nuN = NU * (N - 1);
    const double rho = bv.rho, nrho = -bv.rho;

    // ---------------- Tensor Memory: the whole 512-column file of this SM (one CTA per SM)
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "l"(static_cast<uint64_t>(__cvta_generic_to_shared(&s_tmem_base))) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }

    // ---------------- stage the shared constants once per CTA
    double* sc = smem + dp.off_consts;
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

    // box bounds as (lo, hi) pairs; a disabled side is (-inf, +inf), for which the clamp is the identity
    double* s_xb = smem + dp.off_bounds;              // [N][NX][2]
    double* s_ub = s_xb + 2 * nxN;                    // [N-1][NU][2]
    {
        const bool bx = st.en_state_bound && bv.x_min != nullptr, bu = st.en_input_bound && bv.u_min != nullptr;
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) {
            s_xb[2 * e] = bx ? bv.x_min[e] : -INFINITY; s_xb[2 * e + 1] = bx ? bv.x_max[e] : INFINITY;
        }
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) {
            s_ub[2 * e] = bu ? bv.u_min[e] : -INFINITY; s_ub[2 * e + 1] = bu ? bv.u_max[e] : INFINITY;
        }
    }
    // shared references, stored as the terms the linear cost uses: cq = -(Xref .* Qd), cr = -(Uref .* Rd),
    // tq = -(Xref_{N-1}' Pinf)'   (reference admm.cpp:218, :228, :238)
    double* s_cq = smem + dp.off_refs;                // [N][NX]
    double* s_cr = s_cq + nxN;                        // [N-1][NU]
    double* s_tq = s_cr + nuN;                        // [NX]
    if (REFS == 1) {
        for (int e = threadIdx.x; e < nxN; e += blockDim.x) s_cq[e] = -dmul(bv.Xref ? bv.Xref[e] : 0.0, bv.Qd[e % NX]);
        for (int e = threadIdx.x; e < nuN; e += blockDim.x) s_cr[e] = -dmul(bv.Uref ? bv.Uref[e] : 0.0, bv.Rd[e % NU]);
        for (int row = threadIdx.x; row < NX; row += blockDim.x) {
            double s = 0.0;
            if (bv.Xref) for (int k = 0; k < NX; ++k) s = fma(bv.Xref[(N - 1) * NX + k], bv.Pinf[row * NX + k], s);
            s_tq[row] = -s;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");

    // ---------------- per-warp carve-up
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inst = lane & 7, sub = lane >> 3;
    const bool urow = (NU == 4) || (sub < NU);        // this lane owns an input row
    const int usub = urow ? sub : 0;
    const uint32_t tbase = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) +
                           static_cast<uint32_t>((warp >> 2) * kDTmemColsPerWarp);
    double* wbase = smem + dp.off_warp0 + static_cast<size_t>(warp) * dp.warp_stride;
    double* s_tu = wbase;                             // [N-1][32 lanes][group] : pre-clamp t = u + y_old of the lane's input row
    double* s_rt = s_tu + (N - 1) * 64;               // [N-1][group][2 halves][8 inst][2] : r_i (after a forward sweep) / t_i (after
                                                      // a backward sweep); rows (0,1) and (2,3) of the 8 instances of a group are
                                                      // each 128 contiguous bytes: the 4-vector is two conflict-free LDS.128
    double* s_bc = s_rt + (N - 1) * 64;               // 2 x group x 8 x XP broadcast slots

    // 32-bit shared addresses of everything the sweeps touch (byte strides per knot point in the comments)
    const uint32_t a_tu = smem_addr(s_tu) + lane * 16;                               // + i * 512
    const uint32_t a_rt4 = smem_addr(s_rt) + inst * 16;                              // + i * 512 ; group: + 256 ; rows 2,3: + 128
    const uint32_t a_myrt = smem_addr(s_rt) + ((sub >> 1) * 16 + inst * 2 + (sub & 1)) * 8;   // + i * 512 ; group: + 256
    const uint32_t a_slotA = smem_addr(s_bc) + inst * XP * 8;                        // group: + GS
    const uint32_t a_slot_sum = 2 * a_slotA + kG * GS;                               // slot B = sum - slot A
    const uint32_t a_ub = smem_addr(s_ub) + usub * 16;                               // + i * NU * 16
    const uint32_t a_xb = smem_addr(s_xb) + sub * 16;                                // + i * NX * 16 ; row r: + r * 64
    const uint32_t a_sc = smem_addr(sc);
    const uint32_t a_Mb = a_sc + (CL::B_rm + sub * NU) * 8;                          // row r: + r * 4 * NU * 8
    const uint32_t a_Mq = a_sc + (CL::Quu_rm + usub * NU) * 8;
    const uint32_t a_Mkt = a_sc + (CL::Kinf_cm + sub * NU) * 8;                      // row r: + r * 4 * NU * 8
    const uint32_t a_cq = smem_addr(s_cq) + sub * 8;                                 // + i * NX * 8 ; row r: + r * 32
    const uint32_t a_cr = smem_addr(s_cr) + usub * 8;                                // + i * NU * 8
    uint32_t a_slot = a_slotA;

    const double fr_BPf = urow ? sc[CL::BPf + usub] : 0.0;
    const double fr_Rd = urow ? sc[CL::Rd + usub] : 0.0;
    double fr_f[RX], fr_APf[RX], fr_Qd[RX];
#pragma unroll
    for (int r = 0; r < RX; ++r) { fr_f[r] = sc[CL::f + r * 4 + sub]; fr_APf[r] = sc[CL::APf + r * 4 + sub]; fr_Qd[r] = sc[CL::Qd + r * 4 + sub]; }

    // ---------------- per-lane instance state, one per group
    int b[kG], iter[kG], status[kG];
    bool valid[kG], first[kG], fin[kG];               // fin = true: the sub-warp wants a (new) instance in this group
    double x0r[kG][RX], p_term[kG][RX];
    // carried state of an instance: one contiguous block (instance-major):
    //   knot i < N    : + 3 NX i : g(NX), v(NX), q(NX)
    //   knot i < N - 1: IB + 3 NU i : y(NU), z(NU), r(NU)          PB: terminal p (NX)
    const int IB = 3 * NX * N, PB = IB + 3 * NU * (N - 1);
    double* sblk[kG];                                 // block of the lane's instance, + sub
    const double* xref_g[kG];                         // REFS == 2 only
    const double* uref_g[kG];
#pragma unroll
    for (int g = 0; g < kG; ++g) {
        b[g] = 0; iter[g] = 0; status[g] = 1; valid[g] = false; first[g] = true; fin[g] = true;
        sblk[g] = bv.state_im + sub; xref_g[g] = nullptr; uref_g[g] = nullptr;
#pragma unroll
        for (int r = 0; r < RX; ++r) { x0r[g][r] = 0.0; p_term[g][r] = 0.0; }
    }

    for (;;) {
        // =====================================================================================================
        // write-out of finished instances + refill from the queue.  The Tensor Memory part is warp-uniform
        // (tcgen05.ld/st are .aligned): every lane walks the knot points; lanes whose instance is not finished
        // ignore what they loaded (cold) or store it back unchanged (warm).
        // =====================================================================================================
        if (__any_sync(kAll, fin[0] || fin[1])) {
            bool wr[kG], unconv[kG], nvalid[kG], warm[kG];
            int nb[kG];
            double* nblk[kG];
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                wr[g] = fin[g] && valid[g];               // results to write
                unconv[g] = (status[g] != 0);
                nb[g] = 0x7fffffff;
                if (fin[g] && sub == 0) nb[g] = atomicAdd(dp.counter, 1);
                nb[g] = __shfl_sync(kAll, nb[g], inst);
                nvalid[g] = fin[g] && nb[g] < bv.B;
                warm[g] = !COLD && nvalid[g];
                nblk[g] = bv.state_im + (nvalid[g] ? static_cast<size_t>(nb[g]) : 0) * static_cast<size_t>(bv.state_im_stride) + sub;
            }
            const bool any_wr = __any_sync(kAll, wr[0] || wr[1]);

            // ---- state rows: solution, duals, slacks and q of the finished instances; carried g of the new ones
            if (any_wr || !COLD) {
                uint32_t w[W];
                tm_load<W>(tbase, w);
                uint32_t o_xb = 0;
                size_t o_x = 0;                           // NX * i
                for (int i = 0; i < N; ++i) {
                    double lo[RX], hi[RX];
#pragma unroll
                    for (int r = 0; r < RX; ++r) xlds2(a_xb + o_xb + r * 64, lo[r], hi[r]);
                    double gn[kG][RX];
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
#pragma unroll
                        for (int r = 0; r < RX; ++r) gn[g][r] = 0.0;
                        if (warm[g]) {
#pragma unroll
                            for (int r = 0; r < RX; ++r) gn[g][r] = nblk[g][3 * o_x + r * 4];
                        }
                    }
                    tm_wait_ld<W>(w);
                    uint32_t wc[W];
#pragma unroll
                    for (int k = 0; k < W; ++k) wc[k] = w[k];
                    if (i + 1 < N) tm_load<W>(tbase + (i + 1) * W, w);
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        if (wr[g]) {
                            double* Xs = bv.Xsol + static_cast<size_t>(b[g]) * nxN + sub + o_x;
                            double* op = sblk[g] + 3 * o_x;       // g +0, v +NX, q +2 NX
#pragma unroll
                            for (int r = 0; r < RX; ++r) {
                                const double t = mk_double(wc[g * WG + 2 * r], wc[g * WG + 2 * r + 1]);
                                const double vv = clampd(t, lo[r], hi[r]);
                                const double gg = dsub(t, vv);
                                Xs[r * 4] = vv;                   // solution = vnew of the last iteration (admm.cpp:370,386)
                                op[r * 4] = gg;
                                if (unconv[g]) op[NX + r * 4] = vv;   // v = vnew happened (admm.cpp:379); else the lagging copy is there
                                if (i < N - 1) {                  // q as update_linear_cost left it (admm.cpp:218-219)
                                    double q;
                                    if (REFS == 0) q = dmul(nrho, dsub(vv, gg));
                                    else {
                                        double cq;
                                        if (REFS == 1) cq = s_cq[i * NX + r * 4 + sub];
                                        else cq = -dmul(xref_g[g] ? __ldg(xref_g[g] + o_x + r * 4) : 0.0, fr_Qd[r]);
                                        q = dsub(cq, dmul(rho, dsub(vv, gg)));
                                    }
                                    op[2 * NX + r * 4] = q;
                                }
                            }
                        }
                        if (!COLD && fin[g]) {                    // a warm-started instance enters with its carried dual g in the t slot
#pragma unroll
                            for (int r = 0; r < RX; ++r) { wc[g * WG + 2 * r] = lo32(gn[g][r]); wc[g * WG + 2 * r + 1] = hi32(gn[g][r]); }
                        }
                    }
                    if (!COLD) tm_store<W>(tbase + i * W, wc);
                    o_xb += NX * 16; o_x += NX;
                }
                if (!COLD) tm_wait_st();
            }

            // ---- input rows (lane-private shared-memory slots) and the scalars
            if (urow && (fin[0] || fin[1])) {
                for (int i = 0; i < N - 1; ++i) {
                    double tu[kG], ulo, uhi;
                    xlds2(a_tu + i * 512, tu[0], tu[1]);
                    xlds2(a_ub + i * NU * 16, ulo, uhi);
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        if (wr[g]) {
                            const double zz = clampd(tu[g], ulo, uhi);
                            const double yy = dsub(tu[g], zz);
                            double* ou = sblk[g] + IB + 3 * NU * i;       // y +0, z +NU, r +2 NU
                            bv.Usol[static_cast<size_t>(b[g]) * nuN + sub + NU * i] = zz;   // solution = znew of the last iteration (admm.cpp:371,387)
                            ou[0] = yy;
                            if (unconv[g]) ou[NU] = zz;           // z = znew happened (admm.cpp:380); else the lagging copy is there
                            ou[2 * NU] = xlds1(a_myrt + i * 512 + g * 256);   // r as update_linear_cost left it (admm.cpp:228-233)
                        }
                        if (fin[g]) {
                            const double* nu_p = nblk[g] + IB + 3 * NU * i;
                            xsts1(a_myrt + i * 512 + g * 256, warm[g] ? nu_p[2 * NU] : 0.0);
                            if (!COLD) tu[g] = warm[g] ? nu_p[0] : 0.0;   // carried dual y in the t slot (first-sweep select)
                        }
                    }
                    if (!COLD) xsts2(a_tu + i * 512, tu[0], tu[1]);
                }
            }
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                if (wr[g]) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) sblk[g][PB + r * 4] = p_term[g][r];
                    if (sub == 0) { bv.status[b[g]] = status[g]; bv.iters[b[g]] = iter[g]; }
                }
                if (fin[g]) {
                    const size_t nbb = nvalid[g] ? static_cast<size_t>(nb[g]) : 0;
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        x0r[g][r] = nvalid[g] ? bv.x0[nbb * NX + r * 4 + sub] : 0.0;
                        p_term[g][r] = warm[g] ? nblk[g][PB + r * 4] : 0.0;
                    }
                    b[g] = nvalid[g] ? nb[g] : 0;
                    sblk[g] = nblk[g];
                    valid[g] = nvalid[g]; iter[g] = 0; first[g] = true; status[g] = 1; fin[g] = false;
                    if (REFS == 2) {
                        xref_g[g] = bv.Xref ? bv.Xref + nbb * bv.xref_stride + sub : nullptr;            // + i NX + 4 r
                        uref_g[g] = (bv.Uref && urow) ? bv.Uref + nbb * bv.uref_stride + sub : nullptr;   // + i NU
                    }
                }
            }
            __syncwarp();
        }
        if (!__any_sync(kAll, valid[0] || valid[1])) break;

        // =====================================================================================================
        // backward sweep: p_i for i = N-2 .. 0, leaving t_i = B' p_{i+1} + r_i + BPf for the forward sweep
        // (reference admm.cpp:13-20; d_i = Quu_inv t_i is formed where it is consumed)
        // =====================================================================================================
        {
            double Mab[RX][NX], Mbt[NX];
#pragma unroll
            for (int r = 0; r < RX; ++r) {
#pragma unroll
                for (int k = 0; k < NX; k += 2) ld2(sc + CL::AmBKt_rm + (r * 4 + sub) * NX + k, Mab[r][k], Mab[r][k + 1]);
            }
#pragma unroll
            for (int k = 0; k < NX; ++k) Mbt[k] = urow ? sc[CL::B_cm + usub * NX + k] : 0.0;

            double p[kG][RX];
            double qcar[kG][RX];                          // carried q of the first sweep of a warm-started solve
            bool ldq[kG];
            double t_def[kG];
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                t_def[g] = 0.0;
                ldq[g] = !COLD && first[g] && valid[g];
#pragma unroll
                for (int r = 0; r < RX; ++r) { p[g][r] = p_term[g][r]; qcar[g][r] = 0.0; }
                if (ldq[g]) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) qcar[g][r] = sblk[g][2 * NX + (N - 2) * 3 * NX + r * 4];
                }
            }
            uint32_t w[W];
            tm_load<W>(tbase + (N - 2) * W, w);
            uint32_t o_rt = (N - 2) * 512;                // byte offset of knot i in the r/t array
            uint32_t o_cq = (N - 2) * NX * 8;
            uint32_t o_xb = (N - 2) * NX * 16;

            for (int i = N - 2; i >= 0; --i) {
                double cq[kG][RX];
                if (REFS == 2) {
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
#pragma unroll
                        for (int r = 0; r < RX; ++r) cq[g][r] = -dmul(xref_g[g] ? __ldg(xref_g[g] + i * NX + r * 4) : 0.0, fr_Qd[r]);
                    }
                }
                // broadcast p_{i+1} to the 4 lanes of each instance
                a_slot = a_slot_sum - a_slot;
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) xsts1(a_slot + sub * 8 + g * GS + r * 32, p[g][r]);
                }
                __syncwarp();
                if (i < N - 2 && urow) {                  // t_{i+1}: every lane has read r_{i+1} by now
#pragma unroll
                    for (int g = 0; g < kG; ++g) xsts1(a_myrt + o_rt + 512 + g * 256, t_def[g]);
                }
                double pb[kG][NX], rr[kG][4], r_own[kG];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int k = 0; k < NX; k += 2) xlds2(a_slot + g * GS + k * 8, pb[g][k], pb[g][k + 1]);
                    xlds2(a_rt4 + o_rt + g * 256, rr[g][0], rr[g][1]);
                    xlds2(a_rt4 + o_rt + g * 256 + 128, rr[g][2], rr[g][3]);
                    r_own[g] = xlds1(a_myrt + o_rt + g * 256);
                }
                double lo[RX], hi[RX], cqs[RX], Mkt[RX][4];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    xlds2(a_xb + o_xb + r * 64, lo[r], hi[r]);
                    if (REFS == 1) cqs[r] = xlds1(a_cq + o_cq + r * 32);
                    if (NU == 4) {
                        xlds2(a_Mkt + r * 4 * NU * 8, Mkt[r][0], Mkt[r][1]);
                        xlds2(a_Mkt + r * 4 * NU * 8 + 16, Mkt[r][2], Mkt[r][3]);
                    } else {
#pragma unroll
                        for (int k = 0; k < NU; ++k) Mkt[r][k] = sc[CL::Kinf_cm + (r * 4 + sub) * NU + k];
                    }
                }

                // q_i (own rows): carried on the first sweep of a solve, otherwise rebuilt from the pre-clamp value
                tm_wait_ld<W>(w);
                double q_own[kG][RX];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const double t = mk_double(w[g * WG + 2 * r], w[g * WG + 2 * r + 1]);
                        const double vv = clampd(t, lo[r], hi[r]);
                        const double gg = dsub(t, vv);
                        double q;
                        if (REFS == 0) q = dmul(nrho, dsub(vv, gg));
                        else q = dsub(REFS == 1 ? cqs[r] : cq[g][r], dmul(rho, dsub(vv, gg)));
                        q_own[g][r] = first[g] ? qcar[g][r] : q;
                    }
                }
                if (i > 0) {
                    tm_load<W>(tbase + (i - 1) * W, w);
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        if (ldq[g]) {
#pragma unroll
                            for (int r = 0; r < RX; ++r) qcar[g][r] = sblk[g][2 * NX + (i - 1) * 3 * NX + r * 4];
                        }
                    }
                }

                // own row of B' p_{i+1} and the lane's rows of AmBKt p_{i+1}: 2 (RX + 1) independent chains, interleaved
                double s[kG], s1[kG][RX];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
                    s[g] = 0.0;
#pragma unroll
                    for (int r = 0; r < RX; ++r) s1[g][r] = 0.0;
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) {
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        s[g] = fma(Mbt[k], pb[g][k], s[g]);
#pragma unroll
                        for (int r = 0; r < RX; ++r) s1[g][r] = fma(Mab[r][k], pb[g][k], s1[g][r]);
                    }
                }
#pragma unroll
                for (int g = 0; g < kG; ++g) {
                    t_def[g] = dadd(dadd(s[g], r_own[g]), fr_BPf);     // t_i = B' p_{i+1} + r_i + BPf
                    // p_i = q_i + AmBKt p_{i+1} - Kinf' r_i + APf
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        double s2 = 0.0;
#pragma unroll
                        for (int k = 0; k < NU; ++k) s2 = fma(Mkt[r][k], rr[g][k], s2);
                        p[g][r] = dadd(dsub(dadd(q_own[g][r], s1[g][r]), s2), fr_APf[r]);
                    }
                }
                o_rt -= 512; o_cq -= NX * 8; o_xb -= NX * 16;
            }
            __syncwarp();
            if (urow) {
#pragma unroll
                for (int g = 0; g < kG; ++g) xsts1(a_myrt + g * 256, t_def[g]);
            }
        }

        // =====================================================================================================
        // forward sweep fused with slack / dual / linear-cost / residual updates
        // (reference admm.cpp:25-32, :84-98, :184-187, :218-239, :257-265)
        // =====================================================================================================
        bool ok[kG];                 // every residual this lane has seen in this sweep is under its tolerance
        {
            // register-resident: the lane's rows of A and Kinf (48 doubles), shared by the two instances of the lane.
            // B and Quu_inv rows are re-read from shared memory every knot point (8 broadcast LDS.128 per PAIR of
            // instances, off the critical path).
            double Ma[RX][NX], Mk[NX];
#pragma unroll
            for (int r = 0; r < RX; ++r) {
#pragma unroll
                for (int k = 0; k < NX; k += 2) ld2(sc + CL::A_rm + (r * 4 + sub) * NX + k, Ma[r][k], Ma[r][k + 1]);
            }
#pragma unroll
            for (int k = 0; k < NX; ++k) Mk[k] = urow ? sc[CL::Kinf_rm + usub * NX + k] : 0.0;

            double x[kG][RX];
            bool lag[kG];
            double r_def[kG], vg[kG][RX];
            // first sweep of a warm-started instance: the carried previous slacks v, z come from the instance's
            // block in HBM, one knot point ahead (the carried duals g, y sit in the t slots, see the refill)
            double vin[kG][RX], zin[kG];
#pragma unroll
            for (int g = 0; g < kG; ++g) {
                ok[g] = true; r_def[g] = 0.0; zin[g] = 0.0;
                // the carried "previous slacks" must lag the solution by one iteration when the solve converges
                // (admm.cpp:375 returns before :379-380): while this sweep can still end in convergence the lane parks
                // them in the instance's carried-state block in HBM ([pointer + offset] stores, whole 32-byte sectors)
                lag[g] = valid[g] && ((iter[g] + 1) % st.check_termination == 0);
#pragma unroll
                for (int r = 0; r < RX; ++r) { x[g][r] = x0r[g][r]; vin[g][r] = 0.0; vg[g][r] = 0.0; }
                if (!COLD && first[g] && valid[g]) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) vin[g][r] = sblk[g][NX + r * 4];
                    if (urow) zin[g] = sblk[g][IB + NU];
                }
            }
            uint32_t w[W];
            tm_load<W>(tbase, w);
            uint32_t o_xb = 0;                                                // byte offset of knot i in the state-bound table
            uint32_t tcol = tbase;
            int o_v = NX, o_z = IB + NU;                                      // v / z of knot i in the instance's block (doubles)

            // Residual tests are evaluated lazily: as soon as every instance of the warp has failed one test of
            // this sweep (warp vote after each knot point), or when this iteration cannot terminate the solve
            // (check_termination), the rest of the sweep runs the variant without tests and without lag stores.
            // The decision per instance is unchanged: a failed lane keeps ok = false.

            // state column: slack, dual, [residuals]; new pre-clamp values back to Tensor Memory
            auto state_column = [&](auto chk, auto term) {
                constexpr bool CHK = decltype(chk)::value;
                constexpr bool TERM = decltype(term)::value;
                double lo[RX], hi[RX];
#pragma unroll
                for (int r = 0; r < RX; ++r) xlds2(a_xb + o_xb + r * 64, lo[r], hi[r]);
                tm_wait_ld<W>(w);
                uint32_t wn[W];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        const double told = mk_double(w[g * WG + 2 * r], w[g * WG + 2 * r + 1]);
                        double vold = clampd(told, lo[r], hi[r]);
                        double gold = dsub(told, vold);
                        if (first[g]) { gold = COLD ? 0.0 : told; vold = COLD ? 0.0 : vin[g][r]; }
                        const double t = dadd(x[g][r], gold);
                        if (CHK || TERM) {
                            const double vnew = clampd(t, lo[r], hi[r]);
                            if (CHK) {
                                const bool p1 = fabs(dsub(x[g][r], vnew)) < dp.pri_thresh;
                                const bool p2 = fabs(dsub(vold, vnew)) <= dp.dua_thresh;
                                ok[g] = ok[g] & p1 & p2;
                                if (lag[g] & ok[g]) sblk[g][o_v + r * 4] = vold;
                            }
                            if (TERM) vg[g][r] = dsub(vnew, dsub(t, vnew));
                        }
                        wn[g * WG + 2 * r] = lo32(t); wn[g * WG + 2 * r + 1] = hi32(t);
                    }
                    if (!COLD && CHK && !TERM && first[g] && valid[g]) {
#pragma unroll
                        for (int r = 0; r < RX; ++r) vin[g][r] = sblk[g][o_v + 3 * NX + r * 4];
                    }
                }
                tm_store<W>(tcol, wn);
            };

            uint32_t o_rt = 0, o_tu = 0, o_ub = 0, o_cr = 0;
            int i = 0;
            // one non-terminal knot point
            auto knot = [&](auto chk) {
                constexpr bool CHK = decltype(chk)::value;
                double cr[kG];
                if (REFS == 2) {
#pragma unroll
                    for (int g = 0; g < kG; ++g) cr[g] = -dmul(uref_g[g] ? __ldg(uref_g[g] + i * NU) : 0.0, fr_Rd);
                }
                a_slot = a_slot_sum - a_slot;
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) xsts1(a_slot + sub * 8 + g * GS + r * 32, x[g][r]);
                }
                __syncwarp();
                if (i > 0 && urow) {                                         // r_{i-1}: every lane has read t_{i-1} by now
#pragma unroll
                    for (int g = 0; g < kG; ++g) xsts1(a_myrt + o_rt - 512 + g * 256, r_def[g]);
                }
                double xb[kG][NX], tt[kG][4], mq[4];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int k = 0; k < NX; k += 2) xlds2(a_slot + g * GS + k * 8, xb[g][k], xb[g][k + 1]);
                    xlds2(a_rt4 + o_rt + g * 256, tt[g][0], tt[g][1]);
                    xlds2(a_rt4 + o_rt + g * 256 + 128, tt[g][2], tt[g][3]);
                }
                if (NU == 4) { xlds2(a_Mq, mq[0], mq[1]); xlds2(a_Mq + 16, mq[2], mq[3]); }
                else {
#pragma unroll
                    for (int k = 0; k < NU; ++k) mq[k] = sc[CL::Quu_rm + usub * NU + k];
                }
                double tu[kG], ulo = 0.0, uhi = 0.0, crs = 0.0;
                tu[0] = 0.0; tu[1] = 0.0;
                if (urow) {
                    xlds2(a_tu + o_tu, tu[0], tu[1]);
                    xlds2(a_ub + o_ub, ulo, uhi);
                    if (REFS == 1) crs = xlds1(a_cr + o_cr);
                }

                // d_i = Quu_inv t_i ;  Kinf x_i (own row) ;  A x_i (own rows): 2 (RX + 2) independent chains, interleaved
                double d[kG], s[kG], s1[kG][RX];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
                    d[g] = 0.0; s[g] = 0.0;
#pragma unroll
                    for (int k = 0; k < NU; ++k) d[g] = fma(mq[k], tt[g][k], d[g]);
#pragma unroll
                    for (int r = 0; r < RX; ++r) s1[g][r] = 0.0;
                }
#pragma unroll
                for (int k = 0; k < NX; ++k) {
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        s[g] = fma(Mk[k], xb[g][k], s[g]);
#pragma unroll
                        for (int r = 0; r < RX; ++r) s1[g][r] = fma(Ma[r][k], xb[g][k], s1[g][r]);
                    }
                }

                SCHED_FENCE();
                // u_i = -Kinf x_i - d_i, broadcast at once; then input slack, dual, [residuals]
                // (admm.cpp:28, :87, :97, :187, :259-260)
                double u_own[kG], ub[kG][NU];
#pragma unroll
                for (int g = 0; g < kG; ++g) {
                    u_own[g] = urow ? dsub(-s[g], d[g]) : 0.0;
#pragma unroll
                    for (int k = 0; k < NU; ++k) ub[g][k] = shfl_dq(u_own[g], k * kDInst + inst);
                }
                // rows of B for the tail of this knot point: issued now, consumed after the state column
                double mb[RX][4];
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (NU == 4) {
                        xlds2(a_Mb + r * 4 * NU * 8, mb[r][0], mb[r][1]);
                        xlds2(a_Mb + r * 4 * NU * 8 + 16, mb[r][2], mb[r][3]);
                    } else {
#pragma unroll
                        for (int k = 0; k < NU; ++k) mb[r][k] = sc[CL::B_rm + (r * 4 + sub) * NU + k];
                    }
                }
                if (urow) {
                    double tn[kG];
#pragma unroll
                    for (int g = 0; g < kG; ++g) {
                        double zold = clampd(tu[g], ulo, uhi);
                        double yold = dsub(tu[g], zold);
                        if (first[g]) { yold = COLD ? 0.0 : tu[g]; zold = COLD ? 0.0 : zin[g]; }
                        const double t = dadd(u_own[g], yold);
                        const double znew = clampd(t, ulo, uhi);
                        const double ynew = dsub(t, znew);                  // y + u - znew
                        if (CHK) {
                            const bool p1 = fabs(dsub(u_own[g], znew)) < dp.pri_thresh;
                            const bool p2 = fabs(dsub(zold, znew)) <= dp.dua_thresh;
                            ok[g] = ok[g] & p1 & p2;
                            if (lag[g] & ok[g]) sblk[g][o_z] = zold;
                            if (!COLD && first[g] && valid[g] && i + 1 < N - 1) zin[g] = sblk[g][o_z + 3 * NU];
                        }
                        tn[g] = t;
                        // r_i = -(Uref .* Rd) - rho (znew - y)    (admm.cpp:228-229)
                        if (REFS == 0) r_def[g] = dmul(nrho, dsub(znew, ynew));
                        else r_def[g] = dsub(REFS == 1 ? crs : cr[g], dmul(rho, dsub(znew, ynew)));
                    }
                    xsts2(a_tu + o_tu, tn[0], tn[1]);
                }

                state_column(chk, std::false_type{});         // independent of u_i: overlaps the shuffles
                tcol += W;
                tm_load<W>(tcol, w);                          // next knot point; consumed after the next broadcast

                // x_{i+1} = A x_i + B u_i + f
#pragma unroll
                for (int g = 0; g < kG; ++g) {
#pragma unroll
                    for (int r = 0; r < RX; ++r) {
                        double s2 = 0.0;
#pragma unroll
                        for (int k = 0; k < NU; ++k) s2 = fma(mb[r][k], ub[g][k], s2);
                        x[g][r] = dadd(dadd(s1[g][r], s2), fr_f[r]);
                    }
                }
                o_v += 3 * NX; o_z += 3 * NU;
                o_rt += 512; o_tu += 512; o_ub += NU * 16; o_cr += NU * 8; o_xb += NX * 16;
            };

            bool checking = __any_sync(kAll, lag[0] || lag[1]);   // no instance of the warp can terminate: no tests at all
            if (!checking) { ok[0] = false; ok[1] = false; }
            while (checking && i < N - 1) {
                knot(std::true_type{});
                ++i;
                const unsigned v0 = __ballot_sync(kAll, ok[0] && lag[0]);
                const unsigned v1 = __ballot_sync(kAll, ok[1] && lag[1]);
                const unsigned vv = (v0 & (v0 >> 8) & (v0 >> 16) & (v0 >> 24)) | (v1 & (v1 >> 8) & (v1 >> 16) & (v1 >> 24));
                checking = (vv & 0xffu) != 0;                 // some instance has all 4 lanes ok
            }
            for (; i < N - 1; ++i) knot(std::false_type{});
            // terminal knot point: p_{N-1} = -(Xref_{N-1}' Pinf)' - rho (vnew - g)   (admm.cpp:238-239)
            if (checking) state_column(std::true_type{}, std::true_type{});
            else state_column(std::false_type{}, std::true_type{});
#pragma unroll
            for (int g = 0; g < kG; ++g) {
#pragma unroll
                for (int r = 0; r < RX; ++r) {
                    if (REFS == 0) p_term[g][r] = dmul(nrho, vg[g][r]);
                    else if (REFS == 1) p_term[g][r] = dsub(s_tq[r * 4 + sub], dmul(rho, vg[g][r]));
                    else {
                        double s = 0.0;
                        if (xref_g[g]) for (int k = 0; k < NX; ++k) s = fma(__ldg(xref_g[g] - sub + (N - 1) * NX + k), sc[CL::Pinf_cm + (r * 4 + sub) * NX + k], s);
                        p_term[g][r] = dsub(-s, dmul(rho, vg[g][r]));
                    }
                }
            }
            tm_wait_st();
            __syncwarp();
            if (urow) {
#pragma unroll
                for (int g = 0; g < kG; ++g) xsts1(a_myrt + o_rt - 512 + g * 256, r_def[g]);   // r_{N-2}
            }
        }

        // =====================================================================================================
        // termination (admm.cpp:253-271, :364-388), per instance
        // =====================================================================================================
        const unsigned mine = 0x01010101u << inst;
#pragma unroll
        for (int g = 0; g < kG; ++g) {
            iter[g] += 1;
            first[g] = false;
            const unsigned votes = __ballot_sync(kAll, ok[g]);
            const bool conv = valid[g] && (iter[g] % st.check_termination == 0) && ((votes & mine) == mine);
            status[g] = conv ? 0 : 1;
            fin[g] = valid[g] && (conv || iter[g] >= st.max_iter);
        }
    }

    __syncthreads();
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(s_tmem_base));
    }
}

// Largest a >= 0 with fl(rho * a) < tol, so that  rho * max|.| < tol  <=>  every |.| <= a.
double dual_threshold_d(double rho, double tol) {
    if (!(tol > 0.0) || !(rho > 0.0)) return -1.0;
    double a = tol / rho;
    while (rho * a < tol) a = std::nextafter(a, INFINITY);
    while (!(rho * a < tol)) a = std::nextafter(a, -INFINITY);
    return a;
}

struct DPlan { size_t smem_bytes; DualParams dp; };

template <int NX, int NU>
bool make_dplan(int N, int refs, DPlan& pl) {
    constexpr int XP = QPad<NX>::value;
    constexpr int W = kG * 2 * (NX / 4);
    if (N < 2 || N * W > kDTmemColsPerWarp) return false;          // t of a lane's two instances must fit its 256 TMEM columns
    const int nxN = NX * N, nuN = NU * (N - 1);
    DualParams dp{};
    int off = 0;
    dp.off_consts = off; off += QConst<NX, NU>::size;
    dp.off_bounds = off; off += 2 * (nxN + nuN);
    dp.off_refs = off; if (refs == 1) off += nxN + nuN + NX;
    off = (off + 1) / 2 * 2;
    dp.off_warp0 = off;
    int per_warp = (N - 1) * 128 + 2 * kG * kDInst * XP;
    per_warp = (per_warp + 1) / 2 * 2;
    dp.warp_stride = per_warp;
    pl.smem_bytes = (static_cast<size_t>(off) + static_cast<size_t>(per_warp) * kDWarps) * sizeof(double);
    pl.dp = dp;
    return pl.smem_bytes <= 227 * 1024;
}

int refs_mode_d(const BatchView& bv) {
    if (bv.Xref == nullptr && bv.Uref == nullptr) return 0;
    const bool xs = bv.Xref == nullptr || bv.xref_stride == 0, us = bv.Uref == nullptr || bv.uref_stride == 0;
    return (xs && us) ? 1 : 2;
}

template <int NX, int NU, bool COLD, int REFS>
cudaError_t launch_d(const BatchView& bv, const AdmmSettings& st, const DPlan& pl, int num_sms, cudaStream_t stream) {
    auto kern = admm_tmem2_kernel<NX, NU, COLD, REFS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(pl.smem_bytes));
    if (e != cudaSuccess) return e;
    const int per_cta = kDWarps * kDInst * kG;
    int grid = (bv.B + per_cta - 1) / per_cta;
    if (grid > num_sms) grid = num_sms;            // persistent: one CTA per SM, sub-warps pull instances from the queue
    kern<<<grid, kDWarps * 32, pl.smem_bytes, stream>>>(bv, st, pl.dp);
    return cudaGetLastError();
}

template <int NX, int NU>
cudaError_t launch_shape_d(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                           cudaStream_t stream) {
    DPlan pl;
    const int refs = refs_mode_d(bv);
    if (!make_dplan<NX, NU>(bv.N, refs, pl)) return cudaErrorInvalidConfiguration;
    pl.dp.counter = counter;
    pl.dp.pri_thresh = st.abs_pri_tol;
    pl.dp.dua_thresh = dual_threshold_d(bv.rho, st.abs_dua_tol);
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    if (cold) {
        if (refs == 0) return launch_d<NX, NU, true, 0>(bv, st, pl, num_sms, stream);
        if (refs == 1) return launch_d<NX, NU, true, 1>(bv, st, pl, num_sms, stream);
        return launch_d<NX, NU, true, 2>(bv, st, pl, num_sms, stream);
    }
    if (refs == 0) return launch_d<NX, NU, false, 0>(bv, st, pl, num_sms, stream);
    if (refs == 1) return launch_d<NX, NU, false, 1>(bv, st, pl, num_sms, stream);
    return launch_d<NX, NU, false, 2>(bv, st, pl, num_sms, stream);
}

}  // namespace

bool tmem2_kernel_supports(const BatchView& bv, const AdmmSettings& st) {
    if (st.en_state_soc || st.en_input_soc || st.en_state_linear || st.en_input_linear) return false;
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return false;
    if (bv.x_min != nullptr && (bv.xb_stride != 0 || bv.ub_stride != 0)) return false;   // per-instance bounds: other kernels
    DPlan pl;
    if (bv.nx == 12 && bv.nu == 4) return make_dplan<12, 4>(bv.N, refs_mode_d(bv), pl);
    if (bv.nx == 4 && bv.nu == 1) return make_dplan<4, 1>(bv.N, refs_mode_d(bv), pl);
    return false;
}

cudaError_t launch_admm_tmem2(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                              cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (bv.nx == 12 && bv.nu == 4) return launch_shape_d<12, 4>(bv, st, cold, counter, num_sms, stream);
    if (bv.nx == 4 && bv.nu == 1) return launch_shape_d<4, 1>(bv, st, cold, counter, num_sms, stream);
    return cudaErrorNotSupported;
}

}  // namespace tmpc
