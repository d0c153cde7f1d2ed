// ADMM solve kernel for problems with second-order cones and half-spaces (config 4: rocket landing nx = 6, nu = 3,
// N = 50), all ADMM state resident on chip.
//
// Why a different mapping than admm_tmem.cu.  With cones and half-spaces update_slack / update_dual / update_linear_cost
// (reference admm.cpp:81-247) dominate: per iteration the rocket runs ~9,400 FP64 instructions of Riccati recursion but
// ~19,000 of elementwise projection work (three slack / dual pairs per element, a square root and two divisions per
// cone).  The elementwise part is independent across knot points, the recursions are sequential in the horizon.  So an
// iteration runs in two phases inside one warp that owns a tile of kCInst = 3 instances:
//
//   phase A  backward_pass_grad + forward_pass (admm.cpp:13-32): lane = matrix row.  10 lanes per instance, lanes 0-5
//            own the state rows, 6-8 the input rows; the row a lane multiplies with is A / Kinf (forward) or AmBKt / B'
//            (backward), so ONE 6-FMA chain per knot point serves both kinds of lane.  Inputs q_i, r_i and outputs
//            x_i, u_i go through a per-warp shared-memory exchange array of one 80-byte record per knot point (x
//            overwrites q, u overwrites the t = B'p + r + BPf that overwrote r); the same records all-gather the
//            vectors inside the sweeps (no shuffles).
//   phase B  update_slack (box clamp, project_soc, sequential project_hyperplane), update_dual, update_linear_cost and
//            the residual tests (admm.cpp:81-271): lane = knot point.  The 3 x N (instance, knot) tasks of the tile are
//            dealt to the 32 lanes; a task holds its whole knot point in registers, so cone and half-space projections
//            need no cross-lane traffic and every lane does useful FP64 work.
//
// Residency.  A task's state (g, v, gc, gl: 4 x 6; y, z, yc, yl: 4 x 3 doubles = 72 32-bit words) lives in the Tensor
// Memory row of the lane that always processes it (tcgen05.ld/st .32x32b as lane-private scratch; 8 warps x 3 slots of
// 72 columns = 96 tasks per warp), the remaining 3 N - 96 tasks in shared memory (37-double pitch): 24 instances per SM stay
// on chip for the whole solve (224 KB shared + 216 of 256 TMEM columns per warp at N = 50), HBM is touched at install and
// write-out only; shared references, knot-invariant bounds and the half-space rows are staged once per CTA.  The
// thread-per-instance kernel streams 31 KB per instance-iteration through L2 / HBM instead.
//
// Arithmetic: the same sequence of correctly rounded operations as admm_thread.cu (ascending-k FMA chains, explicitly
// rounded elementwise updates, project_soc3 with the reference's float mu / norm), so both kernels agree bit for bit
// and carried state (element-major arrays, admm_device.cuh) can be handed from one to the other.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>

#include "admm_device.cuh"
#include "admm_kernels.hpp"
#include "tmem_util.cuh"

namespace tmpc {
namespace {

using namespace tmu;

#ifndef TMPC_CONE_WARPS
#define TMPC_CONE_WARPS 8
#endif
#ifndef TMPC_CONE_INST
#define TMPC_CONE_INST 3
#endif
#ifndef TMPC_CONE_SPEC
#define TMPC_CONE_SPEC 1
#endif
constexpr int kCWarps = TMPC_CONE_WARPS;   // warps per CTA, one CTA per SM: two per scheduler
constexpr int kCInst = TMPC_CONE_INST;     // instances per warp (a tile)
constexpr int kCGroup = 10;         // lanes per instance in phase A
constexpr int kRec = 10;            // doubles per knot record of the exchange array: [x | q (6)] [u | r | t (3)] [pad]
constexpr int kTaskDoubles = 36;    // g v gc gl (4 x 6)  y z yc yl (4 x 3)
constexpr int kTaskWords = 2 * kTaskDoubles;
constexpr int kTaskPitch = kTaskDoubles + 1;   // doubles between consecutive tasks kept in shared memory: an odd pitch makes the
                                               // 64-bit accesses of 16 consecutive lanes hit 16 different bank pairs
constexpr int kTmemCols = (512 / ((kCWarps + 3) / 4)) & ~1;   // Tensor-Memory columns of a warp (4 lane quarters)
constexpr int kTmemSlots = kTmemCols / kTaskWords;            // tasks per lane in Tensor Memory: 3 x 72 = 216 of 256 columns at 8 warps
constexpr int kTmemTasks = 32 * kTmemSlots;

struct ConeParams {
    int* counter;                   // tile queue
    int cold;                       // carried state is logically zero: do not read it
    int smem_tasks;                 // tasks per warp kept in shared memory: max(0, 3 N - 96)
    int off_warp0, warp_stride;     // shared-memory carve-up, in doubles
    // per-CTA tables staged once (what phase B would otherwise fetch through the read-only path for every task):
    int tab_refs;                   // shared references as the products the cost uses: cq [N][NX], cr [N-1][NU], tq [NX]
    int tab_bounds;                 // box bounds that are the same at every knot point: (lo, hi) pairs [NX], [NU]
    int off_refs, off_bounds, off_lin;   // in doubles; half-spaces: rows of [a (len), b, |a|^2]
};

// 3-block of a column starting at run-time index `a`, by selects (a dynamic index would push the column into local memory)
template <int LEN>
__device__ __forceinline__ void project_cones(double (&vc)[LEN], int ncones, const int* A, const double* mu) {
    for (int k = 0; k < ncones; ++k) {
        const int a = A[k];
        double s3[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            s3[c] = vc[c];
#pragma unroll
            for (int o = 1; o + 3 <= LEN; ++o) s3[c] = (a == o) ? vc[c + o] : s3[c];
        }
        project_soc3(s3, mu[k]);
#pragma unroll
        for (int j = 0; j < LEN; ++j) {
            const int c = j - a;
            vc[j] = (c == 0) ? s3[0] : ((c == 1) ? s3[1] : ((c == 2) ? s3[2] : vc[j]));
        }
    }
}

// Sequential (Gauss-Seidel) half-space projections (admm.cpp:148-173) from the staged rows [a (LEN), b, |a|^2]
template <int LEN>
__device__ __forceinline__ void project_halfspaces_tab(double (&col)[LEN], const double* tab, int nrows) {
    for (int k = 0; k < nrows; ++k, tab += LEN + 2) {
        double dot = 0.0;
#pragma unroll
        for (int j = 0; j < LEN; ++j) dot = fma(tab[j], col[j], dot);
        const double bk = tab[LEN];
        if (dot > bk) {
            const double dist = __ddiv_rn(dsub(dot, bk), tab[LEN + 1]);
#pragma unroll
            for (int j = 0; j < LEN; ++j) col[j] = dsub(col[j], dmul(dist, tab[j]));
        }
    }
}

// ---- Tensor Memory, 72 words per task: x32 + x32 + x8 (tcgen05.ld is not scoreboarded: see tie_after_wait)
__device__ __forceinline__ void tm_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
}
__device__ __forceinline__ void tm_st32(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
                 "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                    "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
                    "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
                    "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
                 : "memory");
}
// Registers written by tcgen05.ld are valid only after tcgen05.wait::ld; an empty volatile asm that takes them as
// in/out operands AFTER the wait makes every consumer depend on it, so the compiler cannot hoist one above the wait.
__device__ __forceinline__ void tie24(uint32_t* w) {
    asm volatile("" : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]), "+r"(w[6]), "+r"(w[7]),
                      "+r"(w[8]), "+r"(w[9]), "+r"(w[10]), "+r"(w[11]), "+r"(w[12]), "+r"(w[13]), "+r"(w[14]), "+r"(w[15]),
                      "+r"(w[16]), "+r"(w[17]), "+r"(w[18]), "+r"(w[19]), "+r"(w[20]), "+r"(w[21]), "+r"(w[22]), "+r"(w[23])
                 :: "memory");
}
__device__ __forceinline__ void task_tm_load(uint32_t taddr, uint32_t* w) {
    tm_ld32(taddr, w); tm_ld32(taddr + 32, w + 32); tm_ld8(taddr + 64, w + 64);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    tie24(w); tie24(w + 24); tie24(w + 48);
}
__device__ __forceinline__ void task_tm_store(uint32_t taddr, const uint32_t* w) {
    tm_st32(taddr, w); tm_st32(taddr + 32, w + 32); tm_st8(taddr + 64, w + 64);
}

__device__ __forceinline__ double shfl_d(double v, int src) { return shfl_dq(v, src); }

// SPEC >= 0: the constraint set is fixed at compile time - the box on both sides through the staged bound table (a side
// without a box holds (-inf, inf)), shared references staged as tables, and per bit of SPEC
//   1: ONE cone per side, starting at element 0 (the reference's rocket-landing example, rocket_landing_mpc.cpp:59-94)
//   2: ONE half-space per side;
// RTAB says whether the references are shared (staged as tables) or per instance (read through the read-only path).
// The run-time flags, cone loops and half-space loops then fold away and phase B becomes a few long basic blocks, which
// is what its instruction-level parallelism needs (config 4: 50.0 -> 32.9 ms).  SPEC = -1 is the general form.
template <int NX, int NU, int SPEC, bool RTAB = true>
__global__ void __launch_bounds__(kCWarps * 32, 1)
admm_cone_kernel(const __grid_constant__ BatchView bv, const __grid_constant__ AdmmSettings st,
                 const __grid_constant__ ConeParams cp) {
    static_assert(NX + NU < kCGroup + 1 && kCInst * kCGroup <= 32, "phase A: one lane per state / input row");
    static_assert(4 * NX + 4 * NU == kTaskDoubles, "task layout");
    // a global access in uniform control flow before anything else: ptxas then keeps the global-memory descriptor in a
    // uniform register instead of paying two R2UR per load / store (see admm_tmem.cu); the counter never holds this value
    if (*reinterpret_cast<const volatile int*>(cp.counter) == 0x7ffffff0) return;
    extern __shared__ __align__(16) double smem[];
    __shared__ uint32_t s_tmem_base;
    const int N = bv.N;
    const int T = kCInst * N;                          // tasks of a tile
    const size_t ld = static_cast<size_t>(bv.ld);
    const double rho = bv.rho;
    const double tol_pri = st.abs_pri_tol, tol_dua = st.abs_dua_tol;

    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;"
                     :: "l"(static_cast<uint64_t>(__cvta_generic_to_shared(&s_tmem_base))) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // ---------------- shared constants (column-major as given): A, AmBKt, Pinf, B, Kinf, Quu_inv, f, APf, BPf, Qd, Rd
    double* sA = smem;                     double* sAm = sA + NX * NX;      double* sP = sAm + NX * NX;
    double* sB = sP + NX * NX;             double* sK = sB + NX * NU;       double* sQu = sK + NU * NX;
    double* sf = sQu + NU * NU;            double* sAPf = sf + NX;          double* sBPf = sAPf + NX;
    double* sQd = sBPf + NU;               double* sRd = sQd + NX;
    for (int e = threadIdx.x; e < NX * NX; e += blockDim.x) { sA[e] = bv.A[e]; sAm[e] = bv.AmBKt[e]; sP[e] = bv.Pinf[e]; }
    for (int e = threadIdx.x; e < NX * NU; e += blockDim.x) { sB[e] = bv.Bm[e]; sK[e] = bv.Kinf[e]; }
    for (int e = threadIdx.x; e < NU * NU; e += blockDim.x) sQu[e] = bv.Quu_inv[e];
    for (int e = threadIdx.x; e < NX; e += blockDim.x) { sf[e] = bv.f[e]; sAPf[e] = bv.APf[e]; sQd[e] = bv.Qd[e]; }
    for (int e = threadIdx.x; e < NU; e += blockDim.x) { sBPf[e] = bv.BPf[e]; sRd[e] = bv.Rd[e]; }
    // shared references as the terms the linear cost uses: cq = -(Xref .* Qd), cr = -(Uref .* Rd), tq = Xref_{N-1}' Pinf
    // (admm.cpp:218, :228, :238); knot-invariant box bounds as (lo, hi) pairs; half-space rows with |a|^2
    double* s_cq = smem + cp.off_refs;  double* s_cr = s_cq + NX * N;  double* s_tq = s_cr + NU * (N - 1);
    double* s_xb = smem + cp.off_bounds;  double* s_ub = s_xb + 2 * NX;
    double* s_lx = smem + cp.off_lin;  double* s_lu = s_lx + bv.nlx * (NX + 2);
    if (cp.tab_refs) {
        for (int e = threadIdx.x; e < NX * N; e += blockDim.x) s_cq[e] = -dmul(bv.Xref ? bv.Xref[e] : 0.0, bv.Qd[e % NX]);
        for (int e = threadIdx.x; e < NU * (N - 1); e += blockDim.x) s_cr[e] = -dmul(bv.Uref ? bv.Uref[e] : 0.0, bv.Rd[e % NU]);
        for (int j = threadIdx.x; j < NX; j += blockDim.x) {
            double sp = 0.0;
            if (bv.Xref) for (int k = 0; k < NX; ++k) sp = fma(bv.Xref[(N - 1) * NX + k], bv.Pinf[j * NX + k], sp);
            s_tq[j] = sp;
        }
    }
    if (cp.tab_bounds) {
        const bool bx = st.en_state_bound && bv.x_min != nullptr, bu = st.en_input_bound && bv.u_min != nullptr;
        for (int j = threadIdx.x; j < NX; j += blockDim.x) { s_xb[2 * j] = bx ? bv.x_min[j] : -INFINITY; s_xb[2 * j + 1] = bx ? bv.x_max[j] : INFINITY; }
        for (int j = threadIdx.x; j < NU; j += blockDim.x) { s_ub[2 * j] = bu ? bv.u_min[j] : -INFINITY; s_ub[2 * j + 1] = bu ? bv.u_max[j] : INFINITY; }
    }
    if (st.en_state_linear) for (int k = threadIdx.x; k < bv.nlx; k += blockDim.x) {
        double sq = 0.0;
        for (int j = 0; j < NX; ++j) { const double a = bv.Alin_x[j * bv.nlx + k]; s_lx[k * (NX + 2) + j] = a; sq = fma(a, a, sq); }
        s_lx[k * (NX + 2) + NX] = bv.blin_x[k]; s_lx[k * (NX + 2) + NX + 1] = sq;
    }
    if (st.en_input_linear) for (int k = threadIdx.x; k < bv.nlu; k += blockDim.x) {
        double sq = 0.0;
        for (int j = 0; j < NU; ++j) { const double a = bv.Alin_u[j * bv.nlu + k]; s_lu[k * (NU + 2) + j] = a; sq = fma(a, a, sq); }
        s_lu[k * (NU + 2) + NU] = bv.blin_u[k]; s_lu[k * (NU + 2) + NU + 1] = sq;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tbase = s_tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16) + static_cast<uint32_t>((warp >> 2) * kTmemCols);
    double* ex = smem + cp.off_warp0 + static_cast<size_t>(warp) * cp.warp_stride;   // [T][kRec] knot records
    double* ss = ex + T * kRec;                                                       // [smem_tasks][kTaskPitch] state of tasks >= 96

    // ---------------- phase-A role of the lane
    const int grp = lane / kCGroup;                    // instance of the tile (3 = spare lanes 30, 31)
    const int role = lane - grp * kCGroup;
    const bool is_x = grp < kCInst && role < NX;
    const bool is_u = grp < kCInst && role >= NX && role < NX + NU;
    const int row = is_x ? role : 0, jrow = is_u ? role - NX : 0;
    const int gbase = (grp < kCInst ? grp : 0) * kCGroup;
    double Wf[NX], Vf[NU], Wb[NX], Vb[NU];             // the lane's rows: forward (A | Kinf ; B | Quu_inv), backward (AmBKt | B' ; Kinf')
#pragma unroll
    for (int k = 0; k < NX; ++k) {
        Wf[k] = is_x ? sA[k * NX + row] : (is_u ? sK[k * NU + jrow] : 0.0);
        Wb[k] = is_x ? sAm[k * NX + row] : (is_u ? sB[jrow * NX + k] : 0.0);
    }
#pragma unroll
    for (int k = 0; k < NU; ++k) {
        Vf[k] = is_x ? sB[k * NX + row] : (is_u ? sQu[k * NU + jrow] : 0.0);
        Vb[k] = is_x ? sK[row * NU + k] : 0.0;
    }
    const double cf = is_x ? sf[row] : 0.0;
    const double cb = is_x ? sAPf[row] : (is_u ? sBPf[jrow] : 0.0);

    constexpr bool FIX = SPEC >= 0, FIX_SOC = FIX && (SPEC & 1), FIX_LIN = FIX && (SPEC & 2);
    const bool box_x = FIX || (st.en_state_bound && bv.x_min != nullptr);
    const bool box_u = FIX || (st.en_input_bound && bv.u_min != nullptr);
    const bool soc_x = FIX ? FIX_SOC : (st.en_state_soc && bv.ncx > 0);
    const bool soc_u = FIX ? FIX_SOC : (st.en_input_soc && bv.ncu > 0);
    const bool lin_x = FIX ? FIX_LIN : st.en_state_linear != 0;
    const bool lin_u = FIX ? FIX_LIN : st.en_input_linear != 0;
    const bool tab_bounds = FIX || cp.tab_bounds != 0;
    const bool tab_refs = FIX ? RTAB : cp.tab_refs != 0;      // RTAB = false: per-instance references through the read-only path
    const int nlx = FIX ? 1 : bv.nlx, nlu = FIX ? 1 : bv.nlu;
    const int rounds = (T + 31) / 32;

    for (;;) {
        // ================================================================= next tile
        int tile = 0;
        if (lane == 0) tile = atomicAdd(cp.counter, 1);
        tile = __shfl_sync(kAll, tile, 0);
        const int b_first = tile * kCInst;
        if (b_first >= bv.B) break;
        int iter[kCInst], status[kCInst];
        bool live[kCInst];
        // element gi of a per-instance register array, by selects (a dynamic index would go through local memory)
        // instance of the tile that task tau belongs to (tau / N without the integer division: kCInst - 1 compares)
        auto inst_of = [&](int tau) { int gi = 0;
#pragma unroll
            for (int k = 1; k < kCInst; ++k) gi += (tau >= k * N) ? 1 : 0;
            return gi; };
        auto pick = [&](const auto& arr, int gi) { auto v = arr[0];
#pragma unroll
            for (int k = 1; k < kCInst; ++k) v = (gi == k) ? arr[k] : v;
            return v; };
#pragma unroll
        for (int k = 0; k < kCInst; ++k) { iter[k] = 0; status[k] = 1; live[k] = b_first + k < bv.B; }

        // ---------------- install: zeros (cold) or the carried element-major state of the tile's instances
        for (int rd = 0; rd < rounds; ++rd) {
            const int tau = rd * 32 + lane;
            const bool act = tau < T;
            const int gi = act ? inst_of(tau) : 0, i = act ? tau - gi * N : 0;
            const int b = b_first + gi;
            const bool ldst = act && b < bv.B && !cp.cold;
            uint32_t w[kTaskWords];
            double s[kTaskDoubles];
#pragma unroll
            for (int e = 0; e < kTaskDoubles; ++e) s[e] = 0.0;
            double qr[NX + NU];
#pragma unroll
            for (int e = 0; e < NX + NU; ++e) qr[e] = 0.0;
            if (ldst) {
#pragma unroll
                for (int r = 0; r < NX; ++r) {
                    const size_t o = static_cast<size_t>(i * NX + r) * ld + b;
                    s[r] = bv.g[o]; s[NX + r] = bv.v[0][o];
                    s[2 * NX + r] = bv.gc ? bv.gc[o] : 0.0; s[3 * NX + r] = bv.gl ? bv.gl[o] : 0.0;
                    qr[r] = (i < N - 1) ? bv.q[o] : bv.pN[static_cast<size_t>(r) * ld + b];
                }
                if (i < N - 1) {
#pragma unroll
                    for (int r = 0; r < NU; ++r) {
                        const size_t o = static_cast<size_t>(i * NU + r) * ld + b;
                        s[4 * NX + r] = bv.y[o]; s[4 * NX + NU + r] = bv.z[0][o];
                        s[4 * NX + 2 * NU + r] = bv.yc ? bv.yc[o] : 0.0; s[4 * NX + 3 * NU + r] = bv.yl ? bv.yl[o] : 0.0;
                        qr[NX + r] = bv.r[o];
                    }
                }
            }
            if (act) {
#pragma unroll
                for (int e = 0; e < NX + NU; ++e) ex[tau * kRec + e] = qr[e];
                ex[tau * kRec + kRec - 1] = 0.0;
            }
            if (rd < kTmemSlots) {
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) { w[2 * e] = lo32(s[e]); w[2 * e + 1] = hi32(s[e]); }
                task_tm_store(tbase + rd * kTaskWords, w);
            } else if (act) {
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) ss[(tau - kTmemTasks) * kTaskPitch + e] = s[e];
            }
        }
        tm_wait_st();
        double x0r = 0.0;
        if (is_x && b_first + grp < bv.B) x0r = bv.x0[static_cast<size_t>(b_first + grp) * NX + row];
        __syncwarp();

        // ================================================================= ADMM iterations of the tile
        for (;;) {
            const bool glive = grp < kCInst && pick(live, grp);
            // ------------------------------------------------- phase A: backward sweep (admm.cpp:13-20)
            // Vectors are all-gathered through the knot records themselves: the state lanes leave p_i in the q slot of
            // record i (q_i has just been consumed), everybody reads it back with three 16-byte loads one knot later.
            {
                double* rec = ex + (static_cast<size_t>(grp < kCInst ? grp : 0) * N + N - 1) * kRec;   // record N-1 holds p_{N-1}
                for (int i = N - 2; i >= 0; --i) {
                    double pb[NX], rr[NU];
                    ld2(rec, pb[0], pb[1]); ld2(rec + 2, pb[2], pb[3]); ld2(rec + 4, pb[4], pb[5]);   // p_{i+1}
                    rec -= kRec;
                    ld2(rec + NX, rr[0], rr[1]); rr[2] = rec[NX + 2];                                   // r_i
                    const double q_own = is_x ? rec[row] : 0.0;
                    const double r_own = is_u ? rec[NX + jrow] : 0.0;
                    double c1 = 0.0, c2 = 0.0;
#pragma unroll
                    for (int k = 0; k < NX; ++k) c1 = fma(Wb[k], pb[k], c1);
#pragma unroll
                    for (int k = 0; k < NU; ++k) c2 = fma(Vb[k], rr[k], c2);
                    const double pn = dadd(dsub(dadd(q_own, c1), c2), cb);       // p_i = q_i + AmBKt p - Kinf' r_i + APf
                    const double tn = dadd(dadd(c1, r_own), cb);                 // t_i = B' p + r_i + BPf
                    __syncwarp();                                                // every lane has read q_i and r_i
                    if (glive) {
                        if (is_x) rec[row] = pn;
                        if (is_u) rec[NX + jrow] = tn;
                    }
                    __syncwarp();
                }
            }
            // ------------------------------------------------- phase A: forward sweep (admm.cpp:25-32)
            // x_i goes to record i where phase B wants it and where every lane of the group reads the whole vector back;
            // u_i likewise (over t_i).
            {
                double* rec = ex + static_cast<size_t>(grp < kCInst ? grp : 0) * N * kRec;
                if (is_x && glive) rec[row] = x0r;
                __syncwarp();
                for (int i = 0; i < N - 1; ++i) {
                    double xb[NX], tt[NU], ub[NU];
                    ld2(rec, xb[0], xb[1]); ld2(rec + 2, xb[2], xb[3]); ld2(rec + 4, xb[4], xb[5]);   // x_i
                    ld2(rec + NX, tt[0], tt[1]); tt[2] = rec[NX + 2];                                  // t_i
                    double c1 = 0.0, d = 0.0, c2 = 0.0;
#pragma unroll
                    for (int k = 0; k < NX; ++k) c1 = fma(Wf[k], xb[k], c1);     // A x (state lanes) | Kinf x (input lanes)
#pragma unroll
                    for (int k = 0; k < NU; ++k) d = fma(Vf[k], tt[k], d);       // d_i = Quu_inv t_i (input lanes)
                    const double u = dsub(-c1, d);                               // u_i = -Kinf x_i - d_i
                    __syncwarp();                                                // every lane has read t_i
                    if (is_u && glive) rec[NX + jrow] = u;
                    __syncwarp();
                    ld2(rec + NX, ub[0], ub[1]); ub[2] = rec[NX + 2];
#pragma unroll
                    for (int k = 0; k < NU; ++k) c2 = fma(Vf[k], ub[k], c2);     // B u (state lanes)
                    const double xn = dadd(dadd(c1, c2), cf);                    // x_{i+1} = A x + B u + f
                    rec += kRec;
                    if (is_x && glive) rec[row] = xn;                            // q_{i+1} was consumed by the backward sweep
                    __syncwarp();
                }
            }
            __syncwarp();

            // ------------------------------------------------- phase B: slack / dual / linear cost / residuals per knot point
            unsigned okbits = (1u << kCInst) - 1u;
            bool tst[kCInst];
#pragma unroll
            for (int k = 0; k < kCInst; ++k) tst[k] = ((iter[k] + 1) % st.check_termination) == 0;
            // One task: the whole knot point of one instance in registers.  The body is instantiated twice - for the rounds whose
            // state sits in Tensor Memory and for those in shared memory - so that neither path has to move its registers
            // to where the other one expects them.
            auto task = [&](double (&s)[kTaskDoubles], const int tau, const bool in_tile) {
                const int gi = in_tile ? inst_of(tau) : 0, i = in_tile ? tau - gi * N : 0;
                const int b = b_first + gi;
                const bool act = in_tile && pick(live, gi);
                const bool testing = pick(tst, gi);
                const bool terminal = (i == N - 1);
                if (act) {
                    double* rec = ex + static_cast<size_t>(tau) * kRec;
                    double x[NX], u[NU], acc[NX], racc[NU], vc[NX];
                    ld2(rec, x[0], x[1]); ld2(rec + 2, x[2], x[3]); ld2(rec + 4, x[4], x[5]);
                    ld2(rec + 6, u[0], u[1]); u[2] = rec[8];
                    double* g = s; double* v = s + NX; double* gc = s + 2 * NX; double* gl = s + 3 * NX;
                    double* y = s + 4 * NX; double* z = y + NU; double* yc = z + NU; double* yl = yc + NU;
                    const double* xref = bv.Xref ? bv.Xref + static_cast<size_t>(b) * bv.xref_stride + i * NX : nullptr;
                    bool ok = true;
                    double vold[NX], zold[NU];
                    // ---- input column i (admm.cpp:84, :95-98, :187, :228-233, :259-260)
                    if (!terminal) {
                        const double* uref = bv.Uref ? bv.Uref + static_cast<size_t>(b) * bv.uref_stride + i * NU : nullptr;
                        const double* umin = box_u ? bv.u_min + static_cast<size_t>(b) * bv.ub_stride + i * NU : nullptr;
                        const double* umax = box_u ? bv.u_max + static_cast<size_t>(b) * bv.ub_stride + i * NU : nullptr;
#pragma unroll
                        for (int j = 0; j < NU; ++j) {
                            const double yj = y[j];
                            const double t = dadd(u[j], yj);
                            double znew;
                            if (tab_bounds) znew = clampd(t, s_ub[2 * j], s_ub[2 * j + 1]);           // (-inf, inf) when the box is off
                            else znew = box_u ? clampd(t, __ldg(umin + j), __ldg(umax + j)) : t;
                            const double ynew = dsub(t, znew);
                            zold[j] = z[j];
                            ok = ok && (fabs(dsub(u[j], znew)) < tol_pri) && (dmul(fabs(dsub(zold[j], znew)), rho) < tol_dua);
                            double ra = tab_refs ? s_cr[i * NU + j] : -dmul(uref ? __ldg(uref + j) : 0.0, sRd[j]);
                            ra = dsub(ra, dmul(rho, dsub(znew, ynew)));
                            y[j] = ynew; z[j] = znew; racc[j] = ra;
                        }
                        if (soc_u) {
#pragma unroll
                            for (int j = 0; j < NU; ++j) vc[j] = dadd(u[j], yc[j]);
                            if (FIX) project_soc3(vc, bv.cu[0]);
                            else { double v3[NU] = {vc[0], vc[1], vc[2]}; project_cones<NU>(v3, bv.ncu, bv.Acu, bv.cu); vc[0] = v3[0]; vc[1] = v3[1]; vc[2] = v3[2]; }
#pragma unroll
                            for (int j = 0; j < NU; ++j) {
                                const double ycn = dsub(dadd(yc[j], u[j]), vc[j]);
                                racc[j] = dsub(racc[j], dmul(rho, dsub(vc[j], ycn)));
                                yc[j] = ycn;
                            }
                        }
                        if (lin_u) {
#pragma unroll
                            for (int j = 0; j < NU; ++j) vc[j] = dadd(u[j], yl[j]);
                            { double v3[NU] = {vc[0], vc[1], vc[2]}; project_halfspaces_tab<NU>(v3, s_lu, nlu); vc[0] = v3[0]; vc[1] = v3[1]; vc[2] = v3[2]; }
#pragma unroll
                            for (int j = 0; j < NU; ++j) {
                                const double yln = dsub(dadd(yl[j], u[j]), vc[j]);
                                racc[j] = dsub(racc[j], dmul(rho, dsub(vc[j], yln)));
                                yl[j] = yln;
                            }
                        }
                    }
                    // ---- state column i (admm.cpp:81, :90-93, :184, :218-246, :257-258)
                    {
                        const double* xmin = box_x ? bv.x_min + static_cast<size_t>(b) * bv.xb_stride + i * NX : nullptr;
                        const double* xmax = box_x ? bv.x_max + static_cast<size_t>(b) * bv.xb_stride + i * NX : nullptr;
#pragma unroll
                        for (int j = 0; j < NX; ++j) {
                            const double gj = g[j];
                            const double t = dadd(x[j], gj);
                            double vnew;
                            if (tab_bounds) vnew = clampd(t, s_xb[2 * j], s_xb[2 * j + 1]);
                            else vnew = box_x ? clampd(t, __ldg(xmin + j), __ldg(xmax + j)) : t;
                            const double gnew = dsub(t, vnew);
                            vold[j] = v[j];
                            ok = ok && (fabs(dsub(x[j], vnew)) < tol_pri) && (dmul(fabs(dsub(vold[j], vnew)), rho) < tol_dua);
                            g[j] = gnew; v[j] = vnew;
                            acc[j] = dmul(rho, dsub(vnew, gnew));
                        }
                        if (!terminal) {
#pragma unroll
                            for (int j = 0; j < NX; ++j)
                                acc[j] = dsub(tab_refs ? s_cq[i * NX + j] : -dmul(xref ? __ldg(xref + j) : 0.0, sQd[j]), acc[j]);
                        } else {
#pragma unroll
                            for (int j = 0; j < NX; ++j) {
                                double sp = 0.0;
                                if (tab_refs) sp = s_tq[j];
                                else if (xref) {
#pragma unroll
                                    for (int k = 0; k < NX; ++k) sp = fma(__ldg(xref + k), sP[j * NX + k], sp);   // Xref_{N-1}' Pinf
                                }
                                acc[j] = dsub(-sp, acc[j]);
                            }
                        }
                        if (soc_x) {
#pragma unroll
                            for (int j = 0; j < NX; ++j) vc[j] = dadd(x[j], gc[j]);
                            if (FIX) project_soc3(vc, bv.cx[0]);        // the cone is elements 0..2
                            else project_cones<NX>(vc, bv.ncx, bv.Acx, bv.cx);
#pragma unroll
                            for (int j = 0; j < NX; ++j) {
                                const double gcn = dsub(dadd(gc[j], x[j]), vc[j]);
                                acc[j] = dsub(acc[j], dmul(rho, dsub(vc[j], gcn)));
                                gc[j] = gcn;
                            }
                        }
                        if (lin_x) {
#pragma unroll
                            for (int j = 0; j < NX; ++j) vc[j] = dadd(x[j], gl[j]);
                            project_halfspaces_tab<NX>(vc, s_lx, nlx);
#pragma unroll
                            for (int j = 0; j < NX; ++j) {
                                const double gln = dsub(dadd(gl[j], x[j]), vc[j]);
                                acc[j] = dsub(acc[j], dmul(rho, dsub(vc[j], gln)));
                                gl[j] = gln;
                            }
                        }
                    }
                    // q_i (or the terminal p_{N-1}) and r_i for the next backward sweep
                    st2(rec, acc[0], acc[1]); st2(rec + 2, acc[2], acc[3]); st2(rec + 4, acc[4], acc[5]);
                    if (!terminal) { st2(rec + 6, racc[0], racc[1]); rec[8] = racc[2]; }
                    // the carried previous slacks lag the solution by one iteration when the solve converges
                    // (admm.cpp:375 returns before :379-380): a task whose own tests pass parks them in the carried arrays
                    // (if any task of the instance fails, the instance does not converge and the copies are never read)
                    if (testing && ok) {
#pragma unroll
                        for (int j = 0; j < NX; ++j) bv.v[0][static_cast<size_t>(i * NX + j) * ld + b] = vold[j];
                        if (!terminal) {
#pragma unroll
                            for (int j = 0; j < NU; ++j) bv.z[0][static_cast<size_t>(i * NU + j) * ld + b] = zold[j];
                        }
                    }
                    if (!(testing && ok)) okbits &= ~(1u << gi);
                }
                return act;
            };
            const int rounds_tm = rounds < kTmemSlots ? rounds : kTmemSlots;
            for (int rd = 0; rd < rounds_tm; ++rd) {
                const int tau = rd * 32 + lane;
                uint32_t w[kTaskWords];
                double s[kTaskDoubles];
                task_tm_load(tbase + rd * kTaskWords, w);
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) s[e] = mk_double(w[2 * e], w[2 * e + 1]);
                task(s, tau, tau < T);
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) { w[2 * e] = lo32(s[e]); w[2 * e + 1] = hi32(s[e]); }
                task_tm_store(tbase + rd * kTaskWords, w);
            }
            for (int rd = kTmemSlots; rd < rounds; ++rd) {
                const int tau = rd * 32 + lane;
                const bool in_tile = tau < T;
                double* sp = ss + (in_tile ? tau - kTmemTasks : 0) * kTaskPitch;
                double s[kTaskDoubles];
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) s[e] = sp[e];
                if (task(s, tau, in_tile)) {
#pragma unroll
                    for (int e = 0; e < kTaskDoubles; ++e) sp[e] = s[e];
                }
            }
            tm_wait_st();
            __syncwarp();

            // ------------------------------------------------- termination per instance (admm.cpp:253-271, :364-388)
            bool any_live = false;
#pragma unroll
            for (int k = 0; k < kCInst; ++k) {
                const bool conv = __all_sync(kAll, (okbits >> k) & 1u);
                if (live[k]) {
                    iter[k] += 1;
                    const bool c = conv && (iter[k] % st.check_termination == 0);
                    status[k] = c ? 0 : 1;
                    if (c || iter[k] >= st.max_iter) live[k] = false;
                }
                any_live = any_live || live[k];
            }
            if (!any_live) break;
        }

        // ================================================================= write-out of the tile
        for (int rd = 0; rd < rounds; ++rd) {
            const int tau = rd * 32 + lane;
            const bool in_tile = tau < T;
            const int gi = in_tile ? inst_of(tau) : 0, i = in_tile ? tau - gi * N : 0;
            const int b = b_first + gi;
            const bool act = in_tile && b < bv.B;
            const int st_k = pick(status, gi);
            uint32_t w[kTaskWords];
            double s[kTaskDoubles];
            if (rd < kTmemSlots) {
                task_tm_load(tbase + rd * kTaskWords, w);
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) s[e] = mk_double(w[2 * e], w[2 * e + 1]);
            } else {
#pragma unroll
                for (int e = 0; e < kTaskDoubles; ++e) s[e] = in_tile ? ss[(tau - kTmemTasks) * kTaskPitch + e] : 0.0;
            }
            if (act) {
                const double* rec = ex + static_cast<size_t>(tau) * kRec;
                const bool unconv = st_k != 0;
                double* Xs = bv.Xsol + (static_cast<size_t>(b) * N + i) * NX;
#pragma unroll
                for (int r = 0; r < NX; ++r) {
                    const size_t o = static_cast<size_t>(i * NX + r) * ld + b;
                    Xs[r] = s[NX + r];                                   // solution = vnew of the last iteration (admm.cpp:370,386)
                    bv.g[o] = s[r];
                    if (unconv) bv.v[0][o] = s[NX + r];                  // v = vnew happened (admm.cpp:379); else the lagging copy is there
                    if (bv.gc) bv.gc[o] = s[2 * NX + r];
                    if (bv.gl) bv.gl[o] = s[3 * NX + r];
                    if (i < N - 1) bv.q[o] = rec[r]; else bv.pN[static_cast<size_t>(r) * ld + b] = rec[r];
                }
                if (i < N - 1) {
                    double* Us = bv.Usol + (static_cast<size_t>(b) * (N - 1) + i) * NU;
#pragma unroll
                    for (int r = 0; r < NU; ++r) {
                        const size_t o = static_cast<size_t>(i * NU + r) * ld + b;
                        Us[r] = s[4 * NX + NU + r];
                        bv.y[o] = s[4 * NX + r];
                        if (unconv) bv.z[0][o] = s[4 * NX + NU + r];
                        if (bv.yc) bv.yc[o] = s[4 * NX + 2 * NU + r];
                        if (bv.yl) bv.yl[o] = s[4 * NX + 3 * NU + r];
                        bv.r[o] = rec[NX + r];
                    }
                }
                if (i == 0) { bv.status[b] = st_k; bv.iters[b] = pick(iter, gi); }
            }
        }
        __syncwarp();
    }

    __syncthreads();
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(s_tmem_base));
    }
}

template <int NX, int NU>
bool cone_plan(const BatchView& bv, ConeParams& cp, size_t& smem_bytes) {
    const int N = bv.N;
    if (N < 2) return false;
    const int T = kCInst * N;
    if ((T + 31) / 32 > kTmemSlots + 8) return false;
    if (bv.nlx > kMaxLinear || bv.nlu > kMaxLinear) return false;
    constexpr size_t kLimit = 227 * 1024 - 64;
    cp.smem_tasks = T > kTmemTasks ? T - kTmemTasks : 0;
    int per_warp = T * kRec + kTaskPitch * cp.smem_tasks;
    per_warp = (per_warp + 1) / 2 * 2;
    cp.warp_stride = per_warp;
    const bool shared_refs = (bv.Xref == nullptr || bv.xref_stride == 0) && (bv.Uref == nullptr || bv.uref_stride == 0);
    const bool const_bounds = bv.x_min == nullptr || bv.bounds_const != 0;
    // tables are optional: dropped (the kernel then fetches through the read-only path) when shared memory is short
    for (int attempt = 0; attempt < 3; ++attempt) {
        cp.tab_refs = (attempt == 0) && shared_refs;
        cp.tab_bounds = (attempt <= 1) && const_bounds;
        int off = 3 * NX * NX + 2 * NX * NU + NU * NU + 3 * NX + 2 * NU;
        cp.off_lin = off;    off += bv.nlx * (NX + 2) + bv.nlu * (NU + 2);
        cp.off_bounds = off; off += cp.tab_bounds ? 2 * (NX + NU) : 0;
        cp.off_refs = off;   off += cp.tab_refs ? NX * N + NU * (N - 1) + NX : 0;
        off = (off + 1) / 2 * 2;
        cp.off_warp0 = off;
        smem_bytes = (static_cast<size_t>(off) + static_cast<size_t>(per_warp) * kCWarps) * sizeof(double);
        if (smem_bytes <= kLimit) return true;
    }
    return false;
}

}  // namespace

bool cone_kernel_supports(const BatchView& bv, const AdmmSettings& st) {
    if (st.max_iter <= 0 || !(bv.rho > 0.0)) return false;
    if (bv.nx != 6 || bv.nu != 3) return false;
    if (bv.ncx > kMaxCones || bv.ncu > kMaxCones) return false;
    ConeParams cp{};
    size_t bytes = 0;
    return cone_plan<6, 3>(bv, cp, bytes);
}

cudaError_t launch_admm_cone(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                             cudaStream_t stream) {
    if (bv.B <= 0) return cudaSuccess;
    if (bv.nx != 6 || bv.nu != 3) return cudaErrorNotSupported;
    ConeParams cp{};
    size_t bytes = 0;
    if (!cone_plan<6, 3>(bv, cp, bytes)) return cudaErrorInvalidConfiguration;
    cp.counter = counter;
    cp.cold = cold;
    // constraint sets of the forms listed at the kernel run the variant with everything folded at compile time
    const bool no_soc = !(st.en_state_soc && bv.ncx > 0) && !(st.en_input_soc && bv.ncu > 0);
    const bool one_soc = st.en_state_soc && st.en_input_soc && bv.ncx == 1 && bv.ncu == 1 && bv.Acx[0] == 0 && bv.Acu[0] == 0;
    const bool no_lin = !st.en_state_linear && !st.en_input_linear;
    const bool one_lin = st.en_state_linear && st.en_input_linear && bv.nlx == 1 && bv.nlu == 1;
    int spec = -1;
    if (TMPC_CONE_SPEC && cp.tab_bounds && (no_soc || one_soc) && (no_lin || one_lin)) spec = (one_soc ? 1 : 0) | (one_lin ? 2 : 0);
    auto kern = admm_cone_kernel<6, 3, -1>;
    if (cp.tab_refs) {
        if (spec == 0) kern = admm_cone_kernel<6, 3, 0, true>;
        if (spec == 1) kern = admm_cone_kernel<6, 3, 1, true>;
        if (spec == 2) kern = admm_cone_kernel<6, 3, 2, true>;
        if (spec == 3) kern = admm_cone_kernel<6, 3, 3, true>;
    } else {
        if (spec == 0) kern = admm_cone_kernel<6, 3, 0, false>;
        if (spec == 1) kern = admm_cone_kernel<6, 3, 1, false>;
        if (spec == 2) kern = admm_cone_kernel<6, 3, 2, false>;
        if (spec == 3) kern = admm_cone_kernel<6, 3, 3, false>;
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(counter, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    const int tiles = (bv.B + kCInst - 1) / kCInst;
    int grid = (tiles + kCWarps - 1) / kCWarps;
    if (grid > num_sms) grid = num_sms;
    kern<<<grid, kCWarps * 32, bytes, stream>>>(bv, st, cp);
    return cudaGetLastError();
}

}  // namespace tmpc
