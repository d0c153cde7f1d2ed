// Host-visible launch interface of the ADMM solve kernels.
#pragma once
#include <cuda_runtime.h>

#include <vector>

#include "admm_device.cuh"

namespace tmpc {

constexpr int kThreadKernelBlock = 128;
constexpr int kThreadKernelMaxNx = 32;   // runtime-dimension path of the thread kernel
constexpr int kThreadKernelMaxNu = 16;

// Problem constants on the host, column-major (the same numbers the device copies hold).
struct HostMats {
    int nx = 0, nu = 0;
    std::vector<double> A, B, f, Kinf, Pinf, Quu_inv, AmBKt, APf, BPf, Qd, Rd;
};

// ---- per-instance dynamics: every instance has its own A, B, f, Q, R and therefore its own cache.
// Element-major device arrays [e][ld] (column-major matrix entries e), written by the batched setup kernel
// (lqr_cache_device.cu) and read coalesced by the thread kernel.
constexpr int kPerInstanceMaxNx = 12;
constexpr int kPerInstanceMaxNu = 4;
struct PerInstanceMats {
    double *A, *B, *f, *Kinf, *Pinf, *Quu_inv, *AmBKt, *APf, *BPf, *Qd, *Rd;
};
// Inputs in the Julia layout (nx x nx x B, ...); q_stride / r_stride = elements between instances (0 = shared Q / R).
// *fail is set to 1 when a matrix that must be inverted is singular for some instance.
cudaError_t launch_riccati_cache(int nx, int nu, int B, long long ld, double rho, const double* A, const double* Bdyn,
                                 const double* f, const double* Q, long long q_stride, const double* R,
                                 long long r_stride, const PerInstanceMats& out, int* fail, cudaStream_t stream);

// ---- one thread per instance (admm_thread.cu)
bool thread_kernel_supports(int nx, int nu);
// `precision` = 64 (the reference's default `typedef double tinytype`) or 32 (its float build): the arithmetic type of
// the solve; the arrays in HBM are double either way.
cudaError_t launch_admm_thread_per_instance(const BatchView& bv, const AdmmSettings& st, const PerInstanceMats& pm,
                                            cudaStream_t stream, int precision = 64);
cudaError_t launch_admm_thread(const BatchView& bv, const AdmmSettings& st, const HostMats& h, cudaStream_t stream,
                               int precision = 64);

// ---- adaptive rho on the thread kernel (reference admm.cpp:331-357, rho_benchmark.cpp).  The
// reference adapts rho, Kinf, Pinf (and C1, C2, which the solve never reads) in place in the solver's cache, so for a
// batch they are per-instance carried state: element-major arrays [e][ld], broadcast from the shared cache by
// launch_adaptive_init, updated by the solve, kept across solves.  dK / dP are the shared sensitivity matrices
// (column-major nu x nx, nx x nx) that `codegen_with_sensitivity` stores in the reference's cache (codegen.cpp:91-96).
struct AdaptiveRho {
    int enabled;                 // settings->adaptive_rho; 0 = solve with the adapted cache as it stands (the reference keeps it
                                 // after the flag is cleared)
    int clip;                    // adaptive_rho_enable_clipping
    double rho_min, rho_max;
    double* rho;                 // [ld]
    double* Kinf;                // [nu * nx][ld]
    double* Pinf;                // [nx * nx][ld]
    const double* dK;
    const double* dP;
};
// keep_rho: only Kinf / Pinf are re-broadcast (set_cache_terms on an adapted batch: bindings.cpp:277-280 leaves cache->rho alone)
cudaError_t launch_adaptive_init(const BatchView& bv, const AdaptiveRho& ad, cudaStream_t stream, bool keep_rho = false);
cudaError_t launch_adaptive_gather(const BatchView& bv, const AdaptiveRho& ad, double* Kinf, double* Pinf,
                                   cudaStream_t stream);
cudaError_t launch_admm_thread_adaptive(const BatchView& bv, const AdmmSettings& st, const AdaptiveRho& ad,
                                        cudaStream_t stream);

// ---- one instance per 4-lane sub-warp, shared-memory resident (admm_subwarp.cu)
// Applies to box-constrained problems (no cone / half-space flags) of a compiled-in shape.
bool subwarp_kernel_supports(int nx, int nu, int N, const AdmmSettings& st, int precision = 64);
// `cold` = 1: the carried state is known to be all-zero (fresh / reset batch): it is not read.
// `counter` = device int used as the tile work queue (zeroed by the launcher).
cudaError_t launch_admm_subwarp(const BatchView& bv, const AdmmSettings& st, const HostMats& h, int cold,
                                int* counter, int num_sms, cudaStream_t stream, int precision = 64);

// ---- one warp per instance, built for latency (admm_warp.cu): the recursions with a lane per matrix row, everything else
// over all knot points at once, state in shared memory for the whole solve.  Box-constrained problems of a compiled-in shape,
// double build, shared dynamics; element-major carried state (interchangeable with the thread kernel).  TMPC_KERNEL_AUTO
// picks it for batches of at most warp_kernel_auto_max_b(nx) instances (the B = 1 singleton behind the reference's symbols first
// of all), where the throughput kernels would leave the GPU empty.
constexpr int kWarpKernelAutoMaxB = 2048;            // quadrotor shape: 118 us against 138 on the TMEM kernel at 2,048, 215 against 138 at 4,096
constexpr int kWarpKernelAutoMaxBSmall = 4096;       // cartpole shape (96 registers, 5 KB of shared memory per block): 196 us against 332 at 4,096
inline int warp_kernel_auto_max_b(int nx) { return nx <= 4 ? kWarpKernelAutoMaxBSmall : kWarpKernelAutoMaxB; }
bool warp_kernel_supports(const BatchView& bv, const AdmmSettings& st);
cudaError_t launch_admm_warp(const BatchView& bv, const AdmmSettings& st, int cold, int num_sms, cudaStream_t stream);

// ---- one instance per 4-lane sub-warp, state rows in Tensor Memory + input rows in shared memory, per-instance
// refill from a global queue (admm_tmem.cu).  Box-constrained (or unconstrained) problems with shared bounds.
// `order` (all TMEM launchers): device permutation of 0..B-1 the queue tickets index (queue_order.cu), or nullptr for
// the natural order.
// `precision` = 32: the float instantiation (tinytype = float: 16 warps per SM, carried state as float blocks).
bool tmem_kernel_supports(const BatchView& bv, const AdmmSettings& st, int precision = 64);
// the default form of the kernel applies (what the fused closed loop needs; tmem_kernel_supports also accepts the long-horizon
// form: quadrotor shape in double, N = 22 .. 42, four warps per SM)
bool tmem_kernel_fused_supports(const BatchView& bv, const AdmmSettings& st, int precision = 64);
cudaError_t launch_admm_tmem(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                             cudaStream_t stream, const int* order = nullptr, int precision = 64);

// Per-instance dynamics on the TMEM kernel: one QConst-layout block of constants per instance (packed once from the
// element-major cache); the lane's matrix rows come from its instance's block at every sweep start (L2-resident for
// the instances on chip), the per-knot B / Quu_inv rows from a lane-private shared-memory slot filled at install.
long long tmem_pi_block_doubles(int nx, int nu);      // 0 = shape not compiled
cudaError_t launch_tmem_pi_pack(int nx, int nu, const PerInstanceMats& pm, long long ld, int B, double* blocks,
                                cudaStream_t stream);
bool tmem_pi_kernel_supports(const BatchView& bv, const AdmmSettings& st);
cudaError_t launch_admm_tmem_pi(const BatchView& bv, const AdmmSettings& st, int cold, const double* pi_blocks,
                                int* counter, int num_sms, cudaStream_t stream, const int* order = nullptr);

// Fused closed loop on the TMEM kernel: every instance runs `nsteps` MPC steps (solve -> apply znew(:,0) -> plant step ->
// slide the reference window -> next solve) on the sub-warp that owns it, its ADMM state staying on chip between steps.
// bv.Xref / bv.xref_stride describe the LONG per-instance reference (nx x (N + nsteps - 1)) when `slide` is set.
struct RolloutArgs {
    int nsteps;
    int cold;               // carried state is logically zero (fresh / reset batch)
    int slide;
    double* x0_rw;          // plant state nx x B (the same memory as bv.x0)
    int* status_out;        // [nsteps][B], may be null
    int* iters_out;
    const double* pi_blocks = nullptr;   // per-instance dynamics: the packed operand blocks (launch_tmem_pi_pack)
    const int* order = nullptr;          // queue order (queue_order.cu)
};
cudaError_t launch_admm_tmem_rollout(const BatchView& bv, const AdmmSettings& st, const RolloutArgs& ra, int* counter,
                                     int num_sms, cudaStream_t stream);

// ---- cones and half-spaces on chip (admm_cone.cu): a warp owns a tile of 3 instances; the Riccati recursions run with
// one lane per matrix row, the slack / dual / linear-cost updates with one lane per knot point; state in Tensor Memory
// and shared memory.  Compiled for nx = 6, nu = 3 (rocket landing), N <= 50, every constraint family, element-major
// carried state (interchangeable with the thread kernel).
bool cone_kernel_supports(const BatchView& bv, const AdmmSettings& st);
cudaError_t launch_admm_cone(const BatchView& bv, const AdmmSettings& st, int cold, int* counter, int num_sms,
                             cudaStream_t stream);

// ---- longest-first instance queue (queue_order.cu): a permutation of the instances in which the kOrderTop instances of
// every 256 expected to run longest jump the queue (one kernel, block-local counting sort over 256 key buckets).
// Cold start: key = saturation of the LQR input Kinf x0 against the input box; warm start: key = the previous solve's
// iteration count.
constexpr int kOrderBuckets = 256;
constexpr int kOrderTop = 64;   // of every 256: a quarter of the batch may jump the queue (config 2's iteration counts have a
                                // heavy tail: 12 % of the cartpole instances need more than 1.5x the mean), the rest keeps the natural mix
struct OrderArgs {
    int B, nx, nu;
    const double* x0;                         // nx x B
    const double* Kinf; long long kinf_ld;    // column-major nu x nx shared by all instances (kinf_ld = 0), or per instance [e][ld]
    const double* u_min; const double* u_max; long long ub_stride;   // input bounds, knot 0 (stride 0 = shared)
    const int* prev_iters;                    // non-null: key = iteration count of the previous solve
    int top;                                  // instances of every 256 that jump the queue (1..256)
    int* order;                               // out, B
};
cudaError_t launch_build_order(const OrderArgs& a, cudaStream_t stream);

// ---- carried-state formats (state_layout.cu): element-major [e][ld] arrays <-> instance-major blocks
// The instance-major block, in VALUES (doubles, or floats in the float build).  A lane of the TMEM kernel owns the state
// rows 4 r + sub (r < rx) and the input row sub of its instance.  It stores 8 bytes at a time: one double, or a chunk of
// vw = 2 floats (measured: 128-bit stores of double pairs are 1.2 % slower than 64-bit ones on config 3, 64-bit stores of
// float pairs 6 % faster than 32-bit ones); the chunks of the 4 lanes are interleaved so that one store of the sub-warp
// fills whole sectors:
//   knot i < N      at xk i        : chunk c of lane sub at (4 c + sub) vw; a lane's sequence is g_0..g_{rx-1},
//                                    q_0..q_{rx-1} (written together), padded to whole chunks, then v_0..v_{rx-1}
//                                    (written only when v = vnew happened), padded likewise
//   knot i < N - 1  at ib + uk i   : vw = 1: y_j at j, r_j at nu + j, z_j at 2 nu + j
//                                    vw = 2: (y_j, r_j) at 2 j, z_j at 2 nu + j;  uk = 3 nu rounded up to even
//   terminal p      at pb + row
#if defined(__CUDACC__)
#define TMPC_HD __host__ __device__
#else
#define TMPC_HD
#endif
struct ImLayout {
    int rx, nu, vw, gc, xk, uk, ib, pb;
    long long size;
    TMPC_HD ImLayout(int nx, int nu_, int N, int vw_) {
        rx = (nx + 3) / 4; nu = nu_; vw = vw_;
        gc = (2 * rx + vw - 1) / vw;                      // chunks of the (g, q) sequence
        xk = (gc + (rx + vw - 1) / vw) * 4 * vw;
        uk = vw == 1 ? 3 * nu : ((3 * nu + 1) & ~1);
        ib = xk * N; pb = ib + uk * (N - 1);
        size = pb + 4 * rx;
    }
    // field f: 0 g, 1 v, 2 q
    TMPC_HD int x_off(int f, int i, int row) const {
        const int r = row / 4, sub = row % 4, s = (f == 0 ? 0 : f == 2 ? rx : gc * vw) + r;
        return xk * i + ((s / vw) * 4 + sub) * vw + s % vw;
    }
    // field f: 0 y, 1 z, 2 r
    TMPC_HD int u_off(int f, int i, int j) const {
        if (vw == 1) return ib + uk * i + (f == 0 ? 0 : f == 2 ? nu : 2 * nu) + j;
        return ib + uk * i + (f == 1 ? 2 * nu + j : 2 * j + (f == 2 ? 1 : 0));
    }
    TMPC_HD int p_off(int row) const { return pb + row; }
};
// values per instance block at the given precision (64: doubles, 32: floats); an allocation of state_im_values(.., 64) doubles
// per instance holds either
inline long long state_im_values(int nx, int nu, int N, int precision) { return ImLayout(nx, nu, N, precision == 32 ? 2 : 1).size; }
cudaError_t launch_state_convert(const BatchView& bv, bool to_instance_major, cudaStream_t stream, int precision = 64);

// ---- closed-loop plant step (rollout.cu): x0 <- A x0 + B Usol(:,0) + f, optional reference slide
cudaError_t launch_plant_step(const BatchView& bv, double* x0_rw, cudaStream_t stream);
cudaError_t launch_plant_step_per_instance(const BatchView& bv, const PerInstanceMats& pm, long long ld, double* x0_rw,
                                           cudaStream_t stream);
cudaError_t launch_record_step(const BatchView& bv, int* status_out, int* iters_out, cudaStream_t stream);
// out[nu x B] = Usol(:, 0, b): the first control of every instance, packed for one small device-to-host copy
// with st_out / it_out: also copies the status / iteration words (all three destinations may be mapped host memory)
cudaError_t launch_gather_u0(const double* Usol, long long stride, int nu, int B, double* out, cudaStream_t stream,
                             const int* st_in = nullptr, const int* it_in = nullptr, int* st_out = nullptr,
                             int* it_out = nullptr);

}  // namespace tmpc
