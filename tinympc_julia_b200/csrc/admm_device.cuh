// Device-side description of one batch of TinyMPC problem instances, shared by the solve kernels.
//
// HBM layout
// ----------
// * User-facing arrays keep the Julia layout (instance = slowest dimension): x0 is nx x B,
//   references nx x N [x B], bounds likewise, solutions nx x N x B.  A per-instance array and a
//   shared one differ only by the instance stride (0 when shared).
// * (element-major format, thread and shared-memory sub-warp kernels)  Carried ADMM state (what TinyWorkspace keeps between solves, reference types.hpp:86-187:
//   duals g,y,gc,yc,gl,yl; previous box slacks v,z; linear cost terms q,r and the terminal p)
//   is stored element-major: entry e of instance b lives at [e * ld + b], ld = B rounded up to 32.
//   One thread per instance then reads it fully coalesced, and an 8-instance tile of it is a
//   64-byte-wide 2-D box for the sub-warp kernel.
#pragma once
#include <cstdint>

namespace tmpc {

struct AdmmSettings {        // reference: TinySettings, types.hpp:63-80
    double abs_pri_tol, abs_dua_tol;
    int max_iter, check_termination;
    int en_state_bound, en_input_bound;
    int en_state_soc, en_input_soc;
    int en_state_linear, en_input_linear;
};

constexpr int kMaxCones = 4;     // cones per side
constexpr int kMaxLinear = 16;   // half-spaces per side (runtime-dim kernel reads them from global memory)

struct BatchView {
    int B, ld;
    int nx, nu, N;
    double rho;

    // ---- inputs, Julia layout; stride = elements between consecutive instances (0 = shared)
    const double* x0;                       // nx x B
    const double* Xref; long long xref_stride;   // nullptr => zeros
    const double* Uref; long long uref_stride;
    const double* x_min; const double* x_max; long long xb_stride;   // nullptr => unbounded
    const double* u_min; const double* u_max; long long ub_stride;
    int bounds_const;                       // shared bounds that are the same at every knot point (checked when they are set)

    // ---- shared problem constants, column-major, device memory
    const double *A, *Bm, *f, *Kinf, *Pinf, *Quu_inv, *AmBKt, *APf, *BPf, *Qd, *Rd;

    // ---- second-order cones in the slots the reference core stores them in (size-3 blocks)
    int ncx; int Acx[kMaxCones]; double cx[kMaxCones];
    int ncu; int Acu[kMaxCones]; double cu[kMaxCones];
    // ---- half-spaces a_k . z <= b_k, A column-major (rows = constraints)
    int nlx; const double* Alin_x; const double* blin_x;
    int nlu; const double* Alin_u; const double* blin_u;

    // ---- carried state, element-major [e][ld]
    double *g, *y;          // box duals                       nx*N, nu*(N-1)
    double *v[2], *z[2];    // box slacks, ping-pong: v[0] is the carried "v" between solves
    double *q, *r, *pN;     // linear cost terms, terminal p    nx*N, nu*(N-1), nx
    double *d;              // feed-forward scratch             nu*(N-1)
    double *gc, *yc;        // cone duals   (allocated when a cone flag is on)
    double *gl, *yl;        // linear duals (allocated when a linear flag is on)

    // ---- carried state, instance-major twin used by the on-chip TMEM kernel: one contiguous block per instance,
    //      knot i < N: [3 nx i] g(nx) v(nx) q(nx);  knot i < N-1: [3 nx N + 3 nu i] y(nu) z(nu) r(nu);  then p_N (nx).
    //      The host runtime converts between the two formats when consecutive solves use kernels of different
    //      families (state_layout.cu); exactly one of them is current at any time.
    double* state_im; long long state_im_stride;

    // ---- results
    double* Xsol;           // nx x N x B      (Julia layout)  = vnew of the last iteration
    double* Usol;           // nu x (N-1) x B                  = znew of the last iteration
    int* status;            // 0 converged / 1 max_iter         (what solve() returns, admm.cpp:375,388)
    int* iters;
};

// ---- arithmetic helpers: the elementwise updates are written with explicitly rounded operations
// so that nvcc cannot contract them differently from one kernel to the next; dot products are
// ascending-index FMA chains.
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
// min(hi, max(lo, t)) with the comparison semantics of std::max / std::min (Eigen cwiseMax/cwiseMin,
// reference admm.cpp:92,97): two compares + selects, cheaper than DMNMX on sm_100.
__device__ __forceinline__ double clampd(double t, double lo, double hi) {
    const double w = (lo < t) ? t : lo;      // std::max(lo, t)
    return (w < hi) ? w : hi;                // std::min(hi, w)
}

// ---- the same operations at either value of the reference's `tinytype` (types.hpp:15): kernels that exist in a
// float instantiation take their arithmetic from here.
template <class T> struct Ar;
template <> struct Ar<double> {
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
    static __device__ __forceinline__ double abs(double a) { return ::fabs(a); }
    static __device__ __forceinline__ double max(double a, double b) { return ::fmax(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return ::sqrt(a); }
    static __device__ __forceinline__ double clamp(double t, double lo, double hi) { return clampd(t, lo, hi); }
};
template <> struct Ar<float> {
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return ::fmaf(a, b, c); }
    static __device__ __forceinline__ float abs(float a) { return ::fabsf(a); }
    static __device__ __forceinline__ float max(float a, float b) { return ::fmaxf(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    static __device__ __forceinline__ float clamp(float t, float lo, float hi) {
        const float w = (lo < t) ? t : lo;       // std::max(lo, t)
        return (w < hi) ? w : hi;                // std::min(hi, w)
    }
};

// ---- projections of update_slack shared by the thread kernel and the on-chip cone kernel
// Second-order-cone projection of a 3-block, with the reference's mixed precision: mu and the
// norm are rounded to float (reference admm.cpp:39,42), the last entry comes from a float division (:54).
template <class T>
__device__ __forceinline__ void project_soc3(T* s, double mu_d) {
    using R = Ar<T>;
    const float mu = static_cast<float>(mu_d);
    const T u0 = R::mul(s[2], static_cast<T>(mu));
    const float a = static_cast<float>(R::sqrt(R::add(R::mul(s[0], s[0]), R::mul(s[1], s[1]))));
    const T ad = static_cast<T>(a);
    if (ad <= -u0) {
        s[0] = 0; s[1] = 0; s[2] = 0;
    } else if (ad <= u0) {
        // inside the cone: unchanged
    } else {
        const T scale = R::mul(T(0.5), R::add(T(1), R::div(u0, ad)));
        const T last = static_cast<T>(__fdiv_rn(a, mu));
        s[0] = R::mul(scale, s[0]); s[1] = R::mul(scale, s[1]); s[2] = R::mul(scale, last);
    }
}

// Sequential (Gauss-Seidel) half-space projections, reference admm.cpp:148-173.
template <class T>
__device__ __forceinline__ void project_halfspaces(T* col, int len, const double* __restrict__ Alin,
                                                   const double* __restrict__ blin, int nrows) {
    for (int k = 0; k < nrows; ++k) {
        using R = Ar<T>;
        T dot = 0;
        for (int j = 0; j < len; ++j) dot = R::fma(static_cast<T>(__ldg(Alin + j * nrows + k)), col[j], dot);
        const T bk = static_cast<T>(__ldg(blin + k));
        if (dot > bk) {
            T sq = 0;
            for (int j = 0; j < len; ++j) { const T a = static_cast<T>(__ldg(Alin + j * nrows + k)); sq = R::fma(a, a, sq); }
            const T dist = R::div(R::sub(dot, bk), sq);
            for (int j = 0; j < len; ++j) col[j] = R::sub(col[j], R::mul(dist, static_cast<T>(__ldg(Alin + j * nrows + k))));
        }
    }
}

}  // namespace tmpc
