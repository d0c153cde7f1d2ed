// Handle-based batched C ABI (include/tinympc_b200.h, PART 2).
//
// Host-side runtime of the library: owns the device buffers of one batch, mirrors the reference's
// setter semantics (reference: tinympc/TinyMPC/src/tinympc/tiny_api.cpp:139-242, 375-409 and the
// flag handling of src/bindings.cpp:336-490), picks the solve kernel and launches it.  Nothing here
// computes an ADMM iteration on the host: without a usable CUDA device every call fails.
#include <cuda_runtime.h>

#include <algorithm>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <string>
#include <vector>

#include "../../include/tinympc_b200.h"
#include "admm_device.cuh"
#include "admm_kernels.hpp"
#include "lqr_cache.hpp"

namespace {

thread_local std::string g_last_error;

void set_error(const std::string& s) {
    g_last_error = s;
    std::fprintf(stderr, "tinympc_b200: %s\n", s.c_str());
}

struct CudaFail { cudaError_t e; const char* what; };

#define CUDA_TRY(expr)                                                       \
    do {                                                                     \
        cudaError_t e__ = (expr);                                            \
        if (e__ != cudaSuccess) throw CudaFail{e__, #expr};                  \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    bool mapped = false;      // zero-copy: pinned host memory mapped into the device address space (UVA: p is valid on both sides)
    void alloc(size_t count) {
        release();
        n = count;
        if (count) CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&p), count * sizeof(T)));
    }
    // Result buffers of a tiny batch (the B = 1 singleton of PART 1): the kernel's stores go straight to host memory,
    // so solve_mpc -> get_controls costs one stream synchronisation instead of a copy + synchronisation per getter.
    void alloc_mapped(size_t count) {
        release();
        n = count;
        if (count) {
            CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&p), count * sizeof(T), cudaHostAllocMapped));
            std::memset(p, 0, count * sizeof(T));
            mapped = true;
        }
    }
    void zero(cudaStream_t s) { if (p) CUDA_TRY(cudaMemsetAsync(p, 0, n * sizeof(T), s)); }
    void upload(const T* src, size_t count, cudaStream_t s) {
        if (count > n) alloc(count);
        if (count) CUDA_TRY(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, s));
    }
    void release() { if (p) { if (mapped) cudaFreeHost(p); else cudaFree(p); p = nullptr; } n = 0; mapped = false; }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
};

}  // namespace

struct tmpc_batch {
    int device = 0;
    int nx = 0, nu = 0, N = 0, B = 0, ld = 0;
    double rho = 0.0;
    int precision = 64;                       // arithmetic type of the solve: 64 = `typedef double tinytype`, 32 = float build
    std::vector<double> hQ, hR, hA, hB, hf;   // the problem as given (doubles), kept so that set_precision can redo the setup
    double rho_given = 0.0;
    int num_sms = 148;
    tmpc::HostMats hm;
    tmpc::AdmmSettings st{};
    // adaptive rho (reference admm.cpp:331-357, rho_benchmark.cpp): the settings, the shared
    // sensitivity matrices (what codegen_with_sensitivity stores in the reference's cache, codegen.cpp:91-96) and the
    // adapted cache - rho, Kinf, Pinf - which the reference updates in place and which is therefore per-instance carried
    // state here: broadcast from the shared cache on a fresh / reset batch (ad_valid = false), then kept across solves.
    bool adaptive_rho = false;
    double adaptive_rho_min = 1.0, adaptive_rho_max = 100.0;   // tiny_api.cpp:368-369
    bool adaptive_clip = true;
    bool sens_set = false;
    DevBuf<double> d_dK, d_dP, ad_rho, ad_Kinf, ad_Pinf;
    bool ad_valid = false;
    int kernel_pref = TMPC_KERNEL_AUTO;
    bool rollout_fused = true;                // tmpc_batch_rollout_async: one fused launch where the TMEM kernel applies
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;       // tmpc_batch_solve_host: result chunks go home while the next chunk solves
    cudaStream_t up_stream = nullptr;         // ... and the x0 chunks come up in front of their launches
    std::vector<cudaEvent_t> chunk_done, chunk_up;
    // A solve / rollout enqueued on a CALLER stream is not ordered against the handle's own stream: the event below is
    // recorded behind it, and everything that later works on the handle's stream (getters, synchronize, setters, the
    // next solve) waits on it first.
    cudaEvent_t foreign_done = nullptr, own_done = nullptr;
    bool foreign_pending = false;
    bool own_pending = false;                 // asynchronous work sits on the handle's stream (solve_async / rollout_async with a NULL stream)
    // queue order of the on-chip kernels (queue_order.cu)
    int order_mode = TMPC_ORDER_AUTO;
    bool order_built = false;                 // `order` holds the permutation the last solve used
    DevBuf<int> order, order_user;
    bool results_mapped = false;              // Xsol / Usol / status / iters live in mapped pinned host memory (tiny batch)
    double* x0_stage = nullptr;               // pinned staging of x0 for tiny batches: set_x0 returns without a stream sync
    cudaEvent_t x0_staged = nullptr;
    bool x0_lazy = false;                     // the current x0 is in the staging buffer only; kernels read it there (mapped)
    long long launches = 0;
    bool state_zero = true;   // carried state is LOGICALLY zero (fresh or reset) ...
    bool mem_zero = true;     // ... and the device buffers physically hold those zeros.  tmpc_batch_reset only flips
                              // state_zero: the sub-warp kernel never reads a cold state and rewrites all of it, so
                              // the ~0.9 GB of memsets (quadrotor, B = 65,536) happen only for readers that need them.

    // problem constants
    DevBuf<double> dA, dB, df, dKinf, dPinf, dQuu, dAmBKt, dAPf, dBPf, dQd, dRd;
    // inputs
    DevBuf<double> x0, Xref, Uref, xmin, xmax, umin, umax;
    int xref_batch = 0, uref_batch = 0, bounds_batch = 0;      // 0 = never set
    bool bounds_knot_invariant = false;                        // shared bounds with identical columns (the usual MPC box)
    const double *ext_x0 = nullptr, *ext_xref = nullptr, *ext_uref = nullptr;
    int ext_xref_batch = 0, ext_uref_batch = 0;
    // constraints
    int ncx = 0, ncu = 0, nlx = 0, nlu = 0;
    int Acx[tmpc::kMaxCones]{}, Acu[tmpc::kMaxCones]{};
    double cx[tmpc::kMaxCones]{}, cu[tmpc::kMaxCones]{};
    DevBuf<double> Alin_x, blin_x, Alin_u, blin_u;
    // carried state + scratch
    DevBuf<double> g, y, v0, v1, z0, z1, q, r, pN, d, gc, yc, gl, yl;
    // results
    DevBuf<double> Xsol, Usol, u0_pack;
    DevBuf<int> status, iters, counter;
    // per-instance dynamics (tmpc_batch_set_instance_dynamics): element-major [e][ld] cache of every instance
    bool per_instance = false;
    DevBuf<double> pi_blocks;   // the same, one QConst-layout block per instance, for the TMEM kernel (compiled shapes)
    DevBuf<double> piA, piB, pif, piKinf, piPinf, piQuu, piAmBKt, piAPf, piBPf, piQd, piRd;
    DevBuf<double> state_im;  // instance-major twin of the carried state (TMEM kernel), allocated on first use
    bool state_in_im = false; // which format holds the current carried state (meaningless while state_zero)

    // Every API call starts here: select the device, and make the handle's own stream wait for work a previous call
    // enqueued on a caller stream (a stream-side wait; the host does not block).
    void use_device() {
        CUDA_TRY(cudaSetDevice(device));
        if (foreign_pending) { CUDA_TRY(cudaStreamWaitEvent(stream, foreign_done, 0)); foreign_pending = false; own_pending = true; }
    }
};

namespace {

using tmpc::Mat;

void upload_mats(tmpc_batch* h) {
    auto up = [&](DevBuf<double>& b, const std::vector<double>& v) { b.upload(v.data(), v.size(), h->stream); };
    up(h->dA, h->hm.A); up(h->dB, h->hm.B); up(h->df, h->hm.f); up(h->dKinf, h->hm.Kinf); up(h->dPinf, h->hm.Pinf);
    up(h->dQuu, h->hm.Quu_inv); up(h->dAmBKt, h->hm.AmBKt); up(h->dAPf, h->hm.APf); up(h->dBPf, h->hm.BPf);
    up(h->dQd, h->hm.Qd); up(h->dRd, h->hm.Rd);
    CUDA_TRY(cudaStreamSynchronize(h->stream));   // sources are host vectors that may change
}

// Cone / half-space duals are allocated on first use and zeroed ON THE STREAM THE SOLVE RUNS ON (a memset on the
// handle's stream would race with a solve enqueued on a caller stream).
void ensure_aux_state(tmpc_batch* h, cudaStream_t s) {
    const size_t sx = static_cast<size_t>(h->nx) * h->N * h->ld, su = static_cast<size_t>(h->nu) * (h->N - 1) * h->ld;
    if ((h->st.en_state_soc || h->st.en_input_soc) && !h->gc.p) {
        h->gc.alloc(sx); h->yc.alloc(su); h->gc.zero(s); h->yc.zero(s);
    }
    if ((h->st.en_state_linear || h->st.en_input_linear) && !h->gl.p) {
        h->gl.alloc(sx); h->yl.alloc(su); h->gl.zero(s); h->yl.zero(s);
    }
}

// About to enqueue on `s`: a caller stream first waits for asynchronous work still queued on the handle's stream
// (the other direction is use_device()).
void order_before_enqueue(tmpc_batch* h, cudaStream_t s) {
    if (s == h->stream || !h->own_pending) return;
    if (!h->own_done) CUDA_TRY(cudaEventCreateWithFlags(&h->own_done, cudaEventDisableTiming));
    CUDA_TRY(cudaEventRecord(h->own_done, h->stream));
    CUDA_TRY(cudaStreamWaitEvent(s, h->own_done, 0));
    h->own_pending = false;
}
// Work was enqueued on `s`: when that is a caller stream, leave an event behind it ...
void mark_enqueued(tmpc_batch* h, cudaStream_t s) {
    if (s == h->stream) { h->own_pending = true; return; }
    if (!h->foreign_done) CUDA_TRY(cudaEventCreateWithFlags(&h->foreign_done, cudaEventDisableTiming));
    CUDA_TRY(cudaEventRecord(h->foreign_done, s));
    h->foreign_pending = true;
}
// Block the host until everything this handle enqueued anywhere has finished (use_device() already chained the
// handle's stream behind caller-stream work).
void sync_all(tmpc_batch* h) {
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    h->own_pending = false;
}

// Tiny batch: bring an x0 that lives in the staging buffer only into the device buffer, before something other than a
// solve kernel (which reads the mapped staging buffer directly) needs it there.
void flush_x0(tmpc_batch* h, cudaStream_t s) {
    if (!h->x0_lazy) return;
    CUDA_TRY(cudaMemcpyAsync(h->x0.p, h->x0_stage, sizeof(double) * h->nx * h->B, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaEventRecord(h->x0_staged, s));
    h->x0_lazy = false;
    mark_enqueued(h, s);
}

tmpc::BatchView make_view(tmpc_batch* h) {
    tmpc::BatchView v{};
    v.B = h->B; v.ld = h->ld; v.nx = h->nx; v.nu = h->nu; v.N = h->N; v.rho = h->rho;
    const long long sxN = static_cast<long long>(h->nx) * h->N, suN = static_cast<long long>(h->nu) * (h->N - 1);
    v.x0 = h->ext_x0 ? h->ext_x0 : (h->x0_lazy ? h->x0_stage : h->x0.p);
    if (h->ext_xref) { v.Xref = h->ext_xref; v.xref_stride = h->ext_xref_batch > 1 ? sxN : 0; }
    else if (h->xref_batch) { v.Xref = h->Xref.p; v.xref_stride = h->xref_batch > 1 ? sxN : 0; }
    if (h->ext_uref) { v.Uref = h->ext_uref; v.uref_stride = h->ext_uref_batch > 1 ? suN : 0; }
    else if (h->uref_batch) { v.Uref = h->Uref.p; v.uref_stride = h->uref_batch > 1 ? suN : 0; }
    if (h->bounds_batch) {
        v.x_min = h->xmin.p; v.x_max = h->xmax.p; v.xb_stride = h->bounds_batch > 1 ? sxN : 0;
        v.u_min = h->umin.p; v.u_max = h->umax.p; v.ub_stride = h->bounds_batch > 1 ? suN : 0;
        v.bounds_const = h->bounds_knot_invariant ? 1 : 0;
    }
    v.A = h->dA.p; v.Bm = h->dB.p; v.f = h->df.p; v.Kinf = h->dKinf.p; v.Pinf = h->dPinf.p; v.Quu_inv = h->dQuu.p;
    v.AmBKt = h->dAmBKt.p; v.APf = h->dAPf.p; v.BPf = h->dBPf.p; v.Qd = h->dQd.p; v.Rd = h->dRd.p;
    v.ncx = h->ncx; v.ncu = h->ncu;
    for (int k = 0; k < tmpc::kMaxCones; ++k) { v.Acx[k] = h->Acx[k]; v.cx[k] = h->cx[k]; v.Acu[k] = h->Acu[k]; v.cu[k] = h->cu[k]; }
    v.nlx = h->nlx; v.Alin_x = h->Alin_x.p; v.blin_x = h->blin_x.p;
    v.nlu = h->nlu; v.Alin_u = h->Alin_u.p; v.blin_u = h->blin_u.p;
    v.g = h->g.p; v.y = h->y.p; v.v[0] = h->v0.p; v.v[1] = h->v1.p; v.z[0] = h->z0.p; v.z[1] = h->z1.p;
    v.q = h->q.p; v.r = h->r.p; v.pN = h->pN.p; v.d = h->d.p;
    v.gc = h->gc.p; v.yc = h->yc.p; v.gl = h->gl.p; v.yl = h->yl.p;
    v.Xsol = h->Xsol.p; v.Usol = h->Usol.p; v.status = h->status.p; v.iters = h->iters.p;
    v.state_im = h->state_im.p; v.state_im_stride = tmpc::state_im_values(h->nx, h->nu, h->N, h->precision);
    return v;
}

tmpc::PerInstanceMats per_instance_view(tmpc_batch* h) {
    return tmpc::PerInstanceMats{h->piA.p, h->piB.p, h->pif.p, h->piKinf.p, h->piPinf.p, h->piQuu.p, h->piAmBKt.p,
                                 h->piAPf.p, h->piBPf.p, h->piQd.p, h->piRd.p};
}

// The adapted cache is in use: the flag is on, or an earlier adaptive solve left per-instance rho / Kinf / Pinf behind
// (the reference keeps solving with the adapted cache after the flag is cleared; only a new setup restores it - here:
// reset, set_cache_terms, set_precision).
bool adaptive_path(const tmpc_batch* h) { return h->adaptive_rho || h->ad_valid; }

tmpc::AdaptiveRho adaptive_view(tmpc_batch* h) {
    return tmpc::AdaptiveRho{h->adaptive_rho ? 1 : 0, h->adaptive_clip ? 1 : 0, h->adaptive_rho_min, h->adaptive_rho_max,
                             h->ad_rho.p, h->ad_Kinf.p, h->ad_Pinf.p, h->d_dK.p, h->d_dP.p};
}

// nullptr = fine; else why a solve with adaptive rho cannot run
const char* adaptive_refusal(const tmpc_batch* h) {
    if (!adaptive_path(h)) return nullptr;
    if (h->adaptive_rho && !h->sens_set)
        return "adaptive_rho is set but no sensitivity matrices were given (tmpc_batch_set_sensitivity / codegen_with_sensitivity): "
               "the reference then adds 0 x 0 matrices to its cache and crashes at the first rho update (admm.cpp:331-357, "
               "rho_benchmark.cpp:204-207); the solve is refused instead";
    if (h->precision == 32) return "adaptive rho runs in the 64-bit build only";
    if (h->per_instance) return "adaptive rho needs shared dynamics (one set of sensitivity matrices)";
    if (h->kernel_pref != TMPC_KERNEL_AUTO && h->kernel_pref != TMPC_KERNEL_THREAD)
        return "adaptive rho runs on the thread kernel (TMPC_KERNEL_AUTO or TMPC_KERNEL_THREAD)";
    return nullptr;
}

int resolve_kernel(tmpc_batch* h) {
    if (adaptive_path(h)) return adaptive_refusal(h) ? -1 : TMPC_KERNEL_THREAD;
    if (h->precision == 32) {   // float build: the TMEM, sub-warp (shared-memory) and thread kernels are instantiated for float
        if (h->per_instance) return -1;
        const bool sub32 = tmpc::subwarp_kernel_supports(h->nx, h->nu, h->N, h->st, 32);
        const bool tm32 = tmpc::tmem_kernel_supports(make_view(h), h->st, 32);
        if (h->kernel_pref == TMPC_KERNEL_THREAD) return TMPC_KERNEL_THREAD;
        if (h->kernel_pref == TMPC_KERNEL_SUBWARP) return sub32 ? TMPC_KERNEL_SUBWARP : -1;
        if (h->kernel_pref == TMPC_KERNEL_TMEM) return tm32 ? TMPC_KERNEL_TMEM : -1;
        if (h->kernel_pref != TMPC_KERNEL_AUTO) return -1;
        return tm32 ? TMPC_KERNEL_TMEM : (sub32 ? TMPC_KERNEL_SUBWARP : TMPC_KERNEL_THREAD);
    }
    if (h->per_instance) {   // every instance has its own cache: TMEM kernel reading per-instance blocks, else thread kernel
        const bool tm_pi = h->pi_blocks.p && tmpc::tmem_pi_kernel_supports(make_view(h), h->st);
        if (h->kernel_pref == TMPC_KERNEL_THREAD) return TMPC_KERNEL_THREAD;
        if (h->kernel_pref == TMPC_KERNEL_TMEM) return tm_pi ? TMPC_KERNEL_TMEM : -1;
        if (h->kernel_pref != TMPC_KERNEL_AUTO) return -1;
        return tm_pi ? TMPC_KERNEL_TMEM : TMPC_KERNEL_THREAD;
    }
    const bool sub_ok = tmpc::subwarp_kernel_supports(h->nx, h->nu, h->N, h->st);
    const bool tm_ok = tmpc::tmem_kernel_supports(make_view(h), h->st);
    const bool cone_ok = tmpc::cone_kernel_supports(make_view(h), h->st);
    const bool warp_ok = tmpc::warp_kernel_supports(make_view(h), h->st);
    if (h->kernel_pref == TMPC_KERNEL_WARP) return warp_ok ? TMPC_KERNEL_WARP : -1;
    if (h->kernel_pref == TMPC_KERNEL_CONE) return cone_ok ? TMPC_KERNEL_CONE : -1;
    if (h->kernel_pref == TMPC_KERNEL_TMEM) return tm_ok ? TMPC_KERNEL_TMEM : -1;
    if (h->kernel_pref == TMPC_KERNEL_SUBWARP) return sub_ok ? TMPC_KERNEL_SUBWARP : -1;
    if (h->kernel_pref == TMPC_KERNEL_THREAD) return TMPC_KERNEL_THREAD;
    if (h->kernel_pref != TMPC_KERNEL_AUTO) return -1;
    if (warp_ok && h->B <= tmpc::warp_kernel_auto_max_b(h->nx)) return TMPC_KERNEL_WARP;       // small batch: latency first
    return tm_ok ? TMPC_KERNEL_TMEM : (sub_ok ? TMPC_KERNEL_SUBWARP : (cone_ok ? TMPC_KERNEL_CONE : TMPC_KERNEL_THREAD));
}

void zero_state(tmpc_batch* h, cudaStream_t s) {
    h->g.zero(s); h->y.zero(s); h->v0.zero(s); h->v1.zero(s); h->z0.zero(s); h->z1.zero(s);
    h->q.zero(s); h->r.zero(s); h->pN.zero(s); h->d.zero(s);
    h->gc.zero(s); h->yc.zero(s); h->gl.zero(s); h->yl.zero(s);
    h->Xsol.zero(s); h->Usol.zero(s); h->status.zero(s); h->iters.zero(s);
    h->state_zero = true;
    h->mem_zero = true;
}

// Make a logically-zero state physically zero before something reads it (thread kernel, result getters).
void materialize_zero(tmpc_batch* h) {
    if (h->state_zero && !h->mem_zero) { zero_state(h, h->stream); h->own_pending = true; }
}

// Queue order of the on-chip kernels for the coming solve (queue_order.cu), built on `s` in front of it; nullptr = natural
// order.  AUTO: longest-first by the LQR saturation of x0 against the input box (cold start) or by the previous solve's
// iteration counts (warm start); only when the batch is more than one wave of the persistent grid.
const int* build_order(tmpc_batch* h, const tmpc::BatchView& v, cudaStream_t s, int b0 = 0) {
    // `v` may be a chunk view [b0, b0 + v.B) of the batch: the permutation is then over the chunk's own instances
    const bool whole = (b0 == 0 && v.B == h->B);
    if (whole) h->order_built = false;
    if (h->order_mode == TMPC_ORDER_NATURAL) return nullptr;
    if (h->order_mode == TMPC_ORDER_USER) return whole ? h->order_user.p : nullptr;
    if (v.B <= h->num_sms * 64) return nullptr;
    tmpc::OrderArgs a{};
    a.B = v.B; a.nx = h->nx; a.nu = h->nu; a.x0 = v.x0;
    if (!h->state_zero) {
        a.prev_iters = v.iters;
    } else {
        if (!h->st.en_input_bound || !v.u_min || !v.u_max) return nullptr;
        a.u_min = v.u_min; a.u_max = v.u_max; a.ub_stride = v.ub_stride;
        if (h->per_instance) { a.Kinf = h->piKinf.p + b0; a.kinf_ld = h->ld; }
        else { a.Kinf = v.Kinf; a.kinf_ld = 0; }
    }
    if (!a.prev_iters && sizeof(double) * (256u * (h->nx + 1) + static_cast<size_t>(h->nu) * h->nx + h->nu) > 48 * 1024) return nullptr;
    if (!h->order.p) h->order.alloc(h->B);
    a.order = h->order.p + b0;
    static const int top_env = [] { const char* e = std::getenv("TMPC_ORDER_TOP"); return e ? std::atoi(e) : 0; }();   // experiment aid
    a.top = (top_env >= 1 && top_env <= 256) ? top_env : tmpc::kOrderTop;
    const cudaError_t e = tmpc::launch_build_order(a, s);
    if (e != cudaSuccess) throw CudaFail{e, "queue-order kernel"};
    h->launches += 1;
    if (whole) h->order_built = true;
    return a.order;
}

// Device alias of a host pointer when it is page-locked memory mapped into the device address space (cudaHostAlloc /
// cudaHostRegister under unified addressing: torch.pin_memory, tmpc_host_alloc), else nullptr.
template <class T>
T* mapped_alias(T* p) {
    if (!p) return nullptr;
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return static_cast<T*>(a.devicePointer);
}

// One solve kernel on `s`.  A cold (reset) state is passed to the sub-warp kernel as a flag; every other
// reader gets physically zeroed buffers first, on the SAME stream as the solve: an earlier solve enqueued on a
// caller stream may still be running, and memsets on another stream would race with it.
cudaError_t launch_solve(tmpc_batch* h, int k, const tmpc::BatchView& v_in, cudaStream_t s, bool ordered = true) {
    const bool tm = (k == TMPC_KERNEL_TMEM);
    const bool sub = (k == TMPC_KERNEL_SUBWARP || k == TMPC_KERNEL_CONE || k == TMPC_KERNEL_WARP || tm) && h->st.max_iter > 0 && h->rho > 0.0;
    tmpc::BatchView v = v_in;
    const bool wants_im = sub && tm;
    if (wants_im && !h->state_im.p) {
        h->state_im.alloc(static_cast<size_t>(tmpc::state_im_values(h->nx, h->nu, h->N, 64)) * h->B);
        v.state_im = h->state_im.p;
    }
    if (!h->state_zero && wants_im != h->state_in_im) {      // the previous solve left the other format: convert on `s`
        if (!v.state_im) return cudaErrorInvalidValue;
        cudaError_t ce = tmpc::launch_state_convert(v, wants_im, s, h->precision);
        if (ce != cudaSuccess) return ce;
        h->launches += 1;
    }
    if (!sub && h->state_zero && !h->mem_zero) {
        zero_state(h, s);
    }
    cudaError_t e;
    const int* order = (sub && tm && ordered) ? build_order(h, v, s) : nullptr;
    if (!ordered) h->order_built = false;
    if (sub && k == TMPC_KERNEL_TMEM && h->per_instance)
        e = tmpc::launch_admm_tmem_pi(v, h->st, h->state_zero ? 1 : 0, h->pi_blocks.p, h->counter.p, h->num_sms, s, order);
    else if (sub && k == TMPC_KERNEL_TMEM) e = tmpc::launch_admm_tmem(v, h->st, h->state_zero ? 1 : 0, h->counter.p, h->num_sms, s, order, h->precision);
    else if (sub && k == TMPC_KERNEL_WARP) e = tmpc::launch_admm_warp(v, h->st, h->state_zero ? 1 : 0, h->num_sms, s);
    else if (sub && k == TMPC_KERNEL_CONE) e = tmpc::launch_admm_cone(v, h->st, h->state_zero ? 1 : 0, h->counter.p, h->num_sms, s);
    else if (sub) e = tmpc::launch_admm_subwarp(v, h->st, h->hm, h->state_zero ? 1 : 0, h->counter.p, h->num_sms, s, h->precision);
    else if (adaptive_path(h)) {
        if (!h->ad_rho.p) {
            h->ad_rho.alloc(h->ld); h->ad_Kinf.alloc(static_cast<size_t>(h->nu) * h->nx * h->ld);
            h->ad_Pinf.alloc(static_cast<size_t>(h->nx) * h->nx * h->ld);
        }
        const tmpc::AdaptiveRho ad = adaptive_view(h);
        if (!h->ad_valid) {      // every instance starts from the cache the setup computed
            e = tmpc::launch_adaptive_init(v, ad, s);
            if (e != cudaSuccess) return e;
            h->launches += 1;
            h->ad_valid = true;
        }
        e = tmpc::launch_admm_thread_adaptive(v, h->st, ad, s);
    }
    else if (h->per_instance) e = tmpc::launch_admm_thread_per_instance(v, h->st, per_instance_view(h), s, h->precision);
    else e = tmpc::launch_admm_thread(v, h->st, h->hm, s, h->precision);
    h->launches += 1;
    if (h->st.max_iter > 0) { h->state_zero = false; h->mem_zero = false; h->state_in_im = wants_im; }
    return e;
}

// Enqueue one batched solve on `s` (ordered after everything already queued on the handle's own
// stream when s is that stream; callers passing another stream order it themselves).
int enqueue_solve(tmpc_batch* h, cudaStream_t s) {
    if (h->st.check_termination == 0) { set_error("check_termination = 0 is not supported (the reference divides by it)"); return TMPC_ERR_STATE; }
    if (const char* why = adaptive_refusal(h)) { set_error(why); return TMPC_ERR_STATE; }
    order_before_enqueue(h, s);
    ensure_aux_state(h, s);
    const int k = resolve_kernel(h);
    if (k < 0) { set_error("requested kernel does not support this problem (shape, constraint flags or per-instance bounds)"); return TMPC_ERR_STATE; }
    tmpc::BatchView v = make_view(h);
    cudaError_t e;
    e = launch_solve(h, k, v, s);
    if (e != cudaSuccess) throw CudaFail{e, "solve kernel launch"};
    mark_enqueued(h, s);
    return TMPC_OK;
}

template <class F>
int guarded(F&& body) {
    try {
        return body();
    } catch (const CudaFail& f) {
        set_error(std::string("CUDA error: ") + cudaGetErrorString(f.e) + " in " + f.what);
        return TMPC_ERR_CUDA;
    } catch (const std::bad_alloc&) {
        set_error("out of host memory");
        return TMPC_ERR_EXCEPTION;
    } catch (const std::exception& e) {
        set_error(std::string("exception: ") + e.what());
        return TMPC_ERR_EXCEPTION;
    } catch (...) {
        set_error("unknown exception");
        return TMPC_ERR_EXCEPTION;
    }
}

bool finite_all(const double* p, size_t n) {
    for (size_t i = 0; i < n; ++i) if (!std::isfinite(p[i])) return false;
    return true;
}

}  // namespace

extern "C" {

const char* tmpc_last_error(void) { return g_last_error.c_str(); }
const char* tmpc_version(void) { return "tinympc_b200 0.1 (sm_100a)"; }

static int compute_cache_any(int bits, const double* A, const double* Bdyn, const double* f, const double* Q, const double* R, double rho,
                       int nx, int nu, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                       double* BPf, double* Qd, double* Rd) {
    return guarded([&]() -> int {
        if (!A || !Bdyn || !f || !Q || !R || nx < 1 || nu < 1) { set_error("tmpc_compute_cache: bad argument"); return TMPC_ERR_DIM; }
        tmpc::LqrCache c;
        Mat mA(nx, nx, A), mB(nx, nu, Bdyn), mf(nx, 1, f), mQ(nx, nx, Q), mR(nu, nu, R);
        if ((bits == 32 ? tmpc::compute_lqr_cache_f32(mA, mB, mf, mQ, mR, rho, c) : tmpc::compute_lqr_cache(mA, mB, mf, mQ, mR, rho, c)) != 0) { set_error("tmpc_compute_cache: singular matrix"); return TMPC_ERR_DIM; }
        auto cp = [](double* dst, const std::vector<double>& v) { if (dst) std::memcpy(dst, v.data(), v.size() * sizeof(double)); };
        cp(Kinf, c.Kinf.a); cp(Pinf, c.Pinf.a); cp(Quu_inv, c.Quu_inv.a); cp(AmBKt, c.AmBKt.a); cp(APf, c.APf.a); cp(BPf, c.BPf.a);
        cp(Qd, c.Qd); cp(Rd, c.Rd);
        return c.riccati_sweeps > 0 ? TMPC_OK : TMPC_ERR_DIM;
    });
}

int tmpc_compute_cache(const double* A, const double* Bdyn, const double* f, const double* Q, const double* R, double rho,
                       int nx, int nu, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                       double* BPf, double* Qd, double* Rd) {
    return compute_cache_any(64, A, Bdyn, f, Q, R, rho, nx, nu, Kinf, Pinf, Quu_inv, AmBKt, APf, BPf, Qd, Rd);
}

int tmpc_compute_cache_f32(const double* A, const double* Bdyn, const double* f, const double* Q, const double* R, double rho,
                           int nx, int nu, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                           double* BPf, double* Qd, double* Rd) {
    return compute_cache_any(32, A, Bdyn, f, Q, R, rho, nx, nu, Kinf, Pinf, Quu_inv, AmBKt, APf, BPf, Qd, Rd);
}

int tmpc_batch_create(tmpc_batch** out, const double* A, const double* Bdyn, const double* f, const double* Q,
                      const double* R, double rho, int nx, int nu, int N, int B, int device, int verbose) {
    if (out) *out = nullptr;
    return guarded([&]() -> int {
        if (!out || !A || !Bdyn || !f || !Q || !R) { set_error("tmpc_batch_create: null argument"); return TMPC_ERR_DIM; }
        if (nx < 1 || nu < 1 || N < 2 || B < 1) { set_error("tmpc_batch_create: need nx >= 1, nu >= 1, N >= 2, B >= 1"); return TMPC_ERR_DIM; }
        if (!tmpc::thread_kernel_supports(nx, nu)) {
            set_error("tmpc_batch_create: nx/nu beyond the compiled limits (nx <= 32, nu <= 16)");
            return TMPC_ERR_DIM;
        }
        if (!finite_all(A, static_cast<size_t>(nx) * nx) || !finite_all(Bdyn, static_cast<size_t>(nx) * nu) ||
            !finite_all(f, nx) || !std::isfinite(rho)) { set_error("tmpc_batch_create: non-finite problem data"); return TMPC_ERR_DIM; }
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
            set_error("no CUDA device available: this library has no CPU path");
            return TMPC_ERR_CUDA;
        }
        std::unique_ptr<tmpc_batch> h(new tmpc_batch());
        if (device < 0) CUDA_TRY(cudaGetDevice(&device));
        h->device = device;
        h->use_device();
        cudaDeviceProp prop{};
        CUDA_TRY(cudaGetDeviceProperties(&prop, device));
        h->num_sms = prop.multiProcessorCount;
        h->nx = nx; h->nu = nu; h->N = N; h->B = B; h->rho = rho;
        h->ld = (B + 31) / 32 * 32;

        tmpc::LqrCache c;
        Mat mA(nx, nx, A), mB(nx, nu, Bdyn), mf(nx, 1, f), mQ(nx, nx, Q), mR(nu, nu, R);
        if (tmpc::compute_lqr_cache(mA, mB, mf, mQ, mR, rho, c) != 0) {
            set_error("tmpc_batch_create: singular matrix in the Riccati recursion");
            return TMPC_ERR_DIM;
        }
        if (verbose) std::printf("tinympc_b200: Kinf converged after %d Riccati sweeps (nx=%d nu=%d N=%d B=%d)\n",
                                 c.riccati_sweeps, nx, nu, N, B);
        h->hA = mA.a; h->hB = mB.a; h->hf = mf.a; h->hQ = mQ.a; h->hR = mR.a; h->rho_given = rho;
        h->hm.nx = nx; h->hm.nu = nu;
        h->hm.A = mA.a; h->hm.B = mB.a; h->hm.f = mf.a; h->hm.Kinf = c.Kinf.a; h->hm.Pinf = c.Pinf.a;
        h->hm.Quu_inv = c.Quu_inv.a; h->hm.AmBKt = c.AmBKt.a; h->hm.APf = c.APf.a; h->hm.BPf = c.BPf.a;
        h->hm.Qd = c.Qd; h->hm.Rd = c.Rd;

        // Julia-side defaults (src/TinyMPC.jl:57-61, :94-99)
        h->st.abs_pri_tol = 1e-3; h->st.abs_dua_tol = 1e-3; h->st.max_iter = 100; h->st.check_termination = 1;

        CUDA_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        upload_mats(h.get());
        const size_t sx = static_cast<size_t>(nx) * N * h->ld, su = static_cast<size_t>(nu) * (N - 1) * h->ld;
        h->g.alloc(sx); h->v0.alloc(sx); h->v1.alloc(sx); h->q.alloc(sx);
        h->y.alloc(su); h->z0.alloc(su); h->z1.alloc(su); h->r.alloc(su); h->d.alloc(su);
        h->pN.alloc(static_cast<size_t>(nx) * h->ld);
        h->x0.alloc(static_cast<size_t>(nx) * B); h->x0.zero(h->stream);
        // A tiny batch (the B = 1 singleton behind the reference's own symbols) keeps its results in mapped pinned host
        // memory and stages x0 through a pinned buffer: set_x0 -> solve_mpc -> get_controls then costs ONE stream
        // synchronisation (in solve_mpc) instead of one per call plus pageable copies.
        h->results_mapped = (static_cast<size_t>(nx) * N + static_cast<size_t>(nu) * (N - 1)) * B * sizeof(double) <= 64 * 1024;
        if (h->results_mapped) {
            h->Xsol.alloc_mapped(static_cast<size_t>(nx) * N * B); h->Usol.alloc_mapped(static_cast<size_t>(nu) * (N - 1) * B);
            h->status.alloc_mapped(B); h->iters.alloc_mapped(B);
            CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&h->x0_stage), sizeof(double) * nx * B, cudaHostAllocMapped));
            CUDA_TRY(cudaEventCreateWithFlags(&h->x0_staged, cudaEventDisableTiming));
        } else {
            h->Xsol.alloc(static_cast<size_t>(nx) * N * B); h->Usol.alloc(static_cast<size_t>(nu) * (N - 1) * B);
            h->status.alloc(B); h->iters.alloc(B);
        }
        h->counter.alloc(8); h->counter.zero(h->stream);
        zero_state(h.get(), h->stream);
        sync_all(h.get());
        *out = h.release();
        return TMPC_OK;
    });
}

int tmpc_batch_destroy(tmpc_batch* h) {
    if (!h) return TMPC_OK;
    return guarded([&]() -> int {
        cudaSetDevice(h->device);
        if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
        if (h->copy_stream) { cudaStreamSynchronize(h->copy_stream); cudaStreamDestroy(h->copy_stream); }
        if (h->up_stream) { cudaStreamSynchronize(h->up_stream); cudaStreamDestroy(h->up_stream); }
        for (cudaEvent_t e : h->chunk_done) cudaEventDestroy(e);
        for (cudaEvent_t e : h->chunk_up) cudaEventDestroy(e);
        if (h->foreign_done) { cudaEventSynchronize(h->foreign_done); cudaEventDestroy(h->foreign_done); }
        if (h->own_done) cudaEventDestroy(h->own_done);
        if (h->x0_staged) cudaEventDestroy(h->x0_staged);
        if (h->x0_stage) cudaFreeHost(h->x0_stage);
        delete h;
        return TMPC_OK;
    });
}

int tmpc_batch_update_settings(tmpc_batch* h, double abs_pri_tol, double abs_dua_tol, int max_iter,
                               int check_termination, int en_state_bound, int en_input_bound, int en_state_soc,
                               int en_input_soc, int en_state_linear, int en_input_linear, int adaptive_rho,
                               double adaptive_rho_min, double adaptive_rho_max, int adaptive_rho_enable_clipping) {
    return guarded([&]() -> int {
        if (!h) { set_error("update_settings: null handle"); return TMPC_ERR_DIM; }
        if (check_termination == 0) { set_error("update_settings: check_termination = 0 rejected (reference computes iter % 0)"); return TMPC_ERR_STATE; }
        if (max_iter <= 0) { set_error("update_settings: max_iter must be >= 1 (with 0 the reference returns whatever its workspace held, admm.cpp:384-388; nothing is solved)"); return TMPC_ERR_STATE; }
        h->st.abs_pri_tol = abs_pri_tol; h->st.abs_dua_tol = abs_dua_tol;
        h->st.max_iter = max_iter; h->st.check_termination = check_termination;
        h->st.en_state_bound = en_state_bound != 0; h->st.en_input_bound = en_input_bound != 0;
        h->st.en_state_soc = en_state_soc != 0; h->st.en_input_soc = en_input_soc != 0;
        h->st.en_state_linear = en_state_linear != 0; h->st.en_input_linear = en_input_linear != 0;
        // src/bindings.cpp:362-365.  A solve with the flag on needs the sensitivity matrices (tmpc_batch_set_sensitivity)
        h->adaptive_rho = adaptive_rho != 0;
        h->adaptive_rho_min = adaptive_rho_min; h->adaptive_rho_max = adaptive_rho_max;
        h->adaptive_clip = adaptive_rho_enable_clipping != 0;
        return TMPC_OK;
    });
}

/* Arithmetic type of the solve (reference: `typedef double tinytype`, types.hpp:15; float is its other build).
 * bits = 32 redoes the setup the way the float build of the reference does it — A, B, f, Q, R, rho rounded to float
 * (src/bindings.cpp:50-55), Qd / Rd and the Riccati cache computed in float (tiny_api.cpp:107-108, 244-318) — and
 * makes every later solve run the float instantiation of the thread kernel: inputs are rounded to float on load,
 * results widen to double, the C ABI and the device arrays stay double.  bits = 64 restores the default.  Call it on
 * a fresh or reset batch; a cache installed with set_cache_terms is replaced by the recomputed one. */
int tmpc_batch_set_precision(tmpc_batch* h, int bits) {
    return guarded([&]() -> int {
        if (!h || (bits != 32 && bits != 64)) { set_error("set_precision: bits must be 32 or 64"); return TMPC_ERR_DIM; }
        if (!h->state_zero) { set_error("set_precision: the batch carries solver state; reset it first"); return TMPC_ERR_STATE; }
        if (h->per_instance) { set_error("set_precision: per-instance dynamics are built for 64-bit only"); return TMPC_ERR_STATE; }
        if (bits == h->precision) return TMPC_OK;
        h->use_device();
        const int nx = h->nx, nu = h->nu;
        tmpc::LqrCache c;
        Mat mA(nx, nx, h->hA.data()), mB(nx, nu, h->hB.data()), mf(nx, 1, h->hf.data()), mQ(nx, nx, h->hQ.data()),
            mR(nu, nu, h->hR.data());
        const int rc = bits == 32 ? tmpc::compute_lqr_cache_f32(mA, mB, mf, mQ, mR, h->rho_given, c)
                                  : tmpc::compute_lqr_cache(mA, mB, mf, mQ, mR, h->rho_given, c);
        if (rc != 0) { set_error("set_precision: singular matrix in the Riccati recursion"); return TMPC_ERR_DIM; }
        auto rnd = [&](std::vector<double> v) {
            if (bits == 32) for (double& x : v) x = static_cast<double>(static_cast<float>(x));
            return v;
        };
        h->hm.A = rnd(h->hA); h->hm.B = rnd(h->hB); h->hm.f = rnd(h->hf);
        h->hm.Kinf = c.Kinf.a; h->hm.Pinf = c.Pinf.a; h->hm.Quu_inv = c.Quu_inv.a; h->hm.AmBKt = c.AmBKt.a;
        h->hm.APf = c.APf.a; h->hm.BPf = c.BPf.a; h->hm.Qd = c.Qd; h->hm.Rd = c.Rd;
        h->rho = bits == 32 ? static_cast<double>(static_cast<float>(h->rho_given)) : h->rho_given;
        h->precision = bits;
        sync_all(h);
        upload_mats(h);
        return TMPC_OK;
    });
}

int tmpc_batch_get_precision(tmpc_batch* h) { return h ? h->precision : -1; }

int tmpc_batch_set_kernel(tmpc_batch* h, int kernel) {
    if (!h || kernel < TMPC_KERNEL_AUTO || kernel > TMPC_KERNEL_WARP) return TMPC_ERR_DIM;
    h->kernel_pref = kernel;
    return TMPC_OK;
}

int tmpc_batch_get_kernel(tmpc_batch* h) { return h ? resolve_kernel(h) : -1; }

int tmpc_batch_set_order(tmpc_batch* h, int mode, const int* perm) {
    return guarded([&]() -> int {
        if (!h || mode < TMPC_ORDER_NATURAL || mode > TMPC_ORDER_USER) { set_error("set_order: mode must be TMPC_ORDER_NATURAL, _AUTO or _USER"); return TMPC_ERR_DIM; }
        if (mode == TMPC_ORDER_USER) {
            if (!perm) { set_error("set_order: TMPC_ORDER_USER needs a permutation of 0..B-1"); return TMPC_ERR_DIM; }
            std::vector<char> seen(h->B, 0);
            for (int i = 0; i < h->B; ++i) {
                if (perm[i] < 0 || perm[i] >= h->B || seen[perm[i]]) { set_error("set_order: not a permutation of 0..B-1"); return TMPC_ERR_DIM; }
                seen[perm[i]] = 1;
            }
            h->use_device();
            h->order_user.upload(perm, h->B, h->stream);
            sync_all(h);
        }
        h->order_mode = mode;
        h->order_built = false;
        return TMPC_OK;
    });
}

int tmpc_batch_get_order(tmpc_batch* h, int* perm) {
    return guarded([&]() -> int {
        if (!h || !perm) return TMPC_ERR_DIM;
        const int* src = h->order_mode == TMPC_ORDER_USER ? h->order_user.p : (h->order_built ? h->order.p : nullptr);
        if (!src) { set_error("get_order: the last solve ran in natural order"); return TMPC_ERR_STATE; }
        h->use_device();
        CUDA_TRY(cudaMemcpyAsync(perm, src, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_x0(tmpc_batch* h, const double* x0, int rows, int batch) {
    return guarded([&]() -> int {
        if (!h || !x0) { set_error("set_x0: null argument"); return TMPC_ERR_DIM; }
        if (rows != h->nx || batch != h->B) { set_error("set_x0: expected nx x B"); return TMPC_ERR_DIM; }
        h->use_device();
        h->ext_x0 = nullptr;
        if (h->x0_stage && !h->own_pending && !h->foreign_pending) {
            // tiny batch with nothing in flight (set_x0 -> solve -> get_controls): x0 goes into the mapped staging buffer and the
            // solve kernel reads it there - no CUDA call here, no copy in front of the kernel
            CUDA_TRY(cudaEventSynchronize(h->x0_staged));      // an upload from the staging buffer, if any, has left it
            std::memcpy(h->x0_stage, x0, sizeof(double) * h->nx * h->B);
            h->x0_lazy = true;
            return TMPC_OK;
        }
        if (h->x0_stage) {      // tiny batch: copy into the pinned staging buffer and return; the upload is stream-ordered
            if (h->x0_lazy) sync_all(h);      // a solve still in flight reads the staging buffer itself (use_device() chained caller-stream work)
            h->x0_lazy = false;
            CUDA_TRY(cudaEventSynchronize(h->x0_staged));      // the previous upload has left the staging buffer
            std::memcpy(h->x0_stage, x0, sizeof(double) * h->nx * h->B);
            CUDA_TRY(cudaMemcpyAsync(h->x0.p, h->x0_stage, sizeof(double) * h->nx * h->B, cudaMemcpyHostToDevice, h->stream));
            CUDA_TRY(cudaEventRecord(h->x0_staged, h->stream));
            h->own_pending = true;
            return TMPC_OK;
        }
        h->x0_lazy = false;
        h->x0.upload(x0, static_cast<size_t>(h->nx) * h->B, h->stream);
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_x_ref(tmpc_batch* h, const double* x_ref, int rows, int cols, int batch) {
    return guarded([&]() -> int {
        if (!h || !x_ref) { set_error("set_x_ref: null argument"); return TMPC_ERR_DIM; }
        if (rows != h->nx || cols != h->N || (batch != 1 && batch != h->B)) { set_error("set_x_ref: expected nx x N [x B]"); return TMPC_ERR_DIM; }
        h->use_device();
        h->ext_xref = nullptr;
        h->Xref.upload(x_ref, static_cast<size_t>(rows) * cols * batch, h->stream);
        h->xref_batch = batch;
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_u_ref(tmpc_batch* h, const double* u_ref, int rows, int cols, int batch) {
    return guarded([&]() -> int {
        if (!h || !u_ref) { set_error("set_u_ref: null argument"); return TMPC_ERR_DIM; }
        if (rows != h->nu || cols != h->N - 1 || (batch != 1 && batch != h->B)) { set_error("set_u_ref: expected nu x (N-1) [x B]"); return TMPC_ERR_DIM; }
        h->use_device();
        h->ext_uref = nullptr;
        h->Uref.upload(u_ref, static_cast<size_t>(rows) * cols * batch, h->stream);
        h->uref_batch = batch;
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_bound_constraints(tmpc_batch* h, const double* x_min, const double* x_max, const double* u_min,
                                     const double* u_max, int batch) {
    return guarded([&]() -> int {
        if (!h || !x_min || !x_max || !u_min || !u_max) { set_error("set_bound_constraints: null argument"); return TMPC_ERR_DIM; }
        if (batch != 1 && batch != h->B) { set_error("set_bound_constraints: batch must be 1 or B"); return TMPC_ERR_DIM; }
        h->use_device();
        const size_t sx = static_cast<size_t>(h->nx) * h->N * batch, su = static_cast<size_t>(h->nu) * (h->N - 1) * batch;
        h->xmin.upload(x_min, sx, h->stream); h->xmax.upload(x_max, sx, h->stream);
        h->umin.upload(u_min, su, h->stream); h->umax.upload(u_max, su, h->stream);
        h->bounds_batch = batch;
        h->bounds_knot_invariant = false;
        if (batch == 1) {   // the same box at every knot point: kernels may keep one column instead of the nx x N table
            bool same = true;
            auto cols_equal = [&](const double* m, int rows, int cols) {
                for (int c = 1; c < cols && same; ++c)
                    for (int r = 0; r < rows; ++r) if (!(m[c * rows + r] == m[r])) { same = false; break; }
            };
            cols_equal(x_min, h->nx, h->N); cols_equal(x_max, h->nx, h->N);
            cols_equal(u_min, h->nu, h->N - 1); cols_equal(u_max, h->nu, h->N - 1);
            h->bounds_knot_invariant = same;
        }
        h->st.en_state_bound = 1; h->st.en_input_bound = 1;     // src/bindings.cpp:400-404
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_linear_constraints(tmpc_batch* h, const double* Alin_x, int nlx, const double* blin_x,
                                      const double* Alin_u, int nlu, const double* blin_u) {
    return guarded([&]() -> int {
        if (!h || nlx < 0 || nlu < 0) { set_error("set_linear_constraints: bad argument"); return TMPC_ERR_DIM; }
        if ((nlx > 0 && (!Alin_x || !blin_x)) || (nlu > 0 && (!Alin_u || !blin_u))) { set_error("set_linear_constraints: null matrix"); return TMPC_ERR_DIM; }
        h->use_device();
        h->nlx = nlx; h->nlu = nlu;
        if (nlx) { h->Alin_x.upload(Alin_x, static_cast<size_t>(nlx) * h->nx, h->stream); h->blin_x.upload(blin_x, nlx, h->stream); }
        if (nlu) { h->Alin_u.upload(Alin_u, static_cast<size_t>(nlu) * h->nu, h->stream); h->blin_u.upload(blin_u, nlu, h->stream); }
        if (nlx > 0) h->st.en_state_linear = 1;                 // src/bindings.cpp:438-443
        if (nlu > 0) h->st.en_input_linear = 1;
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_cone_constraints(tmpc_batch* h, const int* Acx, const int* qcx, const double* cx, int ncx,
                                    const int* Acu, const int* qcu, const double* cu, int ncu) {
    return guarded([&]() -> int {
        if (!h || ncx < 0 || ncu < 0 || ncx > tmpc::kMaxCones || ncu > tmpc::kMaxCones) {
            set_error("set_cone_constraints: at most 4 cones per side"); return TMPC_ERR_DIM;
        }
        for (int k = 0; k < ncx; ++k)
            if (qcx[k] != 3 || Acx[k] < 0 || Acx[k] + 3 > h->nx) { set_error("set_cone_constraints: state cone must be a 3-block inside the state"); return TMPC_ERR_DIM; }
        for (int k = 0; k < ncu; ++k)
            if (qcu[k] != 3 || Acu[k] < 0 || Acu[k] + 3 > h->nu) { set_error("set_cone_constraints: input cone must be a 3-block inside the input"); return TMPC_ERR_DIM; }
        h->ncx = ncx; h->ncu = ncu;
        for (int k = 0; k < ncx; ++k) { h->Acx[k] = Acx[k]; h->cx[k] = cx[k]; }
        for (int k = 0; k < ncu; ++k) { h->Acu[k] = Acu[k]; h->cu[k] = cu[k]; }
        if (ncx > 0) h->st.en_state_soc = 1;
        if (ncu > 0) h->st.en_input_soc = 1;
        return TMPC_OK;
    });
}

int tmpc_batch_set_cone_constraints_jl(tmpc_batch* h, const int* Acu, const int* qcu, const double* cu, int ncu_len,
                                       const int* Acx, const int* qcx, const double* cx, int ncx_len) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        const tmpc::AdmmSettings before = h->st;
        // first triple -> state-cone slots, second -> input-cone slots (src/bindings.cpp:472-475 calling
        // tiny_api.cpp:166-168) ...
        const int rc = tmpc_batch_set_cone_constraints(h, Acu, qcu, cu, ncu_len, Acx, qcx, cx, ncx_len);
        if (rc != TMPC_OK) return rc;
        // ... while the flags follow the caller's naming and are only ever switched on (src/bindings.cpp:478-483)
        h->st.en_state_soc = before.en_state_soc || (ncx_len > 0);
        h->st.en_input_soc = before.en_input_soc || (ncu_len > 0);
        return TMPC_OK;
    });
}

int tmpc_batch_get_settings(tmpc_batch* h, double* tols, int* ints) {
    if (!h) return TMPC_ERR_DIM;
    if (tols) { tols[0] = h->st.abs_pri_tol; tols[1] = h->st.abs_dua_tol; }
    if (ints) {
        ints[0] = h->st.max_iter; ints[1] = h->st.check_termination; ints[2] = h->st.en_state_bound; ints[3] = h->st.en_input_bound;
        ints[4] = h->st.en_state_soc; ints[5] = h->st.en_input_soc; ints[6] = h->st.en_state_linear; ints[7] = h->st.en_input_linear;
    }
    return TMPC_OK;
}

/* Per-instance dynamics (the setup the reference runs once per solver, tiny_api.cpp:21-137 and
 * :244-318, run for B different systems on the device).  A nx x nx x B, Bdyn nx x nu x B, f nx x B or NULL (zeros),
 * Q nx x nx [x B], R nu x nu [x B] (qr_batch = 1: shared, B: per instance), Julia layout, host pointers; rho is the
 * handle's.  Replaces the shared cache for every later solve and clears the carried state (as a new setup would). */
int tmpc_batch_set_instance_dynamics(tmpc_batch* h, const double* A, const double* Bdyn, const double* f, const double* Q,
                                     const double* R, int qr_batch) {
    return guarded([&]() -> int {
        if (!h || !A || !Bdyn || !Q || !R) { set_error("set_instance_dynamics: null argument"); return TMPC_ERR_DIM; }
        if (qr_batch != 1 && qr_batch != h->B) { set_error("set_instance_dynamics: qr_batch must be 1 or B"); return TMPC_ERR_DIM; }
        if (h->precision == 32) { set_error("set_instance_dynamics: per-instance dynamics are built for 64-bit only"); return TMPC_ERR_STATE; }
        if (h->nx > tmpc::kPerInstanceMaxNx || h->nu > tmpc::kPerInstanceMaxNu) {
            set_error("set_instance_dynamics: per-instance dynamics are compiled for nx <= 12, nu <= 4"); return TMPC_ERR_DIM;
        }
        h->use_device();
        const size_t n = h->nx, m = h->nu, B = h->B, ld = h->ld;
        if (!finite_all(A, n * n * B) || !finite_all(Bdyn, n * m * B) || (f && !finite_all(f, n * B)) ||
            !finite_all(Q, n * n * (qr_batch > 1 ? B : 1)) || !finite_all(R, m * m * (qr_batch > 1 ? B : 1))) {
            set_error("set_instance_dynamics: non-finite input"); return TMPC_ERR_DIM;
        }
        DevBuf<double> dA, dB, df, dQ, dR;
        DevBuf<int> fail;
        dA.upload(A, n * n * B, h->stream); dB.upload(Bdyn, n * m * B, h->stream);
        if (f) df.upload(f, n * B, h->stream);
        dQ.upload(Q, n * n * (qr_batch > 1 ? B : 1), h->stream); dR.upload(R, m * m * (qr_batch > 1 ? B : 1), h->stream);
        fail.alloc(1); fail.zero(h->stream);
        h->piA.alloc(n * n * ld); h->piB.alloc(n * m * ld); h->pif.alloc(n * ld); h->piKinf.alloc(m * n * ld);
        h->piPinf.alloc(n * n * ld); h->piQuu.alloc(m * m * ld); h->piAmBKt.alloc(n * n * ld); h->piAPf.alloc(n * ld);
        h->piBPf.alloc(m * ld); h->piQd.alloc(n * ld); h->piRd.alloc(m * ld);
        cudaError_t e = tmpc::launch_riccati_cache(h->nx, h->nu, h->B, h->ld, h->rho, dA.p, dB.p, f ? df.p : nullptr, dQ.p,
                                                   qr_batch > 1 ? static_cast<long long>(n * n) : 0, dR.p,
                                                   qr_batch > 1 ? static_cast<long long>(m * m) : 0, per_instance_view(h),
                                                   fail.p, h->stream);
        if (e != cudaSuccess) throw CudaFail{e, "batched cache kernel launch"};
        h->launches += 1;
        int failed = 0;
        CUDA_TRY(cudaMemcpyAsync(&failed, fail.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        if (failed) { h->per_instance = false; set_error("set_instance_dynamics: singular R + B'PB for some instance"); return TMPC_ERR_STATE; }
        const long long blk = tmpc::tmem_pi_block_doubles(h->nx, h->nu);
        h->pi_blocks.release();
        if (blk > 0) {
            h->pi_blocks.alloc(static_cast<size_t>(blk) * B);
            e = tmpc::launch_tmem_pi_pack(h->nx, h->nu, per_instance_view(h), h->ld, h->B, h->pi_blocks.p, h->stream);
            if (e != cudaSuccess) throw CudaFail{e, "per-instance block packing"};
            h->launches += 1;
            sync_all(h);
        }
        h->per_instance = true;
        h->state_zero = true;
        return TMPC_OK;
    });
}

/* Cache of ONE instance of a per-instance batch (column-major; any pointer may be NULL). */
int tmpc_batch_get_instance_cache(tmpc_batch* h, int b, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt,
                                  double* APf, double* BPf) {
    return guarded([&]() -> int {
        if (!h || !h->per_instance || b < 0 || b >= h->B) return TMPC_ERR_DIM;
        h->use_device();
        const size_t n = h->nx, m = h->nu, pitch = sizeof(double) * h->ld;
        auto get = [&](double* dst, const DevBuf<double>& src, size_t elems) {
            if (dst) CUDA_TRY(cudaMemcpy2DAsync(dst, sizeof(double), src.p + b, pitch, sizeof(double), elems,
                                               cudaMemcpyDeviceToHost, h->stream));
        };
        get(Kinf, h->piKinf, m * n); get(Pinf, h->piPinf, n * n); get(Quu_inv, h->piQuu, m * m);
        get(AmBKt, h->piAmBKt, n * n); get(APf, h->piAPf, n); get(BPf, h->piBPf, m);
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_set_cache_terms(tmpc_batch* h, const double* Kinf, const double* Pinf, const double* Quu_inv,
                               const double* AmBKt) {
    return guarded([&]() -> int {
        if (!h || !Kinf || !Pinf || !Quu_inv || !AmBKt) { set_error("set_cache_terms: null argument"); return TMPC_ERR_DIM; }
        if (h->per_instance) { set_error("set_cache_terms: the batch has per-instance dynamics (tmpc_batch_set_instance_dynamics); one shared cache cannot replace B of them"); return TMPC_ERR_STATE; }
        h->use_device();
        const int n = h->nx, m = h->nu;
        h->hm.Kinf.assign(Kinf, Kinf + static_cast<size_t>(m) * n);
        h->hm.Pinf.assign(Pinf, Pinf + static_cast<size_t>(n) * n);
        h->hm.Quu_inv.assign(Quu_inv, Quu_inv + static_cast<size_t>(m) * m);
        h->hm.AmBKt.assign(AmBKt, AmBKt + static_cast<size_t>(n) * n);
        if (h->precision == 32)      // Kinf.cast<tinytype>() ... (src/bindings.cpp:277-280)
            for (auto* vec : {&h->hm.Kinf, &h->hm.Pinf, &h->hm.Quu_inv, &h->hm.AmBKt})
                for (double& x : *vec) x = static_cast<double>(static_cast<float>(x));
        sync_all(h);
        upload_mats(h);      // APf, BPf, rho deliberately untouched
        if (h->ad_valid) {   // an adapted batch: the new Kinf / Pinf replace every instance's, cache->rho stays (bindings.cpp:277-280)
            const cudaError_t e = tmpc::launch_adaptive_init(make_view(h), adaptive_view(h), h->stream, /*keep_rho=*/true);
            if (e != cudaSuccess) throw CudaFail{e, "adapted-cache refresh"};
            h->launches += 1;
            h->own_pending = true;
        }
        return TMPC_OK;
    });
}

int tmpc_batch_get_cache_terms(tmpc_batch* h, double* Kinf, double* Pinf, double* Quu_inv, double* AmBKt, double* APf,
                               double* BPf) {
    if (!h) return TMPC_ERR_DIM;
    auto cp = [](double* dst, const std::vector<double>& v) { if (dst) std::memcpy(dst, v.data(), v.size() * sizeof(double)); };
    cp(Kinf, h->hm.Kinf); cp(Pinf, h->hm.Pinf); cp(Quu_inv, h->hm.Quu_inv); cp(AmBKt, h->hm.AmBKt);
    cp(APf, h->hm.APf); cp(BPf, h->hm.BPf);
    return TMPC_OK;
}

int tmpc_batch_set_sensitivity(tmpc_batch* h, const double* dKinf_drho, const double* dPinf_drho, const double* dC1_drho,
                               const double* dC2_drho) {
    return guarded([&]() -> int {
        if (!h || !dKinf_drho || !dPinf_drho || !dC1_drho || !dC2_drho) { set_error("set_sensitivity: null argument"); return TMPC_ERR_DIM; }
        const size_t n = h->nx, m = h->nu;
        if (!finite_all(dKinf_drho, m * n) || !finite_all(dPinf_drho, n * n) || !finite_all(dC1_drho, m * m) || !finite_all(dC2_drho, n * n)) {
            set_error("set_sensitivity: non-finite entry"); return TMPC_ERR_DIM;
        }
        h->use_device();
        sync_all(h);
        // dC1 / dC2 update C1 / C2 in the reference (rho_benchmark.cpp:206-207), which no part of the solve reads
        // (SURVEY.md 8a a12): checked, not stored
        h->d_dK.upload(dKinf_drho, m * n, h->stream);
        h->d_dP.upload(dPinf_drho, n * n, h->stream);
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        h->sens_set = true;
        return TMPC_OK;
    });
}

int tmpc_batch_get_adapted_cache(tmpc_batch* h, double* rho, double* Kinf, double* Pinf) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        sync_all(h);
        const size_t nk = static_cast<size_t>(h->nu) * h->nx, np = static_cast<size_t>(h->nx) * h->nx, B = h->B;
        if (!h->ad_valid) {      // nothing adapted yet: every instance is on the shared cache
            for (size_t b = 0; b < B; ++b) {
                if (rho) rho[b] = h->rho;
                if (Kinf) std::memcpy(Kinf + b * nk, h->hm.Kinf.data(), sizeof(double) * nk);
                if (Pinf) std::memcpy(Pinf + b * np, h->hm.Pinf.data(), sizeof(double) * np);
            }
            return TMPC_OK;
        }
        DevBuf<double> tk, tp;
        if (Kinf) tk.alloc(nk * B);
        if (Pinf) tp.alloc(np * B);
        const cudaError_t e = tmpc::launch_adaptive_gather(make_view(h), adaptive_view(h), tk.p, tp.p, h->stream);
        if (e != cudaSuccess) throw CudaFail{e, "adapted-cache gather"};
        h->launches += 1;
        if (rho) CUDA_TRY(cudaMemcpyAsync(rho, h->ad_rho.p, sizeof(double) * B, cudaMemcpyDeviceToHost, h->stream));
        if (Kinf) CUDA_TRY(cudaMemcpyAsync(Kinf, tk.p, sizeof(double) * nk * B, cudaMemcpyDeviceToHost, h->stream));
        if (Pinf) CUDA_TRY(cudaMemcpyAsync(Pinf, tp.p, sizeof(double) * np * B, cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(cudaStreamSynchronize(h->stream));
        return TMPC_OK;
    });
}

int tmpc_batch_reset(tmpc_batch* h) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        h->state_zero = true;            // logical: physical zeroing is deferred to the first reader that needs it (stream-ordered there)
        h->order_built = false;
        h->ad_valid = false;             // ... and every instance is back on the cache the setup computed
        return TMPC_OK;
    });
}

int tmpc_batch_solve_async(tmpc_batch* h, void* cuda_stream) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        return enqueue_solve(h, cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->stream);
    });
}

int tmpc_batch_synchronize(tmpc_batch* h) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_solve(tmpc_batch* h) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        const int rc = enqueue_solve(h, h->stream);
        if (rc != TMPC_OK) return rc < 0 ? rc : 100 + rc;
        if (h->results_mapped) {      // the kernel wrote the status words to host memory
            sync_all(h);
            for (int b = 0; b < h->B; ++b) if (h->status.p[b] != 0) return 1;
            return 0;
        }
        std::vector<int> st(h->B);
        CUDA_TRY(cudaMemcpyAsync(st.data(), h->status.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        for (int s : st) if (s != 0) return 1;
        return 0;
    });
}

/* One call, host buffers in and out: x0 up, solve, solution + status down.  The batch is cut into chunks of whole
 * instances (the on-chip kernels own an instance per sub-warp and the carried state is one block per instance, so a
 * chunk is just an offset view); chunk c solves on the handle's stream while chunk c-1 travels home on a second
 * stream and the x0 of the chunks ahead come up on a third, which hides the solve behind the PCIe transfer of the
 * solution (166 MB for the quadrotor at B = 65,536).  Output selector: any of X, U, u0 (the first control of every
 * instance, nu x B - what a closed loop consumes, examples/cartpole_example_mpc.jl:41-42), status, iters may be NULL. */
static int solve_host_impl(tmpc_batch* h, const double* x0, int cold, double* X, double* U, double* u0, int* status,
                           int* iters, int chunks) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        if (h->st.check_termination == 0) { set_error("check_termination = 0 is not supported (the reference divides by it)"); return 100 + TMPC_ERR_STATE; }
        if (const char* why = adaptive_refusal(h)) { set_error(why); return 100 + TMPC_ERR_STATE; }
        if (x0) {   // copied below: whole before a single launch, else chunk by chunk in front of each chunk's launch
            h->ext_x0 = nullptr;
            h->x0_lazy = false;
            if (h->x0.n < static_cast<size_t>(h->nx) * h->B) h->x0.alloc(static_cast<size_t>(h->nx) * h->B);
        }
        if (cold) { h->state_zero = true; h->order_built = false; h->ad_valid = false; }
        ensure_aux_state(h, h->stream);
        const int k = resolve_kernel(h);
        if (k < 0) { set_error("requested kernel does not support this problem"); return 100 + TMPC_ERR_STATE; }
        const bool tm = (k == TMPC_KERNEL_TMEM) && h->st.max_iter > 0 && h->rho > 0.0;
        // chunks are whole waves of the persistent grid (one CTA per SM, 64 resident instances each).  Default with the
        // full solution going home: one wave per chunk, at most 8 chunks - the device->host copy is the long pole
        // (166 MB against 0.8 ms of solve at B = 65,536 quadrotors), so an early first copy is worth more than the
        // refill efficiency of longer chunks (measured: 4 chunks of 2 waves 3.44 ms, 7 chunks of 1 wave 3.37 ms).
        // With only u0 / status going home the solve is the long pole: two chunks, a short one first so that the
        // rest of x0 comes up behind it, the second keeping the queue (and its longest-first order) over 7/8 of the batch.
        const int wave = h->num_sms * 64;
        const size_t out_bytes = sizeof(double) * h->B * ((X ? static_cast<size_t>(h->nx) * h->N : 0) + (U ? static_cast<size_t>(h->nu) * (h->N - 1) : 0));
        // default: a chunk per 16 MB going home (0.3 ms of PCIe), at most 8 and at most one per wave.  When only u0 /
        // status go home the solve is the long pole and the batch stays ONE launch: every cut of the queue adds a drain
        // (a one-wave chunk lasts as long as its longest instance), which costs more than overlapping the 0.1 ms upload
        // of x0 saves (measured: 61 M solves/s as [1 wave | rest] against the single launch).
        int nch = chunks > 0 ? chunks : static_cast<int>(std::min<size_t>(std::min(8, (h->B + wave - 1) / wave), out_bytes / (16u << 20)));
        if (nch < 1) nch = 1;
        if (!tm || h->B < 3 * wave) nch = 1;
        std::vector<int> cut;     // chunk c = [cut[c], cut[c + 1])
        if (nch > 1) {
            int chunk_size = ((h->B + nch - 1) / nch + wave - 1) / wave * wave;
            if (chunk_size < wave) chunk_size = wave;
            for (int b0 = 0; b0 < h->B; b0 += chunk_size) cut.push_back(b0);
            cut.push_back(h->B);
        } else {
            cut = {0, h->B};
        }
        nch = static_cast<int>(cut.size()) - 1;
        if (!h->copy_stream) CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        if (!h->up_stream) CUDA_TRY(cudaStreamCreateWithFlags(&h->up_stream, cudaStreamNonBlocking));
        while (static_cast<int>(h->chunk_done.size()) < nch) {
            cudaEvent_t e; CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->chunk_done.push_back(e);
            CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->chunk_up.push_back(e);
        }
        std::vector<int> st_tmp;
        if (!status) { st_tmp.resize(h->B); status = st_tmp.data(); }
        tmpc::BatchView v = make_view(h);
        const size_t sxN = static_cast<size_t>(h->nx) * h->N, suN = static_cast<size_t>(h->nu) * (h->N - 1);
        const size_t nu = static_cast<size_t>(h->nu);
        if (u0 && h->u0_pack.n < nu * h->B) h->u0_pack.alloc(nu * h->B);
        // [b0, b1) of the results, on stream `cs` (the solve that produced them is already ordered before it)
        auto fetch = [&](int b0, int b1, cudaStream_t cs) {
            const size_t nb = static_cast<size_t>(b1 - b0);
            if (X) CUDA_TRY(cudaMemcpyAsync(X + b0 * sxN, v.Xsol + b0 * sxN, sizeof(double) * sxN * nb, cudaMemcpyDeviceToHost, cs));
            if (U) CUDA_TRY(cudaMemcpyAsync(U + b0 * suN, v.Usol + b0 * suN, sizeof(double) * suN * nb, cudaMemcpyDeviceToHost, cs));
            if (u0) {
                cudaError_t ge = tmpc::launch_gather_u0(v.Usol + b0 * suN, static_cast<long long>(suN), h->nu, b1 - b0, h->u0_pack.p + b0 * nu, cs);
                if (ge != cudaSuccess) throw CudaFail{ge, "u0 gather launch"};
                h->launches += 1;
                CUDA_TRY(cudaMemcpyAsync(u0 + b0 * nu, h->u0_pack.p + b0 * nu, sizeof(double) * nu * nb, cudaMemcpyDeviceToHost, cs));
            }
        };
        if (nch > 1) {
            const bool was_cold = h->state_zero;
            if (!h->state_im.p) { h->state_im.alloc(static_cast<size_t>(tmpc::state_im_values(h->nx, h->nu, h->N, 64)) * h->B); v.state_im = h->state_im.p; }
            if (!was_cold && !h->state_in_im) {
                cudaError_t ce = tmpc::launch_state_convert(v, true, h->stream, h->precision);
                if (ce != cudaSuccess) throw CudaFail{ce, "state conversion"};
                h->launches += 1;
            }
            if (x0) {   // every chunk's x0 on the upload stream, each followed by its event; nothing older may still read h->x0
                CUDA_TRY(cudaEventRecord(h->chunk_done[0], h->stream));
                CUDA_TRY(cudaStreamWaitEvent(h->up_stream, h->chunk_done[0], 0));
                for (int c = 0; c < nch; ++c) {
                    const size_t o = static_cast<size_t>(cut[c]) * h->nx;
                    CUDA_TRY(cudaMemcpyAsync(h->x0.p + o, x0 + o, sizeof(double) * h->nx * (cut[c + 1] - cut[c]), cudaMemcpyHostToDevice, h->up_stream));
                    CUDA_TRY(cudaEventRecord(h->chunk_up[c], h->up_stream));
                }
            }
            for (int c = 0; c < nch; ++c) {
                const int b0 = cut[c], b1 = cut[c + 1];
                tmpc::BatchView vc = v;
                vc.B = b1 - b0;
                if (x0) CUDA_TRY(cudaStreamWaitEvent(h->stream, h->chunk_up[c], 0));
                vc.x0 = v.x0 + static_cast<size_t>(b0) * h->nx;
                if (v.Xref && v.xref_stride) vc.Xref = v.Xref + static_cast<size_t>(b0) * v.xref_stride;
                if (v.Uref && v.uref_stride) vc.Uref = v.Uref + static_cast<size_t>(b0) * v.uref_stride;
                // one block of state_im_stride VALUES per instance: doubles, or floats in the float build
                vc.state_im = reinterpret_cast<double*>(reinterpret_cast<char*>(v.state_im) +
                                                        static_cast<size_t>(b0) * v.state_im_stride * (h->precision == 32 ? 4 : 8));
                vc.Xsol = v.Xsol + b0 * sxN; vc.Usol = v.Usol + b0 * suN;
                vc.status = v.status + b0; vc.iters = v.iters + b0;
                // a chunk longer than one wave goes through the queue: longest-first order over the chunk's own instances
                const int* order = build_order(h, vc, h->stream, b0);
                cudaError_t e = h->per_instance
                    ? tmpc::launch_admm_tmem_pi(vc, h->st, was_cold ? 1 : 0,
                                                h->pi_blocks.p + static_cast<size_t>(b0) * tmpc::tmem_pi_block_doubles(h->nx, h->nu),
                                                h->counter.p, h->num_sms, h->stream, order)
                    : tmpc::launch_admm_tmem(vc, h->st, was_cold ? 1 : 0, h->counter.p, h->num_sms, h->stream, order, h->precision);
                if (e != cudaSuccess) throw CudaFail{e, "solve kernel launch"};
                h->launches += 1;
                CUDA_TRY(cudaEventRecord(h->chunk_done[c], h->stream));
                CUDA_TRY(cudaStreamWaitEvent(h->copy_stream, h->chunk_done[c], 0));
                fetch(b0, b1, h->copy_stream);
                if (c == nch - 1) {   // status / iteration counts of the whole batch behind the last chunk: two copies, not 2 per chunk
                    CUDA_TRY(cudaMemcpyAsync(status, v.status, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->copy_stream));
                    if (iters) CUDA_TRY(cudaMemcpyAsync(iters, v.iters, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->copy_stream));
                }
            }
            h->state_zero = false; h->mem_zero = false; h->state_in_im = true; h->order_built = false;
            CUDA_TRY(cudaStreamSynchronize(h->copy_stream));
            sync_all(h);
        } else if (tm && !X && !U && u0 && mapped_alias(x0) && h->B > wave &&
                   (!h->state_zero || sizeof(double) * h->nx * h->B >= (4u << 20))) {
            // Closed-loop shape of the call (only the first controls and the status come home) on page-locked buffers:
            // the solve is the long pole, so nothing may sit in front of it.  The kernel reads x0 straight from the
            // mapped host buffer - 96 bytes per instance at refill, not needed before the first forward sweep, 6 MB
            // spread over the whole launch - instead of waiting 0.12 ms for an upload.  Warm start: the queue order
            // comes from the previous iteration counts, which are on the device.  Cold start: the longest-first
            // pre-pass would have to pull all of x0 across PCIe first, so the queue keeps its natural order (config 3:
            // 2-3 % against 13 %); a small x0 (config 2: 2 MB, and a heavy-tailed iteration count) takes the upload path;
            // the device copy of x0 that later calls see is refreshed on the upload stream beside the solve; and one small
            // kernel writes u0 / status / iterations into the caller's mapped buffers.
            tmpc::BatchView vz = v;
            vz.x0 = mapped_alias(x0);
            CUDA_TRY(cudaEventRecord(h->chunk_done[0], h->stream));
            CUDA_TRY(cudaStreamWaitEvent(h->up_stream, h->chunk_done[0], 0));
            CUDA_TRY(cudaMemcpyAsync(h->x0.p, x0, sizeof(double) * h->nx * h->B, cudaMemcpyHostToDevice, h->up_stream));
            CUDA_TRY(cudaEventRecord(h->chunk_up[0], h->up_stream));
            cudaError_t e = launch_solve(h, k, vz, h->stream, /*ordered=*/!h->state_zero);
            if (e != cudaSuccess) throw CudaFail{e, "solve kernel launch"};
            double* u0m = mapped_alias(u0); int* stm = mapped_alias(status); int* itm = iters ? mapped_alias(iters) : nullptr;
            if (u0m && stm && (itm || !iters)) {
                e = tmpc::launch_gather_u0(v.Usol, static_cast<long long>(suN), h->nu, h->B, u0m, h->stream, v.status, v.iters, stm, itm);
                if (e != cudaSuccess) throw CudaFail{e, "u0 gather launch"};
                h->launches += 1;
            } else {
                fetch(0, h->B, h->stream);
                CUDA_TRY(cudaMemcpyAsync(status, h->status.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
                if (iters) CUDA_TRY(cudaMemcpyAsync(iters, h->iters.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
            }
            CUDA_TRY(cudaStreamWaitEvent(h->stream, h->chunk_up[0], 0));
            sync_all(h);
        } else {
            if (x0) CUDA_TRY(cudaMemcpyAsync(h->x0.p, x0, sizeof(double) * h->nx * h->B, cudaMemcpyHostToDevice, h->stream));
            cudaError_t e = launch_solve(h, k, v, h->stream);
            if (e != cudaSuccess) throw CudaFail{e, "solve kernel launch"};
            fetch(0, h->B, h->stream);
            CUDA_TRY(cudaMemcpyAsync(status, h->status.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
            if (iters) CUDA_TRY(cudaMemcpyAsync(iters, h->iters.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
            sync_all(h);
        }
        for (int b = 0; b < h->B; ++b) if (status[b] != 0) return 1;
        return 0;
    });
}

int tmpc_batch_solve_host(tmpc_batch* h, const double* x0, int cold, double* X, double* U, int* status, int* iters,
                          int chunks) {
    return solve_host_impl(h, x0, cold, X, U, nullptr, status, iters, chunks);
}

int tmpc_batch_solve_host_select(tmpc_batch* h, const double* x0, int cold, double* X, double* U, double* u0, int* status,
                                 int* iters, int chunks) {
    return solve_host_impl(h, x0, cold, X, U, u0, status, iters, chunks);
}

int tmpc_batch_get_status(tmpc_batch* h, int* status, int* iters) {
    return guarded([&]() -> int {
        if (!h) return TMPC_ERR_DIM;
        h->use_device();
        materialize_zero(h);
        if (h->results_mapped) {
            if (h->own_pending) sync_all(h);
            if (status) std::memcpy(status, h->status.p, sizeof(int) * h->B);
            if (iters) std::memcpy(iters, h->iters.p, sizeof(int) * h->B);
            return TMPC_OK;
        }
        if (status) CUDA_TRY(cudaMemcpyAsync(status, h->status.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
        if (iters) CUDA_TRY(cudaMemcpyAsync(iters, h->iters.p, sizeof(int) * h->B, cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_get_states(tmpc_batch* h, double* buf) {
    return guarded([&]() -> int {
        if (!h || !buf) return TMPC_ERR_DIM;
        h->use_device();
        materialize_zero(h);
        if (h->results_mapped) {
            if (h->own_pending) sync_all(h);
            std::memcpy(buf, h->Xsol.p, sizeof(double) * h->nx * h->N * h->B);
            return TMPC_OK;
        }
        CUDA_TRY(cudaMemcpyAsync(buf, h->Xsol.p, sizeof(double) * h->nx * h->N * h->B, cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_get_controls(tmpc_batch* h, double* buf) {
    return guarded([&]() -> int {
        if (!h || !buf) return TMPC_ERR_DIM;
        h->use_device();
        materialize_zero(h);
        if (h->results_mapped) {
            if (h->own_pending) sync_all(h);
            std::memcpy(buf, h->Usol.p, sizeof(double) * h->nu * (h->N - 1) * h->B);
            return TMPC_OK;
        }
        CUDA_TRY(cudaMemcpyAsync(buf, h->Usol.p, sizeof(double) * h->nu * (h->N - 1) * h->B, cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_bind_x0_device(tmpc_batch* h, const double* d_x0) {
    if (!h) return TMPC_ERR_DIM;
    h->ext_x0 = d_x0;
    return TMPC_OK;
}

int tmpc_batch_bind_x_ref_device(tmpc_batch* h, const double* d_x_ref, int batch) {
    if (!h || (d_x_ref && batch != 1 && batch != h->B)) return TMPC_ERR_DIM;
    h->ext_xref = d_x_ref; h->ext_xref_batch = batch;
    return TMPC_OK;
}

int tmpc_batch_bind_u_ref_device(tmpc_batch* h, const double* d_u_ref, int batch) {
    if (!h || (d_u_ref && batch != 1 && batch != h->B)) return TMPC_ERR_DIM;
    h->ext_uref = d_u_ref; h->ext_uref_batch = batch;
    return TMPC_OK;
}

void* tmpc_batch_device_ptr(tmpc_batch* h, int which) {
    if (!h) return nullptr;
    switch (which) {
        case 0: return h->Xsol.p;
        case 1: return h->Usol.p;
        case 2: return h->status.p;
        case 3: return h->iters.p;
        case 4: try { h->use_device(); flush_x0(h, h->stream); } catch (...) { return nullptr; } return h->x0.p;
        default: return nullptr;
    }
}

int tmpc_batch_rollout_async(tmpc_batch* h, int nsteps, const double* d_x_ref_long, int* d_status_out, int* d_iters_out,
                             void* cuda_stream) {
    return guarded([&]() -> int {
        if (!h || nsteps < 1) return TMPC_ERR_DIM;
        if (h->precision == 32) { set_error("rollout: the device closed loop is built for 64-bit only"); return TMPC_ERR_STATE; }
        h->use_device();
        cudaStream_t s = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->stream;
        flush_x0(h, s);
        if (h->ext_x0) {   // the plant state is advanced in place: work on the internal copy
            CUDA_TRY(cudaMemcpyAsync(h->x0.p, h->ext_x0, sizeof(double) * h->nx * h->B, cudaMemcpyDeviceToDevice, s));
            h->ext_x0 = nullptr;
        }
        if (h->st.check_termination == 0) { set_error("check_termination = 0 is not supported"); return TMPC_ERR_STATE; }
        if (const char* why = adaptive_refusal(h)) { set_error(why); return TMPC_ERR_STATE; }
        order_before_enqueue(h, s);
        ensure_aux_state(h, s);
        int k = resolve_kernel(h);
        if (k < 0) { set_error("requested kernel does not support this problem"); return TMPC_ERR_STATE; }
        // a small batch left to AUTO: the closed loop in ONE launch (TMEM kernel) rather than three launches per step
        if (k == TMPC_KERNEL_WARP && h->kernel_pref == TMPC_KERNEL_AUTO && h->rollout_fused &&
            tmpc::tmem_kernel_fused_supports(make_view(h), h->st)) k = TMPC_KERNEL_TMEM;
        const long long ref_len = static_cast<long long>(h->nx) * (h->N + nsteps - 1);
        if (k == TMPC_KERNEL_TMEM && h->st.max_iter > 0 && h->rho > 0.0 && h->rollout_fused) {
            // one launch: every instance runs all its steps on the sub-warp that owns it (admm_tmem.cu, ROLL)
            tmpc::BatchView v = make_view(h);
            if (d_x_ref_long) { v.Xref = d_x_ref_long; v.xref_stride = ref_len; }
            if (h->per_instance ? tmpc::tmem_pi_kernel_supports(v, h->st) : tmpc::tmem_kernel_fused_supports(v, h->st)) {
                if (!h->state_im.p) { h->state_im.alloc(static_cast<size_t>(tmpc::state_im_values(h->nx, h->nu, h->N, 64)) * h->B); v.state_im = h->state_im.p; }
                if (!h->state_zero && !h->state_in_im) {
                    cudaError_t ce = tmpc::launch_state_convert(v, true, s);
                    if (ce != cudaSuccess) throw CudaFail{ce, "state conversion"};
                    h->launches += 1;
                }
                tmpc::RolloutArgs ra{nsteps, h->state_zero ? 1 : 0, d_x_ref_long ? 1 : 0, h->x0.p, d_status_out, d_iters_out,
                                     h->per_instance ? h->pi_blocks.p : nullptr,
                                     h->order_mode == TMPC_ORDER_USER ? h->order_user.p : nullptr};
                cudaError_t e = tmpc::launch_admm_tmem_rollout(v, h->st, ra, h->counter.p, h->num_sms, s);
                if (e != cudaSuccess) throw CudaFail{e, "fused rollout launch"};
                h->launches += 1;
                h->state_zero = false; h->mem_zero = false; h->state_in_im = true; h->order_built = false;
                mark_enqueued(h, s);
                return TMPC_OK;
            }
        }
        for (int step = 0; step < nsteps; ++step) {
            tmpc::BatchView v = make_view(h);
            if (d_x_ref_long) { v.Xref = d_x_ref_long + static_cast<size_t>(h->nx) * step; v.xref_stride = ref_len; }
            cudaError_t e = launch_solve(h, k, v, s);
            if (e != cudaSuccess) throw CudaFail{e, "rollout solve launch"};
            if (d_status_out || d_iters_out) {
                e = tmpc::launch_record_step(v, d_status_out ? d_status_out + static_cast<size_t>(step) * h->B : nullptr,
                                             d_iters_out ? d_iters_out + static_cast<size_t>(step) * h->B : nullptr, s);
                if (e != cudaSuccess) throw CudaFail{e, "rollout record launch"};
                h->launches += 1;
            }
            e = h->per_instance ? tmpc::launch_plant_step_per_instance(v, per_instance_view(h), h->ld, h->x0.p, s)
                                : tmpc::launch_plant_step(v, h->x0.p, s);
            if (e != cudaSuccess) throw CudaFail{e, "rollout plant-step launch"};
            h->launches += 1;
        }
        mark_enqueued(h, s);
        return TMPC_OK;
    });
}

int tmpc_batch_set_rollout_fused(tmpc_batch* h, int fused) {
    if (!h) return TMPC_ERR_DIM;
    h->rollout_fused = fused != 0;
    return TMPC_OK;
}

int tmpc_batch_get_x0(tmpc_batch* h, void* dst, int device_dst) {
    return guarded([&]() -> int {
        if (!h || !dst) return TMPC_ERR_DIM;
        h->use_device();
        flush_x0(h, h->stream);
        const double* src = h->ext_x0 ? h->ext_x0 : h->x0.p;
        CUDA_TRY(cudaMemcpyAsync(dst, src, sizeof(double) * h->nx * h->B,
                                 device_dst ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
        sync_all(h);
        return TMPC_OK;
    });
}

int tmpc_batch_dims(tmpc_batch* h, int* nx, int* nu, int* N, int* B) {
    if (!h) return TMPC_ERR_DIM;
    if (nx) *nx = h->nx;
    if (nu) *nu = h->nu;
    if (N) *N = h->N;
    if (B) *B = h->B;
    return TMPC_OK;
}

long long tmpc_batch_launch_count(tmpc_batch* h) { return h ? h->launches : 0; }

void* tmpc_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}

void tmpc_host_free(void* p) { if (p) cudaFreeHost(p); }

}  // extern "C"
