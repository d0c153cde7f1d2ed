// Shared device helpers of the Tensor-Memory ADMM kernels (admm_tmem.cu, admm_tmem2.cu): Tensor Memory used as
// lane-private scratch (tcgen05.ld/st .32x32b), shared memory by explicit 32-bit address, pinned 64-bit shuffles,
// and the shared-memory staging layout of the problem constants.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace tmpc {
namespace tmu {

constexpr unsigned kAll = 0xffffffffu;

template <int NX, int BYTES = 8>
struct QPad {   // pitch (values) of one instance's broadcast slot: an odd number of 16-byte chunks, so that the eight
                // instances of a warp read their slots with 128-bit loads from disjoint banks
    static constexpr int per_chunk = 16 / BYTES;
    static constexpr int chunks = (NX + per_chunk - 1) / per_chunk;
    static constexpr int value = (chunks % 2 == 1 ? chunks : chunks + 1) * per_chunk;
};

template <int NX, int NU>
struct QConst {      // shared staging of the problem constants (doubles)
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

// Per-instance dynamics: the constants of ONE instance as the four lanes of its sub-warp consume them.  Every lane
// (sub = 0..3, owning state rows 4 r + sub and input row sub) has a flat array of `per_lane` doubles in the order
// below; the four arrays are interleaved in 16-byte chunks (chunk c of lane sub at double offset (4 c + sub) * 2), so
// one load instruction of the sub-warp touches 64 contiguous bytes of the instance's block instead of four rows 96
// bytes apart (8 cache lines per warp-wide load instead of 32).
template <int NX, int NU>
struct QLane {
    static constexpr int RX = NX / 4;
    static constexpr int A = 0;                          // A(4 r + sub, k)        : r * NX + k
    static constexpr int AmBKt = A + RX * NX;            // AmBKt(4 r + sub, k)
    static constexpr int Pinf = AmBKt + RX * NX;         // Pinf(k, 4 r + sub)
    static constexpr int B_rm = Pinf + RX * NX;          // B(4 r + sub, k)        : r * NU + k
    static constexpr int Kinf_cm = B_rm + RX * NU;       // Kinf(k, 4 r + sub)
    static constexpr int f = Kinf_cm + RX * NU;          // f, APf, Qd (4 r + sub) : r
    static constexpr int APf = f + RX;
    static constexpr int Qd = APf + RX;
    static constexpr int B_cm = Qd + RX;                 // input row `sub` (zeros when sub >= NU): B(k, sub) : k
    static constexpr int Kinf_rm = B_cm + NX;            // Kinf(sub, k)
    static constexpr int Quu = Kinf_rm + NX;             // Quu_inv(sub, k)
    static constexpr int BPf = Quu + NU;
    static constexpr int Rd = BPf + 1;
    static constexpr int per_lane = ((Rd + 1) + 1) / 2 * 2;
    static constexpr int size = 4 * per_lane;            // doubles per instance
    // double offset of element idx of lane sub inside the block
    __host__ __device__ static constexpr int at(int idx, int sub) { return (((idx >> 1) * 4 + sub) << 1) + (idx & 1); }
};

// ---------------------------------------------------------------- Tensor Memory as lane-private scratch
struct TmRow { uint32_t w[12]; };     // one knot point of a lane: g(3 rows), v(3 rows) as 32-bit halves

__device__ __forceinline__ void tm_ld4(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void tm_st4(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
}
__device__ __forceinline__ void tm_ld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
}
__device__ __forceinline__ void tm_st8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tm_ld2(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr));
}
__device__ __forceinline__ void tm_st2(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" :: "r"(taddr), "r"(r[0]), "r"(r[1]) : "memory");
}
// a knot point of a lane is WORDS 32-bit columns: doubles 12 (x8 + x4: each tcgen05 instruction needs its address in a
// uniform register, so fewer, wider instructions also mean fewer R2UR) or 4; floats 6 (x4 + x2) or 2
template <int WORDS>
__device__ __forceinline__ void tm_load(uint32_t taddr, uint32_t* w) {
    if (WORDS == 12) { tm_ld8(taddr, w); tm_ld4(taddr + 8, w + 8); }
    else if (WORDS == 6) { tm_ld4(taddr, w); tm_ld2(taddr + 4, w + 4); }
    else if (WORDS == 2) tm_ld2(taddr, w);
    else {
#pragma unroll
        for (int k = 0; k < WORDS; k += 4) tm_ld4(taddr + k, w + k);
    }
}
template <int WORDS>
__device__ __forceinline__ void tm_store(uint32_t taddr, const uint32_t* w) {
    if (WORDS == 12) { tm_st8(taddr, w); tm_st4(taddr + 8, w + 8); }
    else if (WORDS == 6) { tm_st4(taddr, w); tm_st2(taddr + 4, w + 4); }
    else if (WORDS == 2) tm_st2(taddr, w);
    else {
#pragma unroll
        for (int k = 0; k < WORDS; k += 4) tm_st4(taddr + k, w + k);
    }
}
// tcgen05.ld is asynchronous and NOT covered by the register scoreboard: the destination registers are valid
// only after tcgen05.wait::ld.  Passing them through the wait as in/out operands makes every later use depend
// on it, so the compiler cannot hoist a consumer above the wait.
__device__ __forceinline__ void tm_wait_ld12(uint32_t* w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]),
                   "+r"(w[6]), "+r"(w[7]), "+r"(w[8]), "+r"(w[9]), "+r"(w[10]), "+r"(w[11]) :: "memory");
}
__device__ __forceinline__ void tm_wait_ld4(uint32_t* w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]) :: "memory");
}
__device__ __forceinline__ void tm_wait_ld6(uint32_t* w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]) :: "memory");
}
__device__ __forceinline__ void tm_wait_ld2(uint32_t* w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(w[0]), "+r"(w[1]) :: "memory");
}
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ double mk_double(uint32_t lo, uint32_t hi) { return __hiloint2double(static_cast<int>(hi), static_cast<int>(lo)); }
__device__ __forceinline__ uint32_t lo32(double d) { return static_cast<uint32_t>(__double2loint(d)); }
__device__ __forceinline__ uint32_t hi32(double d) { return static_cast<uint32_t>(__double2hiint(d)); }

__device__ __forceinline__ void ld2(const double* p, double& a, double& b) {
    const double2 t = *reinterpret_cast<const double2*>(p);
    a = t.x; b = t.y;
}
__device__ __forceinline__ void st2(double* p, double a, double b) { *reinterpret_cast<double2*>(p) = make_double2(a, b); }
// 64-bit shuffle as two pinned shfl.sync: issued where written (right after the value exists), so their latency
// overlaps the work that follows instead of being scheduled late by the compiler
__device__ __forceinline__ double shfl_dq(double v, int src_lane) {
    uint32_t lo = lo32(v), hi = hi32(v), rlo, rhi;
    asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(rlo) : "r"(lo), "r"(src_lane));
    asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(rhi) : "r"(hi), "r"(src_lane));
    return mk_double(rlo, rhi);
}

// ---------------------------------------------------------------- shared memory by 32-bit address
// The sweeps address shared memory through explicit .shared instructions on 32-bit addresses computed once per
// kernel: with generic pointers the compiler rebuilt the shared window base (S2UR SR_CgaCtaId + ULEA, ~20 cycles of
// scoreboard wait) twice per knot point.  OFF is a byte offset folded into the instruction.
template <int OFF = 0>
__device__ __forceinline__ void lds2(uint32_t a, double& x, double& y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(x), "=d"(y) : "r"(a), "n"(OFF));
}
template <int OFF = 0>
__device__ __forceinline__ double lds1(uint32_t a) {
    double x;
    asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(x) : "r"(a), "n"(OFF));
    return x;
}
template <int OFF = 0>
__device__ __forceinline__ void sts1(uint32_t a, double x) {
    asm volatile("st.shared.f64 [%0+%1], %2;" :: "r"(a), "n"(OFF), "d"(x) : "memory");
}
template <int OFF = 0>
__device__ __forceinline__ void sts2(uint32_t a, double x, double y) {
    asm volatile("st.shared.v2.f64 [%0+%1], {%2, %3};" :: "r"(a), "n"(OFF), "d"(x), "d"(y) : "memory");
}
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// predicated global store: the write-out passes run with per-lane conditions (finished / not finished) on
// every store; as branches they cost a divergence region (BSSY / BSYNC + branch resolution) per row and knot point
__device__ __forceinline__ void stg_if(double* p, double v, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.s32 p, %2, 0; @p st.global.f64 [%0], %1; }"
                 :: "l"(p), "d"(v), "r"(static_cast<int>(pred)) : "memory");
}

__device__ __forceinline__ void stg2_if(double* p, double a, double b, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.s32 p, %3, 0; @p st.global.v2.f64 [%0], {%1, %2}; }"
                 :: "l"(p), "d"(a), "d"(b), "r"(static_cast<int>(pred)) : "memory");
}

template <int W> __device__ __forceinline__ void tm_wait_ld(uint32_t* w);
template <> __device__ __forceinline__ void tm_wait_ld<12>(uint32_t* w) { tm_wait_ld12(w); }
template <> __device__ __forceinline__ void tm_wait_ld<4>(uint32_t* w) { tm_wait_ld4(w); }
template <> __device__ __forceinline__ void tm_wait_ld<6>(uint32_t* w) { tm_wait_ld6(w); }
template <> __device__ __forceinline__ void tm_wait_ld<2>(uint32_t* w) { tm_wait_ld2(w); }

// ---------------------------------------------------------------- the same helpers at the float value of `tinytype`
// (the TMEM kernel is instantiated for both scalars; a value is two Tensor-Memory words as a double, one as a float)
__device__ __forceinline__ double tm_get(const uint32_t* w, int k, double) { return mk_double(w[2 * k], w[2 * k + 1]); }
__device__ __forceinline__ float tm_get(const uint32_t* w, int k, float) { return __uint_as_float(w[k]); }
__device__ __forceinline__ void tm_put(uint32_t* w, int k, double v) { w[2 * k] = lo32(v); w[2 * k + 1] = hi32(v); }
__device__ __forceinline__ void tm_put(uint32_t* w, int k, float v) { w[k] = __float_as_uint(v); }
// w[k] <- cond ? v : w[k]
__device__ __forceinline__ void tm_put_if(uint32_t* w, int k, double v, bool c) {
    w[2 * k] = c ? lo32(v) : w[2 * k]; w[2 * k + 1] = c ? hi32(v) : w[2 * k + 1];
}
__device__ __forceinline__ void tm_put_if(uint32_t* w, int k, float v, bool c) { w[k] = c ? __float_as_uint(v) : w[k]; }

__device__ __forceinline__ void ld2(const float* p, float& a, float& b) {
    const float2 t = *reinterpret_cast<const float2*>(p);
    a = t.x; b = t.y;
}
__device__ __forceinline__ void st2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
__device__ __forceinline__ float shfl_dq(float v, int src_lane) {
    uint32_t r;
    asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(r) : "r"(__float_as_uint(v)), "r"(src_lane));
    return __uint_as_float(r);
}
template <int OFF = 0>
__device__ __forceinline__ void lds2(uint32_t a, float& x, float& y) {
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2+%3];" : "=f"(x), "=f"(y) : "r"(a), "n"(OFF));
}
template <int OFF = 0>
__device__ __forceinline__ void lds4(uint32_t a, float& x, float& y, float& z, float& w) {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a), "n"(OFF));
}
template <int OFF = 0>
__device__ __forceinline__ void sts1(uint32_t a, float x) {
    asm volatile("st.shared.f32 [%0+%1], %2;" :: "r"(a), "n"(OFF), "f"(x) : "memory");
}
template <int OFF = 0>
__device__ __forceinline__ void sts2(uint32_t a, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0+%1], {%2, %3};" :: "r"(a), "n"(OFF), "f"(x), "f"(y) : "memory");
}
// one scalar from shared memory by 32-bit address, typed by the caller
template <int OFF, class T>
__device__ __forceinline__ T lds1t(uint32_t a) {
    if constexpr (sizeof(T) == 8) return lds1<OFF>(a);
    else {
        float x;
        asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(x) : "r"(a), "n"(OFF));
        return x;
    }
}
__device__ __forceinline__ void stg2_if(float* p, float a, float b, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.s32 p, %3, 0; @p st.global.v2.f32 [%0], {%1, %2}; }"
                 :: "l"(p), "f"(a), "f"(b), "r"(static_cast<int>(pred)) : "memory");
}
__device__ __forceinline__ void stg_if(float* p, float v, bool pred) {
    asm volatile("{ .reg .pred p; setp.ne.s32 p, %2, 0; @p st.global.f32 [%0], %1; }"
                 :: "l"(p), "f"(v), "r"(static_cast<int>(pred)) : "memory");
}

// a chunk of VW values (one double, or two floats: 8 bytes either way) of the instance-major carried state
template <int VW, class T> __device__ __forceinline__ void ldc(const T* p, T* v) {
    if constexpr (VW == 1) v[0] = *p; else ld2(p, v[0], v[1]);
}
template <int VW, class T> __device__ __forceinline__ void stc(T* p, const T* v) {
    if constexpr (VW == 1) *p = v[0]; else st2(p, v[0], v[1]);
}
template <int VW, class T> __device__ __forceinline__ void stgc_if(T* p, const T* v, bool pred) {
    if constexpr (VW == 1) stg_if(p, v[0], pred); else stg2_if(p, v[0], v[1], pred);
}

}  // namespace tmu
}  // namespace tmpc
