// ptx.cuh — the handful of sm_100a PTX wrappers the force kernel needs:
// mbarrier + 1-D TMA bulk copy (cp.async.bulk -> SASS UBLKCP) for the j-stream ring,
// MUFU.RSQ without the denormal fix-up code, and packed f32x2 arithmetic (FFMA2/FADD2/FMUL2).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void fence_mbar_init() {
    // make the initialised barriers visible to the async (TMA) proxy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// 1-D bulk copy global -> shared, completion counted in bytes on `bar`.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ float rsqrt_fast(float x) {
#ifdef NB_UBENCH_NO_MUFU  // tools/ubench only: what-if without the SFU op (wrong numerics)
    return __int_as_float(0x5f3759df - (__float_as_int(x) >> 1));
#endif
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- packed f32x2 (Blackwell FFMA2 / FADD2 / FMUL2) ---------------------------------
struct f2 {
    unsigned long long r;  // one 64-bit register: {lo, hi} floats
};

__device__ __forceinline__ f2 f2_make(float lo, float hi) {
    f2 o;
    asm("mov.b64 %0, {%1, %2};" : "=l"(o.r) : "f"(lo), "f"(hi));
    return o;
}
__device__ __forceinline__ void f2_split(f2 a, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.r));
}
__device__ __forceinline__ f2 f2_fma(f2 a, f2 b, f2 c) {
    f2 o;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(o.r) : "l"(a.r), "l"(b.r), "l"(c.r));
    return o;
}
__device__ __forceinline__ f2 f2_mul(f2 a, f2 b) {
    f2 o;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(o.r) : "l"(a.r), "l"(b.r));
    return o;
}
__device__ __forceinline__ f2 f2_add(f2 a, f2 b) {
    f2 o;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(o.r) : "l"(a.r), "l"(b.r));
    return o;
}
__device__ __forceinline__ f2 f2_sub(f2 a, f2 b) {
    f2 o;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(o.r) : "l"(a.r), "l"(b.r));
    return o;
}

}  // namespace nb
