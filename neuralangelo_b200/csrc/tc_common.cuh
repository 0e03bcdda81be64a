// tcgen05 / TMEM / mbarrier / bulk-copy PTX wrappers and the operand-layout helpers shared by the tensor-core
// MLP kernels (mlp_tcgen05.cu, mlp_chain.cu).  sm_100a only.
#pragma once
#include "na_common.cuh"
#include <cuda_fp16.h>

namespace na_tc {

constexpr int kTileM = 128;
constexpr float kLoScale = 2048.f, kLoInv = 1.f / 2048.f;

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One lane of a CONVERGED warp.  Use this, not `lane == 0`, to pick the thread that issues tcgen05.mma / commit: ptxas
// then knows a single thread runs the branch and emits the uniform-datapath instructions back to back; behind a plain
// lane test it wraps every UTCHMMA in an ELECT / BRA.U.ANY loop over the "possibly many" active lanes and moves each
// descriptor through R2UR (14 instructions per MMA, more than the 64 cycles a 128x128x16 MMA takes).
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug must trap, not hang the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
        "%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, no swizzle: LBO = byte stride between core matrices adjacent in K, SBO = between 8-row groups.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
    return d;                // layout type 0 (no swizzle), base offset 0
}

__device__ __forceinline__ uint32_t make_idesc(int m, int n) {
    // c_format F32 (bits 4-5 = 1), a/b format F16 (0), both K-major (0), N>>3 at bit 17, M>>4 at bit 24
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// Two values -> packed fp16 pairs (first value in the low half): v = hi + 2^-11 lo'.  One F2FP.SATFINITE.PACK_AB per
// pair and piece (the scalar cvt.f16.f32 path is a quarter-rate F2F per element and bounded the staging phases);
// satfinite clamps to +-65504 instead of producing inf.
__device__ __forceinline__ void split2_f16(float v0, float v1, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(v1), "f"(v0));
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"((v1 - f.y) * kLoScale), "f"((v0 - f.x) * kLoScale));
}

// softplus_beta(x) = max(x,0) + log(1 + exp(-|beta x|)) / beta.  Differs from torch's thresholded form
// (x when beta*x > 20) by < exp(-20)/beta; the fast exp/log intrinsics are accurate to ~1e-7 ABSOLUTE on the
// log term, i.e. ~1e-9 on the result for beta = 100 -- far below the fp32 rounding of the sums around it.
__device__ __forceinline__ float softplus(float x, float beta, float inv_beta) {
    // raw ex2 / lg2 (ftz): the range fix-ups of __expf only matter below 2^-126, which vanishes against the 1 it is
    // added to; the constant products are loop-invariant
    float e, l;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fabsf(x) * (beta * -1.4426950408889634f)));
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(1.f + e));
    return fmaf(l, 0.6931471805599453f * inv_beta, fmaxf(x, 0.f));
}

// Same function with the logarithm on the FMA pipe: log1p(u) = u * P7(u) on u = exp(-|beta x|) in (0, 1] (Chebyshev
// interpolant, |error| < 2e-7 absolute evaluated in fp32, i.e. < 2e-9 on the activation for beta = 100 -- the class of
// lg2.approx).  One MUFU per element instead of two: the pipelined SDF kernel's epilogue is MUFU-bound, so it
// alternates the two forms to balance the special-function and FMA pipes.
__device__ __forceinline__ float softplus_poly(float x, float beta, float inv_beta) {
    float u;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(u) : "f"(fabsf(x) * (beta * -1.4426950408889634f)));
    float p = -8.371152549e-03f;
    p = fmaf(p, u, 4.349389982e-02f);
    p = fmaf(p, u, -1.068500286e-01f);
    p = fmaf(p, u, 1.768747670e-01f);
    p = fmaf(p, u, -2.447477418e-01f);
    p = fmaf(p, u, 3.327192838e-01f);
    p = fmaf(p, u, -4.999717452e-01f);
    p = fmaf(p, u, 9.999997808e-01f);
    return fmaf(p * u, inv_beta, fmaxf(x, 0.f));
}

// Byte offset of the 16-byte chunk holding (row, k..k+7) inside an A image of kTileM rows.
__device__ __forceinline__ uint32_t a_chunk_offset(int row, int kchunk) {
    return (uint32_t)kchunk * (kTileM / 8 * 128) + (uint32_t)(row >> 3) * 128 + (uint32_t)(row & 7) * 16;
}


}  // namespace na_tc
