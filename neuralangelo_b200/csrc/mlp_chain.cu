// Generic fused MLP forward on tcgen05: a short "program" of stages executed per 128-row tile, with every
// intermediate activation kept on chip.
//
// Replaces the cuBLAS + eager elementwise chains of
//   MLPwithSkipConnection.forward          projects/nerf/utils/nerf_util.py:186-196
//   NeuralRGB.forward (colour MLP)         projects/neuralangelo/utils/modules.py:209-220
//   BackgroundNeRF.forward (both MLPs)     projects/neuralangelo/utils/modules.py:270-285
//
// A stage is:  [epilogue of the previous product: bias + activation -> next A operand in shared memory and/or a
// global output and/or a saved-activation buffer]  [stage <= 256 columns of a global input into the A operand]
// [product  D (+)= A W^T  on the tensor cores].  That is enough to express inputs wider than the 256-column A
// operand (278-wide colour input: two accumulating products), skip connections (accumulate a second product of
// the re-staged network input), a 257-wide layer (two products of the same A: density column, then features)
// and a second network chained on the first (view-dependent background colour), all in ONE launch.
// Same operand format and split-fp16 arithmetic as the SDF kernel (mlp_tcgen05.cu, tc_common.cuh).
#include "tc_common.cuh"

namespace {

using namespace na_tc;

constexpr int kComputeWarps = 8;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kThreads = kComputeThreads + 64;
constexpr int kStages = 5;      // weight ring
constexpr int kBlockK = 16;
constexpr int kMaxK = 256, kMaxN = 256;
constexpr int kABytes = kTileM * kMaxK * 2;
constexpr int kBStageBytes = 2 * kMaxN * kBlockK * 2;
constexpr int kSmemBytes = 2 * kABytes + kStages * kBStageBytes + 1024;
constexpr int kMaxProgram = 20;

// kReluMask: backward link of a relu layer -- the accumulator (an upstream gradient) is multiplied by the step function
// of the layer's SAVED activation, which is READ from the stage's `save` pointer (no bias is added)
enum Act { kNone = 0, kRelu = 1, kSoftplus = 2, kSigmoid = 3, kReluMask = 4 };

struct Stage {
    // epilogue of the previous stage's product
    int has_epi;
    const float* bias;      // [n_valid]
    int act;
    float act_param;        // softplus beta
    int n_valid;            // columns of the accumulator that are real outputs
    int epi_to_a;           // write the activations into the A operand (chunk 0 onwards) for the next product
    float* out;             // global output rows (stride out_ld) or null
    int64_t out_ld;
    float* save;            // saved activations (for the backward), stride save_ld, or null
    int64_t save_ld;
    // input staging
    const float* in;        // global input rows (stride in_ld) or null
    int64_t in_ld;
    int in_col, in_cols;    // columns [in_col, in_col + in_cols) -> A chunks from 0, zero-padded to a multiple of 16
    // product
    int has_gemm;
    const __half* w_hi;     // packed [k_chunks][n_out/8][8][8]
    const __half* w_lo;
    int k_chunks;           // 8-wide chunks of A consumed (even)
    int n_out;              // accumulator columns (multiple of 16, <= 256)
    int accumulate;         // 1: add to the accumulators of the previous product
};

struct Program {
    int64_t m;
    int n_stages;
    Stage stage[kMaxProgram];
};

__device__ __forceinline__ float activate(float x, int act, float p) {
    switch (act) {
        case kRelu: return fmaxf(x, 0.f);
        case kSoftplus: {  // torch semantics incl. threshold 20
            const float bx = p * x;
            return bx > 20.f ? x : log1pf(expf(bx)) / p;
        }
        case kSigmoid: return 1.f / (1.f + expf(-x));
        default: return x;
    }
}

__global__ void __launch_bounds__(kThreads, 1) mlp_chain_fwd_kernel(const __grid_constant__ Program P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a_hi = smem;
    uint8_t* a_lo = smem + kABytes;
    uint8_t* b_ring = smem + 2 * kABytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(b_ring + kStages * kBStageBytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 2);
    const uint32_t bar0 = smem_u32(bars);
    auto b_full = [&](int s) { return bar0 + 8u * s; };
    auto b_empty = [&](int s) { return bar0 + 8u * (kStages + s); };
    const uint32_t a_ready = bar0 + 8u * (2 * kStages), d_ready = bar0 + 8u * (2 * kStages + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t n_tiles = (P.m + kTileM - 1) / kTileM;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(b_full(s), 1);
            mbar_init(b_empty(s), 1);
        }
        mbar_init(a_ready, kComputeThreads);
        mbar_init(d_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kComputeWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t d_main = tmem_base, d_cross = tmem_base + 256;

    if (warp == kComputeWarps + 1) {
        // ===================== weight loader =====================
        if (elect_one()) {
            uint32_t it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int si = 0; si < P.n_stages; ++si) {
                    const Stage& S = P.stage[si];
                    if (!S.has_gemm) continue;
                    const uint32_t chunk_bytes = (uint32_t)S.n_out * 16;
                    for (int c0 = 0; c0 < S.k_chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        mbar_wait(b_empty(s), ((it / kStages) & 1) ^ 1);
                        const uint32_t bytes = (kBlockK / 8) * chunk_bytes;
                        mbar_expect_tx(b_full(s), 2 * bytes);
                        const uint32_t dst = smem_u32(b_ring + s * kBStageBytes);
                        bulk_copy_g2s(dst, reinterpret_cast<const uint8_t*>(S.w_hi) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                        bulk_copy_g2s(dst + kBStageBytes / 2,
                                      reinterpret_cast<const uint8_t*>(S.w_lo) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                    }
                }
            }
        }
    } else if (warp == kComputeWarps) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t a_lbo = kTileM / 8 * 128;
            uint32_t it = 0, gemm_it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int si = 0; si < P.n_stages; ++si) {
                    const Stage& S = P.stage[si];
                    if (!S.has_gemm) continue;
                    const uint32_t idesc = make_idesc(kTileM, S.n_out);
                    const uint32_t b_lbo = (uint32_t)S.n_out / 8 * 128;
                    mbar_wait(a_ready, gemm_it & 1);
                    tc_fence_after();
                    bool first = !S.accumulate;
                    for (int c0 = 0; c0 < S.k_chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        mbar_wait(b_full(s), (it / kStages) & 1);
                        tc_fence_after();
                        const uint32_t b_hi_addr = smem_u32(b_ring + s * kBStageBytes);
                        const uint32_t b_lo_addr = b_hi_addr + kBStageBytes / 2;
                        const uint32_t a_off = (uint32_t)c0 * a_lbo;
                        const uint64_t ah = make_desc(smem_u32(a_hi) + a_off, a_lbo, 128);
                        const uint64_t al = make_desc(smem_u32(a_lo) + a_off, a_lbo, 128);
                        const uint64_t bh = make_desc(b_hi_addr, b_lbo, 128);
                        const uint64_t bl = make_desc(b_lo_addr, b_lbo, 128);
                        tc_mma_f16(d_cross, al, bh, idesc, first ? 0u : 1u);
                        tc_mma_f16(d_cross, ah, bl, idesc, 1u);
                        tc_mma_f16(d_main, ah, bh, idesc, first ? 0u : 1u);
                        first = false;
                        tc_commit(b_empty(s));
                    }
                    tc_commit(d_ready);
                    ++gemm_it;
                }
            }
        }
    } else {
        // ===================== compute warps =====================
        const int tid = threadIdx.x;
        const int quarter = warp & 3, half = warp >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        uint32_t gemm_it = 0;          // products completed so far (d_ready phase)
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int64_t row0 = tile * kTileM;
            const int64_t grow = row0 + row;
            const bool valid = grow < P.m;
            bool pending = false;       // a product has been issued whose completion has not been waited for
            for (int si = 0; si < P.n_stages; ++si) {
                const Stage& S = P.stage[si];
                if (pending && (S.has_epi || S.in)) {   // anything that reads D or overwrites A needs the product done
                    mbar_wait(d_ready, gemm_it & 1);
                    ++gemm_it;
                    tc_fence_after();
                    pending = false;
                }
                if (S.has_epi) {
                    const int n_cols = (S.n_valid + 31) / 32 * 32;
                    const int chunks = n_cols / 32;
                    const int first_chunk = half == 0 ? 0 : (chunks + 1) / 2;
                    const int last_chunk = half == 0 ? (chunks + 1) / 2 : chunks;
                    for (int ch = first_chunk; ch < last_chunk; ++ch) {
                        const int c0 = ch * 32;
                        uint32_t rm[32], rc[32];
                        tc_ld32(d_main + lane_addr + c0, rm);
                        tc_ld32(d_cross + lane_addr + c0, rc);
                        tc_wait_ld();
                        float act[32];
                        if (S.act == kReluMask) {
                            const float* mk = S.save + grow * S.save_ld + c0;
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const bool real = valid && c0 + j < S.n_valid;
                                const float g = fmaf(__uint_as_float(rc[j]), kLoInv, __uint_as_float(rm[j]));
                                act[j] = (real && __ldg(mk + j) > 0.f) ? g : 0.f;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const bool real = c0 + j < S.n_valid;
                                const float pre = fmaf(__uint_as_float(rc[j]), kLoInv, __uint_as_float(rm[j])) +
                                                  (real ? __ldg(S.bias + c0 + j) : 0.f);
                                act[j] = real ? activate(pre, S.act, S.act_param) : 0.f;
                            }
                        }
                        if (S.epi_to_a) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint32_t hi[4], lo[4];
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    split2_f16(act[q * 8 + j * 2], act[q * 8 + j * 2 + 1], hi[j], lo[j]);
                                }
                                const uint32_t off = a_chunk_offset(row, c0 / 8 + q);
                                *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                                *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                            }
                        }
                        if (valid) {
                            if (S.out) {
                                float* dst = S.out + grow * S.out_ld + c0;
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (c0 + j < S.n_valid) dst[j] = act[j];
                            }
                            if (S.save && S.act != kReluMask) {
                                float* dst = S.save + grow * S.save_ld + c0;
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (c0 + j < S.n_valid) dst[j] = act[j];
                            }
                        }
                    }
                    tc_fence_before();
                }
                if (S.in) {
                    // stage in_cols columns (zero-padded to a multiple of 16) into A chunks 0.. ; consecutive threads take
                    // consecutive 8-column chunks of a row; loads of four chunks are in flight before any conversion
                    const int chunks = (S.in_cols + 15) / 16 * 2;
                    const int total = kTileM * chunks;
                    for (int base = 0; base < total; base += 4 * kComputeThreads) {
                        float v[4][8];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int idx = base + u * kComputeThreads + tid;
                            const int r = idx / chunks, kc = idx - r * chunks;
                            const int64_t gr = row0 + r;
                            const float* src = S.in + gr * S.in_ld + S.in_col + kc * 8;
#pragma unroll
                            for (int e = 0; e < 8; ++e)
                                v[u][e] = (idx < total && gr < P.m && kc * 8 + e < S.in_cols) ? __ldg(src + e) : 0.f;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int idx = base + u * kComputeThreads + tid;
                            if (idx >= total) continue;
                            const int r = idx / chunks, kc = idx - r * chunks;
                            uint32_t hi[4], lo[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                split2_f16(v[u][2 * j], v[u][2 * j + 1], hi[j], lo[j]);
                            }
                            const uint32_t off = a_chunk_offset(r, kc);
                            *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                            *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                        }
                    }
                }
                if (S.has_gemm) {
                    fence_proxy_async();
                    mbar_arrive(a_ready);
                    pending = true;
                }
            }
            if (pending) {   // a program never ends on an un-consumed product, but keep the phases consistent
                mbar_wait(d_ready, gemm_it & 1);
                ++gemm_it;
                tc_fence_after();
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kComputeWarps) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace

extern "C" int na_mlp_chain_fwd(const na_mlp_stage* stages, int n_stages, int64_t m, void* stream) {
    na::clear_error();
    NA_REQUIRE(m >= 0 && stages != nullptr, "mlp_chain_fwd: bad arguments");
    NA_REQUIRE(n_stages >= 1 && n_stages <= kMaxProgram, "mlp_chain_fwd: %d stages (1..%d supported)", n_stages,
               kMaxProgram);
    if (m == 0) return NA_OK;
    Program P{};
    P.m = m;
    P.n_stages = n_stages;
    bool have_product = false, pending = false;
    for (int i = 0; i < n_stages; ++i) {
        const na_mlp_stage& h = stages[i];
        Stage& S = P.stage[i];
        NA_REQUIRE(!pending || h.has_epi || h.in, "mlp_chain_fwd: stage %d follows a product with neither an epilogue "
                   "nor an input load", i);
        S.has_epi = h.has_epi, S.bias = h.bias, S.act = h.act, S.act_param = h.act_param, S.n_valid = h.n_valid;
        S.epi_to_a = h.epi_to_a, S.out = h.out, S.out_ld = h.out_ld, S.save = h.save, S.save_ld = h.save_ld;
        S.in = h.in, S.in_ld = h.in_ld, S.in_col = h.in_col, S.in_cols = h.in_cols;
        S.has_gemm = h.has_gemm, S.w_hi = (const __half*)h.w_hi, S.w_lo = (const __half*)h.w_lo;
        S.k_chunks = h.k_chunks, S.n_out = h.n_out, S.accumulate = h.accumulate;
        if (S.has_epi) {
            NA_REQUIRE(have_product, "mlp_chain_fwd: stage %d has an epilogue but no product precedes it", i);
            NA_REQUIRE(S.bias && S.n_valid >= 1 && S.n_valid <= kMaxN, "mlp_chain_fwd: stage %d epilogue arguments", i);
            NA_REQUIRE(S.act >= kNone && S.act <= kReluMask, "mlp_chain_fwd: stage %d unknown activation %d", i, S.act);
            NA_REQUIRE(S.act != kReluMask || S.save, "mlp_chain_fwd: stage %d relu-mask epilogue needs the saved activation", i);
            NA_REQUIRE(S.act != kSoftplus || S.act_param > 0.f, "mlp_chain_fwd: stage %d softplus beta", i);
        }
        if (S.in) NA_REQUIRE(S.in_cols >= 1 && S.in_cols <= kMaxK && S.in_col >= 0, "mlp_chain_fwd: stage %d input columns", i);
        if (S.has_gemm) {
            NA_REQUIRE(S.w_hi && S.w_lo && ((uintptr_t)S.w_hi | (uintptr_t)S.w_lo) % 16 == 0,
                       "mlp_chain_fwd: stage %d weights null or unaligned", i);
            NA_REQUIRE(S.k_chunks >= 2 && S.k_chunks % 2 == 0 && S.k_chunks * 8 <= kMaxK, "mlp_chain_fwd: stage %d k_chunks %d",
                       i, S.k_chunks);
            NA_REQUIRE(S.n_out >= 16 && S.n_out % 16 == 0 && S.n_out <= kMaxN, "mlp_chain_fwd: stage %d n_out %d", i,
                       S.n_out);
            NA_REQUIRE(!S.accumulate || have_product, "mlp_chain_fwd: stage %d accumulates onto nothing", i);
            have_product = true;
        }
        pending = S.has_gemm != 0;
    }
    NA_REQUIRE(!P.stage[n_stages - 1].has_gemm, "mlp_chain_fwd: the last stage must not issue a product");
    static bool done[64] = {};
    int dev = 0;
    if (int rc = na::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return rc;
    if (dev < 0 || dev >= 64 || !done[dev]) {
        if (int rc = na::check_cuda(cudaFuncSetAttribute(mlp_chain_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         kSmemBytes),
                                    "cudaFuncSetAttribute(mlp_chain_fwd_kernel)"))
            return rc;
        if (dev >= 0 && dev < 64) done[dev] = true;
    }
    const int64_t tiles = (m + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    mlp_chain_fwd_kernel<<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(P);
    NA_CHECK_LAUNCH("mlp_chain_fwd_kernel");
    return NA_OK;
}
