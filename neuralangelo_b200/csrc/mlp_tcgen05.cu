// Fused SDF MLP forward on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
//
// Replaces, for the geometry network, the chain
//   WN-Linear(3+LD -> H) -> softplus(beta) -> [Linear(H -> 1) = SDF]  [-> WN-Linear(H -> H) -> softplus = feature]
// of projects/neuralangelo/utils/mlp.py:55-69 (MLPforNeuralSDF.forward with the default `num_layers: 1`),
// i.e. the cuBLAS + eager elementwise launches behind NeuralSDF.forward / .sdf (modules.py:68-74).
//
// Precision.  The finite-difference normals and curvature (modules.py:132-175) amplify SDF rounding by
// 1/(2 eps) and 1/eps^2, so single-pass TF32/BF16 tensor-core products are not good enough.  Every operand
// is split into two fp16 pieces, a = a_hi + 2^-11 a_lo' with a_lo' = (a - a_hi) * 2^11 (the scaling keeps the
// low piece out of fp16's subnormal range), and three kind::f16 MMAs per k-step accumulate in fp32:
//     D_main  += A_hi B_hi                       D_cross += A_lo' B_hi + A_hi B_lo'
// result = D_main + 2^-11 D_cross (the dropped a_lo b_lo term is 2^-22 relative) -- fp32-grade products at
// 1.5x the cost of one TF32 pass.
//
// Two forward kernels (one CTA per SM, persistent over 128-row tiles; M = 128, cta_group::1):
//  * sdf_mlp_fwd_kernel -- both layers (SDF + feature).  Sixteen compute warps (four column groups x four TMEM lane
//    quarters) stage the input tile into shared memory as fp16 hi/lo in the UMMA core-matrix order, then per layer read
//    the accumulators back (tcgen05.ld), add bias, softplus, take the SDF head's dot product on the fly, and write the
//    activations straight back to shared memory as the next layer's A operand -- intermediate activations never touch
//    HBM; one MMA warp (an elected lane issues tcgen05.mma, tcgen05.commit signals mbarriers); one loader warp
//    streams pre-packed 16-wide K blocks of W_hi / W_lo through a 5-stage cp.async.bulk ring, running ahead across
//    layers and tiles.
//  * sdf_head_fwd_pipe_kernel -- the SDF-only evaluation (first layer + head), pipelined: see its own header below.
// Operands are K-major without swizzle: element (row, k) of a tile with R rows sits at
//     (k / 8) * (R / 8 * 128) + (row / 8) * 128 + (row % 8) * 16 + (k % 8) * 2   bytes,
// so a thread that owns one row writes its eight-column chunks with conflict-free 16-byte stores and the
// host packs the weights once per optimiser step into the exact shared-memory image.
#include "tc_common.cuh"
#include <stdlib.h>

namespace {

using namespace na_tc;

constexpr int kComputeWarps = 8;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kThreads = kComputeThreads + 64;  // + MMA warp + loader warp
constexpr int kStages = 5;
constexpr int kBlockK = 16;                     // K elements per weight-ring stage (one UMMA k-step)
constexpr int kMaxK = 256;                      // widest A operand held in shared memory
constexpr int kMaxN = 256;

constexpr int kABytes = kTileM * kMaxK * 2;             // one fp16 A image (hi or lo): 64 KB
constexpr int kBStageBytes = 2 * kMaxN * kBlockK * 2;   // W_hi block + W_lo block: 32 KB
constexpr int kSmemBytes = 2 * kABytes + kStages * kBStageBytes + 2 * kTileM * 4 + 3 * kMaxN * 4 + 1024;

struct Layer {
    const __half* w_hi;   // packed [K_pad/8][N/8][8][8]
    const __half* w_lo;
    const float* bias;    // [N]
    int k_pad;            // multiple of 16, <= 256
};

struct MlpParams {
    const float* x;       // [M, K0] rows, row stride ldx floats
    int64_t ldx;
    int64_t m;
    int k0;               // valid input columns (<= layer[0].k_pad)
    int n;                // hidden width: multiple of 64, <= 256
    int n_layers;         // 1 (SDF only) or 2 (SDF + feature)
    Layer layer[2];
    const float* w_head;  // [N] SDF head weights (applied to layer 0's activations) or null
    const float* b_head;  // device scalar
    float beta;           // softplus beta
    float* sdf_out;       // [M] or null
    float* feat_out;      // [M, N] rows, stride ld_feat, or null
    int64_t ld_feat;
    float* hidden_out;    // [M, N] layer-0 activations (for the backward), stride ld_hidden, or null
    int64_t ld_hidden;
    // chunk-plane input (pipelined kernel only): column 8 c + e of row r is planes[c * chunk_stride + r * 8 + e]
    const float* planes;
    int64_t chunk_stride;
};

// ---------------------------------------------------------------- the kernel
// The forward kernel runs 16 compute warps (four column groups x four TMEM lane quarters): its time is the softplus /
// split epilogue, and four warps per scheduler hide the MUFU / TMEM-load latencies that two could not.
constexpr int kFwdWarps = 16;
constexpr int kFwdCompute = kFwdWarps * 32;
constexpr int kFwdThreads = kFwdCompute + 64;
constexpr int kFwdGroups = kFwdWarps / 4;
constexpr int kFwdSmemBytes = kSmemBytes + 2 * kTileM * 4;   // head partials for four column groups

__global__ void __launch_bounds__(kFwdThreads, 1) sdf_mlp_fwd_kernel(const __grid_constant__ MlpParams P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a_hi = smem;
    uint8_t* a_lo = smem + kABytes;
    uint8_t* b_ring = smem + 2 * kABytes;
    float* head_partial = reinterpret_cast<float*>(b_ring + kStages * kBStageBytes);  // [kFwdGroups][kTileM]
    float* s_bias = head_partial + kFwdGroups * kTileM;   // [2][kMaxN] per-layer biases
    float* s_head = s_bias + 2 * kMaxN;           // [kMaxN] SDF head weights
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_head + kMaxN);
    // bars: [0,S) b_full, [S,2S) b_empty, 2S a_ready, 2S+1 d_ready; then the TMEM base address word
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 2);
    const uint32_t bar0 = smem_u32(bars);
    auto b_full = [&](int s) { return bar0 + 8u * s; };
    auto b_empty = [&](int s) { return bar0 + 8u * (kStages + s); };
    const uint32_t a_ready = bar0 + 8u * (2 * kStages), d_ready = bar0 + 8u * (2 * kStages + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = P.n;
    const int64_t n_tiles = (P.m + kTileM - 1) / kTileM;

    for (int i = threadIdx.x; i < n; i += kFwdThreads) {
        s_bias[i] = P.layer[0].bias[i];
        s_bias[kMaxN + i] = P.n_layers > 1 ? P.layer[1].bias[i] : 0.f;
        s_head[i] = P.w_head ? P.w_head[i] : 0.f;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(b_full(s), 1);
            mbar_init(b_empty(s), 1);
        }
        mbar_init(a_ready, kFwdCompute);
        mbar_init(d_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kFwdWarps) {  // MMA warp owns the TMEM allocation: D_main [0,256) and D_cross [256,512)
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

    if (warp == kFwdWarps + 1) {
        // ===================== weight loader =====================
        if (elect_one()) {
            uint32_t it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 0; l < P.n_layers; ++l) {
                    const Layer& L = P.layer[l];
                    const int chunks = L.k_pad / 8;                // 8-wide k-chunks in this layer
                    const uint32_t chunk_bytes = (uint32_t)n * 16;  // N/8 core matrices of 128 B
                    for (int c0 = 0; c0 < chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        const uint32_t ph = (it / kStages) & 1;
                        mbar_wait(b_empty(s), ph ^ 1);
                        const int nc = min(kBlockK / 8, chunks - c0);
                        const uint32_t bytes = nc * chunk_bytes;
                        mbar_expect_tx(b_full(s), 2 * bytes);
                        const uint32_t dst = smem_u32(b_ring + s * kBStageBytes);
                        bulk_copy_g2s(dst, reinterpret_cast<const uint8_t*>(L.w_hi) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                        bulk_copy_g2s(dst + kBStageBytes / 2,
                                      reinterpret_cast<const uint8_t*>(L.w_lo) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                    }
                }
            }
        }
    } else if (warp == kFwdWarps) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t idesc = make_idesc(kTileM, n);
            const uint32_t a_lbo = kTileM / 8 * 128, b_lbo = (uint32_t)n / 8 * 128;
            uint32_t it = 0, layer_it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 0; l < P.n_layers; ++l, ++layer_it) {
                    const Layer& L = P.layer[l];
                    mbar_wait(a_ready, layer_it & 1);
                    tc_fence_after();
                    const int chunks = L.k_pad / 8;
                    bool first = true;
                    for (int c0 = 0; c0 < chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        mbar_wait(b_full(s), (it / kStages) & 1);
                        tc_fence_after();
                        const int nc = min(kBlockK / 8, chunks - c0);
                        const uint32_t b_hi_addr = smem_u32(b_ring + s * kBStageBytes);
                        const uint32_t b_lo_addr = b_hi_addr + kBStageBytes / 2;
                        for (int c = 0; c < nc; c += 2) {  // one UMMA k-step = 16 elements = 2 chunks
                            const uint32_t a_off = (uint32_t)(c0 + c) * a_lbo, b_off = (uint32_t)c * b_lbo;
                            const uint64_t ah = make_desc(smem_u32(a_hi) + a_off, a_lbo, 128);
                            const uint64_t al = make_desc(smem_u32(a_lo) + a_off, a_lbo, 128);
                            const uint64_t bh = make_desc(b_hi_addr + b_off, b_lbo, 128);
                            const uint64_t bl = make_desc(b_lo_addr + b_off, b_lbo, 128);
                            tc_mma_f16(d_cross, al, bh, idesc, first ? 0u : 1u);
                            tc_mma_f16(d_cross, ah, bl, idesc, 1u);
                            tc_mma_f16(d_main, ah, bh, idesc, first ? 0u : 1u);
                            first = false;
                        }
                        tc_commit(b_empty(s));  // frees the ring slot once the MMAs above have read it
                    }
                    tc_commit(d_ready);  // accumulators complete (and A no longer being read)
                }
            }
        }
    } else {
        // ===================== compute warps =====================
        const int quarter = warp & 3;               // TMEM lane quarter this warp may access
        const int group = warp >> 2;                // which slice of the N columns this warp handles
        const int cols_per_group = n / kFwdGroups >= 32 ? n / kFwdGroups : 32;
        const int n_groups = n / cols_per_group;    // narrow layers leave the last groups without columns
        const int row = quarter * 32 + lane;        // accumulator row == TMEM lane
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        uint32_t layer_it = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int64_t row0 = tile * kTileM;
            // ---- stage the input tile: fp32 rows -> fp16 hi / lo chunks.  A warp owns 8 rows; lane = (row & 7) +
            // 8 * (k-chunk & 3): a load instruction covers 8 rows x 128 contiguous bytes, and the 16-byte stores of
            // each quarter warp fall into 8 different bank groups (consecutive chunks of ONE row are 2 KB apart --
            // the same banks).  All of a batch's loads are issued before any conversion so their latencies overlap.
            {
                const int chunks = P.layer[0].k_pad / 8;
                const int s_row = warp * 8 + (lane & 7), s_kc = lane >> 3;
                const int64_t gr = row0 + s_row;
                const float* src_row = P.x + gr * P.ldx;
                constexpr int kRounds = kMaxK / 8 / 4;   // 8 rounds of four k-chunks
#pragma unroll 1
                for (int base = 0; base < kRounds && base * 4 < chunks; base += 4) {
                    float v[4][8];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int kc = 4 * (base + u) + s_kc;
#pragma unroll
                        for (int e = 0; e < 8; ++e)
                            v[u][e] = (kc < chunks && gr < P.m && kc * 8 + e < P.k0) ? __ldg(src_row + kc * 8 + e) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int kc = 4 * (base + u) + s_kc;
                        if (kc >= chunks) continue;
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) split2_f16(v[u][2 * j], v[u][2 * j + 1], hi[j], lo[j]);
                        const uint32_t off = a_chunk_offset(s_row, kc);
                        *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
            }
            fence_proxy_async();
            mbar_arrive(a_ready);

            for (int l = 0; l < P.n_layers; ++l, ++layer_it) {
                const bool last = l == P.n_layers - 1;
                const bool head = (l == 0) && P.w_head != nullptr;
                mbar_wait(d_ready, layer_it & 1);
                tc_fence_after();
                const float* bias = s_bias + l * kMaxN;
                const float inv_beta = 1.f / P.beta;
                float dot = 0.f;
                const int col_begin = group < n_groups ? group * cols_per_group : 0;
                const int col_end = group < n_groups ? col_begin + cols_per_group : 0;
                for (int c0 = col_begin; c0 < col_end; c0 += 16) {
                    uint32_t rm[16], rc[16];
                    tc_ld16(d_main + lane_addr + c0, rm);
                    tc_ld16(d_cross + lane_addr + c0, rc);
                    tc_wait_ld();
                    float act[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float pre = fmaf(__uint_as_float(rc[j]), kLoInv, __uint_as_float(rm[j])) + bias[c0 + j];
                        act[j] = softplus(pre, P.beta, inv_beta);
                        if (head) dot = fmaf(act[j], s_head[c0 + j], dot);
                    }
                    if (!last) {  // next layer's A operand, straight back to shared memory
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
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
                    float* dst = nullptr;
                    if (last && l == 1 && P.feat_out) dst = P.feat_out + (row0 + row) * P.ld_feat + c0;
                    if (l == 0 && P.hidden_out) dst = P.hidden_out + (row0 + row) * P.ld_hidden + c0;
                    if (dst && row0 + row < P.m) {
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            *reinterpret_cast<float4*>(dst + j) = make_float4(act[j], act[j + 1], act[j + 2], act[j + 3]);
                    }
                }
                if (head) {
                    head_partial[group * kTileM + row] = dot;    // groups without columns contribute 0
                    asm volatile("bar.sync 1, %0;" ::"n"(kFwdCompute) : "memory");
                    if (group == 0 && row0 + row < P.m && P.sdf_out) {
                        float sum = __ldg(P.b_head);
#pragma unroll
                        for (int gi = 0; gi < kFwdGroups; ++gi) sum += head_partial[gi * kTileM + row];
                        P.sdf_out[row0 + row] = sum;
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(kFwdCompute) : "memory");
                }
                tc_fence_before();
                if (!last) {
                    fence_proxy_async();
                    mbar_arrive(a_ready);
                }
            }
            // all TMEM reads of this tile are done before the next tile's first MMA overwrites the accumulators:
            // the next a_ready arrival (after the input staging above) orders them.
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kFwdWarps) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}


// ==================================================================================================
// Pipelined form of the forward for the SDF-only evaluation (n_layers == 1: finite-difference taps, the no-grad
// evaluations of the hierarchical sampler, the lattice queries of the mesh extractor -- four fifths of all point
// evaluations).  The kernel above runs input staging, MMA and epilogue of a tile one after the other and re-streams
// the weights for every tile; here
//   * W_hi / W_lo of the single layer stay RESIDENT in shared memory (144 x 256 x 2 x 2 B = 144 KB, loaded once per CTA
//     by two bulk copies), so the steady state has no weight traffic at all;
//   * the N columns are processed as two halves ("units"), each with its own D_main / D_cross accumulator pair
//     (2 x 256 TMEM columns), so the MMAs of one unit run under the epilogue of the other;
//   * the compute warps request the NEXT tile's rows from HBM into registers before they start the first epilogue of
//     the current tile and convert them into the split-fp16 A image between the two epilogues, when the MMAs that
//     read the current image have completed (a_free).
// Barriers (one completion per tile each, parity = local tile index & 1):
//   a_ready / a_free       A image:      compute warps <-> MMA warp (a_free = tcgen05.commit after the 2nd unit)
//   d_full[h] / d_empty[h] accumulators: MMA warp <-> compute warps
// Lane mapping of the input staging: a warp owns 8 rows; lane = (row & 7) + 8 * (k-chunk & 3), so a load instruction
// covers 8 rows x 128 contiguous bytes and the 16-byte stores of each quarter warp fall into 8 different bank groups.
// Envelope (host entry point): hidden width 128 or 256, k0 = 32 j + r with r <= 4 (3 + levels * features for every
// even level * feature product that is a multiple of 32), shared memory fits.
// ==================================================================================================
#ifdef NA_MLP_PROF   // developer build (scripts/profile_mlp_phases.py): per-phase clock64 totals of CTA 0
__device__ unsigned long long na_mlp_prof[16];
#define PROF_DECL(cond) const bool prof_on = (cond); long long prof_t = clock64()
#define PROF(i)                                      \
    do {                                             \
        if (prof_on) {                               \
            const long long t1 = clock64();          \
            na_mlp_prof[i] += (unsigned long long)(t1 - prof_t); \
            prof_t = t1;                             \
        }                                            \
    } while (0)
#else
#define PROF_DECL(cond)
#define PROF(i)
#endif

constexpr int kPipeWarps = 16;
constexpr int kPipeCompute = kPipeWarps * 32;
constexpr int kPipeThreads = kPipeCompute + 32;   // + MMA warp (also loads the weights once)
constexpr int kPipeBars = 8;
constexpr int kPipeTail = 4;                      // + up to 4 more

__host__ __device__ inline uint32_t pipe_smem_bytes(int k_pad, int n) {
    return 2u * kTileM * k_pad * 2u + 2u * (uint32_t)k_pad * n * 2u + (2u * 4u * kTileM + 2u * kMaxN) * 4u +
           kPipeBars * 8u + 16u;
}

template <int ROUNDS, bool PLANES>
__global__ void __launch_bounds__(kPipeThreads, 1) sdf_head_fwd_pipe_kernel(const __grid_constant__ MlpParams P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int n = P.n, nh = n / 2, k0 = P.k0, k_pad = P.layer[0].k_pad;
    const int chunks = k_pad / 8;
    const uint32_t a_bytes = (uint32_t)kTileM * k_pad * 2;
    const uint32_t w_bytes = (uint32_t)k_pad * n * 2;
    const uint32_t chunk_bytes = (uint32_t)n * 16;   // one 8-wide k-chunk of all N columns in the packed weights
    uint8_t* a_hi = smem;
    uint8_t* a_lo = a_hi + a_bytes;
    uint8_t* w_hi = a_lo + a_bytes;
    uint8_t* w_lo = w_hi + w_bytes;
    float* head_partial = reinterpret_cast<float*>(w_lo + w_bytes);   // [2][4][kTileM]
    float* s_bias = head_partial + 2 * 4 * kTileM;
    float* s_head = s_bias + kMaxN;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_head + kMaxN);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kPipeBars);
    const uint32_t bar0 = smem_u32(bars);
    const uint32_t w_full = bar0, a_ready = bar0 + 8u, a_free = bar0 + 16u;
    auto d_full = [&](int h) { return bar0 + 24u + 8u * h; };
    auto d_empty = [&](int h) { return bar0 + 40u + 8u * h; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t n_tiles = (P.m + kTileM - 1) / kTileM;
    const bool head = P.w_head != nullptr;

    for (int i = threadIdx.x; i < n; i += kPipeThreads) {
        s_bias[i] = P.layer[0].bias[i];
        s_head[i] = head ? P.w_head[i] : 0.f;
    }
    if (threadIdx.x == 0) {
        mbar_init(w_full, 1);
        mbar_init(a_ready, kPipeCompute);
        mbar_init(a_free, 1);
        for (int h = 0; h < 2; ++h) {
            mbar_init(d_full(h), 1);
            mbar_init(d_empty(h), kPipeCompute);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kPipeWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;   // unit h: D_main at columns [256 h, 256 h + nh), D_cross at + 128

    if (warp == kPipeWarps) {
        // ===================== weights (once) + MMA issuer =====================
        if (elect_one()) {
            mbar_expect_tx(w_full, 2u * w_bytes);
            bulk_copy_g2s(smem_u32(w_hi), P.layer[0].w_hi, w_bytes, w_full);
            bulk_copy_g2s(smem_u32(w_lo), P.layer[0].w_lo, w_bytes, w_full);
            mbar_wait(w_full, 0);
            const uint32_t idesc = make_idesc(kTileM, nh);
            const uint32_t a_lbo = kTileM / 8 * 128;
            uint32_t q = 0;
            PROF_DECL(blockIdx.x == 0);
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
                mbar_wait(a_ready, q & 1);
                tc_fence_after();
                PROF(8);
                for (int h = 0; h < 2; ++h) {
                    if (q > 0) {   // the epilogue of the previous tile has drained this accumulator pair
                        mbar_wait(d_empty(h), (q - 1) & 1);
                        tc_fence_after();
                    }
                    PROF(9 + 2 * h);
                    const uint32_t d_main = tmem_base + 256u * h, d_cross = d_main + 128u;
                    const uint32_t b_half = (uint32_t)h * nh * 16;   // N half inside a k-chunk of the packed weights
                    for (int c = 0; c < chunks; c += 2) {            // one UMMA k-step = 16 elements = 2 chunks
                        const uint32_t a_off = (uint32_t)c * a_lbo, b_off = (uint32_t)c * chunk_bytes + b_half;
                        const uint64_t ah = make_desc(smem_u32(a_hi) + a_off, a_lbo, 128);
                        const uint64_t al = make_desc(smem_u32(a_lo) + a_off, a_lbo, 128);
                        const uint64_t bh = make_desc(smem_u32(w_hi) + b_off, chunk_bytes, 128);
                        const uint64_t bl = make_desc(smem_u32(w_lo) + b_off, chunk_bytes, 128);
                        tc_mma_f16(d_cross, al, bh, idesc, c ? 1u : 0u);
                        tc_mma_f16(d_cross, ah, bl, idesc, 1u);
                        tc_mma_f16(d_main, ah, bh, idesc, c ? 1u : 0u);
                    }
                    tc_commit(d_full(h));
                    PROF(10 + 2 * h);
                }
                tc_commit(a_free);   // every MMA that reads this tile's A image has completed
            }
        }
    } else {
        // ===================== compute warps =====================
        const int quarter = warp & 3, group = warp >> 2;
        const int cpg = nh / 4 >= 16 ? nh / 4 : 16;       // columns of a unit per warp group
        const int n_groups = nh / cpg;
        const int row = quarter * 32 + lane;              // accumulator row == TMEM lane
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        const float inv_beta = 1.f / P.beta;
        // input staging: this thread converts row s_row, k-chunks 4 u + s_kc (u < ROUNDS) and, in lanes 0-7, the
        // remainder columns [32 ROUNDS, k0)
        const int s_row = warp * 8 + (lane & 7), s_kc = lane >> 3;
        const int tail_cols = k0 - 32 * ROUNDS;           // 0..kPipeTail
        float pre[ROUNDS][8], tailv[kPipeTail];

        // load step `step` of tile `tile` into registers: steps 0..ROUNDS-1 are the rounds of four k-chunks, step
        // ROUNDS the remainder columns.
        auto load_step = [&](int64_t tile, int step) {
            const int64_t gr = tile * kTileM + s_row;
            const bool ok = tile < n_tiles && gr < P.m;
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int u = 0; u < ROUNDS; ++u) {
                if (u != step) continue;
                if (PLANES) {   // [chunk][row][8] fp32: two aligned 16-byte loads per chunk, 256 contiguous bytes per 8 lanes
                    const float4* c = reinterpret_cast<const float4*>(P.planes + gr * 8 + (int64_t)(4 * u + s_kc) * P.chunk_stride);
                    const float4 lo4 = ok ? __ldg(c) : z, hi4 = ok ? __ldg(c + 1) : z;
                    pre[u][0] = lo4.x, pre[u][1] = lo4.y, pre[u][2] = lo4.z, pre[u][3] = lo4.w;
                    pre[u][4] = hi4.x, pre[u][5] = hi4.y, pre[u][6] = hi4.z, pre[u][7] = hi4.w;
                } else {
                    const float* src = P.x + gr * P.ldx + (4 * u + s_kc) * 8;
#pragma unroll
                    for (int e = 0; e < 8; ++e) pre[u][e] = ok ? __ldg(src + e) : 0.f;
                }
            }
            if (step == ROUNDS) {
                if (PLANES) {
                    const float4 t4 = (ok && s_kc == 0 && tail_cols > 0)
                                          ? __ldg(reinterpret_cast<const float4*>(P.planes + gr * 8 + (int64_t)(4 * ROUNDS) * P.chunk_stride))
                                          : z;
                    tailv[0] = t4.x, tailv[1] = t4.y, tailv[2] = t4.z, tailv[3] = t4.w;
#pragma unroll
                    for (int e = 0; e < kPipeTail; ++e) tailv[e] = e < tail_cols ? tailv[e] : 0.f;
                } else {
                    const float* src = P.x + gr * P.ldx + 32 * ROUNDS;
#pragma unroll
                    for (int e = 0; e < kPipeTail; ++e) tailv[e] = (ok && s_kc == 0 && e < tail_cols) ? __ldg(src + e) : 0.f;
                }
            }
        };
        auto store_chunk = [&](int kc, const float* v) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                split2_f16(v[2 * j], v[2 * j + 1], hi[j], lo[j]);
            }
            const uint32_t off = a_chunk_offset(s_row, kc);
            *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        };
        auto convert = [&]() {
#pragma unroll
            for (int u = 0; u < ROUNDS; ++u) store_chunk(4 * u + s_kc, pre[u]);
            if (s_kc == 0 && tail_cols > 0) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = e < kPipeTail ? tailv[e] : 0.f;
                store_chunk(4 * ROUNDS, v);
            }
        };

        // Half of the warp groups request a tile's rows one phase earlier than the other half (right after the
        // conversion that frees their registers instead of at the top of the next iteration): the two bursts of
        // load issue then overlap with the other half's epilogue arithmetic instead of stalling all sixteen warps.
        const bool early = group >= 2;
        float dot = 0.f;
        auto epilogue = [&](int h, int64_t row0) {
            const uint32_t d_main = tmem_base + 256u * h + lane_addr, d_cross = d_main + 128u;
            const int col_begin = group < n_groups ? group * cpg : 0;
            const int col_end = group < n_groups ? col_begin + cpg : 0;
            for (int c0 = col_begin; c0 < col_end; c0 += 16) {
                uint32_t rm[16], rc[16];
                tc_ld16(d_main + c0, rm);
                tc_ld16(d_cross + c0, rc);
                tc_wait_ld();
                const int gc = h * nh + c0;   // column of the layer
                float act[16];
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                    const float4 b4 = *reinterpret_cast<const float4*>(s_bias + gc + j);
                    const float4 h4 = *reinterpret_cast<const float4*>(s_head + gc + j);
                    const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float p = fmaf(__uint_as_float(rc[j + e]), kLoInv, __uint_as_float(rm[j + e])) + bb[e];
                        act[j + e] = (e & 1) ? softplus_poly(p, P.beta, inv_beta) : softplus(p, P.beta, inv_beta);
                        dot = fmaf(act[j + e], hh[e], dot);
                    }
                }
                if (P.hidden_out && row0 + row < P.m) {
                    float* dst = P.hidden_out + (row0 + row) * P.ld_hidden + gc;
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        *reinterpret_cast<float4*>(dst + j) = make_float4(act[j], act[j + 1], act[j + 2], act[j + 3]);
                }
            }
        };

        // the zero padding of the A image (k-chunks past the last input column) is written once
        for (int kc = 4 * ROUNDS + (tail_cols > 0 ? 1 : 0) + s_kc; kc < chunks; kc += 4) {
            const uint32_t off = a_chunk_offset(s_row, kc);
            *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(0u, 0u, 0u, 0u);
        }
        uint32_t q = 0;
        if ((int64_t)blockIdx.x < n_tiles) {   // prologue: the first tile's A image
#pragma unroll
            for (int st = 0; st <= ROUNDS; ++st) load_step(blockIdx.x, st);
            convert();
            fence_proxy_async();
            mbar_arrive(a_ready);
            if (early) {
#pragma unroll
                for (int st = 0; st <= ROUNDS; ++st) load_step((int64_t)blockIdx.x + gridDim.x, st);
            }
        }
        PROF_DECL(blockIdx.x == 0 && threadIdx.x == 0);
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
            const int64_t row0 = tile * kTileM;
            const int64_t next = tile + gridDim.x;
            if (!early) {
#pragma unroll
                for (int st = 0; st <= ROUNDS; ++st) load_step(next, st);   // in flight during the first epilogue
            }
            dot = 0.f;
            PROF(0);
            mbar_wait(d_full(0), q & 1);
            tc_fence_after();
            PROF(1);
            epilogue(0, row0);
            tc_fence_before();
            mbar_arrive(d_empty(0));
            PROF(2);
            if (next < n_tiles) {   // next tile's A image, under the second unit's MMAs
                mbar_wait(a_free, q & 1);
                PROF(3);
                convert();
                fence_proxy_async();
                mbar_arrive(a_ready);
                if (early) {   // registers are free again: request the tile after the next one right away
#pragma unroll
                    for (int st = 0; st <= ROUNDS; ++st) load_step(next + gridDim.x, st);
                }
                PROF(4);
            }
            mbar_wait(d_full(1), q & 1);
            tc_fence_after();
            PROF(5);
            epilogue(1, row0);
            tc_fence_before();
            mbar_arrive(d_empty(1));
            PROF(6);
            if (head && P.sdf_out) {
                float* part = head_partial + (q & 1) * 4 * kTileM;
                part[group * kTileM + row] = dot;   // groups without columns contribute 0
                // groups 1-3 only announce their partial sums and move on; group 0 waits for them and writes the SDF.
                // A group can be at most one tile ahead of group 0 (d_empty / a_ready need every compute thread), so
                // both the partial buffer and the named barrier alternate with the tile parity.
                const uint32_t bar_id = 1u + (q & 1u);
                if (group != 0) {
                    asm volatile("bar.arrive %0, %1;" ::"r"(bar_id), "n"(kPipeCompute) : "memory");
                } else {
                    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(kPipeCompute) : "memory");
                    if (row0 + row < P.m)
                        P.sdf_out[row0 + row] = __ldg(P.b_head) + part[row] + part[kTileM + row] + part[2 * kTileM + row] +
                                                part[3 * kTileM + row];
                }
            }
            PROF(7);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kPipeWarps) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

using PipeKernel = void (*)(const MlpParams);
template <bool PLANES>
PipeKernel pipe_kernel_for(int rounds) {
    switch (rounds) {
        case 1: return sdf_head_fwd_pipe_kernel<1, PLANES>;
        case 2: return sdf_head_fwd_pipe_kernel<2, PLANES>;
        case 3: return sdf_head_fwd_pipe_kernel<3, PLANES>;
        case 4: return sdf_head_fwd_pipe_kernel<4, PLANES>;
        default: return nullptr;
    }
}


// ==================================================================================================
// Backward of the same MLP with respect to its INPUT rows (the gradient that the hash-grid backward
// scatters into the table) and the pre-activation gradients dz0 / dz1 from which the host forms the weight
// gradients (dW = dz^T a: reductions over all rows, left to a library GEMM in round 1).
//
//   dz1 = d_feat * s'(z1)                      s'(z) = 1 - exp(-beta * softplus(z))  (from the saved activation)
//   dh0 = dz1 W1 + d_sdf w_head
//   dz0 = dh0 * s'(z0)
//   dx  = dz0 W0
// Same pipeline as the forward: the two products run on tcgen05 with split-fp16 operands; the elementwise
// links run in the accumulator epilogues and hand the next A operand over in shared memory.  Gradients are
// tiny (1e-9 .. 1e-4), far below fp16's range, so every row is multiplied by a power of two derived from a
// cheap upper bound of its magnitude before it is split, and the product is scaled back in the epilogue
// (row scaling commutes with the right-multiplication by W).
// ==================================================================================================
struct BwdParams {
    int64_t m;
    int n;                 // hidden width
    int k0, k0_pad;        // input width (valid, padded to 16)
    int two_layer;         // 1: feature branch present (d_feat / feat / W1), 0: SDF-only rows (taps)
    const float* d_sdf;    // [M] or null
    const float* d_feat;   // [M,N] (stride ld_act) or null
    const float* feat;     // [M,N] layer-1 activations
    const float* hidden;   // [M,N] layer-0 activations
    int64_t ld_act;
    const float* row_bound;  // [M] upper bound of |d_feat| per row (two_layer only)
    Layer w1t;             // packed W1^T: B[n_in][k_out]; k_pad = n
    Layer w0t;             // packed W0^T: B[col][k_hidden] with col padded to k0_pad; k_pad = n
    const float* w_head;   // [N]
    float beta;
    const float* norms;    // device [2]: max_j sum_k |W1[k,j]| (bound of |dz1 W1| per unit |dz1|), max |w_head|
    float* dz1;            // [M,N] out (two_layer), stride ld_act
    float* dz0;            // [M,N] out, stride ld_act
    float* dx;             // [M,k0] out, stride ld_dx
    int64_t ld_dx;
};

__device__ __forceinline__ float pow2_scale(float bound) {
    // largest power of two s with s * bound <= 2^14 (fp16 max is 2^16 - 32); bound == 0 -> 1
    if (!(bound > 0.f)) return 1.f;
    int e;
    frexpf(bound, &e);  // bound = f * 2^e, f in [0.5,1)
    e = 14 - e;
    e = e > 120 ? 120 : (e < -120 ? -120 : e);
    return ldexpf(1.f, e);
}

__global__ void __launch_bounds__(kThreads, 1) sdf_mlp_bwd_kernel(const __grid_constant__ BwdParams P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a_hi = smem;
    uint8_t* a_lo = smem + kABytes;
    uint8_t* b_ring = smem + 2 * kABytes;
    float* s_head = reinterpret_cast<float*>(b_ring + kStages * kBStageBytes);  // [kMaxN]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_head + kMaxN);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 2);
    const uint32_t bar0 = smem_u32(bars);
    auto b_full = [&](int s) { return bar0 + 8u * s; };
    auto b_empty = [&](int s) { return bar0 + 8u * (kStages + s); };
    const uint32_t a_ready = bar0 + 8u * (2 * kStages), d_ready = bar0 + 8u * (2 * kStages + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = P.n;
    const int n_mma = P.two_layer ? 2 : 1;
    const int64_t n_tiles = (P.m + kTileM - 1) / kTileM;

    for (int i = threadIdx.x; i < n; i += kThreads) s_head[i] = P.w_head[i];
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
        // ===================== weight loader: per tile [W1^T] then W0^T =====================
        if (elect_one()) {
            uint32_t it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int g = 0; g < n_mma; ++g) {
                    const bool is_w1 = P.two_layer && g == 0;
                    const Layer& L = is_w1 ? P.w1t : P.w0t;
                    const int n_out = is_w1 ? n : P.k0_pad;
                    const int chunks = n / 8;  // K = hidden width for both products
                    const uint32_t chunk_bytes = (uint32_t)n_out * 16;
                    for (int c0 = 0; c0 < chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        mbar_wait(b_empty(s), ((it / kStages) & 1) ^ 1);
                        const uint32_t bytes = (kBlockK / 8) * chunk_bytes;
                        mbar_expect_tx(b_full(s), 2 * bytes);
                        const uint32_t dst = smem_u32(b_ring + s * kBStageBytes);
                        bulk_copy_g2s(dst, reinterpret_cast<const uint8_t*>(L.w_hi) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                        bulk_copy_g2s(dst + kBStageBytes / 2,
                                      reinterpret_cast<const uint8_t*>(L.w_lo) + (size_t)c0 * chunk_bytes, bytes,
                                      b_full(s));
                    }
                }
            }
        }
    } else if (warp == kComputeWarps) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t a_lbo = kTileM / 8 * 128;
            uint32_t it = 0, mma_it = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int g = 0; g < n_mma; ++g, ++mma_it) {
                    const bool is_w1 = P.two_layer && g == 0;
                    const int n_out = is_w1 ? n : P.k0_pad;
                    const uint32_t idesc = make_idesc(kTileM, n_out);
                    const uint32_t b_lbo = (uint32_t)n_out / 8 * 128;
                    mbar_wait(a_ready, mma_it & 1);
                    tc_fence_after();
                    const int chunks = n / 8;
                    bool first = true;
                    for (int c0 = 0; c0 < chunks; c0 += kBlockK / 8, ++it) {
                        const int s = it % kStages;
                        mbar_wait(b_full(s), (it / kStages) & 1);
                        tc_fence_after();
                        const uint32_t b_hi_addr = smem_u32(b_ring + s * kBStageBytes);
                        const uint32_t b_lo_addr = b_hi_addr + kBStageBytes / 2;
                        for (int c = 0; c < kBlockK / 8; c += 2) {
                            const uint32_t a_off = (uint32_t)(c0 + c) * a_lbo, b_off = (uint32_t)c * b_lbo;
                            const uint64_t ah = make_desc(smem_u32(a_hi) + a_off, a_lbo, 128);
                            const uint64_t al = make_desc(smem_u32(a_lo) + a_off, a_lbo, 128);
                            const uint64_t bh = make_desc(b_hi_addr + b_off, b_lbo, 128);
                            const uint64_t bl = make_desc(b_lo_addr + b_off, b_lbo, 128);
                            tc_mma_f16(d_cross, al, bh, idesc, first ? 0u : 1u);
                            tc_mma_f16(d_cross, ah, bl, idesc, 1u);
                            tc_mma_f16(d_main, ah, bh, idesc, first ? 0u : 1u);
                            first = false;
                        }
                        tc_commit(b_empty(s));
                    }
                    tc_commit(d_ready);
                }
            }
        }
    } else {
        // ===================== compute warps =====================
        const int quarter = warp & 3, half = warp >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        const float beta = P.beta;
        const float w1_colsum_max = __ldg(P.norms), w_head_absmax = __ldg(P.norms + 1);
        uint32_t mma_it = 0;
        // write 32 consecutive values of this thread's row (scaled) as four fp16 hi/lo chunks of the A operand
        auto store_a = [&](const float (&v)[32], int c0, float scale) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    split2_f16(v[q * 8 + j * 2] * scale, v[q * 8 + j * 2 + 1] * scale, hi[j], lo[j]);
                }
                const uint32_t off = a_chunk_offset(row, c0 / 8 + q);
                *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
        };
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int64_t grow = tile * kTileM + row;
            const bool valid = grow < P.m;
            const float dsdf = (valid && P.d_sdf) ? __ldg(P.d_sdf + grow) : 0.f;
            const int col_begin = half * (n / 2), col_end = col_begin + n / 2;
            float scale_a, scale_b;  // row scales of the operands of the first / second product of this tile
            if (P.two_layer) {
                const float b1 = valid ? __ldg(P.row_bound + grow) : 0.f;
                scale_a = pow2_scale(b1);
                scale_b = pow2_scale(b1 * w1_colsum_max + fabsf(dsdf) * w_head_absmax);
                // ---- prologue: dz1 = d_feat * s'(z1) -> HBM (fp32) and A operand (scaled, split)
                for (int c0 = col_begin; c0 < col_end; c0 += 32) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 g = make_float4(0.f, 0.f, 0.f, 0.f), f = g;
                        if (valid) {
                            g = __ldg(reinterpret_cast<const float4*>(P.d_feat + grow * P.ld_act + c0 + j));
                            f = __ldg(reinterpret_cast<const float4*>(P.feat + grow * P.ld_act + c0 + j));
                        }
                        v[j] = g.x * (1.f - __expf(-beta * f.x)), v[j + 1] = g.y * (1.f - __expf(-beta * f.y));
                        v[j + 2] = g.z * (1.f - __expf(-beta * f.z)), v[j + 3] = g.w * (1.f - __expf(-beta * f.w));
                        if (valid)
                            *reinterpret_cast<float4*>(P.dz1 + grow * P.ld_act + c0 + j) =
                                make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    }
                    store_a(v, c0, scale_a);
                }
            } else {
                scale_a = 1.f;
                scale_b = pow2_scale(fabsf(dsdf) * w_head_absmax);
                // ---- prologue (SDF-only rows): dz0 = d_sdf w_head * s'(z0)
                for (int c0 = col_begin; c0 < col_end; c0 += 32) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (valid) h = __ldg(reinterpret_cast<const float4*>(P.hidden + grow * P.ld_act + c0 + j));
                        v[j] = dsdf * s_head[c0 + j] * (1.f - __expf(-beta * h.x));
                        v[j + 1] = dsdf * s_head[c0 + j + 1] * (1.f - __expf(-beta * h.y));
                        v[j + 2] = dsdf * s_head[c0 + j + 2] * (1.f - __expf(-beta * h.z));
                        v[j + 3] = dsdf * s_head[c0 + j + 3] * (1.f - __expf(-beta * h.w));
                        if (valid)
                            *reinterpret_cast<float4*>(P.dz0 + grow * P.ld_act + c0 + j) =
                                make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    }
                    store_a(v, c0, scale_b);
                }
            }
            fence_proxy_async();
            mbar_arrive(a_ready);

            if (P.two_layer) {
                // ---- epilogue of dz1 W1: dh0 = . / scale_a + d_sdf w_head;  dz0 = dh0 * s'(z0) -> HBM and A operand
                mbar_wait(d_ready, mma_it & 1);
                ++mma_it;
                tc_fence_after();
                const float inv_a = 1.f / scale_a;
                for (int c0 = col_begin; c0 < col_end; c0 += 32) {
                    uint32_t rm[32], rc[32];
                    tc_ld32(d_main + lane_addr + c0, rm);
                    tc_ld32(d_cross + lane_addr + c0, rc);
                    tc_wait_ld();
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (valid) h = __ldg(reinterpret_cast<const float4*>(P.hidden + grow * P.ld_act + c0 + j));
                        const float hh[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float dh = fmaf(__uint_as_float(rc[j + e]), kLoInv, __uint_as_float(rm[j + e])) * inv_a +
                                             dsdf * s_head[c0 + j + e];
                            v[j + e] = dh * (1.f - __expf(-beta * hh[e]));
                        }
                        if (valid)
                            *reinterpret_cast<float4*>(P.dz0 + grow * P.ld_act + c0 + j) =
                                make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    }
                    store_a(v, c0, scale_b);
                }
                tc_fence_before();
                fence_proxy_async();
                mbar_arrive(a_ready);
            }
            // ---- epilogue of dz0 W0: dx = . / scale_b  (k0_pad columns, split between the two warp groups)
            mbar_wait(d_ready, mma_it & 1);
            ++mma_it;
            tc_fence_after();
            {
                const float inv_b = 1.f / scale_b;
                const int n_chunks = (P.k0_pad + 31) / 32;
                const int first_chunk = half == 0 ? 0 : (n_chunks + 1) / 2;
                const int last_chunk = half == 0 ? (n_chunks + 1) / 2 : n_chunks;
                for (int ch = first_chunk; ch < last_chunk; ++ch) {
                    const int c0 = ch * 32;
                    uint32_t rm[32], rc[32];
                    tc_ld32(d_main + lane_addr + c0, rm);
                    tc_ld32(d_cross + lane_addr + c0, rc);
                    tc_wait_ld();
                    if (valid) {
                        float* dst = P.dx + grow * P.ld_dx + c0;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c0 + j < P.k0)
                                dst[j] = fmaf(__uint_as_float(rc[j]), kLoInv, __uint_as_float(rm[j])) * inv_b;
                    }
                }
            }
            tc_fence_before();
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kComputeWarps) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// The opt-in to > 48 KB of dynamic shared memory is per device and is not a stream operation: do it once per
// (kernel, device) so that later launches -- possibly under CUDA-graph capture -- make no attribute call.
int ensure_smem_attr(int which, const void* func, const char* what, int smem_bytes) {
    static bool done[10][64] = {};
    int dev = 0;
    if (int rc = na::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return rc;
    if (dev < 0 || dev >= 64 || !done[which][dev]) {
        if (int rc = na::check_cuda(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes),
                                    what))
            return rc;
        if (dev >= 0 && dev < 64) done[which][dev] = true;
    }
    return NA_OK;
}

}  // namespace

#ifdef NA_MLP_PROF
extern "C" int na_dev_mlp_prof(unsigned long long* out16, int reset) {
    if (out16 && cudaMemcpyFromSymbol(out16, na_mlp_prof, sizeof(unsigned long long) * 16) != cudaSuccess) return -1;
    if (reset) {
        unsigned long long z[16] = {};
        if (cudaMemcpyToSymbol(na_mlp_prof, z, sizeof(z)) != cudaSuccess) return -1;
    }
    return 0;
}
#endif

extern "C" int na_sdf_mlp_fwd(const float* x, int64_t ldx, int64_t m, int k0, int n, int n_layers,
                              const void* w0_hi, const void* w0_lo, const float* b0, int k0_pad, const void* w1_hi,
                              const void* w1_lo, const float* b1, const float* w_head, const float* b_head, float beta,
                              float* sdf_out, float* feat_out, int64_t ld_feat, float* hidden_out, int64_t ld_hidden,
                              void* stream) {
    na::clear_error();
    NA_REQUIRE(m >= 0, "sdf_mlp_fwd: negative row count");
    if (m == 0) return NA_OK;
    NA_REQUIRE(x && w0_hi && w0_lo && b0, "sdf_mlp_fwd: null pointer");
    NA_REQUIRE(!w_head || b_head, "sdf_mlp_fwd: w_head given without b_head");
    NA_REQUIRE(n % 64 == 0 && n >= 64 && n <= kMaxN, "sdf_mlp_fwd: hidden width %d unsupported (64, 128, 192, 256)", n);
    NA_REQUIRE(k0 >= 1 && k0 <= k0_pad && k0_pad % 16 == 0 && k0_pad <= kMaxK,
               "sdf_mlp_fwd: k0=%d k0_pad=%d unsupported (k0_pad multiple of 16, <= %d)", k0, k0_pad, kMaxK);
    NA_REQUIRE(n_layers == 1 || (n_layers == 2 && w1_hi && w1_lo && b1), "sdf_mlp_fwd: n_layers must be 1 or 2");
    NA_REQUIRE(beta > 0.f, "sdf_mlp_fwd: softplus beta must be positive");
    NA_REQUIRE(!feat_out || (n_layers == 2 && ld_feat % 4 == 0 && (uintptr_t)feat_out % 16 == 0),
               "sdf_mlp_fwd: feat_out needs n_layers == 2, 16-byte alignment and a stride multiple of 4");
    NA_REQUIRE(!hidden_out || (ld_hidden % 4 == 0 && (uintptr_t)hidden_out % 16 == 0),
               "sdf_mlp_fwd: hidden_out needs 16-byte alignment and a stride multiple of 4");
    NA_REQUIRE(((uintptr_t)w0_hi | (uintptr_t)w0_lo | (uintptr_t)w1_hi | (uintptr_t)w1_lo) % 16 == 0,
               "sdf_mlp_fwd: packed weights must be 16-byte aligned");
    MlpParams P{};
    P.x = x, P.ldx = ldx, P.m = m, P.k0 = k0, P.n = n, P.n_layers = n_layers;
    P.layer[0] = Layer{(const __half*)w0_hi, (const __half*)w0_lo, b0, k0_pad};
    P.layer[1] = Layer{(const __half*)w1_hi, (const __half*)w1_lo, b1, n};
    P.w_head = w_head, P.b_head = b_head, P.beta = beta;
    P.sdf_out = sdf_out, P.feat_out = feat_out, P.ld_feat = ld_feat, P.hidden_out = hidden_out, P.ld_hidden = ld_hidden;
    const int64_t tiles = (m + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    // SDF-only evaluation: the pipelined kernel with resident weights (developer builds with -DNA_MLP_DEV_KNOBS:
    // NA_MLP_NO_PIPE=1 switches it off for A/B runs; the shipped library never reads the environment)
#ifdef NA_MLP_DEV_KNOBS
    static const bool no_pipe = getenv("NA_MLP_NO_PIPE") != nullptr;
#else
    constexpr bool no_pipe = false;
#endif
    const uint32_t pipe_smem = pipe_smem_bytes(k0_pad, n);
    const int rounds = k0 / 32;
    if (!no_pipe && n_layers == 1 && n % 128 == 0 && k0 - 32 * rounds <= kPipeTail && pipe_kernel_for<false>(rounds) &&
        pipe_smem <= 227u * 1024u) {
        const PipeKernel kern = pipe_kernel_for<false>(rounds);
        if (int rc = ensure_smem_attr(1 + rounds, (const void*)kern, "cudaFuncSetAttribute(sdf_head_fwd_pipe_kernel)",
                                      227 * 1024))
            return rc;
        kern<<<grid, kPipeThreads, pipe_smem, (cudaStream_t)stream>>>(P);
        NA_CHECK_LAUNCH("sdf_head_fwd_pipe_kernel");
        return NA_OK;
    }
    if (int rc = ensure_smem_attr(0, (const void*)sdf_mlp_fwd_kernel, "cudaFuncSetAttribute(sdf_mlp_fwd_kernel)", kFwdSmemBytes)) return rc;
    sdf_mlp_fwd_kernel<<<grid, kFwdThreads, kFwdSmemBytes, (cudaStream_t)stream>>>(P);
    NA_CHECK_LAUNCH("sdf_mlp_fwd_kernel");
    return NA_OK;
}

extern "C" int na_sdf_head_fwd_planes(const float* planes, int64_t chunk_stride, int64_t m, int k0, int n,
                                      const void* w0_hi, const void* w0_lo, const float* b0, int k0_pad,
                                      const float* w_head, const float* b_head, float beta, float* sdf_out,
                                      float* hidden_out, int64_t ld_hidden, void* stream) {
    na::clear_error();
    NA_REQUIRE(m >= 0, "sdf_head_fwd_planes: negative row count");
    if (m == 0) return NA_OK;
    NA_REQUIRE(planes && w0_hi && w0_lo && b0, "sdf_head_fwd_planes: null pointer");
    NA_REQUIRE(!w_head || b_head, "sdf_head_fwd_planes: w_head given without b_head");
    NA_REQUIRE((uintptr_t)planes % 16 == 0 && chunk_stride % 4 == 0 && chunk_stride >= m * 8,
               "sdf_head_fwd_planes: planes must be 16-byte aligned with chunk_stride a multiple of 4 and >= 8 m");
    const int rounds = k0 / 32;
    NA_REQUIRE(n % 128 == 0 && n >= 128 && n <= kMaxN, "sdf_head_fwd_planes: hidden width %d unsupported (128, 256)", n);
    NA_REQUIRE(k0 >= 32 && k0 - 32 * rounds <= kPipeTail && pipe_kernel_for<true>(rounds) != nullptr,
               "sdf_head_fwd_planes: k0=%d unsupported (32 j + r, j in 1..4, r <= %d)", k0, kPipeTail);
    NA_REQUIRE(k0 <= k0_pad && k0_pad % 16 == 0 && k0_pad <= kMaxK, "sdf_head_fwd_planes: k0_pad=%d unsupported", k0_pad);
    NA_REQUIRE(beta > 0.f, "sdf_head_fwd_planes: softplus beta must be positive");
    NA_REQUIRE(!hidden_out || (ld_hidden % 4 == 0 && (uintptr_t)hidden_out % 16 == 0),
               "sdf_head_fwd_planes: hidden_out needs 16-byte alignment and a stride multiple of 4");
    NA_REQUIRE(((uintptr_t)w0_hi | (uintptr_t)w0_lo) % 16 == 0, "sdf_head_fwd_planes: packed weights must be 16-byte aligned");
    const uint32_t smem = pipe_smem_bytes(k0_pad, n);
    NA_REQUIRE(smem <= 227u * 1024u, "sdf_head_fwd_planes: k0_pad=%d x n=%d weights do not fit in shared memory", k0_pad, n);
    MlpParams P{};
    P.planes = planes, P.chunk_stride = chunk_stride, P.m = m, P.k0 = k0, P.n = n, P.n_layers = 1;
    P.layer[0] = Layer{(const __half*)w0_hi, (const __half*)w0_lo, b0, k0_pad};
    P.w_head = w_head, P.b_head = b_head, P.beta = beta;
    P.sdf_out = sdf_out, P.hidden_out = hidden_out, P.ld_hidden = ld_hidden;
    const PipeKernel kern = pipe_kernel_for<true>(rounds);
    if (int rc = ensure_smem_attr(5 + rounds, (const void*)kern, "cudaFuncSetAttribute(sdf_head_fwd_pipe_kernel)", 227 * 1024))
        return rc;
    const int64_t tiles = (m + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    kern<<<grid, kPipeThreads, smem, (cudaStream_t)stream>>>(P);
    NA_CHECK_LAUNCH("sdf_head_fwd_pipe_kernel");
    return NA_OK;
}

extern "C" int na_sdf_mlp_bwd(int64_t m, int n, int k0, int k0_pad, int two_layer, const float* d_sdf,
                              const float* d_feat, const float* feat, const float* hidden, int64_t ld_act,
                              const float* row_bound, const void* w1t_hi, const void* w1t_lo, const void* w0t_hi,
                              const void* w0t_lo, const float* w_head, float beta, const float* norms,
                              float* dz1, float* dz0, float* dx, int64_t ld_dx, void* stream) {
    na::clear_error();
    NA_REQUIRE(m >= 0, "sdf_mlp_bwd: negative row count");
    if (m == 0) return NA_OK;
    NA_REQUIRE(n % 64 == 0 && n >= 64 && n <= kMaxN, "sdf_mlp_bwd: hidden width %d unsupported", n);
    NA_REQUIRE(k0 >= 1 && k0 <= k0_pad && k0_pad % 16 == 0 && k0_pad <= kMaxK, "sdf_mlp_bwd: k0=%d k0_pad=%d unsupported",
               k0, k0_pad);
    NA_REQUIRE(hidden && w0t_hi && w0t_lo && w_head && norms && dz0 && dx, "sdf_mlp_bwd: null pointer");
    NA_REQUIRE(!two_layer || (d_feat && feat && row_bound && w1t_hi && w1t_lo && dz1),
               "sdf_mlp_bwd: the two-layer form needs d_feat, feat, row_bound, W1^T and dz1");
    NA_REQUIRE(ld_act % 4 == 0 && ((uintptr_t)hidden | (uintptr_t)dz0 | (uintptr_t)dz1 | (uintptr_t)d_feat |
                                   (uintptr_t)feat) % 16 == 0,
               "sdf_mlp_bwd: activation tensors need 16-byte alignment and a stride multiple of 4");
    BwdParams P{};
    P.m = m, P.n = n, P.k0 = k0, P.k0_pad = k0_pad, P.two_layer = two_layer ? 1 : 0;
    P.d_sdf = d_sdf, P.d_feat = d_feat, P.feat = feat, P.hidden = hidden, P.ld_act = ld_act, P.row_bound = row_bound;
    P.w1t = Layer{(const __half*)w1t_hi, (const __half*)w1t_lo, nullptr, n};
    P.w0t = Layer{(const __half*)w0t_hi, (const __half*)w0t_lo, nullptr, n};
    P.w_head = w_head, P.beta = beta, P.norms = norms;
    P.dz1 = dz1, P.dz0 = dz0, P.dx = dx, P.ld_dx = ld_dx;
    if (int rc = ensure_smem_attr(1, (const void*)sdf_mlp_bwd_kernel, "cudaFuncSetAttribute(sdf_mlp_bwd_kernel)", kSmemBytes)) return rc;
    const int64_t tiles = (m + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    sdf_mlp_bwd_kernel<<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(P);
    NA_CHECK_LAUNCH("sdf_mlp_bwd_kernel");
    return NA_OK;
}
