// na_lattice_sdf: the SDF on a lattice block (extract_mesh's blocked query, projects/neuralangelo/utils/mesh.py:32-35,68-100
// -> NeuralSDF.sdf, modules.py:72-74) as ONE kernel: hash-grid encode -> split-fp16 A image in shared memory -> tcgen05
// hidden layer -> softplus + head dot product in the accumulator epilogue.  No feature planes in HBM.
//
// Tile = 16 lattice columns (consecutive (ix, iy) pairs) x 8 consecutive z = 128 rows of one UMMA.  Along a z run the
// (x, y) part of a cell is fixed, so a level's value is a linear blend of two FACE values (bilinear combinations of
// four table rows each): a run costs 2 faces + 1 per cell crossing instead of 8 rows per point -- at the 2048^3
// lattice of the default model (2 lattice points per finest cell) 23 row loads per point instead of 128.
//
// Warp roles (640 threads, registers re-distributed with setmaxnreg):
//   warps 0-7    gather: thread = (column, level); features of its 8 points are computed into registers while the
//                previous tile is still in the tensor pipe, then stored into the A image of its k-chunk group
//   warps 8-15   epilogue: TMEM -> softplus -> dot with the head row -> SDF
//   warp 16      weights (once, bulk copy) and MMA issue; warps 17-19 only give their registers away
// The A image is single-buffered (W_hi + W_lo take 144 KB of shared memory), but its four k-chunk groups (4 levels
// each; the last one also holds the xyz chunk) are released and re-filled one by one: a_free[g] is committed after the
// LAST pass over group g, a_ready[g] is awaited before the FIRST pass of the next tile over it, so a group's gather
// warps have three group-times of MMA to store their registers without stalling the tensor pipe.
// Operands arrive PRE-SCALED so that the epilogue is seven instructions per hidden unit: W0 and b0 are multiplied by
// beta * log2(e) and b0 is the weight row of a constant-one input column (next to xyz), so the accumulator holds
// t = beta * log2(e) * (W0 x + b0); softplus_beta(z) * beta / ln 2 = max(t, 0) + log2(1 + 2^-|t|); the head row carries the
// remaining ln 2 / beta.
#include "tc_common.cuh"

using namespace na_tc;

namespace {

constexpr int kGatherThreads = 256, kEpiThreads = 256;
constexpr int kThreads = 640;
constexpr int kMmaWarp = 16;
constexpr int kCols = 16, kRun = 8;
constexpr int kLevels = 16, kGroups = 4;
constexpr int kN = 256, kNh = 128;
constexpr int kKPad = 144, kChunks = kKPad / 8;       // 16 levels + xyz chunk + zero chunk
constexpr int kRegsGather = 136, kRegsEpi = 88, kRegsMma = 32;
constexpr int kBars = 13;
constexpr int kSrcLattice = 0, kSrcPlanes = 1, kSrcRows = 2;   // where a tile's 128 input rows come from

struct LevelC {
    float scale;
    uint32_t res, zmul, entries, hashed, mask, offset, pad;
};

struct Params {
    na_hashgrid_desc desc;
    const float* table;
    const float *gx, *gy, *gz;
    int nx, ny, nz;
    na_unit_map map;
    uint32_t active_levels;
    const __half *w_hi, *w_lo;
    const float *w_head, *b_head;
    float out_scale;
    // row sources of the plain SDF-only MLP (kSrcPlanes / kSrcRows): the same pipeline without the gather
    const float* planes;
    int64_t chunk_stride;
    const float* x;
    int64_t ldx, m;
    float* sdf_out;
};

constexpr uint32_t kABytes = kTileM * kKPad * 2;   // one fp16 A image (hi or lo)
constexpr uint32_t kWBytes = kKPad * kN * 2;
constexpr uint32_t kSmemBytes = 2 * kABytes + 2 * kWBytes + (2 * 2 * kTileM + kN) * 4 + kLevels * sizeof(LevelC) +
                                kBars * 8 + 16;

struct R8 {
    float v[8];
};

__device__ __forceinline__ R8 row_load(const float* p, uint64_t pol) {
    R8 r;
    asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]),
                   "=f"(r.v[7])
                 : "l"(p), "l"(pol));
    return r;
}

#ifdef NA_LATTICE_DEV   // developer build (scripts/bench_lattice_sdf.py): 1 = no table loads, 2 = no epilogue arithmetic, 4 = L2 prefetch of the next tile's rows,
                        // 8 = gather warps only shake hands, 16 = no MMA issued
__device__ int na_lattice_dev_mode;
#define DEV_MODE(bit) ((dev_mode & (bit)) != 0)
#else
#define DEV_MODE(bit) false
#endif

// max(t, 0) + log2(1 + 2^-|t|): softplus in the log2 domain (raw ex2 / lg2: ~1e-7 absolute on the log term, which the
// head row scales by ln 2 / beta)
__device__ __forceinline__ float softplus2(float t) {
    float e, l;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-fabsf(t)));
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(1.f + e));
    return fmaxf(t, 0.f) + l;
}

__device__ __noinline__ uint32_t wrap_index(uint32_t idx, uint32_t entries) { return idx % entries; }

template <int REGS>
__device__ __forceinline__ void regs_inc() {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS));
}
template <int REGS>
__device__ __forceinline__ void regs_dec() {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS));
}

template <int SRC>
__global__ void __launch_bounds__(kThreads, 1) lattice_sdf_kernel(const __grid_constant__ Params P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* a_hi = smem;
    uint8_t* a_lo = a_hi + kABytes;
    uint8_t* w_hi = a_lo + kABytes;
    uint8_t* w_lo = w_hi + kWBytes;
    float* head_partial = reinterpret_cast<float*>(w_lo + kWBytes);   // [2][2][kTileM]
    float* s_head = head_partial + 2 * 2 * kTileM;
    LevelC* s_lv = reinterpret_cast<LevelC*>(s_head + kN);
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_lv + kLevels);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kBars);
    const uint32_t bar0 = smem_u32(bars);
    const uint32_t w_full = bar0;
    auto a_ready = [&](int g) { return bar0 + 8u * (1 + g); };
    auto a_free = [&](int g) { return bar0 + 8u * (5 + g); };
    auto d_full = [&](int h) { return bar0 + 8u * (9 + h); };
    auto d_empty = [&](int h) { return bar0 + 8u * (11 + h); };

#ifdef NA_LATTICE_DEV
    const int dev_mode = na_lattice_dev_mode;
#endif
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_zr = SRC == kSrcLattice ? (P.nz + kRun - 1) / kRun : 1;
    const int64_t n_cols = (int64_t)P.nx * P.ny;
    const int64_t n_tiles = SRC == kSrcLattice ? ((n_cols + kCols - 1) / kCols) * n_zr : (P.m + kTileM - 1) / kTileM;

    for (int i = threadIdx.x; i < kN; i += kThreads) {
        s_head[i] = P.w_head[i];
    }
    if (SRC == kSrcLattice && threadIdx.x < kLevels) {
        const int l = threadIdx.x;
        LevelC v;
        v.scale = P.desc.scale[l];
        v.res = P.desc.resolution[l];
        v.zmul = v.res * v.res;
        v.entries = P.desc.entries[l];
        v.hashed = P.desc.hashed[l];
        v.mask = ((v.entries & (v.entries - 1u)) == 0u) ? (v.entries - 1u) : 0u;
        v.offset = P.desc.offset[l];
        v.pad = 0;
        s_lv[l] = v;
    }
    if (threadIdx.x == 0) {
        mbar_init(w_full, 1);
        for (int g = 0; g < kGroups; ++g) {
            mbar_init(a_ready(g), kGatherThreads / kGroups);
            mbar_init(a_free(g), 1);
        }
        for (int h = 0; h < 2; ++h) {
            mbar_init(d_full(h), 1);
            mbar_init(d_empty(h), kEpiThreads);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;   // unit h (columns 128 h .. 128 h + 127 of the layer): D_main at 256 h, D_cross at + 128

    if (warp >= kMmaWarp) {
        regs_dec<kRegsMma>();
        if (warp == kMmaWarp && elect_one()) {
            // ===================== weights (once) + MMA issuer =====================
            mbar_expect_tx(w_full, 2u * kWBytes);
            bulk_copy_g2s(smem_u32(w_hi), P.w_hi, kWBytes, w_full);
            bulk_copy_g2s(smem_u32(w_lo), P.w_lo, kWBytes, w_full);
            mbar_wait(w_full, 0);
            const uint32_t idesc = make_idesc(kTileM, kNh);
            constexpr uint32_t a_lbo = kTileM / 8 * 128, chunk_bytes = kN * 16;
            // descriptors differ only in their 14-bit start-address field: keep the four low words and one high word
            const uint64_t d_ah = make_desc(smem_u32(a_hi), a_lbo, 128), d_bh = make_desc(smem_u32(w_hi), chunk_bytes, 128);
            const uint32_t hi_word = (uint32_t)(d_ah >> 32);
            const uint32_t ah0 = (uint32_t)d_ah, al0 = (uint32_t)make_desc(smem_u32(a_lo), a_lbo, 128);
            const uint32_t bh0 = (uint32_t)d_bh, bl0 = (uint32_t)make_desc(smem_u32(w_lo), chunk_bytes, 128);
            auto desc = [&](uint32_t lo_word) { return ((uint64_t)hi_word << 32) | lo_word; };
            uint32_t q = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
#pragma unroll 1
                for (int h = 0; h < 2; ++h) {
                    if (q > 0) {   // the epilogue of the previous tile has drained this accumulator pair
                        mbar_wait(d_empty(h), (q - 1) & 1);
                        tc_fence_after();
                    }
                    const uint32_t d_main = tmem_base + 256u * h, d_cross = d_main + 128u;
#pragma unroll 1
                    for (int g = 0; g < kGroups; ++g) {
                        if (h == 0) {
                            mbar_wait(a_ready(g), q & 1);
                            tc_fence_after();
                        }
                        const int c_end = g == kGroups - 1 ? kChunks : 4 * g + 4;
#pragma unroll
                        for (int c = 4 * g; c < c_end; c += 2) {   // one UMMA k-step = 16 elements = 2 chunks
                            const uint32_t a_add = (uint32_t)c * (a_lbo >> 4);
                            const uint32_t b_add = (uint32_t)c * (chunk_bytes >> 4) + (uint32_t)h * (kNh * 16 >> 4);
                            if (DEV_MODE(16)) continue;
                            tc_mma_f16(d_cross, desc(al0 + a_add), desc(bh0 + b_add), idesc, c ? 1u : 0u);
                            tc_mma_f16(d_cross, desc(ah0 + a_add), desc(bl0 + b_add), idesc, 1u);
                            tc_mma_f16(d_main, desc(ah0 + a_add), desc(bh0 + b_add), idesc, c ? 1u : 0u);
                        }
                        if (h == 1) tc_commit(a_free(g));   // the last pass over this group's chunks
                    }
                    tc_commit(d_full(h));
                }
            }
        }
    } else if (warp < 8) {
        regs_inc<kRegsGather>();
        // ===================== gather warps =====================
        const int gt = threadIdx.x;
        const int g = gt >> 6, lvl = 4 * g + ((gt & 63) >> 4), c = gt & 15;
        const bool xyz_writer = (gt & 63) < 16 && g == kGroups - 1;
        if (xyz_writer) {   // the zero chunk after xyz is written once
#pragma unroll
            for (int j = 0; j < kRun; ++j) {
                const uint32_t off = a_chunk_offset(j * kCols + c, kLevels + 1);
                *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        if constexpr (SRC == kSrcLattice) {
            const LevelC L = s_lv[lvl];
            const float* base = P.table + (size_t)L.offset * 8;
            const bool active = (uint32_t)lvl < P.active_levels;
            const float vmin = P.map.vol_min, vinv = P.map.inv_extent;
            const uint64_t pol = na::policy_evict_last();
            // tcnn grid_index(): dense stride walk or the coherent prime hash, then modulo the level size -- a mask for the
            // power-of-two hashed levels, otherwise only the wrap-around at the far faces (out of line: a division is ~25
            // instructions and this lambda is inlined once per point of the unrolled run)
            auto row_index = [&](uint32_t bxy, int cz) {
                uint32_t idx = L.hashed ? (bxy ^ ((uint32_t)cz * 805459861u)) : (bxy + (uint32_t)cz * L.zmul);
                if (L.mask) idx &= L.mask;
                else if (idx >= L.entries) idx = wrap_index(idx, L.entries);
                return idx;
            };
            // Every row the thread will need for tile t, requested into the L2 a whole tile ahead (no registers held): the
            // dependent face loads of a run then pay L2-hit latency, not DRAM latency.
            auto prefetch_tile = [&](int64_t t) {
                if (t >= n_tiles || !active || !DEV_MODE(4)) return;
                const int64_t pc = (t / n_zr) * kCols + c;
                if (pc >= n_cols) return;
                const int pz0 = (int)(t % n_zr) * kRun;
                const int pix = (int)(pc / P.ny), piy = (int)(pc - (int64_t)pix * P.ny);
                const float px = fmaf(L.scale, (__ldg(P.gx + pix) - vmin) * vinv, 0.5f);
                const float py = fmaf(L.scale, (__ldg(P.gy + piy) - vmin) * vinv, 0.5f);
                const float za = fmaf(L.scale, (__ldg(P.gz + pz0) - vmin) * vinv, 0.5f);
                const float zb = fmaf(L.scale, (__ldg(P.gz + min(pz0 + kRun - 1, P.nz - 1)) - vmin) * vinv, 0.5f);
                const uint32_t cx = (uint32_t)(int)floorf(px), cy = (uint32_t)(int)floorf(py);
                const int ca = (int)floorf(fminf(za, zb)), cb = min((int)floorf(fmaxf(za, zb)) + 1, ca + 11);
    #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t ux = cx + (k & 1), uy = cy + (k >> 1);
                    const uint32_t b = L.hashed ? (ux ^ (uy * 2654435761u)) : (ux + uy * L.res);
                    for (int cz = ca; cz <= cb; ++cz)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t)row_index(b, cz) * 8));
                }
            };
            prefetch_tile(blockIdx.x);

    uint32_t q = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
                if (DEV_MODE(8)) {
                    if (q > 0) mbar_wait(a_free(g), (q - 1) & 1);
                    fence_proxy_async();
                    mbar_arrive(a_ready(g));
                    continue;
                }
                prefetch_tile(tile + gridDim.x);
                const int zr = (int)(tile % n_zr);
                const int64_t col = (tile / n_zr) * kCols + c;
                const int z0 = zr * kRun;
                const bool col_ok = col < n_cols;
                const int ix = col_ok ? (int)(col / P.ny) : 0, iy = col_ok ? (int)(col - (int64_t)ix * P.ny) : 0;
                const float x = __ldg(P.gx + ix), y = __ldg(P.gy + iy);
                float zs[kRun];
    #pragma unroll
                for (int j = 0; j < kRun; ++j) zs[j] = __ldg(P.gz + min(z0 + j, P.nz - 1));

                uint32_t hi[kRun][4], lo[kRun][4];
                if (active && col_ok) {
                    const float px = fmaf(L.scale, (x - vmin) * vinv, 0.5f), py = fmaf(L.scale, (y - vmin) * vinv, 0.5f);
                    const float fx = floorf(px), fy = floorf(py);
                    const uint32_t cx = (uint32_t)(int)fx, cy = (uint32_t)(int)fy;
                    const float wx = px - fx, wy = py - fy;
                    uint32_t bxy[4];
                    float w4[4];
    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t ux = cx + (k & 1), uy = cy + (k >> 1);
                        bxy[k] = L.hashed ? (ux ^ (uy * 2654435761u)) : (ux + uy * L.res);
                        w4[k] = ((k & 1) ? wx : 1.f - wx) * ((k >> 1) ? wy : 1.f - wy);
                    }
                    // bilinear (x, y) combination of the four rows of the cell face at height cf
                    auto face_into = [&](R8& f, int cf) {
                        R8 r[4];
    #pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (DEV_MODE(1)) {
    #pragma unroll
                                for (int e = 0; e < 8; ++e) r[k].v[e] = 0.f;
                            } else {
                                r[k] = row_load(base + (size_t)row_index(bxy[k], cf) * 8, pol);
                            }
                        }
    #pragma unroll
                        for (int e = 0; e < 8; ++e)
                            f.v[e] = fmaf(w4[3], r[3].v[e], fmaf(w4[2], r[2].v[e], fmaf(w4[1], r[1].v[e], w4[0] * r[0].v[e])));
                    };
                    // fa / fb: the faces below / above the current point.  Both are new at the start of the run or after a jump,
                    // one after a step into the next cell.  ONE inlined copy of face_into() per point keeps the unrolled loop
                    // small, and lanes that need a face at the same time share the instructions.
                    R8 fa{}, fb{};
                    int cur = 0;
    #pragma unroll
                    for (int j = 0; j < kRun; ++j) {
                        const float pz = fmaf(L.scale, (zs[j] - vmin) * vinv, 0.5f);
                        const float fz = floorf(pz);
                        const int cz = (int)fz;
                        const float wz = pz - fz, uz = 1.f - wz;
                        int need = j == 0 ? 2 : (cz == cur ? 0 : (cz == cur + 1 ? 1 : 2));
                        for (; need > 0; --need) {
                            fa = fb;
                            face_into(fb, cz + 2 - need);
                        }
                        cur = cz;
    #pragma unroll
                        for (int e = 0; e < 4; ++e)
                            split2_f16(fmaf(wz, fb.v[2 * e], uz * fa.v[2 * e]), fmaf(wz, fb.v[2 * e + 1], uz * fa.v[2 * e + 1]),
                                       hi[j][e], lo[j][e]);
                    }
                } else {
    #pragma unroll
                    for (int j = 0; j < kRun; ++j)
    #pragma unroll
                        for (int e = 0; e < 4; ++e) hi[j][e] = 0u, lo[j][e] = 0u;
                }

                if (q > 0) mbar_wait(a_free(g), (q - 1) & 1);   // every MMA of the previous tile over this group is done
    #pragma unroll
                for (int j = 0; j < kRun; ++j) {
                    const uint32_t off = a_chunk_offset(j * kCols + c, lvl);
                    *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[j][0], hi[j][1], hi[j][2], hi[j][3]);
                    *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[j][0], lo[j][1], lo[j][2], lo[j][3]);
                }
                if (xyz_writer) {
                    uint32_t xh, xl;
    #pragma unroll
                    for (int j = 0; j < kRun; ++j) {
                        uint32_t zh, zl;
                        split2_f16(x, y, xh, xl);
                        split2_f16(zs[j], 1.f, zh, zl);   // column 131 = 1: its weight row is the (scaled) bias
                        const uint32_t off = a_chunk_offset(j * kCols + c, kLevels);
                        *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(xh, zh, 0u, 0u);
                        *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(xl, zl, 0u, 0u);
                    }
                }
                fence_proxy_async();
                mbar_arrive(a_ready(g));
            }
        } else {
            // ---- plain rows: thread = (row mod 16, k-chunk); its 8 rows are 16 apart, so a load instruction of the 16
            //      lanes of a chunk covers 16 consecutive rows
            uint32_t q = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
                const int64_t r0 = tile * kTileM + c;
                float v[kRun][8];
                float xyz[kRun][3];
#pragma unroll
                for (int j = 0; j < kRun; ++j) {
                    const int64_t gr = r0 + j * kCols;
                    const bool ok = gr < P.m;
                    if constexpr (SRC == kSrcPlanes) {
                        const float* src = P.planes + (int64_t)lvl * P.chunk_stride + gr * 8;
                        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                        const float4 a = ok ? na::ld_stream_f4(src) : z4, b = ok ? na::ld_stream_f4(src + 4) : z4;
                        v[j][0] = a.x, v[j][1] = a.y, v[j][2] = a.z, v[j][3] = a.w;
                        v[j][4] = b.x, v[j][5] = b.y, v[j][6] = b.z, v[j][7] = b.w;
                        if (xyz_writer) {
                            const float4 t = ok ? na::ld_stream_f4(P.planes + (int64_t)kLevels * P.chunk_stride + gr * 8) : z4;
                            xyz[j][0] = t.x, xyz[j][1] = t.y, xyz[j][2] = t.z;
                        }
                    } else {
                        const float* src = P.x + gr * P.ldx;
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[j][e] = ok ? __ldg(src + 3 + 8 * lvl + e) : 0.f;
                        if (xyz_writer) {
#pragma unroll
                            for (int e = 0; e < 3; ++e) xyz[j][e] = ok ? __ldg(src + e) : 0.f;
                        }
                    }
                }
                uint32_t hi[kRun][4], lo[kRun][4];
#pragma unroll
                for (int j = 0; j < kRun; ++j)
#pragma unroll
                    for (int e = 0; e < 4; ++e) split2_f16(v[j][2 * e], v[j][2 * e + 1], hi[j][e], lo[j][e]);
                if (q > 0) mbar_wait(a_free(g), (q - 1) & 1);
#pragma unroll
                for (int j = 0; j < kRun; ++j) {
                    const uint32_t off = a_chunk_offset(j * kCols + c, lvl);
                    *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[j][0], hi[j][1], hi[j][2], hi[j][3]);
                    *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[j][0], lo[j][1], lo[j][2], lo[j][3]);
                }
                if (xyz_writer) {
#pragma unroll
                    for (int j = 0; j < kRun; ++j) {
                        uint32_t xh, xl, zh, zl;
                        split2_f16(xyz[j][0], xyz[j][1], xh, xl);
                        split2_f16(xyz[j][2], 1.f, zh, zl);
                        const uint32_t off = a_chunk_offset(j * kCols + c, kLevels);
                        *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(xh, zh, 0u, 0u);
                        *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(xl, zl, 0u, 0u);
                    }
                }
                fence_proxy_async();
                mbar_arrive(a_ready(g));
            }
        }
    } else {
        regs_dec<kRegsEpi>();
        // ===================== epilogue warps =====================
        const int ew = warp - 8;
        const int quarter = ew & 3, group = ew >> 2;      // TMEM lane quarter, column half of a unit
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        const float b_head = __ldg(P.b_head);
        float dot = 0.f;
        auto epilogue = [&](int h) {
            const uint32_t d_main = tmem_base + 256u * h + lane_addr, d_cross = d_main + 128u;
            const int col_begin = group * (kNh / 2);
            for (int c0 = col_begin; c0 < col_begin + kNh / 2; c0 += 16) {
                if (DEV_MODE(2)) continue;
                uint32_t rm[16], rc[16];
                tc_ld16(d_main + c0, rm);
                tc_ld16(d_cross + c0, rc);
                tc_wait_ld();
                const int gc = h * kNh + c0;
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                    const float4 h4 = *reinterpret_cast<const float4*>(s_head + gc + j);
                    const float hh[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float t = fmaf(__uint_as_float(rc[j + e]), kLoInv, __uint_as_float(rm[j + e]));
                        dot = fmaf(softplus2(t), hh[e], dot);
                    }
                }
            }
        };
        uint32_t q = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++q) {
            dot = 0.f;
            for (int h = 0; h < 2; ++h) {
                mbar_wait(d_full(h), q & 1);
                tc_fence_after();
                epilogue(h);
                tc_fence_before();
                mbar_arrive(d_empty(h));
            }
            float* part = head_partial + (q & 1) * 2 * kTileM;
            part[group * kTileM + row] = dot;
            // group 1 only announces its partial sums; group 0 waits for them and writes the SDF.  A group is at most one
            // tile ahead of the other (d_empty needs every epilogue thread), so buffer and named barrier alternate.
            const uint32_t bar_id = 1u + (q & 1u);
            if (group != 0) {
                asm volatile("bar.arrive %0, %1;" ::"r"(bar_id), "n"(kEpiThreads) : "memory");
            } else {
                asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(kEpiThreads) : "memory");
                if constexpr (SRC == kSrcLattice) {
                    // A-image row r = zi * 16 + column; thread t writes column t >> 3, z t & 7: 8 consecutive floats per column
                    const int oc = row >> 3, oz = row & 7, r = oz * kCols + oc;
                    const int zr = (int)(tile % n_zr);
                    const int64_t col = (tile / n_zr) * kCols + oc;
                    const int z = zr * kRun + oz;
                    if (col < n_cols && z < P.nz) P.sdf_out[col * P.nz + z] = P.out_scale * (b_head + part[r] + part[kTileM + r]);
                } else {
                    const int64_t gr = tile * kTileM + row;
                    if (gr < P.m) P.sdf_out[gr] = P.out_scale * (b_head + part[row] + part[kTileM + row]);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace

namespace {
// once per device and instantiation: shared-memory opt-in, and the register plan checked against the launch allocation
// (setmaxnreg.inc waits for registers that must exist inside the CTA's launch-time allocation)
template <int SRC>
int configure() {
    static int configured[64] = {0};
    int dev = 0;
    if (int rc = na::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return rc;
    if (dev >= 64 || configured[dev]) return NA_OK;
    cudaFuncAttributes attr;
    if (int rc = na::check_cuda(cudaFuncGetAttributes(&attr, (const void*)lattice_sdf_kernel<SRC>), "cudaFuncGetAttributes"))
        return rc;
    NA_REQUIRE(attr.numRegs * kThreads >= kRegsGather * kGatherThreads + kRegsEpi * kEpiThreads + kRegsMma * 128,
               "lattice_sdf: register plan does not fit (%d registers per thread at launch)", attr.numRegs);
    if (int rc = na::check_cuda(cudaFuncSetAttribute((const void*)lattice_sdf_kernel<SRC>,
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes),
                                "cudaFuncSetAttribute(lattice_sdf_kernel)"))
        return rc;
    configured[dev] = 1;
    return NA_OK;
}
}  // namespace

#ifdef NA_LATTICE_DEV
extern "C" int na_lattice_dev_set(int mode) {
    return na::check_cuda(cudaMemcpyToSymbol(na_lattice_dev_mode, &mode, sizeof(int)), "na_lattice_dev_set");
}
#endif

extern "C" int na_lattice_sdf(const na_hashgrid_desc* desc, const float* table, const float* gx, const float* gy,
                              const float* gz, int nx, int ny, int nz, na_unit_map map, uint32_t active_levels,
                              const void* w0s_hi, const void* w0s_lo, int k0_pad, int n, const float* w_head_s,
                              const float* b_head, float out_scale, float* sdf_out, void* stream) {
    na::clear_error();
    NA_REQUIRE(desc && table && gx && gy && gz && w0s_hi && w0s_lo && w_head_s && b_head && sdf_out,
               "lattice_sdf: null pointer");
    NA_REQUIRE(nx > 0 && ny > 0 && nz > 0, "lattice_sdf: empty lattice");
    NA_REQUIRE(desc->n_levels == kLevels && desc->n_features == 8,
               "lattice_sdf: %u levels x %u features unsupported (16 x 8)", desc->n_levels, desc->n_features);
    NA_REQUIRE(n == kN && k0_pad == kKPad, "lattice_sdf: hidden width %d / k0_pad %d unsupported (256 / 144)", n, k0_pad);
    NA_REQUIRE(active_levels <= desc->n_levels, "lattice_sdf: active_levels > n_levels");
    NA_REQUIRE((uintptr_t)table % 32 == 0 && ((uintptr_t)w0s_hi | (uintptr_t)w0s_lo) % 16 == 0,
               "lattice_sdf: table must be 32-byte, packed weights 16-byte aligned");
    if (int rc = configure<kSrcLattice>()) return rc;
    Params P{};
    P.desc = *desc;
    P.table = table, P.gx = gx, P.gy = gy, P.gz = gz, P.nx = nx, P.ny = ny, P.nz = nz;
    P.map = map, P.active_levels = active_levels;
    P.w_hi = (const __half*)w0s_hi, P.w_lo = (const __half*)w0s_lo, P.w_head = w_head_s, P.b_head = b_head;
    P.out_scale = out_scale, P.sdf_out = sdf_out;
    const int64_t tiles = (((int64_t)nx * ny + kCols - 1) / kCols) * ((nz + kRun - 1) / kRun);
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    lattice_sdf_kernel<kSrcLattice><<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(P);
    NA_CHECK_LAUNCH("lattice_sdf_kernel");
    return NA_OK;
}

// The same pipeline fed from rows that already exist: chunk planes [chunk][row][8] (planes != NULL) or [m, 131] rows
// (x != NULL, xyz first as the reference's cat([points, feats])).  Operands as for na_lattice_sdf.
extern "C" int na_sdf_head_ws(const float* planes, int64_t chunk_stride, const float* x, int64_t ldx, int64_t m,
                              const void* w0s_hi, const void* w0s_lo, int k0_pad, int n, const float* w_head_s,
                              const float* b_head, float out_scale, float* sdf_out, void* stream) {
    na::clear_error();
    NA_REQUIRE(m >= 0, "sdf_head_ws: negative row count");
    if (m == 0) return NA_OK;
    NA_REQUIRE((planes != nullptr) != (x != nullptr), "sdf_head_ws: exactly one of planes / x");
    NA_REQUIRE(w0s_hi && w0s_lo && w_head_s && b_head && sdf_out, "sdf_head_ws: null pointer");
    NA_REQUIRE(n == kN && k0_pad == kKPad, "sdf_head_ws: hidden width %d / k0_pad %d unsupported (256 / 144)", n, k0_pad);
    NA_REQUIRE(((uintptr_t)w0s_hi | (uintptr_t)w0s_lo) % 16 == 0, "sdf_head_ws: packed weights must be 16-byte aligned");
    if (planes)
        NA_REQUIRE((uintptr_t)planes % 16 == 0 && chunk_stride % 4 == 0 && chunk_stride >= m * 8,
                   "sdf_head_ws: planes must be 16-byte aligned with chunk_stride a multiple of 4 and >= 8 m");
    else
        NA_REQUIRE(ldx >= 8 * kLevels + 3, "sdf_head_ws: ldx < 131");
    Params P{};
    P.planes = planes, P.chunk_stride = chunk_stride, P.x = x, P.ldx = ldx, P.m = m;
    P.w_hi = (const __half*)w0s_hi, P.w_lo = (const __half*)w0s_lo, P.w_head = w_head_s, P.b_head = b_head;
    P.out_scale = out_scale, P.sdf_out = sdf_out;
    const int64_t tiles = (m + kTileM - 1) / kTileM;
    const unsigned grid = (unsigned)(tiles < na::kNumSMs ? tiles : na::kNumSMs);
    if (planes) {
        if (int rc = configure<kSrcPlanes>()) return rc;
        lattice_sdf_kernel<kSrcPlanes><<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(P);
    } else {
        if (int rc = configure<kSrcRows>()) return rc;
        lattice_sdf_kernel<kSrcRows><<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(P);
    }
    NA_CHECK_LAUNCH("lattice_sdf_kernel (rows)");
    return NA_OK;
}
