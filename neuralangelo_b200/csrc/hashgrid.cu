// Multi-resolution hash-grid encode, forward / backward / fused finite-difference taps.
//
// Replaces tiny-cuda-nn's kernel_grid / kernel_grid_backward behind
// projects/neuralangelo/utils/modules.py:85 (tcnn is an un-vendored pip dependency of the
// reference, requirements.txt:3; the arithmetic restated here is SURVEY.md section 8c items 1-6).
//
// Design (B200): the problem is a gather/scatter of 32-byte table rows, so it is bound by
// L2/HBM sector traffic, not by math.
//   * level-major launch: blockIdx.y = level is the slow grid dimension, so at any instant the
//     resident CTAs work on one or two levels and a 128 MiB hashed level (fp32, F=8) is the L2
//     working set instead of the whole 1.46 GB table.
//   * F/4 lanes per (point, level): each lane owns 4 features = one 16-byte vector, so a warp-wide
//     table load touches whole 32-byte sectors (two lanes per row) and the output store is a
//     16-byte vector per lane.
//   * taps: the 4/6 finite-difference taps of compute_gradients sit within eps <= 1/4 cell of the
//     centre at every level, so the centre's 8 corner rows are kept in registers and re-used by
//     every tap that falls in the same cell (almost all of them); only cell-crossing taps gather
//     again.  The backward pre-sums the taps' corner weights the same way and issues one 16-byte
//     vector reduction (red.global.add.v4.f32) per corner instead of one per tap.
#include "na_common.cuh"
#include <math.h>
#include <stdlib.h>

namespace {

using na::ldg_f4;

constexpr int kThreads = 256;
constexpr int kMaxTaps = 8;

struct TapParams {
    float off[kMaxTaps][3];
    uint32_t n;
};

struct Level {
    float scale;
    uint32_t res, entries, hashed, pow2mask;
    const float* base;  // table + offset*F
};

template <int F>
__device__ __forceinline__ Level load_level(const na_hashgrid_desc& d, uint32_t l, const float* table) {
    Level L;
    L.scale = d.scale[l];
    L.res = d.resolution[l];
    L.entries = d.entries[l];
    L.hashed = d.hashed[l];
    L.pow2mask = ((L.entries & (L.entries - 1u)) == 0u) ? (L.entries - 1u) : 0u;
    L.base = table + (size_t)d.offset[l] * F;
    return L;
}

// tcnn grid_index(): dense stride walk or the coherent prime hash, then modulo the level size.
__device__ __forceinline__ uint32_t entry_index(const Level& L, uint32_t cx, uint32_t cy, uint32_t cz) {
    uint32_t idx;
    if (L.hashed) {
        idx = cx ^ (cy * 2654435761u) ^ (cz * 805459861u);
        idx = L.pow2mask ? (idx & L.pow2mask) : (idx % L.entries);
    } else {
        idx = cx + cy * L.res + cz * L.res * L.res;
        if (idx >= L.entries) idx %= L.entries;  // only the documented wrap-around at the far faces
    }
    return idx;
}

struct Cell {
    uint32_t gx, gy, gz;
    float wx, wy, wz;
};

// tcnn pos_fract(): pos = fma(scale, x, 0.5); cell = floor(pos); w = pos - cell.
__device__ __forceinline__ Cell locate(const Level& L, float px, float py, float pz) {
    Cell c;
    float fx = fmaf(L.scale, px, 0.5f), fy = fmaf(L.scale, py, 0.5f), fz = fmaf(L.scale, pz, 0.5f);
    float ix = floorf(fx), iy = floorf(fy), iz = floorf(fz);
    c.gx = (uint32_t)(int)ix;
    c.gy = (uint32_t)(int)iy;
    c.gz = (uint32_t)(int)iz;
    c.wx = fx - ix;
    c.wy = fy - iy;
    c.wz = fz - iz;
    return c;
}

__device__ __forceinline__ float corner_weight(const Cell& c, int k) {
    float w = (k & 1) ? c.wx : 1.f - c.wx;
    w *= (k & 2) ? c.wy : 1.f - c.wy;
    w *= (k & 4) ? c.wz : 1.f - c.wz;
    return w;
}

__device__ __forceinline__ void axpy4(float4& a, float w, const float4& v) {
    a.x = fmaf(w, v.x, a.x);
    a.y = fmaf(w, v.y, a.y);
    a.z = fmaf(w, v.z, a.z);
    a.w = fmaf(w, v.w, a.w);
}

// V floats per lane (V = 4 for F >= 4, V = 2 for F = 2); F/V lanes per point.
template <int V>
struct Vec;
template <>
struct Vec<4> {
    static __device__ __forceinline__ float4 load(const float* p) { return ldg_f4(p); }
    template <bool ALIGNED>
    static __device__ __forceinline__ float4 load_stream(const float* p) {
        if (ALIGNED) return na::ld_stream_f4(p);
        return make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3));
    }
    template <bool ALIGNED>
    static __device__ __forceinline__ void store(float* p, float4 v) {
        if (ALIGNED) {
            *reinterpret_cast<float4*>(p) = v;
        } else {
            p[0] = v.x, p[1] = v.y, p[2] = v.z, p[3] = v.w;
        }
    }
    static __device__ __forceinline__ void red(float* p, float4 v) { na::red_add_f4(p, v); }
    static __device__ __forceinline__ float4 load_keep(const float* p, uint64_t pol) { return na::ldg_f4_hint(p, pol); }
    static __device__ __forceinline__ void red_keep(float* p, float4 v, uint64_t pol) { na::red_add_f4_hint(p, v, pol); }
    template <bool ALIGNED>
    static __device__ __forceinline__ void store_stream(float* p, float4 v, uint64_t pol) {
        if (ALIGNED) {
            na::st_f4_hint(p, v, pol);
        } else {
            p[0] = v.x, p[1] = v.y, p[2] = v.z, p[3] = v.w;
        }
    }
    template <bool ALIGNED>
    static __device__ __forceinline__ float4 load_stream_hint(const float* p, uint64_t pol) {
        if (ALIGNED) return na::ldg_f4_stream_hint(p, pol);
        return make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3));
    }
};
template <>
struct Vec<2> {
    static __device__ __forceinline__ float4 load_keep(const float* p, uint64_t) { return load(p); }
    static __device__ __forceinline__ void red_keep(float* p, float4 v, uint64_t) { red(p, v); }
    template <bool ALIGNED>
    static __device__ __forceinline__ void store_stream(float* p, float4 v, uint64_t) { store<ALIGNED>(p, v); }
    template <bool ALIGNED>
    static __device__ __forceinline__ float4 load_stream_hint(const float* p, uint64_t) { return load_stream<ALIGNED>(p); }
    static __device__ __forceinline__ float4 load(const float* p) {
        float2 t = __ldg(reinterpret_cast<const float2*>(p));
        return make_float4(t.x, t.y, 0.f, 0.f);
    }
    template <bool ALIGNED>
    static __device__ __forceinline__ float4 load_stream(const float* p) {
        return make_float4(__ldg(p), __ldg(p + 1), 0.f, 0.f);
    }
    template <bool ALIGNED>
    static __device__ __forceinline__ void store(float* p, float4 v) {
        if (ALIGNED) {
            *reinterpret_cast<float2*>(p) = make_float2(v.x, v.y);
        } else {
            p[0] = v.x, p[1] = v.y;
        }
    }
    static __device__ __forceinline__ void red(float* p, float4 v) { na::red_add_f2(p, make_float2(v.x, v.y)); }
};

template <int F>
struct Shape {
    static constexpr int V = F >= 4 ? 4 : 2;
    static constexpr int LPP = F / V;            // lanes per point
};

// ---------------------------------------------------------------------------------------------
// Thread mapping (both directions): one warp walks whole feature rows.  Lane -> (point-in-warp,
// level, 16-byte feature slice); for the default 16 levels x 8 features a warp is exactly one
// point: lanes (2l, 2l+1) own level l, so every output / dout access of the warp is one
// contiguous 512-byte row (full 128-byte lines to HBM instead of lone 32-byte sectors) and the
// point's coordinates are a broadcast load.  A CTA owns a contiguous run of points and its warps
// interleave over it, so spatially coherent inputs (samples along a ray) re-use table rows in L1.
// ---------------------------------------------------------------------------------------------
struct Mapping {
    int lanes_per_level;   // F / V
    int levels_per_pass;   // levels handled by one warp pass
    int points_per_warp;   // >1 only when n_levels * lanes_per_level <= 16
    int passes;            // ceil(n_levels / levels_per_pass)
    int points_per_cta;
};

struct LevelS {
    float scale;
    uint32_t res, entries, hashed, pow2mask, offset;
};

__device__ __forceinline__ void stage_levels(const na_hashgrid_desc& desc, LevelS* sl) {
    if (threadIdx.x < desc.n_levels) {
        const uint32_t l = threadIdx.x;
        LevelS v;
        v.scale = desc.scale[l];
        v.res = desc.resolution[l];
        v.entries = desc.entries[l];
        v.hashed = desc.hashed[l];
        v.pow2mask = ((v.entries & (v.entries - 1u)) == 0u) ? (v.entries - 1u) : 0u;
        v.offset = desc.offset[l];
        sl[l] = v;
    }
    __syncthreads();
}

template <int F>
__device__ __forceinline__ Level make_level(const LevelS& v, const float* table) {
    Level L;
    L.scale = v.scale;
    L.res = v.res;
    L.entries = v.entries;
    L.hashed = v.hashed;
    L.pow2mask = v.pow2mask;
    L.base = table + (size_t)v.offset * F;
    return L;
}

__device__ __forceinline__ Cell locate_tap(const Level& L, const na_unit_map& map, const TapParams& taps, uint32_t t,
                                           float x0, float x1, float x2) {
    return locate(L, (x0 + taps.off[t][0] - map.vol_min) * map.inv_extent,
                  (x1 + taps.off[t][1] - map.vol_min) * map.inv_extent,
                  (x2 + taps.off[t][2] - map.vol_min) * map.inv_extent);
}

__device__ __forceinline__ bool same_cell(const Cell& a, const Cell& b) {
    return a.gx == b.gx && a.gy == b.gy && a.gz == b.gz;
}

template <int F, bool ALIGNED>
__global__ void __launch_bounds__(kThreads, 3)
hashgrid_taps_fwd_kernel(const __grid_constant__ na_hashgrid_desc desc, const float* __restrict__ table,
                         const float* __restrict__ x, int64_t n_points, na_unit_map map, uint32_t active_levels,
                         const __grid_constant__ TapParams taps, float* __restrict__ out, int64_t out_stride,
                         int64_t out_tap_stride, int64_t out_level_stride, const Mapping m) {
    using S = Shape<F>;
    using VT = Vec<S::V>;
    __shared__ LevelS sl[NA_MAX_LEVELS];
    stage_levels(desc, sl);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % S::LPP;
    const int lvl_in_pass = (lane / S::LPP) % m.levels_per_pass;
    const int pt_in_warp = lane / (S::LPP * m.levels_per_pass);
    const uint32_t level = blockIdx.y * m.levels_per_pass + lvl_in_pass;
    if (pt_in_warp >= m.points_per_warp || level >= desc.n_levels) return;
    const bool active = level < active_levels;
    const Level L = make_level<F>(sl[level], table);
    const float* tb = L.base + sub * S::V;
    const uint64_t keep = na::policy_evict_last(), stream = na::policy_evict_first();
    const int64_t cta_base = (int64_t)blockIdx.x * m.points_per_cta;
    const int step = (kThreads / 32) * m.points_per_warp;

    for (int i = warp * m.points_per_warp + pt_in_warp; i < m.points_per_cta; i += step) {
        const int64_t n = cta_base + i;
        if (n >= n_points) break;
        float* o = out + n * out_stride + level * out_level_stride + sub * S::V;
        if (!active) {  // coarse-to-fine mask: inactive levels are exactly zero
            for (uint32_t t = 0; t < taps.n; ++t)
                VT::template store_stream<ALIGNED>(o + t * out_tap_stride, make_float4(0.f, 0.f, 0.f, 0.f), stream);
            continue;
        }
        const float x0 = __ldg(x + n * 3 + 0), x1 = __ldg(x + n * 3 + 1), x2 = __ldg(x + n * 3 + 2);
        // tap 0: gather the 8 corner rows of its cell and keep them in registers.
        const Cell c0 = locate_tap(L, map, taps, 0, x0, x1, x2);
        float4 corner[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t idx = entry_index(L, c0.gx + (k & 1), c0.gy + ((k >> 1) & 1), c0.gz + ((k >> 2) & 1));
            corner[k] = VT::load_keep(tb + (size_t)idx * F, keep);
        }
        {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 8; ++k) axpy4(acc, corner_weight(c0, k), corner[k]);
            VT::template store_stream<ALIGNED>(o, acc, stream);
        }
        // taps that stay in tap 0's cell re-use the cached rows; the others are remembered in `pending`.
        uint32_t pending = 0;
#pragma unroll 1
        for (uint32_t t = 1; t < taps.n; ++t) {
            const Cell c = locate_tap(L, map, taps, t, x0, x1, x2);
            if (same_cell(c, c0)) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < 8; ++k) axpy4(acc, corner_weight(c, k), corner[k]);
                VT::template store_stream<ALIGNED>(o + t * out_tap_stride, acc, stream);
            } else {
                pending |= 1u << t;
            }
        }
        // cell-crossing taps: every lane works through its own list, so the warp needs
        // max(#crossing taps per lane) gather rounds rather than one per tap index.
        while (pending) {
            const uint32_t t = __ffs(pending) - 1;
            pending &= pending - 1;
            const Cell c = locate_tap(L, map, taps, t, x0, x1, x2);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            float4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t idx = entry_index(L, c.gx + (k & 1), c.gy + ((k >> 1) & 1), c.gz + ((k >> 2) & 1));
                v[k] = VT::load_keep(tb + (size_t)idx * F, keep);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) axpy4(acc, corner_weight(c, k), v[k]);
            VT::template store_stream<ALIGNED>(o + t * out_tap_stride, acc, stream);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// backward: scatter-add into the table gradient (same mapping; grid.y covers active levels only)
// ---------------------------------------------------------------------------------------------
template <int F, bool ALIGNED>
__global__ void __launch_bounds__(kThreads, 2)
hashgrid_taps_bwd_kernel(const __grid_constant__ na_hashgrid_desc desc, const float* __restrict__ x,
                         int64_t n_points, na_unit_map map, uint32_t active_levels,
                         const __grid_constant__ TapParams taps, const float* __restrict__ dout,
                         int64_t dout_stride, int64_t dout_tap_stride, int64_t dout_level_stride,
                         float* __restrict__ dtable, const Mapping m) {
    using S = Shape<F>;
    using VT = Vec<S::V>;
    __shared__ LevelS sl[NA_MAX_LEVELS];
    stage_levels(desc, sl);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % S::LPP;
    const int lvl_in_pass = (lane / S::LPP) % m.levels_per_pass;
    const int pt_in_warp = lane / (S::LPP * m.levels_per_pass);
    const uint32_t level = blockIdx.y * m.levels_per_pass + lvl_in_pass;
    if (pt_in_warp >= m.points_per_warp || level >= active_levels) return;
    const Level L = make_level<F>(sl[level], dtable);
    float* tb = const_cast<float*>(L.base) + sub * S::V;
    const uint64_t keep = na::policy_evict_last(), stream = na::policy_evict_first();
    const int64_t cta_base = (int64_t)blockIdx.x * m.points_per_cta;
    const int step = (kThreads / 32) * m.points_per_warp;

    // Software pipeline over the CTA's points: the coordinates and all taps' upstream gradients of point i+1 are
    // requested before point i is processed, so two points' worth of loads are in flight per lane (the kernel is
    // bound by memory latency at two CTAs per SM, not by bytes: profiles/r01_hashgrid_planes_ncu_summary.md).
    float nx0 = 0.f, nx1 = 0.f, nx2 = 0.f;
    float4 ndv[kMaxTaps];
    auto prefetch = [&](int i) {
        const int64_t n = cta_base + i;
        const bool ok = i < m.points_per_cta && n < n_points;
#pragma unroll
        for (int t = 0; t < kMaxTaps; ++t) ndv[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) {
            nx0 = __ldg(x + n * 3 + 0), nx1 = __ldg(x + n * 3 + 1), nx2 = __ldg(x + n * 3 + 2);
            const float* gn = dout + n * dout_stride + level * dout_level_stride + sub * S::V;
#pragma unroll
            for (int t = 0; t < kMaxTaps; ++t)
                if ((uint32_t)t < taps.n) ndv[t] = VT::template load_stream_hint<ALIGNED>(gn + t * dout_tap_stride, stream);
        }
    };
    const int i0 = warp * m.points_per_warp + pt_in_warp;
    prefetch(i0);
    for (int i = i0; i < m.points_per_cta; i += step) {
        const int64_t n = cta_base + i;
        if (n >= n_points) break;
        const float* g = dout + n * dout_stride + level * dout_level_stride + sub * S::V;
        const float x0 = nx0, x1 = nx1, x2 = nx2;
        float4 dv[kMaxTaps];
#pragma unroll
        for (int t = 0; t < kMaxTaps; ++t) dv[t] = ndv[t];
        prefetch(i + step);
        const Cell c0 = locate_tap(L, map, taps, 0, x0, x1, x2);
        float4 acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float w = corner_weight(c0, k);
            acc[k] = make_float4(w * dv[0].x, w * dv[0].y, w * dv[0].z, w * dv[0].w);
        }
        uint32_t pending = 0;
#pragma unroll
        for (int t = 1; t < kMaxTaps; ++t) {
            if ((uint32_t)t < taps.n) {
                const Cell c = locate_tap(L, map, taps, t, x0, x1, x2);
                if (same_cell(c, c0)) {  // pre-sum: one reduction per corner for all taps of this cell
#pragma unroll
                    for (int k = 0; k < 8; ++k) axpy4(acc[k], corner_weight(c, k), dv[t]);
                } else {
                    pending |= 1u << t;
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t idx = entry_index(L, c0.gx + (k & 1), c0.gy + ((k >> 1) & 1), c0.gz + ((k >> 2) & 1));
            VT::red_keep(tb + (size_t)idx * F, acc[k], keep);
        }
        while (pending) {
            const uint32_t t = __ffs(pending) - 1;
            pending &= pending - 1;
            const Cell c = locate_tap(L, map, taps, t, x0, x1, x2);
            const float4 d = VT::template load_stream_hint<ALIGNED>(g + t * dout_tap_stride, stream);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float w = corner_weight(c, k);
                const uint32_t idx = entry_index(L, c.gx + (k & 1), c.gy + ((k >> 1) & 1), c.gz + ((k >> 2) & 1));
                VT::red_keep(tb + (size_t)idx * F, make_float4(w * d.x, w * d.y, w * d.z, w * d.w), keep);
            }
        }
    }
}

// levels_per_pass trades HBM write/read efficiency of the feature rows (more levels per warp ->
// longer contiguous runs per row) against L2 residency of the table (fewer levels in flight ->
// one 128 MiB hashed level is the working set).  Measured on B200, 2^22 random points, default
// table (profiles/r01_hashgrid_mapping_sweep.md): 1 level/pass is best for a single evaluation,
// 2 levels/pass for the 7-evaluation tap kernel whose traffic is dominated by feature rows.
// Developer knobs of the sweep above.  Compiled in only with -DNA_HG_DEV_KNOBS (scripts/microbench_hashgrid.py
// documents the build); read once, clamped to valid values.  The shipped library uses the measured defaults.
struct DevKnobs {
    int group = 0;   // 0 = default mapping
    int iters = 16;  // points per warp and CTA
};
const DevKnobs& dev_knobs() {
    static const DevKnobs k = [] {
        DevKnobs v;
#ifdef NA_HG_DEV_KNOBS
        if (const char* e = getenv("NA_HG_GROUP")) v.group = atoi(e) > 0 ? atoi(e) : 0;
        if (const char* e = getenv("NA_HG_ITERS")) v.iters = atoi(e) >= 1 ? (atoi(e) > 4096 ? 4096 : atoi(e)) : 16;
#endif
        return v;
    }();
    return k;
}

Mapping make_mapping(const na_hashgrid_desc* desc, uint32_t levels, uint32_t n_taps, bool planes) {
    Mapping m;
    m.lanes_per_level = desc->n_features >= 4 ? desc->n_features / 4 : 1;
    const int max_levels_per_warp = 32 / m.lanes_per_level;
    // level-major feature planes: consecutive points are consecutive in memory, so one level per pass is both
    // the coalesced and the L2-friendly choice
    int cap = (n_taps > 1 && !planes) ? 2 : 1;
    if (dev_knobs().group > 0) cap = dev_knobs().group;
    if (cap > max_levels_per_warp) cap = max_levels_per_warp;
    int lp = 1;  // levels per pass: power of two so that points-per-warp divides the warp
    while (lp < (int)levels && lp < cap) lp *= 2;
    m.levels_per_pass = lp;
    m.points_per_warp = 32 / (lp * m.lanes_per_level);
    m.passes = ((int)levels + lp - 1) / lp;
    m.points_per_cta = 8 * m.points_per_warp * dev_knobs().iters;
    return m;
}

int validate(const na_hashgrid_desc* desc, const void* a, const void* b, int64_t n_points, uint32_t active_levels) {
    NA_REQUIRE(desc != nullptr, "hashgrid: desc is null");
    NA_REQUIRE(desc->n_levels >= 1 && desc->n_levels <= NA_MAX_LEVELS, "hashgrid: n_levels %u out of range [1,%d]",
               desc->n_levels, NA_MAX_LEVELS);
    NA_REQUIRE(desc->n_features == 2 || desc->n_features == 4 || desc->n_features == 8,
               "hashgrid: n_features %u unsupported (2, 4 or 8)", desc->n_features);
    NA_REQUIRE(active_levels <= desc->n_levels, "hashgrid: active_levels %u > n_levels %u", active_levels,
               desc->n_levels);
    NA_REQUIRE(n_points >= 0, "hashgrid: negative n_points");
    NA_REQUIRE(n_points == 0 || (a != nullptr && b != nullptr), "hashgrid: null device pointer");
    for (uint32_t l = 0; l < desc->n_levels; ++l)
        NA_REQUIRE(desc->entries[l] > 0, "hashgrid: level %u has no entries", l);
    return NA_OK;
}

bool vec_aligned(const void* p, int64_t stride, int64_t tap_stride, int64_t level_stride, int n_features) {
    const int v = n_features >= 4 ? 4 : 2;
    return ((uintptr_t)p % (v * sizeof(float)) == 0) && (stride % v == 0) && (tap_stride % v == 0) &&
           (level_stride % v == 0);
}

int fill_taps(TapParams& tp, const float* tap_offsets, uint32_t n_taps) {
    NA_REQUIRE(n_taps >= 1 && n_taps <= (uint32_t)kMaxTaps, "hashgrid: n_taps %u out of range [1,%d]", n_taps,
               kMaxTaps);
    NA_REQUIRE(tap_offsets != nullptr, "hashgrid: tap_offsets is null");
    tp.n = n_taps;
    for (uint32_t t = 0; t < (uint32_t)kMaxTaps; ++t)
        for (int d = 0; d < 3; ++d) tp.off[t][d] = t < n_taps ? tap_offsets[t * 3 + d] : 0.f;
    return NA_OK;
}

template <int F>
int launch_fwd(const na_hashgrid_desc* desc, const float* table, const float* x, int64_t n_points, na_unit_map map,
               uint32_t active_levels, const TapParams& tp, float* out, int64_t out_stride, int64_t out_tap_stride,
               int64_t out_level_stride, cudaStream_t stream) {
    const Mapping m = make_mapping(desc, desc->n_levels, tp.n, out_level_stride != F);
    const int64_t blocks = (n_points + m.points_per_cta - 1) / m.points_per_cta;
    NA_REQUIRE(blocks <= 0x7fffffffLL, "hashgrid: n_points too large for one launch");
    dim3 grid((unsigned)blocks, m.passes);
    if (vec_aligned(out, out_stride, out_tap_stride, out_level_stride, F))
        hashgrid_taps_fwd_kernel<F, true><<<grid, kThreads, 0, stream>>>(
            *desc, table, x, n_points, map, active_levels, tp, out, out_stride, out_tap_stride, out_level_stride, m);
    else
        hashgrid_taps_fwd_kernel<F, false><<<grid, kThreads, 0, stream>>>(
            *desc, table, x, n_points, map, active_levels, tp, out, out_stride, out_tap_stride, out_level_stride, m);
    NA_CHECK_LAUNCH("hashgrid_taps_fwd_kernel");
    return NA_OK;
}

template <int F>
int launch_bwd(const na_hashgrid_desc* desc, const float* x, int64_t n_points, na_unit_map map,
               uint32_t active_levels, const TapParams& tp, const float* dout, int64_t dout_stride,
               int64_t dout_tap_stride, int64_t dout_level_stride, float* dtable, cudaStream_t stream) {
    const Mapping m = make_mapping(desc, active_levels, tp.n, dout_level_stride != F);
    const int64_t blocks = (n_points + m.points_per_cta - 1) / m.points_per_cta;
    NA_REQUIRE(blocks <= 0x7fffffffLL, "hashgrid: n_points too large for one launch");
    dim3 grid((unsigned)blocks, m.passes);
    if (vec_aligned(dout, dout_stride, dout_tap_stride, dout_level_stride, F))
        hashgrid_taps_bwd_kernel<F, true><<<grid, kThreads, 0, stream>>>(
            *desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, m);
    else
        hashgrid_taps_bwd_kernel<F, false><<<grid, kThreads, 0, stream>>>(
            *desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, m);
    NA_CHECK_LAUNCH("hashgrid_taps_bwd_kernel");
    return NA_OK;
}

}  // namespace

extern "C" int na_hashgrid_taps_fwd_strided(const na_hashgrid_desc*, const float*, const float*, int64_t, na_unit_map,
                                            uint32_t, const float*, uint32_t, float*, int64_t, int64_t, int64_t, void*);
extern "C" int na_hashgrid_taps_bwd_strided(const na_hashgrid_desc*, const float*, int64_t, na_unit_map, uint32_t,
                                            const float*, uint32_t, const float*, int64_t, int64_t, int64_t, float*,
                                            void*);

extern "C" int na_hashgrid_make_desc(uint32_t n_levels, uint32_t n_features, uint32_t log2_hashmap_size,
                                     uint32_t base_resolution, double per_level_scale, na_hashgrid_desc* desc) {
    na::clear_error();
    NA_REQUIRE(desc != nullptr, "make_desc: desc is null");
    NA_REQUIRE(n_levels >= 1 && n_levels <= NA_MAX_LEVELS, "make_desc: n_levels %u out of range", n_levels);
    NA_REQUIRE(n_features == 2 || n_features == 4 || n_features == 8, "make_desc: n_features %u unsupported",
               n_features);
    NA_REQUIRE(log2_hashmap_size >= 3 && log2_hashmap_size <= 30, "make_desc: log2_hashmap_size %u out of range",
               log2_hashmap_size);
    NA_REQUIRE(base_resolution >= 1 && per_level_scale >= 1.0, "make_desc: bad base_resolution/per_level_scale");
    *desc = na_hashgrid_desc{};
    desc->n_levels = n_levels;
    desc->n_features = n_features;
    // tcnn stores per_level_scale as float and hands log2(per_level_scale) (float) to the kernel, which
    // evaluates exp2f(level * log2_scale) * base - 1.  Both transcendental steps are done here in double
    // and rounded once to float (= correctly rounded; within 1 ulp of any libm / CUDA exp2f) so that this
    // host table, the kernels and the CPU oracle agree bit for bit.
    const float pls = (float)per_level_scale;
    const float log2_scale = (float)log2((double)pls);
    const uint32_t hashmap = 1u << log2_hashmap_size;
    uint64_t offset = 0;
    for (uint32_t l = 0; l < n_levels; ++l) {
        const float e = (float)exp2((double)((float)l * log2_scale));
        const float scale = e * (float)base_resolution - 1.0f;
        const uint32_t res = (uint32_t)ceilf(scale) + 1u;
        const uint32_t max_params = 0xffffffffu / 2u;
        const double dense = (double)res * (double)res * (double)res;
        uint32_t params = dense > (double)max_params ? max_params : (uint32_t)dense;
        params = (params + 7u) / 8u * 8u;  // next multiple of 8
        const uint32_t entries = params < hashmap ? params : hashmap;
        desc->scale[l] = scale;
        desc->resolution[l] = res;
        desc->entries[l] = entries;
        desc->offset[l] = (uint32_t)offset;
        desc->hashed[l] = dense > (double)entries ? 1u : 0u;
        offset += entries;
        NA_REQUIRE(offset <= 0xffffffffull, "make_desc: table exceeds 2^32 entries");
    }
    desc->total_entries = (uint32_t)offset;
    return NA_OK;
}

extern "C" int na_hashgrid_taps_fwd(const na_hashgrid_desc* desc, const float* table, const float* x,
                                    int64_t n_points, na_unit_map map, uint32_t active_levels,
                                    const float* tap_offsets, uint32_t n_taps, float* out, int64_t out_stride,
                                    int64_t out_tap_stride, void* stream) {
    return na_hashgrid_taps_fwd_strided(desc, table, x, n_points, map, active_levels, tap_offsets, n_taps, out,
                                        out_stride, out_tap_stride, desc ? (int64_t)desc->n_features : 0, stream);
}

extern "C" int na_hashgrid_taps_fwd_strided(const na_hashgrid_desc* desc, const float* table, const float* x,
                                            int64_t n_points, na_unit_map map, uint32_t active_levels,
                                            const float* tap_offsets, uint32_t n_taps, float* out,
                                            int64_t out_stride, int64_t out_tap_stride, int64_t out_level_stride,
                                            void* stream) {
    na::clear_error();
    if (int rc = validate(desc, table, x, n_points, active_levels)) return rc;
    NA_REQUIRE(n_points == 0 || out != nullptr, "hashgrid fwd: out is null");
    NA_REQUIRE(out_level_stride >= (int64_t)desc->n_features && out_stride >= (int64_t)desc->n_features,
               "hashgrid fwd: point / level strides must be >= F");
    TapParams tp;
    if (int rc = fill_taps(tp, tap_offsets, n_taps)) return rc;
    if (n_points == 0) return NA_OK;
    cudaStream_t s = (cudaStream_t)stream;
    switch (desc->n_features) {
        case 8: return launch_fwd<8>(desc, table, x, n_points, map, active_levels, tp, out, out_stride, out_tap_stride, out_level_stride, s);
        case 4: return launch_fwd<4>(desc, table, x, n_points, map, active_levels, tp, out, out_stride, out_tap_stride, out_level_stride, s);
        default: return launch_fwd<2>(desc, table, x, n_points, map, active_levels, tp, out, out_stride, out_tap_stride, out_level_stride, s);
    }
}

extern "C" int na_hashgrid_taps_bwd(const na_hashgrid_desc* desc, const float* x, int64_t n_points, na_unit_map map,
                                    uint32_t active_levels, const float* tap_offsets, uint32_t n_taps,
                                    const float* dout, int64_t dout_stride, int64_t dout_tap_stride, float* dtable,
                                    void* stream) {
    return na_hashgrid_taps_bwd_strided(desc, x, n_points, map, active_levels, tap_offsets, n_taps, dout, dout_stride,
                                        dout_tap_stride, desc ? (int64_t)desc->n_features : 0, dtable, stream);
}

extern "C" int na_hashgrid_taps_bwd_strided(const na_hashgrid_desc* desc, const float* x, int64_t n_points,
                                            na_unit_map map, uint32_t active_levels, const float* tap_offsets,
                                            uint32_t n_taps, const float* dout, int64_t dout_stride,
                                            int64_t dout_tap_stride, int64_t dout_level_stride, float* dtable,
                                            void* stream) {
    na::clear_error();
    if (int rc = validate(desc, dtable, x, n_points, active_levels)) return rc;
    NA_REQUIRE(n_points == 0 || dout != nullptr, "hashgrid bwd: dout is null");
    NA_REQUIRE(dout_level_stride >= (int64_t)desc->n_features && dout_stride >= (int64_t)desc->n_features,
               "hashgrid bwd: point / level strides must be >= F");
    NA_REQUIRE((uintptr_t)dtable % 16 == 0, "hashgrid bwd: dtable must be 16-byte aligned");
    TapParams tp;
    if (int rc = fill_taps(tp, tap_offsets, n_taps)) return rc;
    if (n_points == 0 || active_levels == 0) return NA_OK;
    cudaStream_t s = (cudaStream_t)stream;
    switch (desc->n_features) {
        case 8: return launch_bwd<8>(desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, s);
        case 4: return launch_bwd<4>(desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, s);
        default: return launch_bwd<2>(desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, s);
    }
}

extern "C" int na_hashgrid_fwd(const na_hashgrid_desc* desc, const float* table, const float* x, int64_t n_points,
                               na_unit_map map, uint32_t active_levels, float* out, int64_t out_stride,
                               void* stream) {
    const float zero[3] = {0.f, 0.f, 0.f};
    return na_hashgrid_taps_fwd(desc, table, x, n_points, map, active_levels, zero, 1, out, out_stride, 0, stream);
}

extern "C" int na_hashgrid_bwd(const na_hashgrid_desc* desc, const float* x, int64_t n_points, na_unit_map map,
                               uint32_t active_levels, const float* dout, int64_t dout_stride, float* dtable,
                               void* stream) {
    const float zero[3] = {0.f, 0.f, 0.f};
    return na_hashgrid_taps_bwd(desc, x, n_points, map, active_levels, zero, 1, dout, dout_stride, 0, dtable, stream);
}
