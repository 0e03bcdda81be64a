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
#include <type_traits>

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

// The modulo proper lives out of line: a 32-bit division is ~25 instructions, entry_index() is inlined once per table
// row of the unrolled kernels (104 times in the 7-tap backward), and it is taken only by non-power-of-two levels at the
// wrap-around of the far faces.  Inline, those divisions were most of the kernels' code (the backward's 6 320
// instructions = 100 KB of SASS streamed through the instruction caches by every warp).
__device__ __noinline__ uint32_t wrap_index(uint32_t idx, uint32_t entries) { return idx % entries; }

// tcnn grid_index(): dense stride walk or the coherent prime hash, then modulo the level size.
__device__ __forceinline__ uint32_t entry_index(const Level& L, uint32_t cx, uint32_t cy, uint32_t cz) {
    uint32_t idx = L.hashed ? (cx ^ (cy * 2654435761u) ^ (cz * 805459861u)) : (cx + cy * L.res + cz * L.res * L.res);
    if (L.pow2mask) idx &= L.pow2mask;
    else if (idx >= L.entries) idx = wrap_index(idx, L.entries);
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
                         float* __restrict__ dtable, const Mapping m, const uint32_t level_base) {
    using S = Shape<F>;
    using VT = Vec<S::V>;
    __shared__ LevelS sl[NA_MAX_LEVELS];
    stage_levels(desc, sl);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % S::LPP;
    const int lvl_in_pass = (lane / S::LPP) % m.levels_per_pass;
    const int pt_in_warp = lane / (S::LPP * m.levels_per_pass);
    const uint32_t level = level_base + blockIdx.y * m.levels_per_pass + lvl_in_pass;
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

// =============================================================================================
// Axis-factorised kernels (8 features per level; one evaluation, or centre + six axis-aligned taps in the order
// +x -x +y -y +z -z that NeuralSDF.compute_gradients uses, modules.py:135-143).
//
// The kernels above spend their time issuing instructions, not moving bytes (4.2 G warp instructions per 7-tap
// call: two lanes per point each recompute seven cells and 56 corner weights).  Here the trilinear weights W(k) = X[kx] Y[ky] Z[kz] are used in factorised form:
// an axis tap changes a single factor, so with P_x[s] = sum_{ky,kz} Y[ky] Z[kz] row(s,ky,kz) (s = slot along x;
// P_y, P_z alike) every x-tap that stays in the cell is X'[0] P_x[0] + X'[1] P_x[1]: 16 FMAs instead of 64 + the
// weights.  A tap that crosses into the neighbouring cell along its axis needs ONE more slot (s = 2 or s = -1):
// four new rows, the other four are the shared face -- which is also what halves the extra reductions of the
// backward on fine levels.  Taps further away (epsilon larger than a cell: the start of training on fine levels)
// fall back to a full eight-corner evaluation.  The backward is the exact transpose: per-axis slot sums of the
// upstream gradients, then one pair of 16-byte reductions per row.
// =============================================================================================
constexpr int kAxisPoints = 1024;   // points per CTA and level (8 per lane pair)

struct AxisTaps {
    float off0[3];   // offset of tap 0
    float d[6];      // signed offset of tap t = 1..6 along axis (t - 1) / 2
};

// Two lanes per point, each with one 16-byte half of the level's 8 features: the pair's loads / stores / reductions
// of a row are then ONE full 32-byte sector request (a single lane issuing both halves in two instructions doubles
// the L2 request count, which is what bounds the scattered accesses -- measured, profiles/r02_hashgrid.md).
using F8 = float4;
__device__ __forceinline__ F8 f8_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ void f8_axpy(F8& a, float w, const F8& v) { axpy4(a, w, v); }
__device__ __forceinline__ F8 f8_scale(float w, const F8& v) { return make_float4(w * v.x, w * v.y, w * v.z, w * v.w); }
__device__ __forceinline__ F8 f8_load_keep(const float* p, uint64_t pol) { return na::ldg_f4_hint(p, pol); }
__device__ __forceinline__ F8 f8_load_stream(const float* p, uint64_t pol) { return na::ldg_f4_stream_hint(p, pol); }
__device__ __forceinline__ void f8_store_stream(float* p, const F8& v, uint64_t pol) { na::st_f4_hint(p, v, pol); }
__device__ __forceinline__ void f8_red(float* p, const F8& v, uint64_t pol) { na::red_add_f4_hint(p, v, pol); }
// predicated in PTX (no branch: the eight corner reductions of a point stay one straight-line sequence)
__device__ __forceinline__ void f8_red_if(float* p, const F8& v, uint64_t pol, bool on) {
    asm volatile(
        "{\n.reg .pred q;\nsetp.ne.b32 q, %6, 0;\n@q red.relaxed.gpu.global.add.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;\n}" ::"l"(p),
        "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol), "r"((uint32_t)on)
        : "memory");
}

struct AxisCell {
    uint32_t c[3];   // cell of tap 0
    int ci[3];       // the same as signed integers (for the cell distance of a tap)
    float W[3][2];   // {1 - w, w} per axis
};

__device__ __forceinline__ AxisCell axis_locate(const Level& L, const na_unit_map& map, const AxisTaps& taps,
                                                const float (&xs)[3]) {
    AxisCell A;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(L.scale, (xs[d] + taps.off0[d] - map.vol_min) * map.inv_extent, 0.5f);
        const float fl = floorf(p);
        A.ci[d] = (int)fl;
        A.c[d] = (uint32_t)A.ci[d];
        const float w = p - fl;
        A.W[d][0] = 1.f - w, A.W[d][1] = w;
    }
    return A;
}

// position of tap (axis a, signed offset d) along its axis: cell distance from tap 0's cell and weight in its own cell
__device__ __forceinline__ void axis_tap(const Level& L, const na_unit_map& map, float xa, float d, int ci, int& delta,
                                         float& w, uint32_t& cell) {
    const float p = fmaf(L.scale, (xa + d - map.vol_min) * map.inv_extent, 0.5f);
    const float fl = floorf(p);
    const int c = (int)fl;
    delta = c - ci;
    cell = (uint32_t)c;
    w = p - fl;
}

// row index for cell offsets (along axis A: sa; the two other axes B, C: sb, sc) from tap 0's cell
template <int A>
__device__ __forceinline__ uint32_t axis_row(const Level& L, const AxisCell& C, uint32_t ca, int sb, int sc) {
    constexpr int B = (A + 1) % 3, Cc = (A + 2) % 3;
    uint32_t cc[3];
    cc[A] = ca;
    cc[B] = C.c[B] + (uint32_t)sb;
    cc[Cc] = C.c[Cc] + (uint32_t)sc;
    return entry_index(L, cc[0], cc[1], cc[2]);
}

// ---- forward: ONE lane per point and level, whole 32-byte rows per lane -------------------------------------
// sm_100 has 256-bit global loads and stores (LDG.256 / STG.256): a lane fetches a table row, or writes a tap's 8
// features, as one full-sector request -- no second lane that would have to repeat the cell arithmetic.
struct R8 {
    float v[8];
};
__device__ __forceinline__ R8 r8_zero() {
    R8 r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = 0.f;
    return r;
}
__device__ __forceinline__ void r8_axpy(R8& a, float w, const R8& b) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a.v[i] = fmaf(w, b.v[i], a.v[i]);
}
__device__ __forceinline__ R8 r8_scale(float w, const R8& b) {
    R8 r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = w * b.v[i];
    return r;
}
__device__ __forceinline__ R8 r8_load_keep(const float* p, uint64_t pol) {
    R8 r;
    asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]),
                   "=f"(r.v[7])
                 : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ void r8_store_stream(float* p, const R8& r, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8}, %9;" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]),
                 "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]), "l"(pol)
                 : "memory");
}

// optional point source: the points of a lattice block given by its three axis arrays instead of an [N,3] array
struct LatticeSource {
    const float *gx, *gy, *gz;
    int ny, nz;
    float* xyz_out;       // optional: xyz of every point (4 floats written per point), e.g. the MLP's xyz chunk plane
    int64_t xyz_stride;   // floats between consecutive points in xyz_out
};

template <int NT>
__global__ void __launch_bounds__(kThreads, 2)
hashgrid_axis_fwd_kernel(const __grid_constant__ na_hashgrid_desc desc, const float* __restrict__ table,
                         const float* __restrict__ x, int64_t n_points, na_unit_map map, uint32_t active_levels,
                         const AxisTaps taps, float* __restrict__ out, int64_t out_stride, int64_t out_tap_stride,
                         int64_t out_level_stride, const LatticeSource lat) {
    constexpr int F = 8;
    __shared__ LevelS sl[NA_MAX_LEVELS];
    stage_levels(desc, sl);
    const uint32_t level = blockIdx.y;
    const uint64_t keep = na::policy_evict_last(), stream = na::policy_evict_first();
    const Level L = make_level<F>(sl[level], table);
    // a CTA owns kAxisPoints consecutive points of one level; its 256 lanes stride over them
    const int64_t cta_base = (int64_t)blockIdx.x * kAxisPoints;
#pragma unroll 1
    for (int it = 0; it < kAxisPoints / kThreads; ++it) {
        const int64_t n = cta_base + it * kThreads + threadIdx.x;
        if (n >= n_points) break;
        float* o = out + n * out_stride + level * out_level_stride;
        float xs[3];
        if (lat.gx) {   // point n of an axis-aligned lattice block, x slowest (mesh.py:96-99 meshgrid order): no xyz array
            const int64_t plane = (int64_t)lat.ny * lat.nz;
            const int i = (int)(n / plane), j = (int)((n / lat.nz) % lat.ny), k = (int)(n % lat.nz);
            xs[0] = __ldg(lat.gx + i), xs[1] = __ldg(lat.gy + j), xs[2] = __ldg(lat.gz + k);
            if (level == 0 && lat.xyz_out)
                *reinterpret_cast<float4*>(lat.xyz_out + n * lat.xyz_stride) = make_float4(xs[0], xs[1], xs[2], 0.f);
        } else {
            xs[0] = __ldg(x + n * 3 + 0), xs[1] = __ldg(x + n * 3 + 1), xs[2] = __ldg(x + n * 3 + 2);
        }
        if (level >= active_levels) {   // coarse-to-fine mask: inactive levels are exactly zero
#pragma unroll
            for (int t = 0; t < NT; ++t) r8_store_stream(o + t * out_tap_stride, r8_zero(), stream);
            continue;
        }
        const AxisCell C = axis_locate(L, map, taps, xs);
        // the eight rows of the cell, folded into the per-axis slot sums (two batches of four rows in flight)
        R8 P[3][2];
#pragma unroll
        for (int a = 0; a < 3; ++a) P[a][0] = r8_zero(), P[a][1] = r8_zero();
#pragma unroll
        for (int kz = 0; kz < 2; ++kz) {
            R8 r[2][2];
#pragma unroll
            for (int ky = 0; ky < 2; ++ky)
#pragma unroll
                for (int kx = 0; kx < 2; ++kx) {
                    const uint32_t idx = entry_index(L, C.c[0] + kx, C.c[1] + ky, C.c[2] + kz);
                    r[ky][kx] = r8_load_keep(L.base + (size_t)idx * F, keep);
                }
#pragma unroll
            for (int ky = 0; ky < 2; ++ky)
#pragma unroll
                for (int kx = 0; kx < 2; ++kx) {
                    r8_axpy(P[0][kx], C.W[1][ky] * C.W[2][kz], r[ky][kx]);
                    r8_axpy(P[1][ky], C.W[0][kx] * C.W[2][kz], r[ky][kx]);
                    r8_axpy(P[2][kz], C.W[0][kx] * C.W[1][ky], r[ky][kx]);
                }
        }
        {
            R8 v = r8_scale(C.W[0][0], P[0][0]);
            r8_axpy(v, C.W[0][1], P[0][1]);
            r8_store_stream(o, v, stream);
        }
        if (NT == 1) continue;
        // Taps that stay in the cell are written at once; the others are collected and served one per round so that a
        // warp needs max-over-lanes rounds (a dependent memory round trip each), not one per tap index.
        uint32_t crossing = 0;
#pragma unroll
        for (int t = 1; t < 7; ++t) {
            const int A = (t - 1) >> 1;
            int delta;
            float w;
            uint32_t cell;
            axis_tap(L, map, xs[A], taps.d[t - 1], C.ci[A], delta, w, cell);
            if (delta == 0) {
                R8 v = r8_scale(1.f - w, P[A][0]);
                r8_axpy(v, w, P[A][1]);
                r8_store_stream(o + t * out_tap_stride, v, stream);
            } else {
                crossing |= 1u << t;
            }
        }
        while (crossing) {
            const int t = __ffs(crossing) - 1;
            crossing &= crossing - 1;
            const int a = (t - 1) >> 1;
            const float xa = a == 0 ? xs[0] : (a == 1 ? xs[1] : xs[2]);
            const int cia = a == 0 ? C.ci[0] : (a == 1 ? C.ci[1] : C.ci[2]);
            int delta;
            float w;
            uint32_t cell;
            axis_tap(L, map, xa, taps.d[t - 1], cia, delta, w, cell);
            float wb[2], wc[2];   // weights of the two other axes in cyclic order b = a + 1, c = a + 2
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                wb[k] = a == 0 ? C.W[1][k] : (a == 1 ? C.W[2][k] : C.W[0][k]);
                wc[k] = a == 0 ? C.W[2][k] : (a == 1 ? C.W[0][k] : C.W[1][k]);
            }
            auto face = [&](uint32_t ca) {   // sum over the four rows of the face at coordinate ca along axis a
                R8 r[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint32_t cc[3];
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        const int rel = (d - a + 3) % 3;
                        cc[d] = rel == 0 ? ca : C.c[d] + (uint32_t)(rel == 1 ? (q & 1) : (q >> 1));
                    }
                    r[q] = r8_load_keep(L.base + (size_t)entry_index(L, cc[0], cc[1], cc[2]) * F, keep);
                }
                R8 e = r8_zero();
#pragma unroll
                for (int q = 0; q < 4; ++q) r8_axpy(e, wb[q & 1] * wc[q >> 1], r[q]);
                return e;
            };
            R8 v;
            if (delta == 1 || delta == -1) {   // neighbouring cell: one face is shared with tap 0's cell (slot sums P)
                const bool up = delta > 0;
                const R8 s0 = up ? P[0][1] : P[0][0], s1 = up ? P[1][1] : P[1][0], s2 = up ? P[2][1] : P[2][0];
                const R8 shared = a == 0 ? s0 : (a == 1 ? s1 : s2);
                const R8 e = face(up ? cell + 1u : cell);
                v = r8_scale(up ? 1.f - w : w, shared);
                r8_axpy(v, up ? w : 1.f - w, e);
            } else {                           // further away (epsilon larger than this level's cells): both faces
                const R8 e0 = face(cell), e1 = face(cell + 1u);
                v = r8_scale(1.f - w, e0);
                r8_axpy(v, w, e1);
            }
            r8_store_stream(o + t * out_tap_stride, v, stream);
        }
    }
}

// SPLIT: only rows in [row_lo, row_hi) of the level receive their reductions (the caller covers the level in two passes, so
// that the half being written -- 64 MiB of a 2^22-entry level -- stays in the L2 while 2^22 random points scatter into it).
#ifndef NA_HG_BWD_CTAS
#define NA_HG_BWD_CTAS 3
#endif
template <int NT, bool SPLIT>
__global__ void __launch_bounds__(kThreads, NA_HG_BWD_CTAS)
hashgrid_axis_bwd_kernel(const __grid_constant__ na_hashgrid_desc desc, const float* __restrict__ x, int64_t n_points,
                         na_unit_map map, const AxisTaps taps, const float* __restrict__ dout, int64_t dout_stride,
                         int64_t dout_tap_stride, int64_t dout_level_stride, float* __restrict__ dtable,
                         uint32_t level_base, uint32_t row_lo, uint32_t row_hi, uint32_t interleave) {
    constexpr int F = 8;
    __shared__ LevelS sl[NA_MAX_LEVELS];
    stage_levels(desc, sl);
    const int sub = threadIdx.x & 1;
    // interleave > 0: consecutive CTAs work on `interleave` different levels (the small dense levels, whose few rows are
    // hammered by every point: spreading the resident CTAs over several levels spreads the L2's same-address reductions)
    const uint32_t level = level_base + (interleave ? blockIdx.x % interleave : blockIdx.y);
    const uint32_t point_block = interleave ? blockIdx.x / interleave : blockIdx.x;
    const uint64_t keep = na::policy_evict_last(), stream = na::policy_evict_first();
    const Level L = make_level<F>(sl[level], dtable);
    float* tb = const_cast<float*>(L.base) + sub * 4;
    const int64_t cta_base = (int64_t)point_block * kAxisPoints;
    // (measured and rejected, profiles/r02_hashgrid.md: prefetching the next point's loads, and a fourth resident CTA,
    //  both slow this kernel down -- it is bound by the L2's reduction throughput, more requests in flight only queue)
#pragma unroll 1
    for (int it = 0; it < kAxisPoints / (kThreads / 2); ++it) {
    const int64_t n = cta_base + it * (kThreads / 2) + (threadIdx.x >> 1);
    if (n >= n_points) break;
    const float* g = dout + n * dout_stride + level * dout_level_stride + sub * 4;
    F8 gv[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) gv[t] = f8_load_stream(g + t * dout_tap_stride, stream);
    const float xs[3] = {__ldg(x + n * 3 + 0), __ldg(x + n * 3 + 1), __ldg(x + n * 3 + 2)};
    const AxisCell C = axis_locate(L, map, taps, xs);
    F8 S[3][2];   // per-axis slot sums of the upstream gradients; tap 0 is booked on the x axis
    S[0][0] = f8_scale(C.W[0][0], gv[0]), S[0][1] = f8_scale(C.W[0][1], gv[0]);
#pragma unroll
    for (int a = 1; a < 3; ++a) S[a][0] = f8_zero(), S[a][1] = f8_zero();
    if (NT > 1) {
        // four rows of the face at cell coordinate `ca` along axis A receive e, weighted by the other two axes
        auto red_row = [&](uint32_t idx, const F8& v) {
            if (SPLIT) f8_red_if(tb + (size_t)idx * F, v, keep, idx >= row_lo && idx < row_hi);
            else f8_red(tb + (size_t)idx * F, v, keep);
        };
        // four rows of the face at cell coordinate `ca` along axis A receive e, weighted by the other two axes
        auto face = [&](auto axis_tag, uint32_t ca, const F8& e) {
            constexpr int A = decltype(axis_tag)::value, B = (A + 1) % 3, Cc = (A + 2) % 3;
#pragma unroll
            for (int sc = 0; sc < 2; ++sc)
#pragma unroll
                for (int sb = 0; sb < 2; ++sb)
                    red_row(axis_row<A>(L, C, ca, sb, sc), f8_scale(C.W[B][sb] * C.W[Cc][sc], e));
        };
        // a tap adds (1 - w) g to the face at its cell and w g to the next one; a face that belongs to tap 0's cell is
        // booked on that cell's slot sum, any other one is scattered on its own (ONE face() site per half: the kernel's
        // code size is what the instruction caches see)
        auto axis_taps = [&](auto axis_tag) {
            constexpr int A = decltype(axis_tag)::value;
#pragma unroll
            for (int sgn = 0; sgn < 2; ++sgn) {
                const int t = 1 + 2 * A + sgn;
                int delta;
                float w;
                uint32_t cell;
                axis_tap(L, map, xs[A], taps.d[t - 1], C.ci[A], delta, w, cell);
                const F8& gt = gv[t < NT ? t : 0];
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int rel = delta + hf;
                    const float wh = hf ? w : 1.f - w;
                    if (rel == 0) f8_axpy(S[A][0], wh, gt);
                    else if (rel == 1) f8_axpy(S[A][1], wh, gt);
                    else face(axis_tag, cell + (uint32_t)hf, f8_scale(wh, gt));
                }
            }
        };
        axis_taps(std::integral_constant<int, 0>{});
        axis_taps(std::integral_constant<int, 1>{});
        axis_taps(std::integral_constant<int, 2>{});
    }
#pragma unroll
    for (int kz = 0; kz < 2; ++kz)
#pragma unroll
        for (int ky = 0; ky < 2; ++ky)
#pragma unroll
            for (int kx = 0; kx < 2; ++kx) {
                F8 v = f8_scale(C.W[1][ky] * C.W[2][kz], S[0][kx]);
                if (NT > 1) {
                    f8_axpy(v, C.W[0][kx] * C.W[2][kz], S[1][ky]);
                    f8_axpy(v, C.W[0][kx] * C.W[1][ky], S[2][kz]);
                }
                const uint32_t idx = entry_index(L, C.c[0] + kx, C.c[1] + ky, C.c[2] + kz);
                if (SPLIT) f8_red_if(tb + (size_t)idx * F, v, keep, idx >= row_lo && idx < row_hi);
                else f8_red(tb + (size_t)idx * F, v, keep);
            }
    }
}

// The axis kernels apply when the level has 8 features, every vector access is 16-byte aligned, and the taps are
// one evaluation or exactly {tap 0, +-x, +-x, +-y, +-y, +-z, +-z}: tap t >= 1 differs from tap 0 along axis (t-1)/2 only.
bool axis_taps_from(const TapParams& tp, uint32_t n_features, AxisTaps& at) {
    if (n_features != 8 || (tp.n != 1 && tp.n != 7)) return false;
    for (int d = 0; d < 3; ++d) at.off0[d] = tp.off[0][d];
    for (int t = 1; t < 7; ++t) at.d[t - 1] = 0.f;
    if (tp.n == 7) {
        for (int t = 1; t < 7; ++t) {
            const int a = (t - 1) / 2;
            for (int d = 0; d < 3; ++d)
                if (d != a && tp.off[t][d] != tp.off[0][d]) return false;
            at.d[t - 1] = tp.off[t][a];
        }
        // tap 0 sits on the axes the other taps move along: its own offset must not differ from theirs off-axis
        for (int d = 0; d < 3; ++d)
            if (tp.off[0][d] != 0.f) return false;
    }
    return true;
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
    int generic = 0; // 1 = never take the axis-factorised kernels (A/B measurements)
};
const DevKnobs& dev_knobs() {
    static const DevKnobs k = [] {
        DevKnobs v;
#ifdef NA_HG_DEV_KNOBS
        if (const char* e = getenv("NA_HG_GROUP")) v.group = atoi(e) > 0 ? atoi(e) : 0;
        if (const char* e = getenv("NA_HG_ITERS")) v.iters = atoi(e) >= 1 ? (atoi(e) > 4096 ? 4096 : atoi(e)) : 16;
        if (const char* e = getenv("NA_HG_GENERIC")) v.generic = atoi(e) != 0;
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

// whole 8-float rows at 32-byte alignment (the 256-bit accesses of the one-lane-per-point forward)
bool row_aligned(const void* p, int64_t stride, int64_t tap_stride, int64_t level_stride) {
    return ((uintptr_t)p % 32 == 0) && (stride % 8 == 0) && (tap_stride % 8 == 0) && (level_stride % 8 == 0);
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
    if (F == 8) {
        AxisTaps at;
        if (axis_taps_from(tp, F, at) && row_aligned(out, out_stride, out_tap_stride, out_level_stride) &&
            row_aligned(table, 8, 8, 8) && !dev_knobs().generic) {
            const int64_t nb = (n_points + kAxisPoints - 1) / kAxisPoints;
            NA_REQUIRE(nb <= 0x7fffffffLL, "hashgrid: n_points too large for one launch");
            dim3 g((unsigned)nb, desc->n_levels);
            if (tp.n == 1)
                hashgrid_axis_fwd_kernel<1><<<g, kThreads, 0, stream>>>(*desc, table, x, n_points, map, active_levels, at,
                                                                        out, out_stride, out_tap_stride, out_level_stride,
                                                                        LatticeSource{});
            else
                hashgrid_axis_fwd_kernel<7><<<g, kThreads, 0, stream>>>(*desc, table, x, n_points, map, active_levels, at,
                                                                        out, out_stride, out_tap_stride, out_level_stride,
                                                                        LatticeSource{});
            NA_CHECK_LAUNCH("hashgrid_axis_fwd_kernel");
            return NA_OK;
        }
    }
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
               int64_t dout_tap_stride, int64_t dout_level_stride, float* dtable, cudaStream_t stream,
               uint32_t level_base = 0) {
    if (level_base >= active_levels) return NA_OK;
    if (F == 8) {
        AxisTaps at;
        if (axis_taps_from(tp, F, at) && vec_aligned(dout, dout_stride, dout_tap_stride, dout_level_stride, F) &&
            !dev_knobs().generic) {
            const int64_t nb = (n_points + kAxisPoints - 1) / kAxisPoints;
            NA_REQUIRE(nb <= 0x7fffffffLL, "hashgrid: n_points too large for one launch");
            // Large batches: the finest levels (most cell-crossing taps, 128 MiB of gradient each) are covered in two passes
            // over disjoint halves of their rows; measured on 2^22 points x 7 taps: 7.73 ms unsplit, 7.39 / 7.29 / 7.33 ms
            // with the last 2 / 4 / 6 levels split (profiles/r02_hashgrid.md).
            uint32_t split = active_levels;
            if (n_points >= (1 << 21) && tp.n == 7)
                while (split > level_base && split + 4 > active_levels && desc->hashed[split - 1] &&
                       (uint64_t)desc->entries[split - 1] * F * 4 > (96ull << 20))
                    --split;
#ifdef NA_HG_DEV_KNOBS
            if (const char* es = getenv("NA_HG_SPLIT")) split = (uint32_t)atoi(es);
            if (split < level_base) split = level_base;
            if (split > active_levels) split = active_levels;
#endif
            // Large batches: the leading dense levels (a few MB of gradient each, every row hit hundreds of times) are
            // interleaved across consecutive CTAs instead of processed one after the other: 7.28 -> 7.13 ms at 2^22 x 7
            // with levels 0-5 (4 / 6 / 8 / 12 interleaved levels: 7.14 / 7.13 / 7.16 / 9.75 ms -- hashed levels thrash the L2).
            uint32_t il = 0;
            if (n_points >= (1 << 21))
                while (level_base + il < split && !desc->hashed[level_base + il] && nb * (int64_t)(il + 1) <= 0x7fffffffLL) ++il;
            auto launch = [&](uint32_t l0, uint32_t l1, bool part, uint32_t lo, uint32_t hi, uint32_t inter = 0) {
                if (l1 <= l0) return;
                dim3 g(inter ? (unsigned)(nb * inter) : (unsigned)nb, inter ? 1u : l1 - l0);
                if (tp.n == 1)
                    hashgrid_axis_bwd_kernel<1, false><<<g, kThreads, 0, stream>>>(*desc, x, n_points, map, at, dout, dout_stride,
                                                                                   dout_tap_stride, dout_level_stride, dtable, l0, lo, hi, inter);
                else if (!part)
                    hashgrid_axis_bwd_kernel<7, false><<<g, kThreads, 0, stream>>>(*desc, x, n_points, map, at, dout, dout_stride,
                                                                                   dout_tap_stride, dout_level_stride, dtable, l0, lo, hi, inter);
                else
                    hashgrid_axis_bwd_kernel<7, true><<<g, kThreads, 0, stream>>>(*desc, x, n_points, map, at, dout, dout_stride,
                                                                                  dout_tap_stride, dout_level_stride, dtable, l0, lo, hi, inter);
            };
            if (il > 1) launch(level_base, level_base + il, false, 0u, 0xffffffffu, il);
            else il = 0;
            launch(level_base + il, split, false, 0u, 0xffffffffu);
            for (uint32_t l = split; l < active_levels; ++l) {
                const uint32_t half = (desc->entries[l] + 1) / 2;
                launch(l, l + 1, true, 0u, half);
                launch(l, l + 1, true, half, 0xffffffffu);
            }
            NA_CHECK_LAUNCH("hashgrid_axis_bwd_kernel");
            return NA_OK;
        }
    }
    const Mapping m = make_mapping(desc, active_levels - level_base, tp.n, dout_level_stride != F);
    const int64_t blocks = (n_points + m.points_per_cta - 1) / m.points_per_cta;
    NA_REQUIRE(blocks <= 0x7fffffffLL, "hashgrid: n_points too large for one launch");
    dim3 grid((unsigned)blocks, m.passes);
    if (vec_aligned(dout, dout_stride, dout_tap_stride, dout_level_stride, F))
        hashgrid_taps_bwd_kernel<F, true><<<grid, kThreads, 0, stream>>>(
            *desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, m,
            level_base);
    else
        hashgrid_taps_bwd_kernel<F, false><<<grid, kThreads, 0, stream>>>(
            *desc, x, n_points, map, active_levels, tp, dout, dout_stride, dout_tap_stride, dout_level_stride, dtable, m,
            level_base);
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

// Lattice block query: encode the (nx, ny, nz) points of an axis-aligned lattice given by its three axis arrays (device
// pointers), x slowest -- the blocks of extract_mesh -- without materialising their [N,3] coordinates; optionally the
// coordinates are written as the MLP's xyz chunk (4 floats per point at xyz_out + n * xyz_stride).
extern "C" int na_hashgrid_lattice_fwd_strided(const na_hashgrid_desc* desc, const float* table, const float* gx,
                                               const float* gy, const float* gz, int nx, int ny, int nz, na_unit_map map,
                                               uint32_t active_levels, float* out, int64_t out_stride,
                                               int64_t out_level_stride, float* xyz_out, int64_t xyz_stride,
                                               void* stream) {
    na::clear_error();
    const int64_t n_points = (int64_t)nx * ny * nz;
    if (int rc = validate(desc, table, gx, n_points, active_levels)) return rc;
    NA_REQUIRE(nx > 0 && ny > 0 && nz > 0 && gy != nullptr && gz != nullptr && out != nullptr, "lattice fwd: bad arguments");
    NA_REQUIRE(desc->n_features == 8, "lattice fwd: 8 features per level only");
    NA_REQUIRE(row_aligned(out, out_stride, 8, out_level_stride) && row_aligned(table, 8, 8, 8),
               "lattice fwd: out / table must be 32-byte aligned with strides that are multiples of 8 floats");
    NA_REQUIRE(xyz_out == nullptr || ((uintptr_t)xyz_out % 16 == 0 && xyz_stride % 4 == 0 && xyz_stride >= 4),
               "lattice fwd: xyz_out must be 16-byte aligned with a stride that is a multiple of 4 floats");
    AxisTaps at{};
    LatticeSource lat{gx, gy, gz, ny, nz, xyz_out, xyz_stride};
    const int64_t nb = (n_points + kAxisPoints - 1) / kAxisPoints;
    dim3 g((unsigned)nb, desc->n_levels);
    hashgrid_axis_fwd_kernel<1><<<g, kThreads, 0, (cudaStream_t)stream>>>(*desc, table, nullptr, n_points, map, active_levels, at,
                                                                           out, out_stride, 0, out_level_stride, lat);
    NA_CHECK_LAUNCH("hashgrid_axis_fwd_kernel (lattice)");
    return NA_OK;
}
