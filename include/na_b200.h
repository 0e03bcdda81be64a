/*
 * na_b200.h -- C ABI of libna_b200.so: hand-written sm_100a CUDA kernels for the
 * Neuralangelo per-ray field-evaluation / NeuS-rendering / isosurface hot path.
 *
 * Conventions (all entry points):
 *   - plain C types only: raw DEVICE pointers + explicit sizes; no torch types.
 *   - return 0 on success, a negative NA_ERR_* code otherwise; the message is
 *     retrievable (thread-local) with na_last_error().  Nothing throws or aborts.
 *   - `stream` is a cudaStream_t passed as void*.  Calls are asynchronous on that
 *     stream: no implicit device synchronisation, no allocation (the *_host
 *     convenience entry points, which take HOST buffers, are the documented
 *     exception: they allocate, copy, run and synchronise).
 *   - the caller has made the right device current (cudaSetDevice) beforehand.
 *
 * "Replaces" cites the reference interface (path:line under the reference tree)
 * whose arithmetic the entry point takes over.  The reference has no FFI layer of
 * its own; its native surface is the Python call surface of tiny-cuda-nn and
 * PyMCubes plus eager ATen ops.
 */
#ifndef NA_B200_H
#define NA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NA_OK 0
#define NA_ERR_INVALID (-1)  /* bad argument (null pointer, size, alignment, unsupported width) */
#define NA_ERR_CUDA (-2)     /* a CUDA runtime call or launch failed */
#define NA_ERR_CAPACITY (-3) /* caller-provided output/scratch capacity too small */

#define NA_MAX_LEVELS 32

/* Version of this ABI; bumped when a signature changes. */
int na_abi_version(void);
/* Thread-local message of the last failing call on this thread ("" if none). */
const char* na_last_error(void);

/* ------------------------------------------------------------------------------------------
 * Multi-resolution hash grid (tiny-cuda-nn "HashGrid" encoding, 3-D inputs).
 * Replaces: tcnn.Encoding(3, {otype: HashGrid, ...}) constructed at
 *   projects/neuralangelo/utils/modules.py:42-50 and called at modules.py:85.
 * Table layout = tcnn's: level-major, entry-major, feature-minor, fp32:
 *   table[(offset[l] + idx) * n_features + f].
 * ---------------------------------------------------------------------------------------- */
typedef struct na_hashgrid_desc {
    uint32_t n_levels;
    uint32_t n_features;              /* per level: 2, 4 or 8 */
    uint32_t total_entries;           /* sum of entries[] */
    uint32_t reserved;
    float scale[NA_MAX_LEVELS];       /* grid scale: pos = fma(scale, x, 0.5) */
    uint32_t resolution[NA_MAX_LEVELS]; /* ceil(scale) + 1 */
    uint32_t entries[NA_MAX_LEVELS];  /* "hashmap size" of the level (dense levels: res^3 rounded up to 8) */
    uint32_t offset[NA_MAX_LEVELS];   /* running sum of entries[], in entries */
    uint32_t hashed[NA_MAX_LEVELS];   /* 1 when entries < res^3 (spatial hash), 0 when dense-indexed */
} na_hashgrid_desc;

/* Fill `desc` the way tcnn's GridEncoding constructor does (host side only, no CUDA).
 * Replaces: tcnn grid_scale()/grid_resolution()/offset table [upstream, not vendored];
 * parameters are the config keys built at modules.py:42-49. */
int na_hashgrid_make_desc(uint32_t n_levels, uint32_t n_features, uint32_t log2_hashmap_size,
                          uint32_t base_resolution, double per_level_scale, na_hashgrid_desc* desc);

/* Affine map from caller coordinates to the unit cube, applied per component before
 * the encode: xn = (x - vol_min) * inv_extent.  modules.py:81-82 (vol_min/vol_max).
 * Identity = {0, 1}. */
typedef struct na_unit_map {
    float vol_min;
    float inv_extent;
} na_unit_map;

/* out[n*out_stride + l*F + f] = trilinear interpolation at level l, for l < active_levels;
 * exactly 0 for active_levels <= l < n_levels (the coarse-to-fine mask of
 * modules.py:91-93,109-113 folded in).  x: [n_points,3] fp32 row-major.
 * Replaces: tcnn kernel_grid forward (call site modules.py:85) + mask multiply. */
int na_hashgrid_fwd(const na_hashgrid_desc* desc, const float* table, const float* x,
                    int64_t n_points, na_unit_map map, uint32_t active_levels,
                    float* out, int64_t out_stride, void* stream);

/* dtable[(offset[l]+idx)*F+f] += w_corner * dout[n*dout_stride + l*F + f]  (atomic, fp32;
 * dtable is NOT cleared: it accumulates like autograd's .grad).  Levels >= active_levels
 * receive nothing.  Replaces: tcnn kernel_grid_backward (autograd of modules.py:85). */
int na_hashgrid_bwd(const na_hashgrid_desc* desc, const float* x, int64_t n_points,
                    na_unit_map map, uint32_t active_levels, const float* dout,
                    int64_t dout_stride, float* dtable, void* stream);

/* Fused finite-difference taps: evaluates the encode at x + tap_offsets[t] for t < n_taps
 * (n_taps <= 8; offsets in caller coordinates, added in fp32 BEFORE the unit map exactly as
 * modules.py:138-143,164-167 do), re-using the 8 corner rows of tap 0's cell whenever a tap
 * falls in the same cell.  out[t*out_tap_stride + n*out_stride + l*F + f].
 * Replaces: the 4/6 extra self.sdf() encodes of NeuralSDF.compute_gradients, modules.py:132-175. */
int na_hashgrid_taps_fwd(const na_hashgrid_desc* desc, const float* table, const float* x,
                         int64_t n_points, na_unit_map map, uint32_t active_levels,
                         const float* tap_offsets /* host, [n_taps,3] */, uint32_t n_taps,
                         float* out, int64_t out_stride, int64_t out_tap_stride, void* stream);

int na_hashgrid_taps_bwd(const na_hashgrid_desc* desc, const float* x, int64_t n_points,
                         na_unit_map map, uint32_t active_levels,
                         const float* tap_offsets /* host, [n_taps,3] */, uint32_t n_taps,
                         const float* dout, int64_t dout_stride, int64_t dout_tap_stride,
                         float* dtable, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NA_B200_H */
