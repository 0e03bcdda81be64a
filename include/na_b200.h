/*
 * na_b200.h -- C ABI of libna_b200.so: hand-written sm_100a CUDA kernels for the
 * Neuralangelo per-ray field-evaluation / NeuS-rendering / isosurface hot path.
 *
 * Conventions (all entry points):
 *   - plain C types only: raw DEVICE pointers + explicit sizes; no torch types.
 *   - return 0 on success, a negative NA_ERR_* code otherwise; the message is
 *     retrievable (thread-local) with na_last_error().  Nothing throws or aborts.
 *   - `stream` is a cudaStream_t passed as void*.  Calls are asynchronous on that
 *     stream: no implicit device synchronisation, no allocation.
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

/* Same kernels with a free feature layout: element (tap t, point n, level l, feature f) lives at
 *   t*tap_stride + n*point_stride + l*level_stride + f.
 * The row layout of na_hashgrid_taps_fwd is (point_stride = row stride, level_stride = F); the LEVEL-MAJOR
 * PLANE layout [taps][levels][points][F] is (point_stride = F, level_stride = n_points*F, tap_stride =
 * n_levels*n_points*F): every warp then writes (reads) 512 contiguous bytes instead of sixteen lone 32-byte
 * sectors, and a level's plane is exactly the K-chunk-major operand image the tcgen05 MLP stages. */
int na_hashgrid_taps_fwd_strided(const na_hashgrid_desc* desc, const float* table, const float* x,
                                 int64_t n_points, na_unit_map map, uint32_t active_levels,
                                 const float* tap_offsets, uint32_t n_taps, float* out, int64_t point_stride,
                                 int64_t tap_stride, int64_t level_stride, void* stream);
int na_hashgrid_taps_bwd_strided(const na_hashgrid_desc* desc, const float* x, int64_t n_points,
                                 na_unit_map map, uint32_t active_levels, const float* tap_offsets,
                                 uint32_t n_taps, const float* dout, int64_t point_stride, int64_t tap_stride,
                                 int64_t level_stride, float* dtable, void* stream);

/* Lattice block query: the encode of the nx*ny*nz points (x slowest) of an axis-aligned lattice given by its three
 * axis arrays (device pointers) -- one block of extract_mesh -- without materialising the [N,3] coordinates.  Features
 * go to out[n*point_stride + l*level_stride + f]; when xyz_out is given, (x, y, z, 0) of point n is written at
 * xyz_out + n*xyz_stride (the xyz chunk plane of na_sdf_head_fwd_planes).  8 features per level, 32-byte aligned out.
 * Replaces: LatticeGrid.__getitem__'s meshgrid/stack + the H2D copy + the encode of
 * projects/neuralangelo/utils/mesh.py:32-35,96-99. */
int na_hashgrid_lattice_fwd_strided(const na_hashgrid_desc* desc, const float* table, const float* gx,
                                    const float* gy, const float* gz, int nx, int ny, int nz, na_unit_map map,
                                    uint32_t active_levels, float* out, int64_t point_stride, int64_t level_stride,
                                    float* xyz_out, int64_t xyz_stride, void* stream);

/* The whole block query in ONE kernel: sdf_out[(ix*ny + iy)*nz + iz] = out_scale * SDF head of the one-hidden-layer softplus MLP
 * applied to [xyz | encode(xyz)] of lattice point (gx[ix], gy[iy], gz[iz]).  The encode is gathered along runs of 8 z
 * neighbours (two bilinear face values + one per cell crossing instead of 8 rows per point) straight into the
 * split-fp16 tcgen05 operand in shared memory: no feature planes in HBM.  Envelope: 16 levels x 8 features, hidden
 * width n = 256, k0_pad = 144.  Operands are pre-scaled (PackedSdfMlp.lattice_operands): w0s_hi / w0s_lo = the packed
 * split-fp16 image (layout of na_sdf_head_fwd_planes, feature columns first) of c * [W0 | b0 | 0 ...] with
 * c = beta * log2(e) -- column 131 multiplies a constant-one input, i.e. carries the bias; w_head_s = w_head * ln2 / beta.
 * Replaces: LatticeGrid.__getitem__ + sdf_func(xyz) per block, projects/neuralangelo/utils/mesh.py:32-35,68-100
 * (-> NeuralSDF.sdf, utils/modules.py:72-74). */
int na_lattice_sdf(const na_hashgrid_desc* desc, const float* table, const float* gx, const float* gy,
                   const float* gz, int nx, int ny, int nz, na_unit_map map, uint32_t active_levels,
                   const void* w0s_hi, const void* w0s_lo, int k0_pad, int n, const float* w_head_s,
                   const float* b_head, float out_scale, float* sdf_out, void* stream);

/* The same warp-specialised pipeline (loader warps -> tcgen05 -> epilogue warps, operands pre-scaled as above) for rows
 * that already exist: sdf_out[r] = out_scale * SDF head, r < m, from chunk planes [chunk][row][8] (planes != NULL: 16
 * level chunks + the xyz chunk, chunk_stride floats apart) or from [m, 131] rows with stride ldx (x != NULL, xyz first).
 * Exactly one of planes / x.  Envelope: 16 levels x 8 features + xyz, n = 256, k0_pad = 144.
 * Replaces: NeuralSDF.sdf for the finite-difference taps / sampler evaluations, utils/modules.py:72-74,132-156. */
int na_sdf_head_ws(const float* planes, int64_t chunk_stride, const float* x, int64_t ldx, int64_t m,
                   const void* w0s_hi, const void* w0s_lo, int k0_pad, int n, const float* w_head_s,
                   const float* b_head, float out_scale, float* sdf_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * NeuS volume rendering and hierarchical resampling (one warp per ray).
 * All tensors fp32 contiguous, [n_rays, n_samples(, k)].  `dist_far` is [n_rays] or NULL (= 1e10,
 * the reference's default).  `s_var` is a DEVICE scalar (the learnable log inverse std-dev).
 * ---------------------------------------------------------------------------------------- */

/* weights[i] = alphas[i] * prod_{j<i}(1 - alphas[j]).
 * Replaces: render.alpha_compositing_weights, projects/nerf/utils/render.py:87-99. */
int na_alpha_weights_fwd(const float* alphas, int64_t n_rays, int n_samples, float* weights, void* stream);
/* Adjoint by a reverse affine scan (no division by 1-alpha). */
int na_alpha_weights_bwd(const float* alphas, const float* d_weights, int64_t n_rays, int n_samples,
                         float* d_alphas, void* stream);

/* out[r,c] = sum_i q[r,i,c] * w[r,i].  Replaces: render.composite, render.py:102-112. */
int na_composite_fwd(const float* q, const float* w, int64_t n_rays, int n_samples, int k, float* out,
                     void* stream);
/* d_q and/or d_w may be NULL when not needed. */
int na_composite_bwd(const float* q, const float* w, const float* d_out, int64_t n_rays, int n_samples, int k,
                     float* d_q, float* d_w, void* stream);

/* alpha = 1 - exp(-density * interval).  Replaces: render.volume_rendering_alphas_dist, render.py:68-84. */
int na_density_alphas_fwd(const float* dens, const float* dists, const float* dist_far, int64_t n_rays,
                          int n_samples, float* alphas, void* stream);
int na_density_alphas_bwd(const float* dens, const float* dists, const float* dist_far, const float* d_alphas,
                          int64_t n_rays, int n_samples, float* d_dens, void* stream);

/* NeuS opacity from SDF value + SDF gradient along the ray; anneal_ratio = DEVICE scalar holding
 * min(progress/anneal_end, 1) (device-resident so that a captured CUDA graph follows the schedule).
 * Replaces: Model.compute_neus_alphas + _get_iter_cos, projects/neuralangelo/model.py:275-298. */
int na_neus_alphas_fwd(const float* sdf, const float* grad, const float* dists, const float* ray_unit,
                       const float* dist_far, const float* s_var, const float* anneal_ratio, int64_t n_rays,
                       int n_samples, float* alphas, void* stream);
/* d_s_var (device scalar) is accumulated atomically; d_sdf [R,N], d_grad [R,N,3] are overwritten. */
int na_neus_alphas_bwd(const float* sdf, const float* grad, const float* dists, const float* ray_unit,
                       const float* dist_far, const float* s_var, const float* anneal_ratio, const float* d_alphas,
                       int64_t n_rays, int n_samples, float* d_sdf, float* d_grad, float* d_s_var, void* stream);

/* Inverse-CDF sampling: bins [R,n_bins], weights [R,n_bins-1], unif [n_fine] (device; the u values,
 * normally the midpoints of linspace(0,1,n_fine+1)) -> out [R,n_fine].  2 <= n_bins <= 1024.
 * Replaces: nerf_util.sample_dists_from_pdf, projects/nerf/utils/nerf_util.py:41-68. */
int na_sample_pdf(const float* bins, const float* weights, const float* unif, int64_t n_rays, int n_bins,
                  int n_fine, float* out, void* stream);

/* NeuS hierarchical up-sampling: section opacities from consecutive SDF samples at sharpness inv_s,
 * compositing weights, inverse-CDF sampling.  2 <= n_samples <= 768.
 * Replaces: Model.sample_dists_hierarchical, model.py:250-267. */
int na_neus_upsample(const float* dists, const float* sdfs, const float* unif, float inv_s, int64_t n_rays,
                     int n_samples, int n_fine, float* dists_fine, void* stream);

/* Merge two per-ray sorted rows (and optionally the SDF values riding along).
 * Replaces: torch.cat + sort + gather at model.py:241-247. */
int na_merge_sorted(const float* dists, const float* sdfs, const float* dists_fine, const float* sdfs_fine,
                    int64_t n_rays, int n_samples, int n_fine, float* dists_out, float* sdfs_out, void* stream);

/* Fused renderer: object NeuS alphas ++ background density alphas -> compositing weights -> RGB.
 * rgb [R,3]; weights [R,n_obj+n_bg] and opacity [R] optional (NULL).  n_bg may be 0 (no background).
 * n_obj + n_bg <= 512.  Replaces: model.py:132-141 with the alphas of :168-169 and :198. */
int na_neus_render_fwd(const float* sdf, const float* grad, const float* dists_obj, const float* ray_unit,
                       const float* dist_far, const float* s_var, const float* anneal_ratio, const float* rgb_obj,
                       int n_obj, const float* dens_bg, const float* dists_bg, const float* rgb_bg, int n_bg,
                       int64_t n_rays, float* rgb, float* weights, float* opacity, void* stream);
/* d_weights optional (NULL).  Outputs overwritten except d_s_var (atomic accumulate). */
int na_neus_render_bwd(const float* sdf, const float* grad, const float* dists_obj, const float* ray_unit,
                       const float* dist_far, const float* s_var, const float* anneal_ratio, const float* rgb_obj,
                       int n_obj, const float* dens_bg, const float* dists_bg, const float* rgb_bg, int n_bg,
                       int64_t n_rays, const float* d_rgb, const float* d_weights, float* d_sdf, float* d_grad,
                       float* d_s_var, float* d_rgb_obj, float* d_dens_bg, float* d_rgb_bg, void* stream);

/* ------------------------------------------------------------------------------------------
 * Marching cubes on a device-resident fp32 volume [nx,ny,nz] (C order, x slowest), shared
 * vertices, output identical -- values AND ordering -- to the serial PyMCubes algorithm.
 * Replaces: mcubes.marching_cubes(sdf, 0.) at projects/neuralangelo/utils/mesh.py:121 and the
 * `V * intv + xyz[0,0,0]` that follows it (:123).
 * Two phases because the output size is data dependent:
 *   na_mc_count  classifies the cells and scans; writes counts[0] = #vertices, counts[1] = #triangles
 *                (DEVICE uint32[2]); the caller reads them back and allocates;
 *   na_mc_emit   writes vertices (float64 [nV,3] = index-space position * scale + origin) and
 *                triangles (int32 [nF,3]); entries beyond the given capacities are dropped.
 * `scratch` is caller-owned device memory of na_mc_scratch_bytes() bytes, 256-byte aligned, and
 * must be left untouched between the two calls.
 * ---------------------------------------------------------------------------------------- */
int64_t na_mc_scratch_bytes(int nx, int ny, int nz);
int na_mc_count(const float* volume, int nx, int ny, int nz, double iso, void* scratch, int64_t scratch_size,
                uint32_t* counts, void* stream);
int na_mc_emit(const float* volume, int nx, int ny, int nz, double iso, void* scratch, int64_t scratch_size,
               double scale, const double* origin /* host [3] or NULL */, double* verts, int64_t cap_verts,
               int32_t* faces, int64_t cap_faces, void* stream);

/* Marching cubes for callers that cannot allocate device memory (fixed memory budgets, code running
 * inside another component's allocator, threads that must not block on cudaMalloc).  No counterpart in
 * the reference.  Same result as na_mc_count + na_mc_emit, but the scratch is a fixed array built into
 * the library (about 1.3 MB held on every GPU that loads it, until the process exits), so the volume is
 * limited to na_mc_static_max_cells() = 64^3 cells (e.g. 65^3 samples); larger volumes fail with
 * NA_ERR_CAPACITY.  The caller passes the outputs: verts [cap_verts,3] float64, faces [cap_faces,3] int32
 * (entries beyond the capacities are dropped) and counts (DEVICE uint32[2]: #vertices, #triangles found).
 * Calls no cudaMalloc/cudaFree.  All calls on one GPU share the array: calls that may overlap on the GPU
 * race and must be serialised by the caller.  The first call on each device looks up the array's device
 * address; issue that first call outside of any stream capture. */
int64_t na_mc_static_max_cells(void);
int na_mc_run_static(const float* volume, int nx, int ny, int nz, double iso, double scale,
                     const double* origin /* host [3] or NULL */, double* verts, int64_t cap_verts,
                     int32_t* faces, int64_t cap_faces, uint32_t* counts, void* stream);

/* Weight and bias gradient of a Linear layer, y = x W^T + b, on tcgen05 (csrc/wgrad_tcgen05.cu):
 *   dw[n*ld_w + k] += sum_m dy[m*ld_dy + n] * x[m*ld_x + k]   (k < k_in),      dw[n*ld_w + k_in] += sum_m dy[m*ld_dy + n]
 * i.e. the bias gradient arrives as column k_in of the same buffer.  `dw` must be ZERO-FILLED by the caller (row slices
 * add their partial sums with vector reductions), 16-byte aligned, ld_w >= k_in + 1 and a multiple of 4.  `amax`: device
 * scalar holding max|dy|; dy is multiplied by 2^floor(log2(4096 / amax)) before its conversion to fp16 and the result
 * divided by it again (NULL: every CTA takes the maximum over its own row slice in a first sweep).  Operands of dW in fp16 (10-bit mantissa, as the TF32 GEMMs the reference
 * runs), accumulation in fp32; the bias gradient is an fp32 sum of the unconverted dy.  k_in <= 352, any n_out, any m.
 * Replaces: autograd's weight-gradient GEMM + bias-gradient reduction of every nn.Linear on the path
 * (projects/neuralangelo/utils/mlp.py:55-69, projects/nerf/utils/nerf_util.py:186-196). */
int na_linear_wgrad(const float* dy, int64_t ld_dy, const float* x, int64_t ld_x, int64_t m, int n_out, int k_in,
                    const float* amax, float* dw, int64_t ld_w, void* stream);

/* ------------------------------------------------------------------------------------------
 * Optimizer step.
 * One fused pass of AdamW (decoupled weight decay, bias-corrected, no amsgrad -- torch.optim.AdamW's
 * update rule) over a flat fp32 buffer: p, exp_avg, exp_avg_sq updated in place; grads are multiplied
 * by grad_scale before use (1/world_size when they hold a cross-rank SUM) and, if zero_grad != 0, cleared.
 * `step` is the 1-based step count used for the bias corrections.  Buffers 16-byte aligned.
 * Replaces: torch.optim.AdamW.step + optimizer.zero_grad as wired at
 *   imaginaire/trainers/utils/get_trainer.py:106-150 and imaginaire/trainers/base.py:454-457.
 * ---------------------------------------------------------------------------------------- */
int na_adamw_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int64_t step, float grad_scale, int zero_grad,
                  void* stream);

/* The data-parallel table step as ONE kernel over NVLink multicast (NVSwitch in-fabric reduction): for the slice
 * [offset, offset + n) this rank owns, gradients of all ranks are summed by multimem.ld_reduce through `mc_grads`,
 * AdamW runs in registers (same arithmetic as na_adamw_step, grad_scale = 1 / world), the new parameters are written
 * to every rank with multimem.st through `mc_params`.  Gradients are left as they are: every rank clears its own
 * buffer after the closing barrier (all ranks have read it by then).  `mc_*` are
 * MULTICAST virtual addresses of symmetric allocations (e.g. torch.distributed._symmetric_memory rendezvous:
 * handle.multicast_ptr); params_local / exp_avg / exp_avg_sq are this rank's ordinary device pointers (whole-buffer
 * bases; the kernel adds `offset`).  The caller must separate the launch from the producers of the gradients and from
 * the consumers of the parameters with cross-rank barriers on the stream.  max_ctas <= 0 picks the grid; unroll
 * (2, 4 or 8; 0 = default 4) is the number of multicast loads a thread keeps in flight.
 * Replaces: DDP's bucketed gradient all-reduce (imaginaire/trainers/utils/get_trainer.py:81-88) + AdamW
 * (get_trainer.py:106-150) + zero_grad (imaginaire/trainers/base.py:454-457) for the hash table. */
int na_adamw_step_multimem(float* mc_grads, float* mc_params, const float* params_local, float* exp_avg,
                           float* exp_avg_sq, int64_t offset, int64_t n, float lr, float beta1, float beta2,
                           float eps, float weight_decay, int64_t step, float grad_scale, int max_ctas, int unroll,
                           void* stream);
/* params[i] *= factor: applies the accumulated weight decay of hash-grid levels that were inactive
 * (zero gradient, zero moments: AdamW reduces to p *= 1 - lr*wd per step) when they become active. */
int na_scale_inplace(float* params, int64_t n, float factor, void* stream);

/* Weight-norm reparametrisation of one layer, w[r,:] = g[r] * v[r,:] / ||v[r,:]|| (fp32, row-major [rows, cols]), and
 * its adjoint: dg[r] = <dw[r], v[r]> / ||v[r]||, dv[r] = (g[r] / ||v[r]||) (dw[r] - v[r] dg[r] / ||v[r]||).
 * inv_norm [rows] is written by the forward and read by the backward.  One launch each.
 * Replaces: torch.nn.utils.weight_norm on every MLP layer, projects/neuralangelo/utils/mlp.py:37-38 and
 *   projects/nerf/utils/nerf_util.py:176-177 (three ATen kernels per layer forward, about six backward). */
int na_weight_norm_fwd(const float* v, const float* g, int rows, int cols, float* w, float* inv_norm, void* stream);
int na_weight_norm_bwd(const float* v, const float* g, const float* inv_norm, const float* dw, int rows, int cols,
                       float* dv, float* dg, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused SDF MLP forward on tcgen05 tensor cores (accumulators in TMEM, activations stay on chip):
 *   h0 = softplus_beta(x W0^T + b0);  sdf = h0 . w_head + b_head;  [feat = softplus_beta(h0 W1^T + b1)]
 * x: [m, k0] fp32 rows with stride ldx (e.g. the [N,131] rows of the encode).  n = hidden width
 * (64/128/192/256).  Weights are the EFFECTIVE (weight-norm already applied) matrices split into two fp16
 * pieces and packed K-chunk-major: w_hi/w_lo[(k/8)][(j/8)][j%8][k%8] for W[j,k], K zero-padded to
 * k0_pad (multiple of 16, <= 256) for layer 0 and to n for layer 1; w_lo = (W - float(w_hi)) * 2048.
 * Outputs are optional (NULL): sdf_out [m]; feat_out [m,n] (needs n_layers == 2); hidden_out [m,n] = h0.
 * Replaces: MLPforNeuralSDF.forward, projects/neuralangelo/utils/mlp.py:55-69 (num_layers = 1), behind
 *   NeuralSDF.forward / .sdf, projects/neuralangelo/utils/modules.py:68-74.
 * ---------------------------------------------------------------------------------------- */
int na_sdf_mlp_fwd(const float* x, int64_t ldx, int64_t m, int k0, int n, int n_layers, const void* w0_hi,
                   const void* w0_lo, const float* b0, int k0_pad, const void* w1_hi, const void* w1_lo,
                   const float* b1, const float* w_head, const float* b_head /* device scalar */, float beta, float* sdf_out,
                   float* feat_out,
                   int64_t ld_feat, float* hidden_out, int64_t ld_hidden, void* stream);

/* SDF-only form of na_sdf_mlp_fwd (h0 and the SDF head; no feature layer) reading its input as CHUNK PLANES:
 * column 8 c + e of row r is planes[c * chunk_stride + r * 8 + e] -- for 8 features per level exactly the level-major
 * planes na_hashgrid_taps_fwd_strided writes (chunk = level), followed by one chunk that holds the row's xyz.  The
 * weights are packed for THAT column order.  k0 = 32 j + r (j in 1..4, r <= 4 columns in the last chunk); n = 128 or
 * 256.  One launch of the pipelined kernel: W_hi / W_lo resident in shared memory, two accumulator units in TMEM.
 * Replaces: the same reference code as na_sdf_mlp_fwd with with_feat=False (NeuralSDF.sdf, modules.py:72-74). */
int na_sdf_head_fwd_planes(const float* planes, int64_t chunk_stride, int64_t m, int k0, int n, const void* w0_hi,
                           const void* w0_lo, const float* b0, int k0_pad, const float* w_head,
                           const float* b_head /* device scalar */, float beta, float* sdf_out, float* hidden_out,
                           int64_t ld_hidden, void* stream);

/* Backward of na_sdf_mlp_fwd with respect to the input rows, plus the pre-activation gradients from which
 * the caller forms the weight gradients (dW1 = dz1^T h0, dW0 = dz0^T x, db = column sums):
 *   dz1 = d_feat * s'(z1);  dh0 = dz1 W1 + d_sdf w_head;  dz0 = dh0 * s'(z0);  dx = dz0 W0.
 * two_layer = 0 is the SDF-only form (finite-difference taps): dz0 = d_sdf w_head * s'(z0), dx = dz0 W0.
 * hidden / feat are the activations saved by the forward ([m,n], stride ld_act); row_bound[m] >= max_j |d_feat[r,j]|
 * (used to scale each row into fp16 range); w1t / w0t are W1^T and W0^T (rows zero-padded to k0_pad) packed
 * like the forward weights; norms = DEVICE float[2] {max_j sum_k |W1[k,j]|, max |w_head|} (scaling bounds).
 * Replaces: the autograd backward of MLPforNeuralSDF.forward (projects/neuralangelo/utils/mlp.py:55-69). */
int na_sdf_mlp_bwd(int64_t m, int n, int k0, int k0_pad, int two_layer, const float* d_sdf, const float* d_feat,
                   const float* feat, const float* hidden, int64_t ld_act, const float* row_bound,
                   const void* w1t_hi, const void* w1t_lo, const void* w0t_hi, const void* w0t_lo,
                   const float* w_head, float beta, const float* norms, float* dz1, float* dz0, float* dx,
                   int64_t ld_dx, void* stream);

/* ------------------------------------------------------------------------------------------
 * Generic fused MLP chain on tcgen05: a program of stages run per 128-row tile with all intermediate
 * activations on chip; the same kernel runs the forward of a network and the input-gradient chain of its backward.
 * Stage = [epilogue of the previous product: + bias, activation (0 none, 1 relu, 2 softplus(act_param), 3 sigmoid,
 * 4 relu-mask: no bias, the accumulator -- an upstream gradient -- is passed where the layer's saved activation,
 * READ from `save`, is positive) -> next A operand (epi_to_a) and/or global `out` and/or (acts 0-3) `save`]
 * [load in_cols (<= 256) columns of a global input into the A operand] [product D (+)= A W^T, n_out <= 256,
 * k_chunks*8 <= 256 columns of A; weights packed like na_sdf_mlp_fwd's].  Wide inputs, skip connections,
 * 257-wide layers and chained networks are sequences of such stages (see neuralangelo_b200/fused_chain.py).
 * Replaces: MLPwithSkipConnection.forward, projects/nerf/utils/nerf_util.py:186-196, as used by
 *   NeuralRGB.forward (projects/neuralangelo/utils/modules.py:209-220) and BackgroundNeRF.forward (:270-285),
 *   and autograd's dX = dY W chain through the same layers.
 * ---------------------------------------------------------------------------------------- */
typedef struct na_mlp_stage {
    int32_t has_epi, act, n_valid, epi_to_a;
    float act_param;
    int32_t in_col, in_cols, has_gemm, k_chunks, n_out, accumulate, reserved;
    const float* bias;
    float* out;
    int64_t out_ld;
    float* save;
    int64_t save_ld;
    const float* in;
    int64_t in_ld;
    const void* w_hi;
    const void* w_lo;
} na_mlp_stage;
int na_mlp_chain_fwd(const na_mlp_stage* stages, int n_stages, int64_t m, void* stream);

/* One-launch operand packing (csrc/pack.cu): logical W[j, c] = w[j*ld_w + c] (transposed: w[c*ld_w + j]) for j < n_valid,
 * c < k; rows n_valid..n-1 are zero.  Destination column kd < k takes scale * W[j, (kd + col_rot) mod k], column k takes
 * scale * extra_col[j] when extra_col is given, the rest up to k_pad is zero.  hi = fp16(v), lo = fp16((v - hi) * 2048),
 * both in the K-chunk-major layout of the tensor-core kernels: [(k/8)][(j/8)][j%8][k%8].  n multiple of 8, k_pad of 16.
 * Replaces: the host-side torch packing of the effective weights (weight norm applied) behind mlp.py:55-69. */
int na_pack_split_f16(const float* w, int64_t ld_w, int n, int n_valid, int k, int transposed, int col_rot, double scale,
                      const float* extra_col, int k_pad, void* hi, void* lo, void* stream);

/* ------------------------------------------------------------------------------------------
 * Element-wise glue of the per-ray pipeline, one launch each (csrc/glue.cu).  All tensors fp32 contiguous unless a
 * leading dimension is given; `rows` = rays x samples, `samples_per_ray` maps a row to its ray.
 * ---------------------------------------------------------------------------------------- */

/* out[row] = [xyz(3) | SH_levels(ray_unit of the row's ray) | grad / max(|grad|, 1e-12) (3) | feat(n_feat) | app(n_app)]:
 * the colour network's input.  Replaces: NeuralRGB.forward's encode_view + cat (projects/neuralangelo/utils/modules.py:
 * 209-229, spherical_harmonics.py:47-84) and the F.normalize of projects/neuralangelo/model.py:181. */
int na_rgb_input_fwd(const float* points, const float* ray_unit, const float* grad, const float* feat, int64_t ld_feat,
                     const float* app, int64_t ld_app, int64_t rows, int samples_per_ray, int sh_levels, int n_feat,
                     int n_app, float* out, int64_t ld_out, void* stream);
/* d_grad[row] from the three normal columns (first at `normal_col`) of d_out: the adjoint of the normalisation.  The
 * gradient of feat is the corresponding column block of d_out itself. */
int na_rgb_input_bwd(const float* grad, const float* d_out, int64_t ld_out, int normal_col, int64_t rows, float* d_grad,
                     void* stream);

/* Finite-difference gradient and curvature from tap SDF values taps[T][rows] and the centre sdf[rows]:
 * T = 6 (+x -x +y -y +z -z): grad_a = (s_2a - s_2a+1) / den_grad, hess_a = (s_2a + s_2a+1 - 2 sdf) / den_hess;
 * T = 4 (tetrahedral stencil): grad = sum_i k_i s_i / den_grad, hess = ((sum_i s_i) / 2 - 2 sdf) / den_hess / 3 (x3).
 * hess may be NULL.  Replaces: the tensor arithmetic of NeuralSDF.compute_gradients, modules.py:144-175. */
int na_tap_combine_fwd(const float* taps, const float* sdf, int64_t rows, int n_taps, float den_grad, float den_hess,
                       float* grad, float* hess, void* stream);
int na_tap_combine_bwd(const float* d_grad, const float* d_hess, int64_t rows, int n_taps, float den_grad,
                       float den_hess, float* d_taps, float* d_sdf, void* stream);

/* out2[0] += mean over rows of nan_to_num((|grad| - 1)^2), out2[1] += mean of nan_to_num(|hess_x + hess_y + hess_z|)
 * (hess may be NULL), rows of rays with outside[ray] != 0 contributing zero; out2 zero-filled by the caller.
 * Replaces: eikonal_loss / curvature_loss, projects/neuralangelo/utils/misc.py:74-89. */
int na_sdf_losses_fwd(const float* grad, const float* hess, const uint8_t* outside, int64_t rows, int samples_per_ray,
                      float* out2, void* stream);
int na_sdf_losses_bwd(const float* grad, const float* hess, const uint8_t* outside, int64_t rows, int samples_per_ray,
                      const float* d_out2, float* d_grad, float* d_hess, void* stream);

/* Ray set-up for the sampled pixels only, one launch: pixel ray_idx[b, r] of the row-major H x W grid (pixel centres at
 * +0.5) -> center / ray (depth 1) / ray_unit [n_images, rays_per_image, 3] and, when near / far / outside are given, the
 * unit-sphere bounds with the reference's miss handling (near 1, far 1.2, outside 1).  pose [n_images, 3, 4] world-to-
 * camera, intr [n_images, 3, 3].
 * Replaces: camera.get_center_and_ray + slice_by_ray_idx + normalize + get_dist_bounds,
 * projects/nerf/utils/camera.py:283-311, projects/neuralangelo/model.py:117-121,207-213. */
int na_ray_setup(const float* pose, const float* intr, const int64_t* ray_idx, int64_t n_images, int rays_per_image,
                 int width, float* center, float* ray, float* ray_unit, float* near, float* far, uint8_t* outside,
                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NA_B200_H */
