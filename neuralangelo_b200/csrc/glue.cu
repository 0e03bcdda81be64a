// Element-wise glue of the per-ray pipeline as single launches (each replaces 10-40 few-microsecond ATen kernels per
// training step, forward and backward):
//   na_rgb_input_fwd/bwd     the colour network's input row  [xyz | SH(view) | normalize(grad) | feat (| appearance)]
//                            NeuralRGB.forward / encode_view, projects/neuralangelo/utils/modules.py:209-229,
//                            spherical_harmonics.py:47-84, and the F.normalize of model.py:181
//   na_tap_combine_fwd/bwd   gradient / curvature from the centre and tap SDF values, modules.py:144-175
//   na_sdf_losses_fwd/bwd    eikonal and curvature losses, projects/neuralangelo/utils/misc.py:74-89
#include "na_common.cuh"
#include <math.h>

namespace {

constexpr unsigned kFullMask = 0xffffffffu;

// real spherical harmonics up to degree `levels` (<= 4), same constants and ordering as spherical_harmonics.py:47-84
__device__ __forceinline__ int sh_eval(float x, float y, float z, int levels, float* o) {
    int n = 0;
    o[n++] = 0.28209479177387814f;
    if (levels >= 1) {
        o[n++] = -0.4886025119029199f * y;
        o[n++] = 0.4886025119029199f * z;
        o[n++] = -0.4886025119029199f * x;
    }
    if (levels >= 2) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        o[n++] = 1.0925484305920792f * xy;
        o[n++] = -1.0925484305920792f * yz;
        o[n++] = 0.31539156525252005f * (2.0f * zz - xx - yy);
        o[n++] = -1.0925484305920792f * xz;
        o[n++] = 0.5462742152960396f * (xx - yy);
        if (levels >= 3) {
            o[n++] = -0.5900435899266435f * y * (3 * xx - yy);
            o[n++] = 2.890611442640554f * xy * z;
            o[n++] = -0.4570457994644658f * y * (4 * zz - xx - yy);
            o[n++] = 0.3731763325901154f * z * (2 * zz - 3 * xx - 3 * yy);
            o[n++] = -0.4570457994644658f * x * (4 * zz - xx - yy);
            o[n++] = 1.445305721320277f * z * (xx - yy);
            o[n++] = -0.5900435899266435f * x * (xx - 3 * yy);
        }
        if (levels >= 4) {
            o[n++] = 2.5033429417967046f * xy * (xx - yy);
            o[n++] = -1.7701307697799304f * yz * (3 * xx - yy);
            o[n++] = 0.9461746957575601f * xy * (7 * zz - 1);
            o[n++] = -0.6690465435572892f * yz * (7 * zz - 3);
            o[n++] = 0.10578554691520431f * (zz * (35 * zz - 30) + 3);
            o[n++] = -0.6690465435572892f * xz * (7 * zz - 3);
            o[n++] = 0.47308734787878004f * (xx - yy) * (7 * zz - 1);
            o[n++] = -1.7701307697799304f * xz * (xx - 3 * yy);
            o[n++] = 0.6258357354491761f * (xx * (xx - 3 * yy) - yy * (3 * xx - yy));
        }
    }
    return n;
}

// One warp per row.  Row layout: [xyz 3 | SH (levels+1)^2 | normal 3 | feat F | app A].  `ray_unit` is per RAY
// ([rows / samples_per_ray, 3]); grad / feat / points are per row.
__global__ void __launch_bounds__(256)
rgb_input_fwd_kernel(const float* __restrict__ points, const float* __restrict__ ray_unit, const float* __restrict__ grad,
                     const float* __restrict__ feat, int64_t ld_feat, const float* __restrict__ app, int64_t ld_app,
                     int64_t rows, int samples_per_ray, int levels, int n_feat, int n_app, float* __restrict__ out,
                     int64_t ld_out) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    float* o = out + row * ld_out;
    const int n_sh = (levels + 1) * (levels + 1);
    if (lane == 0) {
        o[0] = points[row * 3 + 0], o[1] = points[row * 3 + 1], o[2] = points[row * 3 + 2];
        const int64_t ray = row / samples_per_ray;
        float sh[25];
        sh_eval(ray_unit[ray * 3 + 0], ray_unit[ray * 3 + 1], ray_unit[ray * 3 + 2], levels, sh);
        for (int i = 0; i < n_sh; ++i) o[3 + i] = sh[i];
    } else if (lane == 1) {   // F.normalize(grad, dim=-1): g / max(|g|, 1e-12)
        const float g0 = grad[row * 3 + 0], g1 = grad[row * 3 + 1], g2 = grad[row * 3 + 2];
        const float inv = 1.f / fmaxf(sqrtf(g0 * g0 + g1 * g1 + g2 * g2), 1e-12f);
        float* on = o + 3 + n_sh;
        on[0] = g0 * inv, on[1] = g1 * inv, on[2] = g2 * inv;
    }
    float* of = o + 6 + n_sh;
    const float* f = feat + row * ld_feat;
    for (int i = lane; i < n_feat; i += 32) of[i] = f[i];
    if (n_app) {
        const float* a = app + row * ld_app;
        for (int i = lane; i < n_app; i += 32) of[n_feat + i] = a[i];
    }
    // columns up to the leading dimension are zero: the caller may hand the padded rows to a GEMM whose K must be a
    // multiple of four for the library's vectorised kernels (278 -> 280)
    for (int64_t i = 6 + n_sh + n_feat + n_app + lane; i < ld_out; i += 32) o[i] = 0.f;
}

// d_grad from the normal columns of d_out: n = g / r  =>  dg = (dn - n (n . dn)) / r   (r = max(|g|, 1e-12)).
__global__ void __launch_bounds__(256)
rgb_input_bwd_kernel(const float* __restrict__ grad, const float* __restrict__ d_out, int64_t ld_out, int normal_col,
                     int64_t rows, float* __restrict__ d_grad) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const float g0 = grad[row * 3 + 0], g1 = grad[row * 3 + 1], g2 = grad[row * 3 + 2];
    const float nrm = sqrtf(g0 * g0 + g1 * g1 + g2 * g2);
    const float* dn = d_out + row * ld_out + normal_col;
    const float d0 = dn[0], d1 = dn[1], d2 = dn[2];
    float* dg = d_grad + row * 3;
    if (nrm > 1e-12f) {
        const float inv = 1.f / nrm;
        const float n0 = g0 * inv, n1 = g1 * inv, n2 = g2 * inv;
        const float dot = n0 * d0 + n1 * d1 + n2 * d2;
        dg[0] = (d0 - n0 * dot) * inv, dg[1] = (d1 - n1 * dot) * inv, dg[2] = (d2 - n2 * dot) * inv;
    } else {   // clamped denominator: n = g / 1e-12, a plain scaling
        dg[0] = d0 * 1e12f, dg[1] = d1 * 1e12f, dg[2] = d2 * 1e12f;
    }
}

// ---------------------------------------------------------------------------------------------
// gradient / curvature from tap SDF values (modules.py:144-175).  taps: [T, rows] (T = 6: +x -x +y -y +z -z; T = 4: the
// tetrahedral stencil k = (1,-1,-1), (-1,-1,1), (-1,1,-1), (1,1,1)); den_grad = 2 eps (6) or 4 eps (4), den_hess = eps^2,
// both already rounded to fp32 by the caller exactly as torch rounds the python scalars it divides by.
// ---------------------------------------------------------------------------------------------
template <int T>
__global__ void __launch_bounds__(256)
tap_combine_fwd_kernel(const float* __restrict__ taps, const float* __restrict__ sdf, int64_t rows, float den_grad,
                       float den_hess, float* __restrict__ grad, float* __restrict__ hess) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    float s[T];
#pragma unroll
    for (int t = 0; t < T; ++t) s[t] = taps[(int64_t)t * rows + r];
    if (T == 6) {
#pragma unroll
        for (int a = 0; a < 3; ++a) grad[r * 3 + a] = (s[2 * a] - s[2 * a + 1]) / den_grad;
        if (hess) {
            const float c = 2.f * sdf[r];
#pragma unroll
            for (int a = 0; a < 3; ++a) hess[r * 3 + a] = (s[2 * a] + s[2 * a + 1] - c) / den_hess;
        }
    } else {
        grad[r * 3 + 0] = (s[0] - s[1] - s[2] + s[3]) / den_grad;
        grad[r * 3 + 1] = (-s[0] - s[1] + s[2] + s[3]) / den_grad;
        grad[r * 3 + 2] = (-s[0] + s[1] - s[2] + s[3]) / den_grad;
        if (hess) {
            const float h = ((s[0] + s[1] + s[2] + s[3]) / 2.0f - 2.f * sdf[r]) / den_hess / 3.0f;
            hess[r * 3 + 0] = h, hess[r * 3 + 1] = h, hess[r * 3 + 2] = h;
        }
    }
}

template <int T>
__global__ void __launch_bounds__(256)
tap_combine_bwd_kernel(const float* __restrict__ d_grad, const float* __restrict__ d_hess, int64_t rows, float den_grad,
                       float den_hess, float* __restrict__ d_taps, float* __restrict__ d_sdf) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    float g[3] = {0.f, 0.f, 0.f}, h[3] = {0.f, 0.f, 0.f};
    if (d_grad) g[0] = d_grad[r * 3] / den_grad, g[1] = d_grad[r * 3 + 1] / den_grad, g[2] = d_grad[r * 3 + 2] / den_grad;
    if (d_hess) h[0] = d_hess[r * 3] / den_hess, h[1] = d_hess[r * 3 + 1] / den_hess, h[2] = d_hess[r * 3 + 2] / den_hess;
    if (T == 6) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            d_taps[(int64_t)(2 * a) * rows + r] = g[a] + h[a];
            d_taps[(int64_t)(2 * a + 1) * rows + r] = -g[a] + h[a];
        }
        if (d_sdf) d_sdf[r] = -2.f * (h[0] + h[1] + h[2]);
    } else {
        const float hs = (h[0] + h[1] + h[2]) / 3.0f;
        d_taps[0 * rows + r] = g[0] - g[1] - g[2] + hs * 0.5f;
        d_taps[1 * rows + r] = -g[0] - g[1] + g[2] + hs * 0.5f;
        d_taps[2 * rows + r] = -g[0] + g[1] - g[2] + hs * 0.5f;
        d_taps[3 * rows + r] = g[0] + g[1] + g[2] + hs * 0.5f;
        if (d_sdf) d_sdf[r] = -2.f * hs;
    }
}

// ---------------------------------------------------------------------------------------------
// eikonal / curvature losses (misc.py:74-89): means over ALL rows of nan_to_num((|g| - 1)^2) and nan_to_num(|sum h|),
// rows of rays flagged `outside` contributing zero.  out[0] += eikonal, out[1] += curvature (out zero-filled by the caller).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float finite_or_zero(float v) { return (v == v && fabsf(v) <= 3.4028234e38f) ? v : 0.f; }

__global__ void __launch_bounds__(256)
sdf_losses_fwd_kernel(const float* __restrict__ grad, const float* __restrict__ hess, const uint8_t* __restrict__ outside,
                      int64_t rows, int samples_per_ray, float inv_rows, float* __restrict__ out) {
    float e = 0.f, c = 0.f;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
        if (outside && outside[r / samples_per_ray]) continue;
        const float g0 = grad[r * 3], g1 = grad[r * 3 + 1], g2 = grad[r * 3 + 2];
        const float d = sqrtf(g0 * g0 + g1 * g1 + g2 * g2) - 1.f;
        e += finite_or_zero(d * d);
        if (hess) c += finite_or_zero(fabsf(hess[r * 3] + hess[r * 3 + 1] + hess[r * 3 + 2]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(kFullMask, e, o), c += __shfl_xor_sync(kFullMask, c, o);
    __shared__ float se[8], sc[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) se[warp] = e, sc[warp] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        float te = 0.f, tc = 0.f;
        for (int w = 0; w < 8; ++w) te += se[w], tc += sc[w];
        atomicAdd(out, te * inv_rows);
        if (hess) atomicAdd(out + 1, tc * inv_rows);
    }
}

__global__ void __launch_bounds__(256)
sdf_losses_bwd_kernel(const float* __restrict__ grad, const float* __restrict__ hess, const uint8_t* __restrict__ outside,
                      int64_t rows, int samples_per_ray, float inv_rows, const float* __restrict__ d_out,
                      float* __restrict__ d_grad, float* __restrict__ d_hess) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const bool masked = outside && outside[r / samples_per_ray];
    const float de = __ldg(d_out) * inv_rows, dc = hess ? __ldg(d_out + 1) * inv_rows : 0.f;
    const float g0 = grad[r * 3], g1 = grad[r * 3 + 1], g2 = grad[r * 3 + 2];
    const float nrm = sqrtf(g0 * g0 + g1 * g1 + g2 * g2);
    const float d = nrm - 1.f, sq = d * d;
    float k = 0.f;   // d eik / d g = 2 (|g| - 1) g / |g| where the term is finite (nan_to_num passes no gradient otherwise)
    if (!masked && sq == sq && fabsf(sq) <= 3.4028234e38f && nrm > 0.f) k = de * 2.f * d / nrm;
    const bool live = k != 0.f;   // (0 * NaN would leak the NaN of a non-finite row into its gradient)
    d_grad[r * 3] = live ? k * g0 : 0.f, d_grad[r * 3 + 1] = live ? k * g1 : 0.f, d_grad[r * 3 + 2] = live ? k * g2 : 0.f;
    if (d_hess) {
        const float s = hess[r * 3] + hess[r * 3 + 1] + hess[r * 3 + 2];
        float v = 0.f;
        if (!masked && s == s && fabsf(s) <= 3.4028234e38f) v = dc * (s > 0.f ? 1.f : (s < 0.f ? -1.f : 0.f));
        d_hess[r * 3] = v, d_hess[r * 3 + 1] = v, d_hess[r * 3 + 2] = v;
    }
}

}  // namespace

#define NA_GLUE_PROLOGUE(who)                                        \
    na::clear_error();                                               \
    NA_REQUIRE(rows >= 0, who ": negative row count");               \
    if (rows == 0) return NA_OK;                                     \
    cudaStream_t s = (cudaStream_t)stream;

extern "C" int na_rgb_input_fwd(const float* points, const float* ray_unit, const float* grad, const float* feat,
                                int64_t ld_feat, const float* app, int64_t ld_app, int64_t rows, int samples_per_ray,
                                int sh_levels, int n_feat, int n_app, float* out, int64_t ld_out, void* stream) {
    NA_GLUE_PROLOGUE("rgb_input_fwd")
    NA_REQUIRE(points && ray_unit && grad && feat && out, "rgb_input_fwd: null pointer");
    NA_REQUIRE(samples_per_ray >= 1 && sh_levels >= 0 && sh_levels <= 4 && n_feat >= 0 && n_app >= 0 && (n_app == 0 || app),
               "rgb_input_fwd: bad arguments");
    NA_REQUIRE(ld_out >= 6 + (sh_levels + 1) * (sh_levels + 1) + n_feat + n_app, "rgb_input_fwd: ld_out too small");
    rgb_input_fwd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(points, ray_unit, grad, feat, ld_feat, app, ld_app, rows,
                                                                   samples_per_ray, sh_levels, n_feat, n_app, out, ld_out);
    NA_CHECK_LAUNCH("rgb_input_fwd_kernel");
    return NA_OK;
}

extern "C" int na_rgb_input_bwd(const float* grad, const float* d_out, int64_t ld_out, int normal_col, int64_t rows,
                                float* d_grad, void* stream) {
    NA_GLUE_PROLOGUE("rgb_input_bwd")
    NA_REQUIRE(grad && d_out && d_grad && normal_col >= 0, "rgb_input_bwd: bad arguments");
    rgb_input_bwd_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(grad, d_out, ld_out, normal_col, rows, d_grad);
    NA_CHECK_LAUNCH("rgb_input_bwd_kernel");
    return NA_OK;
}

extern "C" int na_tap_combine_fwd(const float* taps, const float* sdf, int64_t rows, int n_taps, float den_grad,
                                  float den_hess, float* grad, float* hess, void* stream) {
    NA_GLUE_PROLOGUE("tap_combine_fwd")
    NA_REQUIRE(taps && grad && (n_taps == 4 || n_taps == 6) && (!hess || sdf), "tap_combine_fwd: bad arguments");
    const unsigned blocks = (unsigned)((rows + 255) / 256);
    if (n_taps == 6) tap_combine_fwd_kernel<6><<<blocks, 256, 0, s>>>(taps, sdf, rows, den_grad, den_hess, grad, hess);
    else tap_combine_fwd_kernel<4><<<blocks, 256, 0, s>>>(taps, sdf, rows, den_grad, den_hess, grad, hess);
    NA_CHECK_LAUNCH("tap_combine_fwd_kernel");
    return NA_OK;
}

extern "C" int na_tap_combine_bwd(const float* d_grad, const float* d_hess, int64_t rows, int n_taps, float den_grad,
                                  float den_hess, float* d_taps, float* d_sdf, void* stream) {
    NA_GLUE_PROLOGUE("tap_combine_bwd")
    NA_REQUIRE(d_taps && (n_taps == 4 || n_taps == 6), "tap_combine_bwd: bad arguments");
    const unsigned blocks = (unsigned)((rows + 255) / 256);
    if (n_taps == 6) tap_combine_bwd_kernel<6><<<blocks, 256, 0, s>>>(d_grad, d_hess, rows, den_grad, den_hess, d_taps, d_sdf);
    else tap_combine_bwd_kernel<4><<<blocks, 256, 0, s>>>(d_grad, d_hess, rows, den_grad, den_hess, d_taps, d_sdf);
    NA_CHECK_LAUNCH("tap_combine_bwd_kernel");
    return NA_OK;
}

extern "C" int na_sdf_losses_fwd(const float* grad, const float* hess, const uint8_t* outside, int64_t rows,
                                 int samples_per_ray, float* out2, void* stream) {
    NA_GLUE_PROLOGUE("sdf_losses_fwd")
    NA_REQUIRE(grad && out2 && samples_per_ray >= 1, "sdf_losses_fwd: bad arguments");
    int64_t blocks = (rows + 255) / 256;
    if (blocks > na::kNumSMs * 4) blocks = na::kNumSMs * 4;
    sdf_losses_fwd_kernel<<<(unsigned)blocks, 256, 0, s>>>(grad, hess, outside, rows, samples_per_ray, 1.0f / (float)rows, out2);
    NA_CHECK_LAUNCH("sdf_losses_fwd_kernel");
    return NA_OK;
}

extern "C" int na_sdf_losses_bwd(const float* grad, const float* hess, const uint8_t* outside, int64_t rows,
                                 int samples_per_ray, const float* d_out2, float* d_grad, float* d_hess, void* stream) {
    NA_GLUE_PROLOGUE("sdf_losses_bwd")
    NA_REQUIRE(grad && d_out2 && d_grad && samples_per_ray >= 1 && (!d_hess || hess), "sdf_losses_bwd: bad arguments");
    sdf_losses_bwd_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(grad, hess, outside, rows, samples_per_ray,
                                                                        1.0f / (float)rows, d_out2, d_grad, d_hess);
    NA_CHECK_LAUNCH("sdf_losses_bwd_kernel");
    return NA_OK;
}

// ---- ray set-up for selected pixels: pixel index -> centre, ray, unit ray, unit-sphere bounds --------------------------
// Follows the reference's order of operations: img2cam with the closed-form inverse intrinsics, cam2world with
// [R^T | -R^T t] applied to homogeneous coordinates (sum over j = 0..3), ray = world point - centre
// (projects/nerf/utils/camera.py:283-311), normalize (eps 1e-12), sphere intersection with radius 1 and the miss
// handling of Model.get_dist_bounds (projects/neuralangelo/model.py:207-213).
namespace {
__global__ void ray_setup_kernel(const float* __restrict__ pose, const float* __restrict__ intr,
                                 const int64_t* __restrict__ ray_idx, int64_t n, int rays_per_image, int width,
                                 float* __restrict__ center, float* __restrict__ ray, float* __restrict__ ray_unit,
                                 float* __restrict__ near, float* __restrict__ far, uint8_t* __restrict__ outside) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t b = i / rays_per_image;
    const float* P = pose + b * 12;
    const float* K = intr + b * 9;
    const int64_t idx = ray_idx[i];
    const float xy1[3] = {(float)(idx % width) + 0.5f, (float)(idx / width) + 0.5f, 1.f};
    // inverse intrinsics: adjugate / determinant
    const float a = K[0], bb = K[1], c = K[2], d = K[3], e = K[4], f = K[5], g = K[6], h = K[7], k = K[8];
    const float A = __fsub_rn(__fmul_rn(e, k), __fmul_rn(f, h)), B = __fsub_rn(__fmul_rn(c, h), __fmul_rn(bb, k)),
                C = __fsub_rn(__fmul_rn(bb, f), __fmul_rn(c, e));
    const float D = __fsub_rn(__fmul_rn(f, g), __fmul_rn(d, k)), E = __fsub_rn(__fmul_rn(a, k), __fmul_rn(c, g)),
                F = __fsub_rn(__fmul_rn(c, d), __fmul_rn(a, f));
    const float G = __fsub_rn(__fmul_rn(d, h), __fmul_rn(e, g)), H = __fsub_rn(__fmul_rn(bb, g), __fmul_rn(a, h)),
                I = __fsub_rn(__fmul_rn(a, e), __fmul_rn(bb, d));
    const float det = __fadd_rn(__fadd_rn(__fmul_rn(a, A), __fmul_rn(bb, D)), __fmul_rn(c, G));
    const float Ki[3][3] = {{A / det, B / det, C / det}, {D / det, E / det, F / det}, {G / det, H / det, I / det}};
    float cam[4];
#pragma unroll
    for (int r = 0; r < 3; ++r) cam[r] = fmaf(Ki[r][2], xy1[2], fmaf(Ki[r][1], xy1[1], Ki[r][0] * xy1[0]));
    cam[3] = 1.f;
    // pose inverse [R^T | -R^T t]
    float Ri[3][4];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        Ri[r][0] = P[0 * 4 + r], Ri[r][1] = P[1 * 4 + r], Ri[r][2] = P[2 * 4 + r];
        Ri[r][3] = fmaf(-Ri[r][2], P[2 * 4 + 3], fmaf(-Ri[r][1], P[1 * 4 + 3], -Ri[r][0] * P[0 * 4 + 3]));
    }
    float cen[3], dir[3], nrm2 = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        cen[r] = Ri[r][3];   // [0, 0, 0, 1] . row
        const float world = fmaf(Ri[r][3], cam[3], fmaf(Ri[r][2], cam[2], fmaf(Ri[r][1], cam[1], Ri[r][0] * cam[0])));
        dir[r] = world - cen[r];
        nrm2 = fmaf(dir[r], dir[r], nrm2);
    }
    const float den = fmaxf(sqrtf(nrm2), 1e-12f);
    float u[3], ctc = 0.f, ctv = 0.f;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        u[r] = dir[r] / den;
        ctc = fmaf(cen[r], cen[r], ctc);
        ctv = fmaf(cen[r], u[r], ctv);
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        center[i * 3 + r] = cen[r];
        ray[i * 3 + r] = dir[r];
        ray_unit[i * 3 + r] = u[r];
    }
    if (near) {
        const float disc = sqrtf(ctv * ctv - (ctc - 1.f));   // NaN = the ray misses the unit sphere
        const float dn = fmaxf(-ctv - disc, 0.f), df = -ctv + disc;
        const bool miss = disc != disc;
        near[i] = miss ? 1.f : dn;
        far[i] = miss ? 1.2f : df;
        outside[i] = miss ? 1 : 0;
    }
}
}  // namespace

extern "C" int na_ray_setup(const float* pose, const float* intr, const int64_t* ray_idx, int64_t n_images,
                            int rays_per_image, int width, float* center, float* ray, float* ray_unit, float* near,
                            float* far, uint8_t* outside, void* stream) {
    na::clear_error();
    const int64_t n = n_images * rays_per_image;
    NA_REQUIRE(n >= 0 && rays_per_image >= 0, "ray_setup: negative sizes");
    if (n == 0) return NA_OK;
    NA_REQUIRE(pose && intr && ray_idx && center && ray && ray_unit && width > 0, "ray_setup: bad arguments");
    NA_REQUIRE((near == nullptr) == (far == nullptr) && (near == nullptr) == (outside == nullptr),
               "ray_setup: near / far / outside go together");
    ray_setup_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pose, intr, ray_idx, n, rays_per_image, width,
                                                                                     center, ray, ray_unit, near, far, outside);
    NA_CHECK_LAUNCH("ray_setup_kernel");
    return NA_OK;
}
