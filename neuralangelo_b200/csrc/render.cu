// NeuS volume rendering and hierarchical resampling: one warp per ray, warp-shuffle scans.
//
// Replaces the eager ATen chains of
//   projects/nerf/utils/render.py:68-112   (volume_rendering_alphas_dist, alpha_compositing_weights, composite)
//   projects/nerf/utils/nerf_util.py:41-68 (sample_dists_from_pdf)
//   projects/neuralangelo/model.py:232-298 (sample_dists_all / sample_dists_hierarchical /
//                                            compute_neus_alphas / _get_iter_cos)
//
// Layout: every per-sample tensor is [rays, samples(, k)] fp32 contiguous.  Lane j of the ray's warp
// owns samples j, j+32, ... so all global accesses are coalesced; the exclusive transmittance product
// and its adjoint are warp scans with a running carry between 32-sample chunks.  The adjoint of the
// transmittance product is a reverse affine scan (no division by 1-alpha, which the clip to [0,1]
// at model.py:290 makes exactly zero in practice).
#include "na_common.cuh"
#include "tc_common.cuh"
#include <math.h>

namespace {

constexpr int kWarpsPerCta = 8;
constexpr unsigned kFull = 0xffffffffu;

// ex2.approx / rcp.approx forms: with the rows staged in shared memory the renderer is bound by instruction issue, and
// the IEEE expf / division sequences were a third of its instructions.  Relative error ~1e-6 (argument rounding of
// x * log2(e) for |x| up to ~60), against the 2e-5 tolerance of the parity tests.
__device__ __forceinline__ float fast_exp(float x) { return __expf(x); }
__device__ __forceinline__ float sigmoidf(float x) { return __fdividef(1.f, 1.f + fast_exp(-x)); }
// IEEE form for the hierarchical sampler: its inverse-CDF step is ill-conditioned where the CDF is flat, so it keeps the
// reference's arithmetic to the last bit it can
__device__ __forceinline__ float sigmoid_ieee(float x) { return 1.f / (1.f + expf(-x)); }

// Inclusive product scan over the warp's lanes.
__device__ __forceinline__ float warp_scan_prod(float v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float t = __shfl_up_sync(kFull, v, d);
        if (lane >= d) v *= t;
    }
    return v;
}
__device__ __forceinline__ float warp_scan_sum(float v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float t = __shfl_up_sync(kFull, v, d);
        if (lane >= d) v += t;
    }
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
    return v;
}

// Forward transmittance chunk: given this lane's alpha (0 for padding lanes) and the running carry
// (product of 1-alpha over all earlier chunks) returns T for this lane and updates the carry.
__device__ __forceinline__ float transmittance_chunk(float alpha, int lane, float& carry) {
    const float incl = warp_scan_prod(1.f - alpha, lane);
    float excl = __shfl_up_sync(kFull, incl, 1);
    if (lane == 0) excl = 1.f;
    const float T = carry * excl;
    carry *= __shfl_sync(kFull, incl, 31);
    return T;
}

// Reverse affine scan chunk for the adjoint of w_i = alpha_i * prod_{j<i}(1-alpha_j).
// With Q_i = sum_{k>i} dw_k alpha_k prod_{i<j<k}(1-alpha_j):  Q_{i-1} = dw_i alpha_i + (1-alpha_i) Q_i,
// d alpha_i = T_i (dw_i - Q_i).  Chunks are visited last to first; `carry` holds Q at the chunk's
// last sample on entry (0 for the last chunk) and Q just before the chunk's first sample on exit.
__device__ __forceinline__ float adjoint_chunk(float alpha, float dw, int lane, float& carry) {
    float a = 1.f - alpha, b = dw * alpha;  // f(Q) = b + a Q ; padding lanes: alpha=0, dw=0 -> identity
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float a2 = __shfl_down_sync(kFull, a, d), b2 = __shfl_down_sync(kFull, b, d);
        if (lane + d < 32) {
            b = fmaf(a, b2, b);
            a *= a2;
        }
    }
    // (a,b) = f_lane o ... o f_31 ; Q just before this lane's sample = b + a*carry
    const float q_before = fmaf(a, carry, b);
    float q_here = __shfl_down_sync(kFull, q_before, 1);  // Q_i = "Q before sample i+1"
    if (lane == 31) q_here = carry;
    carry = __shfl_sync(kFull, q_before, 0);
    return q_here;
}

// ------------------------------------------------------------------------------------------
// NeuS alpha of one sample (model.py:275-298) and its partial derivatives.
// ------------------------------------------------------------------------------------------
struct NeusAlpha {
    float alpha;
    // d alpha / d sdf, / d true_cos, / d inv_s  (already masked by the clip)
    float d_sdf, d_cos, d_invs;
};

template <bool WITH_GRAD>
__device__ __forceinline__ NeusAlpha neus_alpha(float sdf, float true_cos, float dist_intv, float inv_s,
                                                float anneal_ratio) {
    const float h0 = fmaxf(-true_cos * 0.5f + 0.5f, 0.f), h1 = fmaxf(-true_cos, 0.f);
    const float iter_cos = -(h0 * (1.f - anneal_ratio) + h1 * anneal_ratio);
    const float half = iter_cos * dist_intv * 0.5f;
    const float est_prev = sdf - half, est_next = sdf + half;
    const float p = sigmoidf(est_prev * inv_s), q = sigmoidf(est_next * inv_s);
    const float denom = p + 1e-5f;
    const float raw = __fdividef(p - q, denom);
    NeusAlpha r;
    r.alpha = fminf(fmaxf(raw, 0.f), 1.f);
    if (WITH_GRAD) {
        const bool pass = (raw >= 0.f) && (raw <= 1.f);  // clamp backward is inclusive at the bounds
        const float rden = __fdividef(1.f, denom);
        const float da_dp = (q + 1e-5f) * rden * rden, da_dq = -rden;
        const float gp = pass ? da_dp * p * (1.f - p) : 0.f;  // d alpha / d (est_prev*inv_s)
        const float gq = pass ? da_dq * q * (1.f - q) : 0.f;
        r.d_sdf = (gp + gq) * inv_s;
        const float d_itercos = (gq - gp) * dist_intv * 0.5f * inv_s;
        const float dcos_dtrue = 0.5f * (1.f - anneal_ratio) * (h0 > 0.f ? 1.f : 0.f) +
                                 anneal_ratio * (h1 > 0.f ? 1.f : 0.f);
        r.d_cos = d_itercos * dcos_dtrue;
        r.d_invs = gp * est_prev + gq * est_next;
    } else {
        r.d_sdf = r.d_cos = r.d_invs = 0.f;
    }
    return r;
}

__device__ __forceinline__ float interval(const float* d, int i, int n, float far) {
    return (i + 1 < n ? d[i + 1] : far) - d[i];
}

// ------------------------------------------------------------------------------------------
// standalone pieces (API parity with the reference's functions)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarpsPerCta * 32)
alpha_weights_fwd_kernel(const float* __restrict__ alphas, int64_t n_rays, int n, float* __restrict__ weights) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float* a = alphas + ray * n;
    float* w = weights + ray * n;
    float carry = 1.f;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const float al = i < n ? a[i] : 0.f;
        const float T = transmittance_chunk(al, lane, carry);
        if (i < n) w[i] = al * T;
    }
}

// d_alphas doubles as the scratch row for T between the forward and the reverse sweep.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
alpha_weights_bwd_kernel(const float* __restrict__ alphas, const float* __restrict__ d_weights, int64_t n_rays,
                         int n, float* __restrict__ d_alphas) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float* a = alphas + ray * n;
    const float* dw = d_weights + ray * n;
    float* da = d_alphas + ray * n;
    float carry = 1.f;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const float T = transmittance_chunk(i < n ? a[i] : 0.f, lane, carry);
        if (i < n) da[i] = T;
    }
    __syncwarp();
    float q = 0.f;
    for (int base = ((n - 1) / 32) * 32; base >= 0; base -= 32) {
        const int i = base + lane;
        const float al = i < n ? a[i] : 0.f, g = i < n ? dw[i] : 0.f;
        const float qi = adjoint_chunk(al, g, lane, q);
        if (i < n) da[i] = da[i] * (g - qi);
    }
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
composite_fwd_kernel(const float* __restrict__ q, const float* __restrict__ w, int64_t n_rays, int n, int k,
                     float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float* qr = q + ray * n * k;
    const float* wr = w + ray * n;
    for (int c = 0; c < k; ++c) {
        float acc = 0.f;
        for (int i = lane; i < n; i += 32) acc = fmaf(qr[(int64_t)i * k + c], wr[i], acc);
        acc = warp_sum(acc);
        if (lane == 0) out[ray * k + c] = acc;
    }
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
composite_bwd_kernel(const float* __restrict__ q, const float* __restrict__ w, const float* __restrict__ d_out,
                     int64_t n_rays, int n, int k, float* __restrict__ d_q, float* __restrict__ d_w) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float* g = d_out + ray * k;
    for (int i = lane; i < n; i += 32) {
        const float wi = w[ray * n + i];
        float acc = 0.f;
        for (int c = 0; c < k; ++c) {
            const float gc = g[c];
            if (d_q) d_q[(ray * n + i) * k + c] = gc * wi;
            acc = fmaf(gc, q[(ray * n + i) * k + c], acc);
        }
        if (d_w) d_w[ray * n + i] = acc;
    }
}

template <bool BWD>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
density_alphas_kernel(const float* __restrict__ dens, const float* __restrict__ dists,
                      const float* __restrict__ dist_far, const float* __restrict__ d_alphas, int64_t n_rays, int n,
                      float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float far = dist_far ? dist_far[ray] : 1e10f;
    const float* d = dists + ray * n;
    for (int i = lane; i < n; i += 32) {
        const float dt = interval(d, i, n, far);
        const float e = expf(-dens[ray * n + i] * dt);
        out[ray * n + i] = BWD ? d_alphas[ray * n + i] * dt * e : 1.f - e;
    }
}

template <bool BWD>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
neus_alphas_kernel(const float* __restrict__ sdf, const float* __restrict__ grad, const float* __restrict__ dists,
                   const float* __restrict__ ray_unit, const float* __restrict__ dist_far,
                   const float* __restrict__ s_var, const float* __restrict__ anneal_ptr,
                   const float* __restrict__ d_alphas,
                   int64_t n_rays, int n, float* __restrict__ alphas, float* __restrict__ d_sdf,
                   float* __restrict__ d_grad, float* __restrict__ d_s_var) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float inv_s = expf(__ldg(s_var));
    const float anneal_ratio = __ldg(anneal_ptr);
    const float far = dist_far ? dist_far[ray] : 1e10f;
    const float r0 = ray_unit[ray * 3 + 0], r1 = ray_unit[ray * 3 + 1], r2 = ray_unit[ray * 3 + 2];
    const float* d = dists + ray * n;
    float ds_acc = 0.f;
    for (int i = lane; i < n; i += 32) {
        const int64_t s = ray * n + i;
        const float tc = r0 * grad[s * 3 + 0] + r1 * grad[s * 3 + 1] + r2 * grad[s * 3 + 2];
        const NeusAlpha na_ = neus_alpha<BWD>(sdf[s], tc, interval(d, i, n, far), inv_s, anneal_ratio);
        if (!BWD) {
            alphas[s] = na_.alpha;
        } else {
            const float g = d_alphas[s];
            d_sdf[s] = g * na_.d_sdf;
            const float gc = g * na_.d_cos;
            d_grad[s * 3 + 0] = gc * r0;
            d_grad[s * 3 + 1] = gc * r1;
            d_grad[s * 3 + 2] = gc * r2;
            ds_acc = fmaf(g, na_.d_invs, ds_acc);
        }
    }
    if (BWD) {
        ds_acc = warp_sum(ds_acc) * inv_s;  // d inv_s / d s_var = inv_s
        if (lane == 0 && ds_acc != 0.f) atomicAdd(d_s_var, ds_acc);
    }
}

// ------------------------------------------------------------------------------------------
// inverse-CDF sampling (nerf_util.py:41-68).  cdf row lives in shared memory (n <= kMaxBins).
// ------------------------------------------------------------------------------------------
constexpr int kMaxBins = 1024;

// Builds cdf[0..n-1] (cdf[0] = 0) from n-1 weights held as w(i); returns nothing, leaves cdf in smem.
template <typename WeightFn>
__device__ __forceinline__ void build_cdf(WeightFn w, int n, int lane, float* cdf) {
    float l1 = 0.f;
    for (int i = lane; i < n - 1; i += 32) l1 += fabsf(w(i));
    l1 = warp_sum(l1);
    const float inv = 1.f / fmaxf(l1, 1e-12f);  // F.normalize(p=1, eps=1e-12)
    float carry = 0.f;
    if (lane == 0) cdf[0] = 0.f;
    for (int base = 0; base < n - 1; base += 32) {
        const int i = base + lane;
        const float v = i < n - 1 ? w(i) * inv : 0.f;
        const float incl = warp_scan_sum(v, lane) + carry;
        if (i < n - 1) cdf[i + 1] = incl;
        carry = __shfl_sync(kFull, incl, 31);
    }
    __syncwarp();
}

__device__ __forceinline__ float sample_cdf(const float* cdf, const float* bins, int n, float u) {
    // searchsorted(right=True): number of cdf entries <= u
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] <= u) lo = mid + 1;
        else hi = mid;
    }
    const int low = max(lo - 1, 0), high = min(lo, n - 1);
    const float c0 = cdf[low], c1 = cdf[high];
    const float t = (u - c0) / (c1 - c0 + 1e-8f);
    const float b0 = bins[low], b1 = bins[high];
    return b0 + t * (b1 - b0);
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
sample_pdf_kernel(const float* __restrict__ bins, const float* __restrict__ weights, const float* __restrict__ unif,
                  int64_t n_rays, int n, int n_fine, float* __restrict__ out) {
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + warp;
    if (ray >= n_rays) return;
    float* cdf = smem + warp * n;
    const float* wr = weights + ray * (n - 1);
    build_cdf([&](int i) { return wr[i]; }, n, lane, cdf);
    for (int j = lane; j < n_fine; j += 32) out[ray * n_fine + j] = sample_cdf(cdf, bins + ray * n, n, unif[j]);
}

// sample_dists_hierarchical (model.py:250-267): interval alphas from consecutive SDF samples,
// compositing weights, then inverse-CDF sampling -- all within the ray's warp.
__global__ void __launch_bounds__(kWarpsPerCta * 32)
neus_upsample_kernel(const float* __restrict__ dists, const float* __restrict__ sdfs,
                     const float* __restrict__ unif, float inv_s, int64_t n_rays, int n, int n_fine,
                     float* __restrict__ out) {
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + warp;
    if (ray >= n_rays) return;
    float* cdf = smem + warp * 2 * n;
    float* wts = cdf + n;
    const float* d = dists + ray * n;
    const float* s = sdfs + ray * n;
    float carry = 1.f;
    for (int base = 0; base < n - 1; base += 32) {
        const int i = base + lane;
        float alpha = 0.f;
        if (i < n - 1) {
            const float s0 = s[i], s1 = s[i + 1], d0 = d[i], d1 = d[i + 1];
            const float mid = (s0 + s1) * 0.5f;
            float cosv = (s1 - s0) / (d1 - d0 + 1e-5f);
            const float prev = i > 0 ? (s0 - s[i - 1]) / (d0 - d[i - 1] + 1e-5f) : 0.f;
            cosv = fminf(prev, cosv);
            const float half = cosv * (d1 - d0) * 0.5f;
            const float p = sigmoid_ieee((mid - half) * inv_s), q = sigmoid_ieee((mid + half) * inv_s);
            alpha = fminf(fmaxf((p - q) / (p + 1e-5f), 0.f), 1.f);
        }
        const float T = transmittance_chunk(alpha, lane, carry);
        if (i < n - 1) wts[i] = alpha * T;
    }
    __syncwarp();
    build_cdf([&](int i) { return wts[i]; }, n, lane, cdf);
    for (int j = lane; j < n_fine; j += 32) out[ray * n_fine + j] = sample_cdf(cdf, d, n, unif[j]);
}

// cat + sort + gather of model.py:241-247 as a rank merge of two sorted rows (stable: coarse first on ties).
__global__ void __launch_bounds__(kWarpsPerCta * 32)
merge_sorted_kernel(const float* __restrict__ dists, const float* __restrict__ sdfs,
                    const float* __restrict__ dists_fine, const float* __restrict__ sdfs_fine, int64_t n_rays, int n,
                    int n_fine, float* __restrict__ dists_out, float* __restrict__ sdfs_out) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= n_rays) return;
    const float* a = dists + ray * n;
    const float* b = dists_fine + ray * n_fine;
    float* o = dists_out + ray * (n + n_fine);
    float* so = sdfs_out ? sdfs_out + ray * (n + n_fine) : nullptr;
    for (int i = lane; i < n; i += 32) {  // rank among fine: count of b < a[i]
        const float v = a[i];
        int lo = 0, hi = n_fine;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (b[mid] < v) lo = mid + 1;
            else hi = mid;
        }
        o[i + lo] = v;
        if (so) so[i + lo] = sdfs[ray * n + i];
    }
    for (int j = lane; j < n_fine; j += 32) {  // rank among coarse: count of a <= b[j]
        const float v = b[j];
        int lo = 0, hi = n;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (a[mid] <= v) lo = mid + 1;
            else hi = mid;
        }
        o[j + lo] = v;
        if (so) so[j + lo] = sdfs_fine[ray * n_fine + j];
    }
}

// ------------------------------------------------------------------------------------------
// fused renderer: NeuS alphas (object) ++ density alphas (background) -> weights -> composite RGB.
// model.py:132-141 with the alphas of render_rays_object (:168-169) and render_rays_background (:198).
// ------------------------------------------------------------------------------------------
struct RenderArgs {
    const float *sdf, *grad, *dists_obj, *ray_unit, *dist_far, *s_var, *rgb_obj;  // object: [R,No(,3)]
    const float *dens_bg, *dists_bg, *rgb_bg;                                       // background: [R,Nb(,3)] or null
    const float* anneal_ratio;  // device scalar: min(progress / anneal_end, 1)
    int n_obj, n_bg;
    int64_t n_rays;
};

template <int MAXC, bool BWD>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
neus_render_kernel(const RenderArgs A, float* __restrict__ rgb_out, float* __restrict__ weights_out,
                   float* __restrict__ opacity_out, const float* __restrict__ d_rgb,
                   const float* __restrict__ d_weights, float* __restrict__ d_sdf, float* __restrict__ d_grad,
                   float* __restrict__ d_s_var, float* __restrict__ d_rgb_obj, float* __restrict__ d_dens_bg,
                   float* __restrict__ d_rgb_bg) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = (int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (ray >= A.n_rays) return;
    const int n_obj = A.n_obj, n_bg = A.n_bg, n = n_obj + n_bg;
    const float inv_s = expf(__ldg(A.s_var));
    const float anneal = __ldg(A.anneal_ratio);
    const float far = A.dist_far ? A.dist_far[ray] : 1e10f;
    const float r0 = A.ray_unit[ray * 3 + 0], r1 = A.ray_unit[ray * 3 + 1], r2 = A.ray_unit[ray * 3 + 2];
    const float* dob = A.dists_obj + ray * n_obj;
    const float* dbg = n_bg ? A.dists_bg + ray * n_bg : nullptr;

    float alpha[MAXC], T[MAXC];
    float gsdf[MAXC], gcos[MAXC], ginv[MAXC];  // alpha partials (BWD only; object: sdf/cos/inv_s, bg: gsdf = d/d dens)
    float carry = 1.f;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, accw = 0.f;
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
        const int i = c * 32 + lane;
        float al = 0.f;
        gsdf[c] = gcos[c] = ginv[c] = 0.f;
        if (i < n_obj) {
            const int64_t s = ray * n_obj + i;
            const float tc = r0 * A.grad[s * 3 + 0] + r1 * A.grad[s * 3 + 1] + r2 * A.grad[s * 3 + 2];
            const NeusAlpha na_ = neus_alpha<BWD>(A.sdf[s], tc, interval(dob, i, n_obj, far), inv_s, anneal);
            al = na_.alpha;
            gsdf[c] = na_.d_sdf, gcos[c] = na_.d_cos, ginv[c] = na_.d_invs;
        } else if (i < n) {
            const int ib = i - n_obj;
            const float dt = interval(dbg, ib, n_bg, 1e10f);
            const float e = fast_exp(-A.dens_bg[ray * n_bg + ib] * dt);
            al = 1.f - e;
            gsdf[c] = dt * e;
        }
        alpha[c] = al;
        if (c * 32 < n) {  // warp-uniform
            T[c] = transmittance_chunk(al, lane, carry);
        } else {
            T[c] = 0.f;
        }
        if (!BWD && i < n) {
            const float w = al * T[c];
            const float* col = i < n_obj ? A.rgb_obj + (ray * n_obj + i) * 3 : A.rgb_bg + (ray * n_bg + i - n_obj) * 3;
            acc0 = fmaf(w, col[0], acc0), acc1 = fmaf(w, col[1], acc1), acc2 = fmaf(w, col[2], acc2);
            accw += w;
            if (weights_out) weights_out[ray * n + i] = w;
        }
    }
    if (!BWD) {
        acc0 = warp_sum(acc0), acc1 = warp_sum(acc1), acc2 = warp_sum(acc2), accw = warp_sum(accw);
        if (lane == 0) {
            rgb_out[ray * 3 + 0] = acc0, rgb_out[ray * 3 + 1] = acc1, rgb_out[ray * 3 + 2] = acc2;
            if (opacity_out) opacity_out[ray] = accw;
        }
        return;
    }
    // ---- backward: dw_i = <d_rgb, c_i> (+ d_weights_i); reverse affine scan; chain into the alpha inputs.
    const float g0 = d_rgb[ray * 3 + 0], g1 = d_rgb[ray * 3 + 1], g2 = d_rgb[ray * 3 + 2];
    float q = 0.f, ds_acc = 0.f;
#pragma unroll
    for (int c = MAXC - 1; c >= 0; --c) {
        if (c * 32 >= n) continue;  // warp-uniform
        const int i = c * 32 + lane;
        float dw = 0.f;
        if (i < n) {
            const bool obj = i < n_obj;
            const int64_t s = obj ? ray * n_obj + i : ray * n_bg + (i - n_obj);
            const float* col = (obj ? A.rgb_obj : A.rgb_bg) + s * 3;
            dw = g0 * col[0] + g1 * col[1] + g2 * col[2];
            if (d_weights) dw += d_weights[ray * n + i];
            const float w = alpha[c] * T[c];
            float* dcol = (obj ? d_rgb_obj : d_rgb_bg) + s * 3;
            dcol[0] = g0 * w, dcol[1] = g1 * w, dcol[2] = g2 * w;
        }
        const float qi = adjoint_chunk(alpha[c], dw, lane, q);
        if (i < n) {
            const float da = T[c] * (dw - qi);
            if (i < n_obj) {
                const int64_t s = ray * n_obj + i;
                d_sdf[s] = da * gsdf[c];
                const float gc = da * gcos[c];
                d_grad[s * 3 + 0] = gc * r0, d_grad[s * 3 + 1] = gc * r1, d_grad[s * 3 + 2] = gc * r2;
                ds_acc = fmaf(da, ginv[c], ds_acc);
            } else {
                d_dens_bg[ray * n_bg + (i - n_obj)] = da * gsdf[c];
            }
        }
    }
    ds_acc = warp_sum(ds_acc) * inv_s;
    if (lane == 0 && ds_acc != 0.f) atomicAdd(d_s_var, ds_acc);
}

// ------------------------------------------------------------------------------------------
// The same renderer with every ray's rows STAGED IN SHARED MEMORY by bulk asynchronous copies (cp.async.bulk,
// mbarrier completion): the direct kernel above reads nine lane-strided 4-byte streams per warp (gradient and colour
// rows at a 12-byte lane stride) and reached 0.48 / 0.39 of the HBM peak (profiles/r01_render_ncu_summary.md: DRAM 43 %
// busy, long_scoreboard).  Here one lane requests the ray's rows as a handful of contiguous bulk copies, the next
// ray's rows are in flight while the current ray is scanned (two buffers per warp), the per-sample gradients of the
// backward are collected in shared memory as well and leave as bulk stores.  Persistent CTAs; a warp owns rays
// warp_id, warp_id + total_warps, ...  Applies when every row is a multiple of 16 bytes (samples % 4 == 0) and the
// tensors are 16-byte aligned; the direct kernel remains for everything else.
// ------------------------------------------------------------------------------------------
struct StagedLayout {   // float offsets inside one warp's input buffer / output buffer
    int sdf, grad, dob, rgb, dens, dbg, rgbb, dw, in_floats;
    int o_sdf, o_grad, o_rgb, o_dens, o_rgbb, out_floats;
};

__host__ __device__ inline StagedLayout staged_layout(int n_obj, int n_bg, bool bwd, bool with_dw) {
    StagedLayout L;
    int o = 0;
    L.sdf = o, o += n_obj;
    L.grad = o, o += 3 * n_obj;
    L.dob = o, o += n_obj;
    L.rgb = o, o += 3 * n_obj;
    L.dens = o, o += n_bg;
    L.dbg = o, o += n_bg;
    L.rgbb = o, o += 3 * n_bg;
    L.dw = o, o += (bwd && with_dw) ? n_obj + n_bg : 0;
    L.in_floats = o;
    // the backward's per-sample gradients overwrite the rows they are the gradients of (same lane, read before write)
    L.o_sdf = L.sdf, L.o_grad = L.grad, L.o_rgb = L.rgb, L.o_dens = L.dens, L.o_rgbb = L.rgbb;
    L.out_floats = 0;
    return L;
}

__device__ __forceinline__ void bulk_store_s2g(void* dst, uint32_t src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes)
                 : "memory");
}

template <int MAXC, bool BWD>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
neus_render_staged_kernel(const RenderArgs A, float* __restrict__ rgb_out, float* __restrict__ weights_out,
                          float* __restrict__ opacity_out, const float* __restrict__ d_rgb,
                          const float* __restrict__ d_weights, float* __restrict__ d_sdf, float* __restrict__ d_grad,
                          float* __restrict__ d_s_var, float* __restrict__ d_rgb_obj, float* __restrict__ d_dens_bg,
                          float* __restrict__ d_rgb_bg) {
    using namespace na_tc;
    extern __shared__ __align__(128) float smem_f[];
    __shared__ uint64_t bars[kWarpsPerCta][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_obj = A.n_obj, n_bg = A.n_bg, n = n_obj + n_bg;
    const StagedLayout L = staged_layout(n_obj, n_bg, BWD, d_weights != nullptr);
    float* in_buf[2] = {smem_f + (size_t)warp * (2 * L.in_floats + L.out_floats),
                        smem_f + (size_t)warp * (2 * L.in_floats + L.out_floats) + L.in_floats};
    const uint32_t bar[2] = {smem_u32(&bars[warp][0]), smem_u32(&bars[warp][1])};
    if (lane == 0) {
        mbar_init(bar[0], 1);
        mbar_init(bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const float inv_s = expf(__ldg(A.s_var));
    const float anneal = __ldg(A.anneal_ratio);
    const int64_t warp_stride = (int64_t)gridDim.x * kWarpsPerCta;
    const int64_t first = (int64_t)blockIdx.x * kWarpsPerCta + warp;

    auto request = [&](int64_t ray, int b) {   // lane 0: all rows of `ray` -> buffer b
        if (lane != 0 || ray >= A.n_rays) return;
        const uint32_t bytes = (uint32_t)L.in_floats * 4u;
        mbar_expect_tx(bar[b], bytes);
        const uint32_t base = smem_u32(in_buf[b]);
        bulk_copy_g2s(base + L.sdf * 4, A.sdf + ray * n_obj, n_obj * 4, bar[b]);
        bulk_copy_g2s(base + L.grad * 4, A.grad + ray * n_obj * 3, n_obj * 12, bar[b]);
        bulk_copy_g2s(base + L.dob * 4, A.dists_obj + ray * n_obj, n_obj * 4, bar[b]);
        bulk_copy_g2s(base + L.rgb * 4, A.rgb_obj + ray * n_obj * 3, n_obj * 12, bar[b]);
        if (n_bg) {
            bulk_copy_g2s(base + L.dens * 4, A.dens_bg + ray * n_bg, n_bg * 4, bar[b]);
            bulk_copy_g2s(base + L.dbg * 4, A.dists_bg + ray * n_bg, n_bg * 4, bar[b]);
            bulk_copy_g2s(base + L.rgbb * 4, A.rgb_bg + ray * n_bg * 3, n_bg * 12, bar[b]);
        }
        if (BWD && d_weights) bulk_copy_g2s(base + L.dw * 4, d_weights + ray * n, n * 4, bar[b]);
    };

    request(first, 0);
    uint32_t it = 0;
    float ds_total = 0.f;
    for (int64_t ray = first; ray < A.n_rays; ray += warp_stride, ++it) {
        const int b = it & 1;
        // the other buffer was consumed one iteration ago; in the backward its bulk stores must have read it out too
        if (BWD && lane == 0 && it > 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        request(ray + warp_stride, b ^ 1);
        const float far = A.dist_far ? __ldg(A.dist_far + ray) : 1e10f;
        const float r0 = __ldg(A.ray_unit + ray * 3 + 0), r1 = __ldg(A.ray_unit + ray * 3 + 1),
                    r2 = __ldg(A.ray_unit + ray * 3 + 2);
        mbar_wait(bar[b], (it >> 1) & 1);
        float* S = in_buf[b];
        float* out_buf = S;
        const float* dob = S + L.dob;
        const float* dbg = S + L.dbg;
        float alpha[MAXC], T[MAXC];
        float gsdf[MAXC], gcos[MAXC], ginv[MAXC];
        float carry = 1.f;
        float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, accw = 0.f;
#pragma unroll
        for (int c = 0; c < MAXC; ++c) {
            const int i = c * 32 + lane;
            float al = 0.f;
            gsdf[c] = gcos[c] = ginv[c] = 0.f;
            if (i < n_obj) {
                const float* gr = S + L.grad + i * 3;
                const float tc = r0 * gr[0] + r1 * gr[1] + r2 * gr[2];
                const NeusAlpha na_ = neus_alpha<BWD>(S[L.sdf + i], tc, interval(dob, i, n_obj, far), inv_s, anneal);
                al = na_.alpha;
                gsdf[c] = na_.d_sdf, gcos[c] = na_.d_cos, ginv[c] = na_.d_invs;
            } else if (i < n) {
                const int ib = i - n_obj;
                const float dt = interval(dbg, ib, n_bg, 1e10f);
                const float e = fast_exp(-S[L.dens + ib] * dt);
                al = 1.f - e;
                gsdf[c] = dt * e;
            }
            alpha[c] = al;
            T[c] = (c * 32 < n) ? transmittance_chunk(al, lane, carry) : 0.f;
            if (!BWD && i < n) {
                const float w = al * T[c];
                const float* col = i < n_obj ? S + L.rgb + i * 3 : S + L.rgbb + (i - n_obj) * 3;
                acc0 = fmaf(w, col[0], acc0), acc1 = fmaf(w, col[1], acc1), acc2 = fmaf(w, col[2], acc2);
                accw += w;
                if (weights_out) weights_out[ray * n + i] = w;
            }
        }
        if (!BWD) {
            acc0 = warp_sum(acc0), acc1 = warp_sum(acc1), acc2 = warp_sum(acc2), accw = warp_sum(accw);
            if (lane == 0) {
                rgb_out[ray * 3 + 0] = acc0, rgb_out[ray * 3 + 1] = acc1, rgb_out[ray * 3 + 2] = acc2;
                if (opacity_out) opacity_out[ray] = accw;
            }
            __syncwarp();
            continue;
        }
        // ---- backward ----
        const float g0 = __ldg(d_rgb + ray * 3 + 0), g1 = __ldg(d_rgb + ray * 3 + 1), g2 = __ldg(d_rgb + ray * 3 + 2);
        float q = 0.f, ds_acc = 0.f;
#pragma unroll
        for (int c = MAXC - 1; c >= 0; --c) {
            if (c * 32 >= n) continue;  // warp-uniform
            const int i = c * 32 + lane;
            float dw = 0.f;
            if (i < n) {
                const bool obj = i < n_obj;
                const float* col = obj ? S + L.rgb + i * 3 : S + L.rgbb + (i - n_obj) * 3;
                dw = g0 * col[0] + g1 * col[1] + g2 * col[2];
                if (d_weights) dw += S[L.dw + i];
                const float w = alpha[c] * T[c];
                float* dcol = obj ? out_buf + L.o_rgb + i * 3 : out_buf + L.o_rgbb + (i - n_obj) * 3;
                dcol[0] = g0 * w, dcol[1] = g1 * w, dcol[2] = g2 * w;
            }
            const float qi = adjoint_chunk(alpha[c], dw, lane, q);
            if (i < n) {
                const float da = T[c] * (dw - qi);
                if (i < n_obj) {
                    out_buf[L.o_sdf + i] = da * gsdf[c];
                    const float gc = da * gcos[c];
                    float* dg = out_buf + L.o_grad + i * 3;
                    dg[0] = gc * r0, dg[1] = gc * r1, dg[2] = gc * r2;
                    ds_acc = fmaf(da, ginv[c], ds_acc);
                } else {
                    out_buf[L.o_dens + (i - n_obj)] = da * gsdf[c];
                }
            }
        }
        ds_total += ds_acc;
        fence_proxy_async();        // this lane's shared-memory writes become visible to the bulk-copy engine
        __syncwarp();
        if (lane == 0) {
            const uint32_t ob = smem_u32(out_buf);
            bulk_store_s2g(d_sdf + ray * n_obj, ob + L.o_sdf * 4, n_obj * 4);
            bulk_store_s2g(d_grad + ray * n_obj * 3, ob + L.o_grad * 4, n_obj * 12);
            bulk_store_s2g(d_rgb_obj + ray * n_obj * 3, ob + L.o_rgb * 4, n_obj * 12);
            if (n_bg) {
                bulk_store_s2g(d_dens_bg + ray * n_bg, ob + L.o_dens * 4, n_bg * 4);
                bulk_store_s2g(d_rgb_bg + ray * n_bg * 3, ob + L.o_rgbb * 4, n_bg * 12);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (BWD) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        ds_total = warp_sum(ds_total) * inv_s;
        if (lane == 0 && ds_total != 0.f) atomicAdd(d_s_var, ds_total);
    }
}

inline unsigned ray_blocks(int64_t n_rays) { return (unsigned)((n_rays + kWarpsPerCta - 1) / kWarpsPerCta); }

int check_rays(int64_t n_rays, int n, const char* who) {
    NA_REQUIRE(n_rays >= 0 && n_rays <= (int64_t)0x7fffffff * kWarpsPerCta, "%s: n_rays out of range", who);
    NA_REQUIRE(n >= 1, "%s: n_samples must be >= 1", who);
    return NA_OK;
}

}  // namespace

#define NA_RAYS_PROLOGUE(who, n)                    \
    na::clear_error();                              \
    if (int rc = check_rays(n_rays, n, who)) return rc; \
    if (n_rays == 0) return NA_OK;                  \
    cudaStream_t s = (cudaStream_t)stream;

extern "C" int na_alpha_weights_fwd(const float* alphas, int64_t n_rays, int n_samples, float* weights, void* stream) {
    NA_RAYS_PROLOGUE("alpha_weights_fwd", n_samples)
    NA_REQUIRE(alphas && weights, "alpha_weights_fwd: null pointer");
    alpha_weights_fwd_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(alphas, n_rays, n_samples, weights);
    NA_CHECK_LAUNCH("alpha_weights_fwd_kernel");
    return NA_OK;
}

extern "C" int na_alpha_weights_bwd(const float* alphas, const float* d_weights, int64_t n_rays, int n_samples,
                                    float* d_alphas, void* stream) {
    NA_RAYS_PROLOGUE("alpha_weights_bwd", n_samples)
    NA_REQUIRE(alphas && d_weights && d_alphas, "alpha_weights_bwd: null pointer");
    alpha_weights_bwd_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(alphas, d_weights, n_rays, n_samples,
                                                                             d_alphas);
    NA_CHECK_LAUNCH("alpha_weights_bwd_kernel");
    return NA_OK;
}

extern "C" int na_composite_fwd(const float* q, const float* w, int64_t n_rays, int n_samples, int k, float* out,
                                void* stream) {
    NA_RAYS_PROLOGUE("composite_fwd", n_samples)
    NA_REQUIRE(q && w && out && k >= 1, "composite_fwd: null pointer or k < 1");
    composite_fwd_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(q, w, n_rays, n_samples, k, out);
    NA_CHECK_LAUNCH("composite_fwd_kernel");
    return NA_OK;
}

extern "C" int na_composite_bwd(const float* q, const float* w, const float* d_out, int64_t n_rays, int n_samples,
                                int k, float* d_q, float* d_w, void* stream) {
    NA_RAYS_PROLOGUE("composite_bwd", n_samples)
    NA_REQUIRE(q && w && d_out && k >= 1, "composite_bwd: null pointer or k < 1");
    composite_bwd_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(q, w, d_out, n_rays, n_samples, k, d_q, d_w);
    NA_CHECK_LAUNCH("composite_bwd_kernel");
    return NA_OK;
}

extern "C" int na_density_alphas_fwd(const float* dens, const float* dists, const float* dist_far, int64_t n_rays,
                                     int n_samples, float* alphas, void* stream) {
    NA_RAYS_PROLOGUE("density_alphas_fwd", n_samples)
    NA_REQUIRE(dens && dists && alphas, "density_alphas_fwd: null pointer");
    density_alphas_kernel<false><<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(dens, dists, dist_far, nullptr,
                                                                                 n_rays, n_samples, alphas);
    NA_CHECK_LAUNCH("density_alphas_kernel");
    return NA_OK;
}

extern "C" int na_density_alphas_bwd(const float* dens, const float* dists, const float* dist_far,
                                     const float* d_alphas, int64_t n_rays, int n_samples, float* d_dens,
                                     void* stream) {
    NA_RAYS_PROLOGUE("density_alphas_bwd", n_samples)
    NA_REQUIRE(dens && dists && d_alphas && d_dens, "density_alphas_bwd: null pointer");
    density_alphas_kernel<true><<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(dens, dists, dist_far, d_alphas,
                                                                                n_rays, n_samples, d_dens);
    NA_CHECK_LAUNCH("density_alphas_kernel");
    return NA_OK;
}

extern "C" int na_neus_alphas_fwd(const float* sdf, const float* grad, const float* dists, const float* ray_unit,
                                  const float* dist_far, const float* s_var, const float* anneal_ratio, int64_t n_rays,
                                  int n_samples, float* alphas, void* stream) {
    NA_RAYS_PROLOGUE("neus_alphas_fwd", n_samples)
    NA_REQUIRE(sdf && grad && dists && ray_unit && s_var && anneal_ratio && alphas, "neus_alphas_fwd: null pointer");
    neus_alphas_kernel<false><<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(
        sdf, grad, dists, ray_unit, dist_far, s_var, anneal_ratio, nullptr, n_rays, n_samples, alphas, nullptr,
        nullptr, nullptr);
    NA_CHECK_LAUNCH("neus_alphas_kernel");
    return NA_OK;
}

extern "C" int na_neus_alphas_bwd(const float* sdf, const float* grad, const float* dists, const float* ray_unit,
                                  const float* dist_far, const float* s_var, const float* anneal_ratio,
                                  const float* d_alphas, int64_t n_rays, int n_samples, float* d_sdf, float* d_grad,
                                  float* d_s_var, void* stream) {
    NA_RAYS_PROLOGUE("neus_alphas_bwd", n_samples)
    NA_REQUIRE(sdf && grad && dists && ray_unit && s_var && anneal_ratio && d_alphas && d_sdf && d_grad && d_s_var,
               "neus_alphas_bwd: null pointer");
    neus_alphas_kernel<true><<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(
        sdf, grad, dists, ray_unit, dist_far, s_var, anneal_ratio, d_alphas, n_rays, n_samples, nullptr, d_sdf,
        d_grad, d_s_var);
    NA_CHECK_LAUNCH("neus_alphas_kernel");
    return NA_OK;
}

extern "C" int na_sample_pdf(const float* bins, const float* weights, const float* unif, int64_t n_rays,
                             int n_bins, int n_fine, float* out, void* stream) {
    NA_RAYS_PROLOGUE("sample_pdf", n_bins)
    NA_REQUIRE(bins && weights && unif && out, "sample_pdf: null pointer");
    NA_REQUIRE(n_bins >= 2 && n_bins <= kMaxBins, "sample_pdf: n_bins %d out of range [2,%d]", n_bins, kMaxBins);
    NA_REQUIRE(n_fine >= 1, "sample_pdf: n_fine must be >= 1");
    const size_t smem = (size_t)kWarpsPerCta * n_bins * sizeof(float);
    sample_pdf_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, smem, s>>>(bins, weights, unif, n_rays, n_bins, n_fine,
                                                                         out);
    NA_CHECK_LAUNCH("sample_pdf_kernel");
    return NA_OK;
}

extern "C" int na_neus_upsample(const float* dists, const float* sdfs, const float* unif, float inv_s,
                                int64_t n_rays, int n_samples, int n_fine, float* dists_fine, void* stream) {
    NA_RAYS_PROLOGUE("neus_upsample", n_samples)
    NA_REQUIRE(dists && sdfs && unif && dists_fine, "neus_upsample: null pointer");
    NA_REQUIRE(n_samples >= 2 && n_samples <= kMaxBins / 2 * 2 && n_samples * 2 * kWarpsPerCta * 4 <= 48 * 1024,
               "neus_upsample: n_samples %d out of range [2,%d]", n_samples, 48 * 1024 / (8 * kWarpsPerCta));
    NA_REQUIRE(n_fine >= 1, "neus_upsample: n_fine must be >= 1");
    const size_t smem = (size_t)kWarpsPerCta * 2 * n_samples * sizeof(float);
    neus_upsample_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, smem, s>>>(dists, sdfs, unif, inv_s, n_rays,
                                                                            n_samples, n_fine, dists_fine);
    NA_CHECK_LAUNCH("neus_upsample_kernel");
    return NA_OK;
}

extern "C" int na_merge_sorted(const float* dists, const float* sdfs, const float* dists_fine,
                               const float* sdfs_fine, int64_t n_rays, int n_samples, int n_fine, float* dists_out,
                               float* sdfs_out, void* stream) {
    NA_RAYS_PROLOGUE("merge_sorted", n_samples)
    NA_REQUIRE(dists && dists_fine && dists_out && n_fine >= 1, "merge_sorted: null pointer or n_fine < 1");
    NA_REQUIRE(!sdfs_out || (sdfs && sdfs_fine), "merge_sorted: sdfs_out given without sdfs/sdfs_fine");
    merge_sorted_kernel<<<ray_blocks(n_rays), kWarpsPerCta * 32, 0, s>>>(dists, sdfs, dists_fine, sdfs_fine, n_rays,
                                                                        n_samples, n_fine, dists_out, sdfs_out);
    NA_CHECK_LAUNCH("merge_sorted_kernel");
    return NA_OK;
}

namespace {
template <bool BWD>
int launch_render(const RenderArgs& A, float* rgb, float* weights, float* opacity, const float* d_rgb,
                  const float* d_weights, float* d_sdf, float* d_grad, float* d_s_var, float* d_rgb_obj,
                  float* d_dens_bg, float* d_rgb_bg, cudaStream_t s) {
    const int n = A.n_obj + A.n_bg;
    {   // staged form: every row a multiple of 16 bytes at 16-byte aligned addresses, enough rays to fill the GPU
        auto al16 = [](const void* p) { return p == nullptr || (uintptr_t)p % 16 == 0; };
        const bool rows_ok = A.n_obj % 4 == 0 && A.n_bg % 4 == 0 && n <= 256;
        const bool ptrs_ok = al16(A.sdf) && al16(A.grad) && al16(A.dists_obj) && al16(A.rgb_obj) && al16(A.dens_bg) &&
                             al16(A.dists_bg) && al16(A.rgb_bg) && al16(d_weights) && al16(d_sdf) && al16(d_grad) &&
                             al16(d_rgb_obj) && al16(d_dens_bg) && al16(d_rgb_bg);
        const StagedLayout L = staged_layout(A.n_obj, A.n_bg, BWD, d_weights != nullptr);
        const size_t smem = (size_t)kWarpsPerCta * (2 * L.in_floats + L.out_floats) * sizeof(float);
        if (rows_ok && ptrs_ok && smem <= 200 * 1024 && A.n_rays >= 4096) {
            int per_sm = (int)((220 * 1024) / (smem + 1024));
            if (per_sm > 4) per_sm = 4;
            if (per_sm < 1) per_sm = 1;
            int64_t grid = (int64_t)na::kNumSMs * per_sm;
            const int64_t need = (A.n_rays + kWarpsPerCta - 1) / kWarpsPerCta;
            if (grid > need) grid = need;
#define NA_RENDER_STAGED(MAXC)                                                                                        \
    do {                                                                                                              \
        auto kern = neus_render_staged_kernel<MAXC, BWD>;                                                             \
        if (int rc = na::check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), \
                                    "neus_render_staged smem"))                                                       \
            return rc;                                                                                                \
        kern<<<(unsigned)grid, kWarpsPerCta * 32, smem, s>>>(A, rgb, weights, opacity, d_rgb, d_weights, d_sdf, d_grad, \
                                                             d_s_var, d_rgb_obj, d_dens_bg, d_rgb_bg);                \
    } while (0)
            if (n <= 64) NA_RENDER_STAGED(2);
            else if (n <= 128) NA_RENDER_STAGED(4);
            else if (n <= 192) NA_RENDER_STAGED(6);
            else NA_RENDER_STAGED(8);
#undef NA_RENDER_STAGED
            NA_CHECK_LAUNCH("neus_render_staged_kernel");
            return NA_OK;
        }
    }
    const unsigned blocks = ray_blocks(A.n_rays);
#define NA_RENDER(MAXC)                                                                                           \
    neus_render_kernel<MAXC, BWD><<<blocks, kWarpsPerCta * 32, 0, s>>>(A, rgb, weights, opacity, d_rgb, d_weights, \
                                                                       d_sdf, d_grad, d_s_var, d_rgb_obj,         \
                                                                       d_dens_bg, d_rgb_bg)
    if (n <= 64) NA_RENDER(2);
    else if (n <= 128) NA_RENDER(4);
    else if (n <= 192) NA_RENDER(6);
    else if (n <= 256) NA_RENDER(8);
    else NA_RENDER(16);
#undef NA_RENDER
    NA_CHECK_LAUNCH("neus_render_kernel");
    return NA_OK;
}

int fill_render_args(RenderArgs& A, const float* sdf, const float* grad, const float* dists_obj,
                     const float* ray_unit, const float* dist_far, const float* s_var, const float* anneal_ratio,
                     const float* rgb_obj, int n_obj, const float* dens_bg, const float* dists_bg,
                     const float* rgb_bg, int n_bg, int64_t n_rays) {
    NA_REQUIRE(n_obj >= 1 && n_bg >= 0 && n_obj + n_bg <= 512, "neus_render: n_obj=%d n_bg=%d unsupported (total <= 512)",
               n_obj, n_bg);
    NA_REQUIRE(sdf && grad && dists_obj && ray_unit && s_var && anneal_ratio && rgb_obj, "neus_render: null object pointer");
    NA_REQUIRE(n_bg == 0 || (dens_bg && dists_bg && rgb_bg), "neus_render: null background pointer");
    A = RenderArgs{sdf, grad, dists_obj, ray_unit, dist_far, s_var, rgb_obj, dens_bg, dists_bg, rgb_bg,
                   anneal_ratio, n_obj, n_bg, n_rays};
    return NA_OK;
}
}  // namespace

extern "C" int na_neus_render_fwd(const float* sdf, const float* grad, const float* dists_obj, const float* ray_unit,
                                  const float* dist_far, const float* s_var, const float* anneal_ratio,
                                  const float* rgb_obj, int n_obj, const float* dens_bg, const float* dists_bg,
                                  const float* rgb_bg, int n_bg, int64_t n_rays, float* rgb, float* weights,
                                  float* opacity, void* stream) {
    NA_RAYS_PROLOGUE("neus_render_fwd", n_obj)
    NA_REQUIRE(rgb != nullptr, "neus_render_fwd: rgb is null");
    RenderArgs A;
    if (int rc = fill_render_args(A, sdf, grad, dists_obj, ray_unit, dist_far, s_var, anneal_ratio, rgb_obj, n_obj,
                                  dens_bg, dists_bg, rgb_bg, n_bg, n_rays))
        return rc;
    return launch_render<false>(A, rgb, weights, opacity, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                nullptr, nullptr, s);
}

extern "C" int na_neus_render_bwd(const float* sdf, const float* grad, const float* dists_obj, const float* ray_unit,
                                  const float* dist_far, const float* s_var, const float* anneal_ratio,
                                  const float* rgb_obj, int n_obj, const float* dens_bg, const float* dists_bg,
                                  const float* rgb_bg, int n_bg, int64_t n_rays, const float* d_rgb,
                                  const float* d_weights, float* d_sdf, float* d_grad, float* d_s_var,
                                  float* d_rgb_obj, float* d_dens_bg, float* d_rgb_bg, void* stream) {
    NA_RAYS_PROLOGUE("neus_render_bwd", n_obj)
    NA_REQUIRE(d_rgb && d_sdf && d_grad && d_s_var && d_rgb_obj, "neus_render_bwd: null pointer");
    NA_REQUIRE(n_bg == 0 || (d_dens_bg && d_rgb_bg), "neus_render_bwd: null background gradient pointer");
    RenderArgs A;
    if (int rc = fill_render_args(A, sdf, grad, dists_obj, ray_unit, dist_far, s_var, anneal_ratio, rgb_obj, n_obj,
                                  dens_bg, dists_bg, rgb_bg, n_bg, n_rays))
        return rc;
    return launch_render<true>(A, nullptr, nullptr, nullptr, d_rgb, d_weights, d_sdf, d_grad, d_s_var, d_rgb_obj,
                               d_dens_bg, d_rgb_bg, s);
}
