// Fused AdamW step over a flat fp32 parameter buffer, with gradient averaging and gradient zeroing folded in.
//
// Replaces, for the path's parameters, torch.optim.AdamW as the reference configures it
// (imaginaire/trainers/utils/get_trainer.py:106-150, base.yaml:117-121: lr 1e-3, weight_decay 1e-2) followed by
// optimizer.zero_grad (imaginaire/trainers/base.py:454-457).  The 366 M-entry hash table makes this a pure
// HBM-streaming pass: 4 reads (p, g, m, v) + 4 writes (p, m, v, g = 0) of 4 bytes per parameter; separate
// optimizer + zero_grad passes would read/write 11 arrays.  Each thread handles 4 consecutive floats
// (16-byte vectors), grid-stride over the buffer.
#include "na_common.cuh"

namespace {

struct AdamArgs {
    float lr, beta1, beta2, eps, weight_decay;
    float bias_correction1;       // 1 - beta1^t
    float inv_sqrt_bias_correction2;  // 1 / sqrt(1 - beta2^t)
    float grad_scale;             // multiplies g before use (1/world_size for summed gradients, 1 otherwise)
    int zero_grad;
};

__device__ __forceinline__ void adamw1(float& p, float& g, float& m, float& v, const AdamArgs& a) {
    const float gs = g * a.grad_scale;
    p *= 1.f - a.lr * a.weight_decay;                 // decoupled weight decay
    m = a.beta1 * m + (1.f - a.beta1) * gs;
    v = a.beta2 * v + (1.f - a.beta2) * gs * gs;
    const float denom = sqrtf(v) * a.inv_sqrt_bias_correction2 + a.eps;
    p -= (a.lr / a.bias_correction1) * (m / denom);
    if (a.zero_grad) g = 0.f;
}

__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
             AdamArgs a) {
    const int64_t n4 = n / 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 P = reinterpret_cast<float4*>(p)[i], G = reinterpret_cast<float4*>(g)[i];
        float4 M = reinterpret_cast<float4*>(m)[i], V = reinterpret_cast<float4*>(v)[i];
        adamw1(P.x, G.x, M.x, V.x, a);
        adamw1(P.y, G.y, M.y, V.y, a);
        adamw1(P.z, G.z, M.z, V.z, a);
        adamw1(P.w, G.w, M.w, V.w, a);
        reinterpret_cast<float4*>(p)[i] = P;
        reinterpret_cast<float4*>(m)[i] = M;
        reinterpret_cast<float4*>(v)[i] = V;
        if (a.zero_grad) reinterpret_cast<float4*>(g)[i] = G;
    }
    // tail (n not a multiple of 4)
    const int64_t t = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) adamw1(p[t], g[t], m[t], v[t], a);
}

__global__ void __launch_bounds__(256) scale_kernel(float* __restrict__ p, int64_t n, float factor) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] *= factor;
}

// ---------------------------------------------------------------------------------------------
// Data-parallel table step as ONE kernel over NVLink multicast (NVLS): reduce-scatter + AdamW + all-gather.
// Every rank owns a slice [lo, lo + n) of the active table prefix.  For its slice a rank
//   * reads the SUM of all ranks' gradients with multimem.ld_reduce (the NVSwitch adds the replicas in flight),
//   * updates its parameters / moments (same arithmetic as adamw1 above, grad_scale = 1 / world),
//   * writes the new parameters to EVERY rank with multimem.st.
// Gradients are NOT cleared here (zeroing every replica through the multicast address would double the wire bytes --
// measured: 5.75 ms instead of ~3 ms on two GPUs); each rank clears its own buffer after the closing barrier.
// Wire traffic per rank: 1/W of the gradient in + (W-1)/W out (peers pulling), 1/W of the parameters out +
// (W-1)/W in -- both directions of the links are busy at once, where reduce-scatter followed by all-gather
// uses them one after the other.  The caller brackets the launch with cross-rank barriers (all gradients
// complete before; all parameters landed after).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
__device__ __forceinline__ void mc_st_f4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

// kMcUnroll: multicast loads in flight per thread (the switch round trip is several microseconds)
template <int kMcUnroll>
__global__ void __launch_bounds__(256)
adamw_multimem_kernel(float* __restrict__ mc_g, float* __restrict__ mc_p, const float* __restrict__ p_local,
                      float* __restrict__ m, float* __restrict__ v, int64_t lo, int64_t n, AdamArgs a) {
    const int64_t n4 = n / 4;   // the host guarantees n % 4 == 0 and 16-byte alignment of lo
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * kMcUnroll) {
        float4 G[kMcUnroll], P[kMcUnroll], M[kMcUnroll], V[kMcUnroll];
#pragma unroll
        for (int u = 0; u < kMcUnroll; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) G[u] = mc_ld_reduce_f4(mc_g + lo + i * 4);
        }
#pragma unroll
        for (int u = 0; u < kMcUnroll; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) {
                const int64_t e = lo + i * 4;
                P[u] = *reinterpret_cast<const float4*>(p_local + e);
                M[u] = *reinterpret_cast<float4*>(m + e), V[u] = *reinterpret_cast<float4*>(v + e);
            }
        }
#pragma unroll
        for (int u = 0; u < kMcUnroll; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) {
                const int64_t e = lo + i * 4;
                adamw1(P[u].x, G[u].x, M[u].x, V[u].x, a);
                adamw1(P[u].y, G[u].y, M[u].y, V[u].y, a);
                adamw1(P[u].z, G[u].z, M[u].z, V[u].z, a);
                adamw1(P[u].w, G[u].w, M[u].w, V[u].w, a);
                mc_st_f4(mc_p + e, P[u]);
                *reinterpret_cast<float4*>(m + e) = M[u];
                *reinterpret_cast<float4*>(v + e) = V[u];
            }
        }
    }
}

}  // namespace

extern "C" int na_adamw_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                             float beta1, float beta2, float eps, float weight_decay, int64_t step, float grad_scale,
                             int zero_grad, void* stream) {
    na::clear_error();
    NA_REQUIRE(n >= 0 && step >= 1, "adamw_step: bad n or step");
    if (n == 0) return NA_OK;
    NA_REQUIRE(params && grads && exp_avg && exp_avg_sq, "adamw_step: null pointer");
    NA_REQUIRE(((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) % 16 == 0,
               "adamw_step: buffers must be 16-byte aligned");
    AdamArgs a;
    a.lr = lr, a.beta1 = beta1, a.beta2 = beta2, a.eps = eps, a.weight_decay = weight_decay;
    a.bias_correction1 = (float)(1.0 - pow((double)beta1, (double)step));
    a.inv_sqrt_bias_correction2 = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)step)));
    a.grad_scale = grad_scale;
    a.zero_grad = zero_grad;
    int64_t blocks = (n / 4 + 255) / 256;
    const int64_t cap = (int64_t)na::kNumSMs * 16;  // grid-stride: 16 CTAs of 256 threads per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    adamw_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, n, a);
    NA_CHECK_LAUNCH("adamw_kernel");
    return NA_OK;
}

extern "C" int na_scale_inplace(float* params, int64_t n, float factor, void* stream) {
    na::clear_error();
    NA_REQUIRE(n >= 0 && (n == 0 || params), "scale_inplace: bad arguments");
    if (n == 0) return NA_OK;
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)na::kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    scale_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, n, factor);
    NA_CHECK_LAUNCH("scale_kernel");
    return NA_OK;
}

extern "C" int na_adamw_step_multimem(float* mc_grads, float* mc_params, const float* params_local, float* exp_avg,
                                      float* exp_avg_sq, int64_t offset, int64_t n, float lr, float beta1, float beta2,
                                      float eps, float weight_decay, int64_t step, float grad_scale, int max_ctas,
                                      int unroll, void* stream) {
    na::clear_error();
    NA_REQUIRE(n >= 0 && offset >= 0 && step >= 1, "adamw_step_multimem: bad n, offset or step");
    if (n == 0) return NA_OK;
    NA_REQUIRE(mc_grads && mc_params && params_local && exp_avg && exp_avg_sq, "adamw_step_multimem: null pointer");
    NA_REQUIRE(n % 4 == 0 && offset % 4 == 0, "adamw_step_multimem: slice must be a multiple of 4 floats");
    NA_REQUIRE(((uintptr_t)mc_grads | (uintptr_t)mc_params | (uintptr_t)params_local | (uintptr_t)exp_avg |
                (uintptr_t)exp_avg_sq) % 16 == 0,
               "adamw_step_multimem: buffers must be 16-byte aligned");
    AdamArgs a;
    a.lr = lr, a.beta1 = beta1, a.beta2 = beta2, a.eps = eps, a.weight_decay = weight_decay;
    a.bias_correction1 = (float)(1.0 - pow((double)beta1, (double)step));
    a.inv_sqrt_bias_correction2 = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)step)));
    a.grad_scale = grad_scale;
    a.zero_grad = 0;   // gradients are cleared by the caller after the closing barrier
    int64_t blocks = (n / 4 + 255) / 256;
    // one CTA per SM measured best on 8 GPUs (3.41 ms for the whole default table against 3.77 ms with eight per SM):
    // the NVSwitch, not the SMs, paces the kernel, and more requests in flight only queue
    const int64_t cap = max_ctas > 0 ? max_ctas : (int64_t)na::kNumSMs;
    if (blocks > cap) blocks = cap;
    const dim3 grid((unsigned)blocks);
    cudaStream_t s = (cudaStream_t)stream;
    switch (unroll) {
        case 2: adamw_multimem_kernel<2><<<grid, 256, 0, s>>>(mc_grads, mc_params, params_local, exp_avg, exp_avg_sq, offset, n, a); break;
        case 8: adamw_multimem_kernel<8><<<grid, 256, 0, s>>>(mc_grads, mc_params, params_local, exp_avg, exp_avg_sq, offset, n, a); break;
        default: adamw_multimem_kernel<4><<<grid, 256, 0, s>>>(mc_grads, mc_params, params_local, exp_avg, exp_avg_sq, offset, n, a); break;
    }
    NA_CHECK_LAUNCH("adamw_multimem_kernel");
    return NA_OK;
}
