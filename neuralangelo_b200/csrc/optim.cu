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
