// Weight-norm reparametrisation W = g * v / ||v|| (one norm per output row) and its adjoint, one launch each.
//
// Replaces torch.nn.utils.weight_norm as the reference applies it to every layer of its MLPs
// (projects/neuralangelo/utils/mlp.py:37-38, projects/nerf/utils/nerf_util.py:176-177).  In eager ATen the forward is
// three kernels per layer (row norms, g / norm, v * scale) and the backward about six; the shipped model has 17 such
// layers evaluated once per training step, i.e. ~150 few-microsecond launches -- a measurable slice of a step that is
// bound by launch count, not bytes (profiles/r01_measurements.md).  One warp owns one output row.
#include "na_common.cuh"

namespace {

constexpr int kWnWarps = 8;

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

__global__ void __launch_bounds__(kWnWarps * 32)
weight_norm_fwd_kernel(const float* __restrict__ v, const float* __restrict__ g, int rows, int cols,
                       float* __restrict__ w, float* __restrict__ inv_norm) {
    const int row = blockIdx.x * kWnWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* vr = v + (int64_t)row * cols;
    float ss = 0.f;
    for (int c = lane; c < cols; c += 32) ss = fmaf(vr[c], vr[c], ss);
    const float inv = 1.f / sqrtf(warp_sum_f(ss));
    const float scale = g[row] * inv;
    for (int c = lane; c < cols; c += 32) w[(int64_t)row * cols + c] = vr[c] * scale;
    if (lane == 0) inv_norm[row] = inv;
}

// dg[r] = <dW[r], v[r]> / ||v[r]||;   dv[r] = (g[r] / ||v[r]||) * (dW[r] - v[r] * dg[r] / ||v[r]||)
__global__ void __launch_bounds__(kWnWarps * 32)
weight_norm_bwd_kernel(const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ inv_norm,
                       const float* __restrict__ dw, int rows, int cols, float* __restrict__ dv,
                       float* __restrict__ dg) {
    const int row = blockIdx.x * kWnWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* vr = v + (int64_t)row * cols;
    const float* dr = dw + (int64_t)row * cols;
    float dot = 0.f;
    for (int c = lane; c < cols; c += 32) dot = fmaf(dr[c], vr[c], dot);
    const float inv = inv_norm[row];
    const float dgr = warp_sum_f(dot) * inv;
    const float scale = g[row] * inv, proj = dgr * inv;
    for (int c = lane; c < cols; c += 32) dv[(int64_t)row * cols + c] = scale * (dr[c] - vr[c] * proj);
    if (lane == 0) dg[row] = dgr;
}

}  // namespace

extern "C" int na_weight_norm_fwd(const float* v, const float* g, int rows, int cols, float* w, float* inv_norm,
                                  void* stream) {
    na::clear_error();
    NA_REQUIRE(rows >= 0 && cols >= 1, "weight_norm_fwd: bad shape %d x %d", rows, cols);
    if (rows == 0) return NA_OK;
    NA_REQUIRE(v && g && w && inv_norm, "weight_norm_fwd: null pointer");
    weight_norm_fwd_kernel<<<(rows + kWnWarps - 1) / kWnWarps, kWnWarps * 32, 0, (cudaStream_t)stream>>>(v, g, rows, cols,
                                                                                                      w, inv_norm);
    NA_CHECK_LAUNCH("weight_norm_fwd_kernel");
    return NA_OK;
}

extern "C" int na_weight_norm_bwd(const float* v, const float* g, const float* inv_norm, const float* dw, int rows,
                                  int cols, float* dv, float* dg, void* stream) {
    na::clear_error();
    NA_REQUIRE(rows >= 0 && cols >= 1, "weight_norm_bwd: bad shape %d x %d", rows, cols);
    if (rows == 0) return NA_OK;
    NA_REQUIRE(v && g && inv_norm && dw && dv && dg, "weight_norm_bwd: null pointer");
    weight_norm_bwd_kernel<<<(rows + kWnWarps - 1) / kWnWarps, kWnWarps * 32, 0, (cudaStream_t)stream>>>(
        v, g, inv_norm, dw, rows, cols, dv, dg);
    NA_CHECK_LAUNCH("weight_norm_bwd_kernel");
    return NA_OK;
}
