// One-launch operand packing for the tensor-core MLP kernels: an fp32 matrix -> the two fp16 pieces (hi, lo * 2048) in
// the K-chunk-major UMMA layout  out[(k/8)][(j/8)][j%8][k%8]  for logical W[j, k], j < n, K zero-padded to k_pad.
// Replaces a dozen small ATen kernels per packed operand (zeros / copy / clamp / half / sub / mul / half / permute /
// contiguous), six operands per training step for the SDF network alone.
#include "na_common.cuh"
#include <cuda_fp16.h>

namespace {

struct PackArgs {
    const float* w;
    int64_t ld_w;
    int n, n_valid, k, transposed, col_rot;
    double scale;
    const float* extra_col;
    int k_pad;
    __half *hi, *lo;
};

__global__ void pack_split_f16_kernel(const PackArgs A) {
    const int chunks = A.k_pad / 8;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;   // (k-chunk, row) with the row fastest: 128-byte runs of output
    if (t >= chunks * A.n) return;
    const int kc = t / A.n, j = t - kc * A.n;
    __align__(16) __half h[8], l[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int kd = kc * 8 + e;
        float v = 0.f;
        if (j >= A.n_valid) {
            v = 0.f;
        } else if (kd < A.k) {
            int c = kd + A.col_rot;
            if (c >= A.k) c -= A.k;
            const float s = A.transposed ? A.w[(int64_t)c * A.ld_w + j] : A.w[(int64_t)j * A.ld_w + c];
            v = A.scale == 1.0 ? s : (float)((double)s * A.scale);
        } else if (kd == A.k && A.extra_col) {
            v = (float)((double)A.extra_col[j] * A.scale);
        }
        v = fminf(fmaxf(v, -65504.f), 65504.f);
        h[e] = __float2half_rn(v);
        l[e] = __float2half_rn((v - __half2float(h[e])) * 2048.f);
    }
    const size_t off = (((size_t)kc * (A.n / 8) + j / 8) * 8 + j % 8) * 8;
    *reinterpret_cast<uint4*>(A.hi + off) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(A.lo + off) = *reinterpret_cast<const uint4*>(l);
}

}  // namespace

extern "C" int na_pack_split_f16(const float* w, int64_t ld_w, int n, int n_valid, int k, int transposed, int col_rot, double scale,
                                 const float* extra_col, int k_pad, void* hi, void* lo, void* stream) {
    na::clear_error();
    NA_REQUIRE(w && hi && lo, "pack_split_f16: null pointer");
    NA_REQUIRE(n > 0 && n % 8 == 0 && k > 0 && k_pad % 16 == 0 && k + (extra_col ? 1 : 0) <= k_pad,
               "pack_split_f16: n=%d must be a multiple of 8, k=%d (+1 with an extra column) must fit k_pad=%d (multiple of 16)",
               n, k, k_pad);
    NA_REQUIRE(n_valid >= 0 && n_valid <= n, "pack_split_f16: n_valid out of range");
    NA_REQUIRE(col_rot >= 0 && col_rot < k, "pack_split_f16: col_rot out of range");
    NA_REQUIRE(ld_w >= (transposed ? n_valid : k), "pack_split_f16: ld_w too small");
    NA_REQUIRE(((uintptr_t)hi | (uintptr_t)lo) % 16 == 0, "pack_split_f16: outputs must be 16-byte aligned");
    PackArgs A{w, ld_w, n, n_valid, k, transposed, col_rot, scale, extra_col, k_pad, (__half*)hi, (__half*)lo};
    const int total = k_pad / 8 * n;
    pack_split_f16_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(A);
    NA_CHECK_LAUNCH("pack_split_f16_kernel");
    return NA_OK;
}
