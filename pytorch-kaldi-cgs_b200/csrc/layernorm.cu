// The reference's own LayerNorm (neural_networks.py:40-51), NOT nn.LayerNorm:
//     mean = x.mean(-1); std = x.std(-1)  (unbiased, n-1);  y = gamma * (x - mean) / (std + eps) + beta
// used for *_use_laynorm_inp (the layer input, all rows at once) and dnn_use_laynorm (after a linear).
// One warp per row; a row (<= a few thousand floats) stays in L1 between the passes, so HBM sees one read of x and
// one write of y.  Backward (derived from the formula above, s = std, r = 1/(s+eps), xh = (x-mean)*r, g = dy*gamma):
//     dx = r * (g - mean(g)) - xh * sum(g * xh) / (s * (n-1));   dgamma = sum_rows dy*xh;   dbeta = sum_rows dy
// dgamma / dbeta are accumulated per CTA in shared memory and added to the (zero-initialised) outputs atomically.
#include <algorithm>

#include "common.cuh"

namespace hcgs {

constexpr int LN_ROWS = 8;  // warps per CTA

__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const float* __restrict__ x, int64_t ldx,
                                                            float* __restrict__ y, int64_t ldy, int64_t M, int n,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps,
                                                            float* __restrict__ save_mean, float* __restrict__ save_std) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t stride = (int64_t)gridDim.x * LN_ROWS;
    for (int64_t m = (int64_t)blockIdx.x * LN_ROWS + warp; m < M; m += stride) {
        const float* xr = x + m * ldx;
        float s = 0.f;
        for (int p = lane; p < n; p += 32) s += xr[p];
        const float mean = warp_sum(s) / (float)n;
        float q = 0.f;
        for (int p = lane; p < n; p += 32) {
            const float d = xr[p] - mean;
            q = fmaf(d, d, q);
        }
        const float sd = sqrtf(warp_sum(q) / (float)(n - 1));  // torch.std: unbiased; n == 1 gives nan like the reference
        const float den = sd + eps;
        float* yr = y + m * ldy;
        for (int p = lane; p < n; p += 32) yr[p] = __ldg(gamma + p) * (xr[p] - mean) / den + __ldg(beta + p);
        if (lane == 0 && save_mean) {
            save_mean[m] = mean;
            save_std[m] = sd;
        }
    }
}

__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ dy, int64_t lddy,
                                                            const float* __restrict__ x, int64_t ldx, int64_t M, int n,
                                                            const float* __restrict__ gamma, float eps,
                                                            const float* __restrict__ save_mean,
                                                            const float* __restrict__ save_std, float* __restrict__ dx,
                                                            int64_t lddx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta) {
    extern __shared__ float s_acc[];  // [2][n]: dgamma, dbeta of this CTA
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = threadIdx.x; p < 2 * n; p += 256) s_acc[p] = 0.f;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * LN_ROWS;
    for (int64_t m = (int64_t)blockIdx.x * LN_ROWS + warp; m < M; m += stride) {
        const float* xr = x + m * ldx;
        const float* dr = dy + m * lddy;
        const float mean = save_mean[m], sd = save_std[m];
        const float r = 1.f / (sd + eps);
        float sg = 0.f, sgx = 0.f;
        for (int p = lane; p < n; p += 32) {
            const float g = dr[p] * __ldg(gamma + p);
            const float xh = (xr[p] - mean) * r;
            sg += g;
            sgx = fmaf(g, xh, sgx);
            atomicAdd(s_acc + p, dr[p] * xh);      // lanes hit distinct addresses: no conflicts inside a warp
            atomicAdd(s_acc + n + p, dr[p]);
        }
        sg = warp_sum(sg) / (float)n;
        sgx = warp_sum(sgx);
        const float k2 = sd > 0.f ? sgx / (sd * (float)(n - 1)) : 0.f;
        if (dx) {
            float* o = dx + m * lddx;
            for (int p = lane; p < n; p += 32) {
                const float g = dr[p] * __ldg(gamma + p);
                const float xh = (xr[p] - mean) * r;
                o[p] = r * (g - sg) - xh * k2;
            }
        }
    }
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += 256) {
        if (dgamma) atomicAdd(dgamma + p, s_acc[p]);
        if (dbeta) atomicAdd(dbeta + p, s_acc[n + p]);
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_layernorm_fwd(const float* x, int64_t ldx, float* y, int64_t ldy, int64_t m, int n,
                                  const float* gamma, const float* beta, float eps, float* save_mean,
                                  float* save_std, hcgs_stream stream) {
    HCGS_REQUIRE(x && y && gamma && beta, "hcgs_layernorm_fwd: NULL argument");
    HCGS_REQUIRE((save_mean == nullptr) == (save_std == nullptr), "hcgs_layernorm_fwd: save_mean and save_std go together");
    if (m <= 0 || n <= 0) return 0;
    const int blocks = (int)std::min<int64_t>((m + LN_ROWS - 1) / LN_ROWS, 148 * 8);
    layernorm_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, y, ldy, m, n, gamma, beta, eps, save_mean,
                                                                   save_std);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_layernorm_bwd(const float* dy, int64_t lddy, const float* x, int64_t ldx, int64_t m, int n,
                                  const float* gamma, float eps, const float* save_mean, const float* save_std,
                                  float* dx, int64_t lddx, float* dgamma, float* dbeta, hcgs_stream stream) {
    HCGS_REQUIRE(dy && x && gamma && save_mean && save_std, "hcgs_layernorm_bwd: NULL argument");
    HCGS_REQUIRE(n <= 12000, "hcgs_layernorm_bwd: %d features exceed the shared-memory accumulators", n);
    if (m <= 0 || n <= 0) return 0;
    const int blocks = (int)std::min<int64_t>((m + LN_ROWS - 1) / LN_ROWS, 148 * 4);
    const size_t smem = (size_t)2 * n * sizeof(float);
    if (smem > 48 * 1024)
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    layernorm_bwd_kernel<<<blocks, 256, smem, (cudaStream_t)stream>>>(dy, lddy, x, ldx, m, n, gamma, eps, save_mean,
                                                                      save_std, dx, lddx, dgamma, dbeta);
    HCGS_LAUNCH_CHECK();
    return 0;
}
