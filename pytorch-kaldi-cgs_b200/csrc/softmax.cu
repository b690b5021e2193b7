// Output-layer epilogue: row log-softmax, prior normalisation, fused NLL / error / dlogits (fp32).
//
// One warp per frame (row); a row of P = 1944 .. 3480 log-probabilities stays in L1 between the
// passes, so HBM sees one read of the logits and one write of the result.
#include "common.cuh"

namespace hcgs {

constexpr int ROWS_PER_CTA = 8;  // 8 warps

__global__ void __launch_bounds__(256) logsoftmax_fwd_kernel(const float* __restrict__ z, int ldz,
                                                             float* __restrict__ out, int ldo, int64_t M, int P,
                                                             const float* __restrict__ log_prior) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t m = (int64_t)blockIdx.x * ROWS_PER_CTA + warp;
    if (m >= M) return;
    const float* zr = z + m * ldz;
    float mx = -INFINITY;
    for (int p = lane; p < P; p += 32) mx = fmaxf(mx, zr[p]);
    mx = warp_max(mx);
    float s = 0.f;
    for (int p = lane; p < P; p += 32) s += expf(zr[p] - mx);
    s = warp_sum(s);
    const float lse = mx + logf(s);
    float* orow = out + m * ldo;
    if (log_prior)
        for (int p = lane; p < P; p += 32) orow[p] = (zr[p] - lse) - __ldg(log_prior + p);
    else
        for (int p = lane; p < P; p += 32) orow[p] = zr[p] - lse;
}

__global__ void __launch_bounds__(256) logsoftmax_bwd_kernel(const float* __restrict__ dlogp,
                                                             const float* __restrict__ logp,
                                                             float* __restrict__ dz, int64_t M, int P) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t m = (int64_t)blockIdx.x * ROWS_PER_CTA + warp;
    if (m >= M) return;
    const float* dr = dlogp + m * P;
    const float* lr = logp + m * P;
    float s = 0.f;
    for (int p = lane; p < P; p += 32) s += dr[p];
    s = warp_sum(s);
    float* o = dz + m * P;
    for (int p = lane; p < P; p += 32) o[p] = dr[p] - expf(lr[p]) * s;
}

__global__ void __launch_bounds__(256) nll_fwd_bwd_kernel(const float* __restrict__ logp, int64_t M, int P,
                                                          const float* __restrict__ labels, int ldlab,
                                                          float scale, float* __restrict__ dz,
                                                          float* __restrict__ stats) {
    __shared__ float sh_loss[ROWS_PER_CTA], sh_err[ROWS_PER_CTA];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t m = (int64_t)blockIdx.x * ROWS_PER_CTA + warp;
    float loss = 0.f, err = 0.f;
    if (m < M) {
        const float* lr = logp + m * P;
        const int label = (int)labels[m * ldlab];  // .long() of the float label column
        float best = -INFINITY;
        int besti = 0x7fffffff;
        for (int p = lane; p < P; p += 32) {
            const float v = lr[p];
            if (v > best) {
                best = v;
                besti = p;
            }
        }
        // first index of the maximum (torch.max tie rule on the reference's CPU path)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) {
                best = ob;
                besti = oi;
            }
        }
        if (label >= 0 && label < P) loss = -lr[label];
        err = (besti != label) ? 1.f : 0.f;
        if (dz) {
            float* o = dz + m * P;
            for (int p = lane; p < P; p += 32) o[p] = (expf(lr[p]) - (p == label ? 1.f : 0.f)) * scale;
        }
    }
    if (lane == 0) {
        sh_loss[warp] = loss;
        sh_err[warp] = err;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float l = 0.f, e = 0.f;
#pragma unroll
        for (int w = 0; w < ROWS_PER_CTA; ++w) {
            l += sh_loss[w];
            e += sh_err[w];
        }
        atomicAdd(stats + 0, l);
        atomicAdd(stats + 1, e);
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_logsoftmax_fwd(const float* z, int ldz, float* out, int ldo, int64_t m, int p,
                                   const float* log_prior, hcgs_stream stream) {
    HCGS_REQUIRE(z && out, "hcgs_logsoftmax_fwd: NULL argument");
    if (m <= 0 || p <= 0) return 0;
    const int64_t blocks = (m + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    logsoftmax_fwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(z, ldz, out, ldo, m, p, log_prior);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_logsoftmax_bwd(const float* dlogp, const float* logp, float* dz, int64_t m, int p,
                                   hcgs_stream stream) {
    HCGS_REQUIRE(dlogp && logp && dz, "hcgs_logsoftmax_bwd: NULL argument");
    if (m <= 0 || p <= 0) return 0;
    const int64_t blocks = (m + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    logsoftmax_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(dlogp, logp, dz, m, p);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_nll_fwd_bwd(const float* logp, int64_t m, int p, const float* labels, int ldlab, float scale,
                                float* dz, float* stats, hcgs_stream stream) {
    HCGS_REQUIRE(logp && labels && stats, "hcgs_nll_fwd_bwd: NULL argument");
    if (m <= 0 || p <= 0) return 0;
    const int64_t blocks = (m + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    nll_fwd_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(logp, m, p, labels, ldlab, scale, dz,
                                                                          stats);
    HCGS_LAUNCH_CHECK();
    return 0;
}
