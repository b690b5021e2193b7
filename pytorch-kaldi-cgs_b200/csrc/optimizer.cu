// Fused elementwise optimizer steps on flat fp32 buffers (torch.optim semantics, utils.py:1833-1881).
#include "common.cuh"

namespace hcgs {

struct OptHp {
    float lr, a, b, eps, wd, c, d, step;
};

__global__ void opt_step_kernel(int kind, float* __restrict__ p, const float* __restrict__ g,
                                float* __restrict__ s1, float* __restrict__ s2, int64_t n, OptHp hp,
                                float grad_scale, float bc1, float bc2_sqrt) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        float w = p[i];
        float gr = g[i] * grad_scale;
        if (hp.wd != 0.f) gr = fmaf(hp.wd, w, gr);
        if (kind == 0) {  // SGD: a = momentum, b = dampening, c = nesterov
            if (hp.a != 0.f) {
                float buf = (hp.step <= 1.f) ? gr : hp.a * s1[i] + (1.f - hp.b) * gr;
                s1[i] = buf;
                gr = (hp.c != 0.f) ? gr + hp.a * buf : buf;
            }
            p[i] = w - hp.lr * gr;
        } else if (kind == 1) {  // RMSprop: a = alpha, b = momentum
            const float sq = hp.a * s1[i] + (1.f - hp.a) * gr * gr;
            s1[i] = sq;
            const float avg = sqrtf(sq) + hp.eps;
            if (hp.b > 0.f) {
                const float buf = hp.b * s2[i] + gr / avg;
                s2[i] = buf;
                p[i] = w - hp.lr * buf;
            } else {
                p[i] = w - hp.lr * (gr / avg);
            }
        } else {  // Adam: a = beta1, b = beta2
            const float m = hp.a * s1[i] + (1.f - hp.a) * gr;
            const float v = hp.b * s2[i] + (1.f - hp.b) * gr * gr;
            s1[i] = m;
            s2[i] = v;
            const float denom = sqrtf(v) / bc2_sqrt + hp.eps;
            p[i] = w - (hp.lr / bc1) * (m / denom);
        }
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_opt_step(int kind, float* param, const float* grad, float* state1, float* state2,
                             int64_t numel, const float* hp, float grad_scale, hcgs_stream stream) {
    HCGS_REQUIRE(param && grad && hp, "hcgs_opt_step: NULL argument");
    HCGS_REQUIRE(kind >= 0 && kind <= 2, "hcgs_opt_step: unknown optimizer kind %d", kind);
    if (numel <= 0) return 0;
    OptHp h;
    h.lr = hp[0]; h.a = hp[1]; h.b = hp[2]; h.eps = hp[3]; h.wd = hp[4]; h.c = hp[5]; h.d = hp[6]; h.step = hp[7];
    if (kind == 0) HCGS_REQUIRE(h.a == 0.f || state1, "hcgs_opt_step: SGD momentum needs state1");
    if (kind == 1) HCGS_REQUIRE(state1 && (h.b <= 0.f || state2), "hcgs_opt_step: RMSprop needs state buffers");
    if (kind == 2) HCGS_REQUIRE(state1 && state2 && h.step >= 1.f, "hcgs_opt_step: Adam needs two state buffers and step>=1");
    float bc1 = 1.f, bc2s = 1.f;
    if (kind == 2) {
        bc1 = 1.f - powf(h.a, h.step);
        bc2s = sqrtf(1.f - powf(h.b, h.step));
    }
    int64_t blocks = (numel + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    opt_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(kind, param, grad, state1, state2, numel, h,
                                                                       grad_scale, bc1, bc2s);
    HCGS_LAUNCH_CHECK();
    return 0;
}
