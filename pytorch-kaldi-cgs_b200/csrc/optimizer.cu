// Fused elementwise optimizer steps on flat fp32 buffers (torch.optim semantics, utils.py:1833-1881).
#include <math.h>

#include <algorithm>

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

// one launch for a whole parameter group: blockIdx.y = tensor, blockIdx.x strides over its elements
constexpr int OPT_MULTI_MAX = 48;
struct OptMulti {
    float* p[OPT_MULTI_MAX];
    const float* g[OPT_MULTI_MAX];
    float* s1[OPT_MULTI_MAX];
    float* s2[OPT_MULTI_MAX];
    int64_t n[OPT_MULTI_MAX];
};

// hyper-parameters of the multi-tensor form: derived constants are computed on the host in double precision and rounded
// once, like torch.optim does with its Python floats (1 - beta2 = 0.001 exactly, not 1.f - 0.999f)
struct OptHp2 {
    float lr, a, b, eps, wd, c, one_m_a, one_m_b, step_size, bc2_sqrt, step;
};

__device__ __forceinline__ void opt_update(int kind, float& w, float gr, float* s1, float* s2, int64_t i,
                                           const OptHp2& hp) {
    if (hp.wd != 0.f) gr = fmaf(hp.wd, w, gr);
    if (kind == 0) {  // SGD: a = momentum, one_m_b = 1 - dampening, c = nesterov
        if (hp.a != 0.f) {
            float buf = (hp.step <= 1.f) ? gr : hp.a * s1[i] + hp.one_m_b * gr;
            s1[i] = buf;
            gr = (hp.c != 0.f) ? gr + hp.a * buf : buf;
        }
        w = w - hp.lr * gr;
    } else if (kind == 1) {  // RMSprop: a = alpha, b = momentum
        const float sq = hp.a * s1[i] + hp.one_m_a * gr * gr;
        s1[i] = sq;
        const float avg = sqrtf(sq) + hp.eps;
        if (hp.b > 0.f) {
            const float buf = hp.b * s2[i] + gr / avg;
            s2[i] = buf;
            w = w - hp.lr * buf;
        } else {
            w = w - hp.lr * (gr / avg);
        }
    } else {  // Adam: a = beta1, b = beta2
        const float m = hp.a * s1[i] + hp.one_m_a * gr;
        const float v = hp.b * s2[i] + hp.one_m_b * gr * gr;
        s1[i] = m;
        s2[i] = v;
        const float denom = sqrtf(v) / hp.bc2_sqrt + hp.eps;
        w = w - hp.step_size * (m / denom);
    }
}

__global__ void opt_step_multi_kernel(int kind, const __grid_constant__ OptMulti t, OptHp2 hp, float grad_scale) {
    const int k = blockIdx.y;
    float* __restrict__ p = t.p[k];
    const float* __restrict__ g = t.g[k];
    const int64_t n = t.n[k];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float w = p[i];
        opt_update(kind, w, g[i] * grad_scale, t.s1[k], t.s2[k], i, hp);
        p[i] = w;
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_opt_step_multi(int kind, int n_tensors, float* const* params, const float* const* grads,
                                   float* const* state1, float* const* state2, const int64_t* numels, const double* hp,
                                   float grad_scale, hcgs_stream stream) {
    HCGS_REQUIRE(params && grads && numels && hp, "hcgs_opt_step_multi: NULL argument");
    HCGS_REQUIRE(kind >= 0 && kind <= 2, "hcgs_opt_step_multi: unknown optimizer kind %d", kind);
    // hp = {lr, a, b, eps, weight_decay, c, d, step} as for hcgs_opt_step, in double precision
    OptHp2 h;
    h.lr = (float)hp[0]; h.a = (float)hp[1]; h.b = (float)hp[2]; h.eps = (float)hp[3]; h.wd = (float)hp[4];
    h.c = (float)hp[5]; h.step = (float)hp[7];
    h.one_m_a = (float)(1.0 - hp[1]);
    h.one_m_b = (float)(1.0 - hp[2]);
    h.step_size = h.lr;
    h.bc2_sqrt = 1.f;
    const bool need1 = (kind == 0 && h.a != 0.f) || kind >= 1, need2 = (kind == 1 && h.b > 0.f) || kind == 2;
    HCGS_REQUIRE(kind != 2 || hp[7] >= 1.0, "hcgs_opt_step_multi: Adam needs step >= 1");
    if (kind == 2) {
        h.step_size = (float)(hp[0] / (1.0 - pow(hp[1], hp[7])));
        h.bc2_sqrt = (float)sqrt(1.0 - pow(hp[2], hp[7]));
    }
    for (int base = 0; base < n_tensors; base += OPT_MULTI_MAX) {
        const int cnt = std::min(OPT_MULTI_MAX, n_tensors - base);
        OptMulti t;
        int64_t biggest = 0;
        for (int k = 0; k < cnt; ++k) {
            HCGS_REQUIRE(params[base + k] && grads[base + k], "hcgs_opt_step_multi: NULL tensor %d", base + k);
            HCGS_REQUIRE(!need1 || (state1 && state1[base + k]), "hcgs_opt_step_multi: tensor %d needs state1", base + k);
            HCGS_REQUIRE(!need2 || (state2 && state2[base + k]), "hcgs_opt_step_multi: tensor %d needs state2", base + k);
            t.p[k] = params[base + k];
            t.g[k] = grads[base + k];
            t.s1[k] = state1 ? state1[base + k] : nullptr;
            t.s2[k] = state2 ? state2[base + k] : nullptr;
            t.n[k] = numels[base + k];
            biggest = std::max(biggest, t.n[k]);
        }
        for (int k = cnt; k < OPT_MULTI_MAX; ++k) t.p[k] = nullptr, t.g[k] = nullptr, t.s1[k] = t.s2[k] = nullptr, t.n[k] = 0;
        if (biggest <= 0) continue;
        int64_t bx = (biggest + 256 * 4 - 1) / (256 * 4);
        bx = std::max<int64_t>(1, std::min<int64_t>(bx, 148 * 4));
        opt_step_multi_kernel<<<dim3((unsigned)bx, (unsigned)cnt, 1), 256, 0, (cudaStream_t)stream>>>(kind, t, h,
                                                                                                      grad_scale);
        HCGS_LAUNCH_CHECK();
    }
    return 0;
}

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
