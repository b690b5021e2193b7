// Elementwise weight / activation preparation and MLP pointwise pieces (fp32).
#include "common.cuh"

namespace hcgs {

__global__ void mask_apply_kernel(PtrPack<float> w, int n, const float* __restrict__ mask, int64_t numel) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float m = __ldg(mask + i);
        for (int g = 0; g < n; ++g) w.p[g][i] *= m;
    }
}

// quantized_modules.Quantize(balanced=False): clamp master in place, q = sign*ceil(|w|*2^(b-1))/2^(b-1)
__global__ void quantize_w_kernel(float* __restrict__ master, float* __restrict__ q, int64_t numel, float scale,
                                  float inv_scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        float w = master[i];
        w = fminf(fmaxf(w, -1.f), 1.f);
        master[i] = w;  // q may alias master: written afterwards
        const float sgn = (w > 0.f) ? 1.f : ((w < 0.f) ? -1.f : 0.f);
        q[i] = (ceilf(fabsf(w) * scale) * inv_scale) * sgn;
    }
}

// ---- Quantize_inp: global max|x| then requantise --------------------------------------------
__global__ void absmax_kernel(const float* __restrict__ x, int ldx, int64_t rows, int cols,
                              unsigned int* __restrict__ out_bits) {
    float m = 0.f;
    const int64_t total = rows * cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols;
        const int c = (int)(i - r * cols);
        m = fmaxf(m, fabsf(__ldg(x + r * ldx + c)));
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));  // non-negative floats order as uints
}

__global__ void quantize_act_kernel(const float* __restrict__ x, int ldx, float* __restrict__ y, int ldy,
                                    int64_t rows, int cols, float scale, float inv_scale, int repeat,
                                    const unsigned int* __restrict__ max_bits) {
    const float var = __uint_as_float(*max_bits);
    const int64_t total = rows * cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols;
        const int c = (int)(i - r * cols);
        float v = x[r * ldx + c];
        if (var != 0.f)
            for (int k = 0; k < repeat; ++k) v = quant_act_once(v, var, scale, inv_scale);
        y[r * ldy + c] = v;
    }
}

__global__ void act_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                               const float* __restrict__ drop, int64_t numel, int act) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        float v = act_apply(act, x[i]);
        if (drop) v *= __ldg(drop + i);
        y[i] = v;
    }
}

// y is the saved output (after dropout).  For the activations used here act'(.) is recoverable
// from the sign/value of the pre-dropout output; with dropout the dropped entries have zero
// gradient anyway, the kept ones are y/drop.
__global__ void act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                               const float* __restrict__ drop, float* __restrict__ dx, int64_t numel, int act) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        float d = dy[i];
        float out = y[i];
        if (drop) {
            const float m = __ldg(drop + i);
            d *= m;
            out = (m != 0.f) ? out / m : 0.f;
        }
        dx[i] = d * act_grad_from_out(act, out);
    }
}

static inline int grid_for(int64_t n) {
    int64_t b = (n + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_mask_apply(float* const* w, int n, const float* mask, int64_t numel, hcgs_stream stream) {
    HCGS_REQUIRE(w && mask, "hcgs_mask_apply: NULL argument");
    HCGS_REQUIRE(n >= 1 && n <= HCGS_MAX_GROUPS, "hcgs_mask_apply: n out of range");
    if (numel <= 0) return 0;
    mask_apply_kernel<<<grid_for(numel), 256, 0, (cudaStream_t)stream>>>(make_pack<float>(w, n), n, mask, numel);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_quantize_w(float* master, float* q, int64_t numel, int bits, hcgs_stream stream) {
    HCGS_REQUIRE(master && q, "hcgs_quantize_w: NULL argument");
    HCGS_REQUIRE(bits >= 1 && bits <= 24, "hcgs_quantize_w: bits out of range");
    if (numel <= 0) return 0;
    const float scale = (float)(1 << (bits - 1));
    quantize_w_kernel<<<grid_for(numel), 256, 0, (cudaStream_t)stream>>>(master, q, numel, scale, 1.0f / scale);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_quantize_act(const float* x, int ldx, float* y, int ldy, int64_t rows, int cols, int bits,
                                 int repeat, float* scratch, hcgs_stream stream) {
    HCGS_REQUIRE(x && y && scratch, "hcgs_quantize_act: NULL argument");
    HCGS_REQUIRE(bits >= 1 && bits <= 24, "hcgs_quantize_act: bits out of range");
    if (rows <= 0 || cols <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    HCGS_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(float), st));
    const int64_t total = rows * cols;
    absmax_kernel<<<grid_for(total), 256, 0, st>>>(x, ldx, rows, cols, reinterpret_cast<unsigned int*>(scratch));
    HCGS_LAUNCH_CHECK();
    const float scale = (float)(1 << (bits - 1));
    quantize_act_kernel<<<grid_for(total), 256, 0, st>>>(x, ldx, y, ldy, rows, cols, scale, 1.0f / scale, repeat,
                                                        reinterpret_cast<const unsigned int*>(scratch));
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_act_fwd(const float* x, float* y, const float* drop_mask, int64_t numel, int act,
                            hcgs_stream stream) {
    HCGS_REQUIRE(x && y, "hcgs_act_fwd: NULL argument");
    if (numel <= 0) return 0;
    act_fwd_kernel<<<grid_for(numel), 256, 0, (cudaStream_t)stream>>>(x, y, drop_mask, numel, act);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_act_bwd(const float* dy, const float* y, const float* drop_mask, float* dx, int64_t numel,
                            int act, hcgs_stream stream) {
    HCGS_REQUIRE(dy && y && dx, "hcgs_act_bwd: NULL argument");
    if (numel <= 0) return 0;
    act_bwd_kernel<<<grid_for(numel), 256, 0, (cudaStream_t)stream>>>(dy, y, drop_mask, dx, numel, act);
    HCGS_LAUNCH_CHECK();
    return 0;
}
