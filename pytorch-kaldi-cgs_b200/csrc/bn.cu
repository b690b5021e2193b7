// Batch normalisation over the time-batched rows (nn.BatchNorm1d semantics, fp32).
//
// The reference normalises each gate's [T*B, H] projection with its own BatchNorm1d
// (neural_networks.py:957-976, momentum 0.05, eps 1e-5); padded frames are part of the statistics.
// Here the four gates are one [M, 4H] matrix with per-group parameter pointers.
//
// Statistics: rows are split over S CTAs per 32-column strip; every split computes a local
// (mean, M2) pair with a two-pass scan of its rows and the splits are merged in a fixed order with
// Chan's formula, so the result is deterministic and free of E[x^2]-E[x]^2 cancellation.
#include <mutex>
#include <cuda_bf16.h>

#include "common.cuh"

namespace hcgs {

// column -> (parameter group, index inside the group).  Gate-major: col = g*C + c.  Gate-interleaved (il, the
// tensor-core build): col = c*G + g over a padded unit count; pad units (c >= C) have no parameters.
__device__ __forceinline__ void bn_map(int col, int C, int G, int il, int& g, int& c, bool& valid) {
    if (il) {
        g = col % G;
        c = col / G;
        valid = c < C;
    } else {
        g = col / C;
        c = col % C;
        valid = true;
    }
}

__device__ __forceinline__ float block_col_reduce(float v, float (*red)[33], int warp, int lane) {
    red[warp][lane] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][lane];
    __syncthreads();
    return t;
}

// partial[(split*Ctot + col)*2 + {0,1}] = {mean, M2} over this split's rows
__global__ void __launch_bounds__(256) bn_stats_partial_kernel(const float* __restrict__ x, int ldx, int64_t M,
                                                               int Ctot, int64_t rows_per_split,
                                                               float* __restrict__ partial) {
    __shared__ float red[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    const bool ok = col < Ctot;
    const float n = (float)(mend - mbeg);
    float s = 0.f;
    if (ok)
        for (int64_t m = mbeg + warp; m < mend; m += 8) s += __ldg(x + m * ldx + col);
    const float mean = block_col_reduce(s, red, warp, lane) / n;
    float q = 0.f;
    if (ok)
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            const float d = __ldg(x + m * ldx + col) - mean;
            q = fmaf(d, d, q);
        }
    const float m2 = block_col_reduce(q, red, warp, lane);
    if (warp == 0 && ok) {
        partial[((int64_t)blockIdx.y * Ctot + col) * 2 + 0] = mean;
        partial[((int64_t)blockIdx.y * Ctot + col) * 2 + 1] = m2;
    }
}

__global__ void bn_stats_finalize_kernel(const float* __restrict__ partial, int splits, int64_t M,
                                         int64_t rows_per_split, int Ctot, int C, int G, int il, float eps,
                                         float momentum,
                                         PtrPack<float> rmean, PtrPack<float> rvar,
                                         float* __restrict__ save_mean, float* __restrict__ save_invstd) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= Ctot) return;
    float n = 0.f, mean = 0.f, m2 = 0.f;
    for (int s = 0; s < splits; ++s) {
        const int64_t mbeg = (int64_t)s * rows_per_split;
        const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
        const float nb = (float)(mend - mbeg);
        const float mb = partial[((int64_t)s * Ctot + col) * 2 + 0];
        const float qb = partial[((int64_t)s * Ctot + col) * 2 + 1];
        const float nt = n + nb;
        const float delta = mb - mean;
        mean += delta * (nb / nt);
        m2 += qb + delta * delta * (n * nb / nt);
        n = nt;
    }
    const float var_b = m2 / n;
    save_mean[col] = mean;
    save_invstd[col] = rsqrtf(var_b + eps);
    // 1/sqrt via rsqrtf is ~1 ulp; refine once to match a correctly rounded 1/sqrt
    {
        const float v = var_b + eps;
        float r = save_invstd[col];
        r = r * (1.5f - 0.5f * v * r * r);
        save_invstd[col] = r;
    }
    int g, c;
    bool valid;
    bn_map(col, C, G, il, g, c, valid);
    if (!valid) return;
    if (rmean.p[g]) rmean.p[g][c] = (1.f - momentum) * rmean.p[g][c] + momentum * mean;
    if (rvar.p[g]) {
        const float var_u = (n > 1.f) ? m2 / (n - 1.f) : var_b;
        rvar.p[g][c] = (1.f - momentum) * rvar.p[g][c] + momentum * var_u;
    }
}

// Same merge for many splits (the per-128-row-tile partials written by the tensor-core GEMM epilogue):
// block = 32 columns x 8 warps; warp w merges splits w, w+8, ... in order, then the eight partial
// results are merged in warp order -- deterministic.  training == 0: the saved statistics are the
// running ones (eval mode), nothing is merged or updated.
__global__ void __launch_bounds__(256) bn_stats_merge_kernel(const float* __restrict__ partial, int64_t splits,
                                                             int64_t M, int64_t rows_per_split, int Ctot, int C,
                                                             int G, int il, float eps, float momentum, int training,
                                                             PtrPack<float> rmean, PtrPack<float> rvar,
                                                             float* __restrict__ save_mean,
                                                             float* __restrict__ save_invstd) {
    __shared__ float sn[8][33], sm[8][33], sq[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    const bool ok = col < Ctot;
    int g = 0, c = 0;
    bool valid = false;
    if (ok) bn_map(col, C, G, il, g, c, valid);
    if (!training) {
        if (warp == 0 && ok) {
            float mean = 0.f, invstd = 0.f;
            if (valid) {
                mean = rmean.p[g][c];
                invstd = 1.0f / sqrtf(rvar.p[g][c] + eps);
            }
            save_mean[col] = mean;
            save_invstd[col] = invstd;
        }
        return;
    }
    float n = 0.f, mean = 0.f, m2 = 0.f;
    if (ok) {
        for (int64_t s0 = warp; s0 < splits; s0 += 32) {  // four independent loads in flight per trip
            float2 v[4];
            float nb[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t sidx = s0 + 8 * q;
                nb[q] = 0.f;
                v[q] = make_float2(0.f, 0.f);
                if (sidx < splits) {
                    const int64_t mbeg = sidx * rows_per_split;
                    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
                    nb[q] = (float)(mend - mbeg);
                    v[q] = __ldg(reinterpret_cast<const float2*>(partial + (sidx * Ctot + col) * 2));
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (nb[q] > 0.f) {
                    const float nt = n + nb[q];
                    const float delta = v[q].x - mean;
                    mean += delta * (nb[q] / nt);
                    m2 += v[q].y + delta * delta * (n * nb[q] / nt);
                    n = nt;
                }
            }
        }
    }
    sn[warp][lane] = n;
    sm[warp][lane] = mean;
    sq[warp][lane] = m2;
    __syncthreads();
    if (warp != 0 || !ok) return;
    n = 0.f, mean = 0.f, m2 = 0.f;
    for (int w = 0; w < 8; ++w) {
        const float nb = sn[w][lane];
        if (nb > 0.f) {
            const float nt = n + nb;
            const float delta = sm[w][lane] - mean;
            mean += delta * (nb / nt);
            m2 += sq[w][lane] + delta * delta * (n * nb / nt);
            n = nt;
        }
    }
    const float var_b = m2 / n;
    const float v = var_b + eps;
    float r = rsqrtf(v);
    r = r * (1.5f - 0.5f * v * r * r);
    save_mean[col] = mean;
    save_invstd[col] = r;
    if (!valid) return;
    if (rmean.p[g]) rmean.p[g][c] = (1.f - momentum) * rmean.p[g][c] + momentum * mean;
    if (rvar.p[g]) {
        const float var_u = (n > 1.f) ? m2 / (n - 1.f) : var_b;
        rvar.p[g][c] = (1.f - momentum) * rvar.p[g][c] + momentum * var_u;
    }
}

// y = (x - mean) * invstd * gamma + beta ; in eval mode mean/invstd come from the running stats
__global__ void __launch_bounds__(256) bn_apply_kernel(const float* __restrict__ x, int ldx,
                                                       float* __restrict__ y, int ldy, int64_t M, int Ctot,
                                                       int C, int G, int il, PtrPack<const float> gamma,
                                                       PtrPack<const float> beta, PtrPack<float> rmean,
                                                       PtrPack<float> rvar, float eps, int training,
                                                       const float* __restrict__ save_mean,
                                                       const float* __restrict__ save_invstd,
                                                       int rows_per_block) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    if (col >= Ctot) return;
    int g, c;
    bool valid;
    bn_map(col, C, G, il, g, c, valid);
    float mean = 0.f, invstd = 0.f, ga = 0.f, be = 0.f;  // pad columns: y = 0
    if (valid) {
        if (training) {
            mean = save_mean[col];
            invstd = save_invstd[col];
        } else {
            mean = rmean.p[g][c];
            invstd = 1.0f / sqrtf(rvar.p[g][c] + eps);
        }
        ga = gamma.p[g] ? __ldg(gamma.p[g] + c) : 1.f;
        be = beta.p[g] ? __ldg(beta.p[g] + c) : 0.f;
    }
    const float a = invstd * ga;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_block;
    const int64_t mend = (mbeg + rows_per_block < M) ? mbeg + rows_per_block : M;
    for (int64_t m = mbeg + warp; m < mend; m += 8) y[m * ldy + col] = (x[m * ldx + col] - mean) * a + be;
}

// partial[(split*Ctot+col)*2] = {sum dy, sum dy*xhat}
__global__ void __launch_bounds__(256) bn_bwd_partial_kernel(const float* __restrict__ dy, int lddy,
                                                             const float* __restrict__ x, int ldx, int64_t M,
                                                             int Ctot, int64_t rows_per_split,
                                                             const float* __restrict__ save_mean,
                                                             const float* __restrict__ save_invstd,
                                                             float* __restrict__ partial) {
    __shared__ float red[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    const bool ok = col < Ctot;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    const float mean = ok ? save_mean[col] : 0.f, invstd = ok ? save_invstd[col] : 0.f;
    float s1 = 0.f, s2 = 0.f;
    if (ok)
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            const float d = __ldg(dy + m * lddy + col);
            const float xh = (__ldg(x + m * ldx + col) - mean) * invstd;
            s1 += d;
            s2 = fmaf(d, xh, s2);
        }
    const float t1 = block_col_reduce(s1, red, warp, lane);
    const float t2 = block_col_reduce(s2, red, warp, lane);
    if (warp == 0 && ok) {
        partial[((int64_t)blockIdx.y * Ctot + col) * 2 + 0] = t1;
        partial[((int64_t)blockIdx.y * Ctot + col) * 2 + 1] = t2;
    }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float* __restrict__ dy, int lddy,
                                                           const float* __restrict__ x, int ldx, int64_t M,
                                                           int Ctot, int C, int G, int il,
                                                           PtrPack<const float> gamma,
                                                           const float* __restrict__ save_mean,
                                                           const float* __restrict__ save_invstd,
                                                           const float* __restrict__ partial, int splits,
                                                           float* __restrict__ dx, int lddx,
                                                           __nv_bfloat16* __restrict__ dxb, int lddxb,
                                                           PtrPack<float> dgamma, PtrPack<float> dbeta,
                                                           int rows_per_block) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + lane;
    if (col >= Ctot) return;
    int g, c;
    bool valid;
    bn_map(col, C, G, il, g, c, valid);
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_block;
    const int64_t mend = (mbeg + rows_per_block < M) ? mbeg + rows_per_block : M;
    if (!valid) {  // pad columns of the padded tensors stay zero
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            if (dx) dx[m * lddx + col] = 0.f;
            if (dxb) dxb[m * lddxb + col] = __float2bfloat16_rn(0.f);
        }
        return;
    }
    float s1 = 0.f, s2 = 0.f;
    for (int s = 0; s < splits; ++s) {
        s1 += partial[((int64_t)s * Ctot + col) * 2 + 0];
        s2 += partial[((int64_t)s * Ctot + col) * 2 + 1];
    }
    if (blockIdx.y == 0 && warp == 0) {
        if (dgamma.p[g]) dgamma.p[g][c] = s2;
        if (dbeta.p[g]) dbeta.p[g][c] = s1;
    }
    const float mean = save_mean[col], invstd = save_invstd[col];
    const float ga = gamma.p[g] ? __ldg(gamma.p[g] + c) : 1.f;
    const float inv_m = 1.0f / (float)M;
    const float k1 = s1 * inv_m, k2 = s2 * inv_m, a = ga * invstd;
    for (int64_t m = mbeg + warp; m < mend; m += 8) {
        const float xh = (x[m * ldx + col] - mean) * invstd;
        const float o = a * (dy[m * lddy + col] - k1 - xh * k2);
        if (dx) dx[m * lddx + col] = o;
        if (dxb) dxb[m * lddxb + col] = __float2bfloat16_rn(o);
    }
}

static int pick_splits(int64_t M, int strips, int64_t ws_floats, int Ctot, int64_t* rows_out) {
    int splits = 1;
    const int target = 148 * 4;
    if (strips < target) splits = (target + strips - 1) / strips;
    int64_t max_s = ws_floats / (2 * (int64_t)Ctot);
    if (splits > max_s) splits = (int)max_s;
    if (splits > 256) splits = 256;
    if (splits < 1) splits = 1;
    int64_t rows = (M + splits - 1) / splits;
    if (rows < 64) rows = 64;
    splits = (int)((M + rows - 1) / rows);
    *rows_out = rows;
    return splits;
}

}  // namespace hcgs

using namespace hcgs;

// ---- gate-interleaved backward, two adjacent columns per thread (8-byte x loads, 4- or 8-byte dy loads, 4-byte
// bf16 stores: full 32-byte sectors on every access).  DY = float or __nv_bfloat16: the tensor-core build hands
// over the bf16 dpre the recurrent backward writes for the GEMMs, so the fp32 copy never exists.
template <typename DY>
__device__ __forceinline__ float2 bn_ld2(const DY* p);
template <>
__device__ __forceinline__ float2 bn_ld2<float>(const float* p) {
    return __ldg(reinterpret_cast<const float2*>(p));
}
template <>
__device__ __forceinline__ float2 bn_ld2<__nv_bfloat16>(const __nv_bfloat16* p) {
    const unsigned v = __ldg(reinterpret_cast<const unsigned*>(p));
    return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}

template <typename DY>
__global__ void __launch_bounds__(256) bn_bwd_il_partial_kernel(const DY* __restrict__ dy, const float* __restrict__ x,
                                                                int ld, int64_t M, int64_t rows_per_split,
                                                                const float* __restrict__ save_mean,
                                                                const float* __restrict__ save_invstd,
                                                                float* __restrict__ partial) {
    __shared__ float red[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 64 + lane * 2;
    const bool ok = col < ld;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    float2 mean = make_float2(0.f, 0.f), invstd = mean;
    if (ok) {
        mean = *reinterpret_cast<const float2*>(save_mean + col);
        invstd = *reinterpret_cast<const float2*>(save_invstd + col);
    }
    float s1a = 0.f, s1b = 0.f, s2a = 0.f, s2b = 0.f;
    if (ok) {
#pragma unroll 4
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            const float2 d = bn_ld2<DY>(dy + m * ld + col);
            const float2 xv = __ldg(reinterpret_cast<const float2*>(x + m * ld + col));
            s1a += d.x;
            s1b += d.y;
            s2a = fmaf(d.x, (xv.x - mean.x) * invstd.x, s2a);
            s2b = fmaf(d.y, (xv.y - mean.y) * invstd.y, s2b);
        }
    }
    const float t1a = block_col_reduce(s1a, red, warp, lane);
    const float t1b = block_col_reduce(s1b, red, warp, lane);
    const float t2a = block_col_reduce(s2a, red, warp, lane);
    const float t2b = block_col_reduce(s2b, red, warp, lane);
    if (warp == 0 && ok)
        *reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * ld + col) * 2) = make_float4(t1a, t2a, t1b, t2b);
}

template <typename DY>
__global__ void __launch_bounds__(256) bn_bwd_il_apply_kernel(const DY* __restrict__ dy, const float* __restrict__ x,
                                                              int ld, int64_t M, int C, int G,
                                                              PtrPack<const float> gamma,
                                                              const float* __restrict__ save_mean,
                                                              const float* __restrict__ save_invstd,
                                                              const float* __restrict__ partial, int splits,
                                                              float* __restrict__ dx, __nv_bfloat16* __restrict__ dxb,
                                                              PtrPack<float> dgamma, PtrPack<float> dbeta,
                                                              int rows_per_block) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 64 + lane * 2;
    if (col >= ld) return;
    const int g = col % G, c = col / G;  // G is even: both columns belong to unit c, gates g and g+1
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_block;
    const int64_t mend = (mbeg + rows_per_block < M) ? mbeg + rows_per_block : M;
    if (c >= C) {  // pad units of the padded tensors stay zero
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            if (dx) *reinterpret_cast<float2*>(dx + m * ld + col) = make_float2(0.f, 0.f);
            if (dxb) *reinterpret_cast<unsigned*>(dxb + m * ld + col) = 0u;
        }
        return;
    }
    float s1a = 0.f, s2a = 0.f, s1b = 0.f, s2b = 0.f;
    for (int s = 0; s < splits; ++s) {
        const float4 p = *reinterpret_cast<const float4*>(partial + ((int64_t)s * ld + col) * 2);
        s1a += p.x;
        s2a += p.y;
        s1b += p.z;
        s2b += p.w;
    }
    if (blockIdx.y == 0 && warp == 0) {
        if (dgamma.p[g]) dgamma.p[g][c] = s2a;
        if (dbeta.p[g]) dbeta.p[g][c] = s1a;
        if (dgamma.p[g + 1]) dgamma.p[g + 1][c] = s2b;
        if (dbeta.p[g + 1]) dbeta.p[g + 1][c] = s1b;
    }
    const float2 mean = *reinterpret_cast<const float2*>(save_mean + col);
    const float2 invstd = *reinterpret_cast<const float2*>(save_invstd + col);
    const float ga = gamma.p[g] ? __ldg(gamma.p[g] + c) : 1.f;
    const float gb = gamma.p[g + 1] ? __ldg(gamma.p[g + 1] + c) : 1.f;
    const float inv_m = 1.0f / (float)M;
    const float k1a = s1a * inv_m, k2a = s2a * inv_m, aa = ga * invstd.x;
    const float k1b = s1b * inv_m, k2b = s2b * inv_m, ab = gb * invstd.y;
#pragma unroll 4
    for (int64_t m = mbeg + warp; m < mend; m += 8) {
        const float2 d = bn_ld2<DY>(dy + m * ld + col);
        const float2 xv = __ldg(reinterpret_cast<const float2*>(x + m * ld + col));
        const float oa = aa * (d.x - k1a - (xv.x - mean.x) * invstd.x * k2a);
        const float ob = ab * (d.y - k1b - (xv.y - mean.y) * invstd.y * k2b);
        if (dx) *reinterpret_cast<float2*>(dx + m * ld + col) = make_float2(oa, ob);
        if (dxb) {
            const __nv_bfloat162 o2 = __floats2bfloat162_rn(oa, ob);
            *reinterpret_cast<__nv_bfloat162*>(dxb + m * ld + col) = o2;
        }
    }
}

// ---- the same two passes for n_groups = 4 with ONE UNIT (its four gate columns) per thread: 16-byte x loads, 8- or
// 16-byte dy loads, 8-byte bf16 stores, 512 contiguous bytes of x per warp and row (the two-column version moved 256)
template <typename DY>
__device__ __forceinline__ float4 bn_ld4(const DY* p);
template <>
__device__ __forceinline__ float4 bn_ld4<float>(const float* p) {
    return __ldg(reinterpret_cast<const float4*>(p));
}
template <>
__device__ __forceinline__ float4 bn_ld4<__nv_bfloat16>(const __nv_bfloat16* p) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    return make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xffff0000u), __uint_as_float(v.y << 16),
                       __uint_as_float(v.y & 0xffff0000u));
}

template <typename DY>
__global__ void __launch_bounds__(256) bn_bwd_il4_partial_kernel(const DY* __restrict__ dy, const float* __restrict__ x,
                                                                 int ld, int64_t M, int64_t rows_per_split,
                                                                 const float* __restrict__ save_mean,
                                                                 const float* __restrict__ save_invstd,
                                                                 float* __restrict__ partial) {
    __shared__ float red[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 128 + lane * 4;
    const bool ok = col < ld;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    float4 mean = make_float4(0.f, 0.f, 0.f, 0.f), invstd = mean;
    if (ok) {
        mean = *reinterpret_cast<const float4*>(save_mean + col);
        invstd = *reinterpret_cast<const float4*>(save_invstd + col);
    }
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    if (ok) {
#pragma unroll 4
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            const float4 d = bn_ld4<DY>(dy + m * ld + col);
            const float4 xv = __ldg(reinterpret_cast<const float4*>(x + m * ld + col));
            s1[0] += d.x; s1[1] += d.y; s1[2] += d.z; s1[3] += d.w;
            s2[0] = fmaf(d.x, (xv.x - mean.x) * invstd.x, s2[0]);
            s2[1] = fmaf(d.y, (xv.y - mean.y) * invstd.y, s2[1]);
            s2[2] = fmaf(d.z, (xv.z - mean.z) * invstd.z, s2[2]);
            s2[3] = fmaf(d.w, (xv.w - mean.w) * invstd.w, s2[3]);
        }
    }
    float t1[4], t2[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        t1[q] = block_col_reduce(s1[q], red, warp, lane);
        t2[q] = block_col_reduce(s2[q], red, warp, lane);
    }
    if (warp == 0 && ok) {
        float4* dst = reinterpret_cast<float4*>(partial + ((int64_t)blockIdx.y * ld + col) * 2);
        dst[0] = make_float4(t1[0], t2[0], t1[1], t2[1]);
        dst[1] = make_float4(t1[2], t2[2], t1[3], t2[3]);
    }
}

template <typename DY>
__global__ void __launch_bounds__(256) bn_bwd_il4_apply_kernel(const DY* __restrict__ dy, const float* __restrict__ x,
                                                               int ld, int64_t M, int C, PtrPack<const float> gamma,
                                                               const float* __restrict__ save_mean,
                                                               const float* __restrict__ save_invstd,
                                                               const float* __restrict__ partial, int splits,
                                                               float* __restrict__ dx, __nv_bfloat16* __restrict__ dxb,
                                                               PtrPack<float> dgamma, PtrPack<float> dbeta,
                                                               int rows_per_block) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 128 + lane * 4;
    if (col >= ld) return;
    const int c = col >> 2;  // the unit; the four columns are its gates 0..3
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_block;
    const int64_t mend = (mbeg + rows_per_block < M) ? mbeg + rows_per_block : M;
    if (c >= C) {  // pad units of the padded tensors stay zero
        for (int64_t m = mbeg + warp; m < mend; m += 8) {
            if (dx) *reinterpret_cast<float4*>(dx + m * ld + col) = make_float4(0.f, 0.f, 0.f, 0.f);
            if (dxb) *reinterpret_cast<uint2*>(dxb + m * ld + col) = make_uint2(0u, 0u);
        }
        return;
    }
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < splits; ++s) {
        const float4* p = reinterpret_cast<const float4*>(partial + ((int64_t)s * ld + col) * 2);
        const float4 p0 = p[0], p1 = p[1];
        s1[0] += p0.x; s2[0] += p0.y; s1[1] += p0.z; s2[1] += p0.w;
        s1[2] += p1.x; s2[2] += p1.y; s1[3] += p1.z; s2[3] += p1.w;
    }
    if (blockIdx.y == 0 && warp == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (dgamma.p[q]) dgamma.p[q][c] = s2[q];
            if (dbeta.p[q]) dbeta.p[q][c] = s1[q];
        }
    }
    const float4 mean = *reinterpret_cast<const float4*>(save_mean + col);
    const float4 invstd = *reinterpret_cast<const float4*>(save_invstd + col);
    const float mu[4] = {mean.x, mean.y, mean.z, mean.w}, is[4] = {invstd.x, invstd.y, invstd.z, invstd.w};
    const float inv_m = 1.0f / (float)M;
    float k1[4], k2[4], aq[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        k1[q] = s1[q] * inv_m;
        k2[q] = s2[q] * inv_m;
        aq[q] = (gamma.p[q] ? __ldg(gamma.p[q] + c) : 1.f) * is[q];
    }
#pragma unroll 4
    for (int64_t m = mbeg + warp; m < mend; m += 8) {
        // (evict-first loads here were measured 48 % slower: the pass lives on what the reduction pass left in L2)
        const float4 d = bn_ld4<DY>(dy + m * ld + col);
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x + m * ld + col));
        const float dv[4] = {d.x, d.y, d.z, d.w}, xq[4] = {xv.x, xv.y, xv.z, xv.w};
        float o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) o[q] = aq[q] * (dv[q] - k1[q] - (xq[q] - mu[q]) * is[q] * k2[q]);
        if (dx) *reinterpret_cast<float4*>(dx + m * ld + col) = make_float4(o[0], o[1], o[2], o[3]);
        if (dxb) {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(o[0], o[1]), hi = __floats2bfloat162_rn(o[2], o[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<const unsigned*>(&lo);
            pk.y = *reinterpret_cast<const unsigned*>(&hi);
            __stcs(reinterpret_cast<uint2*>(dxb + m * ld + col), pk);   // streaming store: 0.6 GB nobody re-reads from L2
        }
    }
}

template <typename DY>
static int bn_bwd_il_launch(const DY* dy, const float* x, int64_t m, int n_groups, int units, int ld_units,
                            const float* const* gamma, const float* save_mean, const float* save_invstd, float* dx,
                            void* dx_bf16, float* const* dgamma, float* const* dbeta, float* workspace,
                            int64_t workspace_floats, cudaStream_t st) {
    const int ld = n_groups * ld_units;
    const bool aligned16 = (((uintptr_t)x | (uintptr_t)save_mean | (uintptr_t)save_invstd | (uintptr_t)workspace |
                             (uintptr_t)dx) & 15) == 0 && ((uintptr_t)dy & (sizeof(DY) == 4 ? 15 : 7)) == 0 && ((uintptr_t)dx_bf16 & 7) == 0;
    if (n_groups == 4 && aligned16) {  // one unit (four gate columns) per thread
        const int strips = (ld + 127) / 128;
        int64_t rows;
        const int splits = pick_splits(m, strips, workspace_floats, ld, &rows);
        bn_bwd_il4_partial_kernel<DY><<<dim3(strips, splits), 256, 0, st>>>(dy, x, ld, m, rows, save_mean, save_invstd,
                                                                            workspace);
        HCGS_LAUNCH_CHECK();
        int rows_per_block = 256;
        int64_t yb = (m + rows_per_block - 1) / rows_per_block;
        while (yb > 65535) {
            rows_per_block *= 2;
            yb = (m + rows_per_block - 1) / rows_per_block;
        }
        bn_bwd_il4_apply_kernel<DY><<<dim3(strips, (unsigned)yb), 256, 0, st>>>(
            dy, x, ld, m, units, make_pack<const float>(gamma, n_groups), save_mean, save_invstd, workspace, splits, dx,
            (__nv_bfloat16*)dx_bf16, make_pack<float>(dgamma, n_groups), make_pack<float>(dbeta, n_groups),
            rows_per_block);
        HCGS_LAUNCH_CHECK();
        return 0;
    }
    const int strips = (ld + 63) / 64;
    int64_t rows;
    const int splits = pick_splits(m, strips, workspace_floats, ld, &rows);
    bn_bwd_il_partial_kernel<DY><<<dim3(strips, splits), 256, 0, st>>>(dy, x, ld, m, rows, save_mean, save_invstd,
                                                                       workspace);
    HCGS_LAUNCH_CHECK();
    int rows_per_block = 256;
    int64_t yb = (m + rows_per_block - 1) / rows_per_block;
    while (yb > 65535) {
        rows_per_block *= 2;
        yb = (m + rows_per_block - 1) / rows_per_block;
    }
    bn_bwd_il_apply_kernel<DY><<<dim3(strips, (unsigned)yb), 256, 0, st>>>(
        dy, x, ld, m, units, n_groups, make_pack<const float>(gamma, n_groups), save_mean, save_invstd, workspace,
        splits, dx, (__nv_bfloat16*)dx_bf16, make_pack<float>(dgamma, n_groups), make_pack<float>(dbeta, n_groups),
        rows_per_block);
    HCGS_LAUNCH_CHECK();
    return 0;
}

static int bn_fwd_impl(const float* x, int ldx, float* y, int ldy, int64_t m, int n_groups, int group_cols,
                       int ld_units, int il, const float* const* gamma, const float* const* beta,
                       float* const* running_mean, float* const* running_var, float eps, float momentum,
                       int training, float* save_mean, float* save_invstd, float* workspace,
                       int64_t workspace_floats, hcgs_stream stream) {
    HCGS_REQUIRE(x && y, "hcgs_bn_fwd: NULL argument");
    HCGS_REQUIRE(n_groups >= 1 && n_groups <= HCGS_MAX_GROUPS, "hcgs_bn_fwd: n_groups out of range");
    if (m <= 0) return 0;
    const int Ctot = n_groups * (il ? ld_units : group_cols);
    const int strips = (Ctot + 31) / 32;
    cudaStream_t st = (cudaStream_t)stream;
    if (training) {
        HCGS_REQUIRE(save_mean && save_invstd && workspace, "hcgs_bn_fwd: training needs save buffers + workspace");
        HCGS_REQUIRE(workspace_floats >= 2 * (int64_t)Ctot, "hcgs_bn_fwd: workspace too small");
        int64_t rows;
        const int splits = pick_splits(m, strips, workspace_floats, Ctot, &rows);
        bn_stats_partial_kernel<<<dim3(strips, splits), 256, 0, st>>>(x, ldx, m, Ctot, rows, workspace);
        HCGS_LAUNCH_CHECK();
        bn_stats_finalize_kernel<<<(Ctot + 127) / 128, 128, 0, st>>>(
            workspace, splits, m, rows, Ctot, group_cols, n_groups, il, eps, momentum,
            make_pack<float>(running_mean, n_groups),
            make_pack<float>(running_var, n_groups), save_mean, save_invstd);
        HCGS_LAUNCH_CHECK();
    } else {
        HCGS_REQUIRE(running_mean && running_var, "hcgs_bn_fwd: eval needs running statistics");
    }
    int rows_per_block = 256;
    int64_t yb = (m + rows_per_block - 1) / rows_per_block;
    while (yb > 65535) {
        rows_per_block *= 2;
        yb = (m + rows_per_block - 1) / rows_per_block;
    }
    bn_apply_kernel<<<dim3(strips, (unsigned)yb), 256, 0, st>>>(
        x, ldx, y, ldy, m, Ctot, group_cols, n_groups, il, make_pack<const float>(gamma, n_groups),
        make_pack<const float>(beta, n_groups), make_pack<float>(running_mean, n_groups),
        make_pack<float>(running_var, n_groups), eps, training, save_mean, save_invstd, rows_per_block);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_bn_fwd(const float* x, int ldx, float* y, int ldy, int64_t m, int n_groups, int group_cols,
                           const float* const* gamma, const float* const* beta, float* const* running_mean,
                           float* const* running_var, float eps, float momentum, int training, float* save_mean,
                           float* save_invstd, float* workspace, int64_t workspace_floats, hcgs_stream stream) {
    return bn_fwd_impl(x, ldx, y, ldy, m, n_groups, group_cols, group_cols, 0, gamma, beta, running_mean,
                       running_var, eps, momentum, training, save_mean, save_invstd, workspace, workspace_floats,
                       stream);
}

// ---- allocation-free variant of hcgs_bn_fwd ------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The statistics scratch of
// hcgs_bn_fwd (per row split and column {mean, M2}) lives in a file-scope device array instead of a caller buffer:
// HCGS_BN_STATIC_MAX_COLS columns x 64 row splits x 2 floats = 2 MiB, held by every GPU that loads the library until
// the process exits.  Every slot the finalize kernel reads is written by the partial kernel of the SAME call, so
// nothing has to start at zero and the previous call's contents are never read; but all calls share the array:
// two calls that overlap on the GPU (different streams, no dependency) race on it.
static __device__ __align__(256) float g_bn_static_ws[2 * HCGS_BN_STATIC_MAX_COLS * 64];

namespace {
constexpr int kBnStaticMaxDevices = 64;
std::mutex g_bn_static_mu;
float* g_bn_static_ptr[kBnStaticMaxDevices] = {};   // device address of g_bn_static_ws, per device ordinal
}  // namespace

// In host code `g_bn_static_ws` is only a stand-in for the symbol, not a device address: resolve it once per device,
// with that device current, and hand the resolved pointer to the kernels.
static int bn_static_scratch(float** out) {
    int dev = -1;
    HCGS_CHECK_CUDA(cudaGetDevice(&dev));
    HCGS_REQUIRE(dev >= 0 && dev < kBnStaticMaxDevices, "hcgs_bn_fwd_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_bn_static_mu);
    if (g_bn_static_ptr[dev] == nullptr) {
        float* p = nullptr;
        HCGS_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_bn_static_ws));
        HCGS_REQUIRE(p != nullptr, "hcgs_bn_fwd_static: cudaGetSymbolAddress returned NULL");
        g_bn_static_ptr[dev] = p;
    }
    *out = g_bn_static_ptr[dev];
    return 0;
}

extern "C" int hcgs_bn_fwd_static(const float* x, int ldx, float* y, int ldy, int64_t m, int n_groups, int group_cols,
                                  const float* const* gamma, const float* const* beta, float* const* running_mean,
                                  float* const* running_var, float eps, float momentum, int training,
                                  float* save_mean, float* save_invstd, hcgs_stream stream) {
    HCGS_REQUIRE(n_groups >= 1 && n_groups <= HCGS_MAX_GROUPS, "hcgs_bn_fwd_static: n_groups out of range");
    HCGS_REQUIRE(group_cols >= 1, "hcgs_bn_fwd_static: group_cols must be positive");
    const int64_t Ctot = (int64_t)n_groups * group_cols;
    HCGS_REQUIRE(Ctot <= HCGS_BN_STATIC_MAX_COLS,
                 "hcgs_bn_fwd_static: %lld columns exceed the static scratch (HCGS_BN_STATIC_MAX_COLS = %d); use "
                 "hcgs_bn_fwd with a caller workspace", (long long)Ctot, HCGS_BN_STATIC_MAX_COLS);
    float* ws = nullptr;
    if (int rc = bn_static_scratch(&ws)) return rc;
    // the same 64 row splits per column that the plugin's own workspace (2 * C * 64 floats) allows
    return bn_fwd_impl(x, ldx, y, ldy, m, n_groups, group_cols, group_cols, 0, gamma, beta, running_mean,
                       running_var, eps, momentum, training, save_mean, save_invstd, ws, 2 * Ctot * 64, stream);
}

extern "C" int hcgs_bn_fwd_il(const float* x, float* y, int64_t m, int n_groups, int units, int ld_units,
                              const float* const* gamma, const float* const* beta, float* const* running_mean,
                              float* const* running_var, float eps, float momentum, int training,
                              float* save_mean, float* save_invstd, float* workspace, int64_t workspace_floats,
                              hcgs_stream stream) {
    HCGS_REQUIRE(ld_units >= units, "hcgs_bn_fwd_il: ld_units < units");
    return bn_fwd_impl(x, n_groups * ld_units, y, n_groups * ld_units, m, n_groups, units, ld_units, 1, gamma, beta,
                       running_mean, running_var, eps, momentum, training, save_mean, save_invstd, workspace,
                       workspace_floats, stream);
}

static int bn_bwd_impl(const float* dy, int lddy, const float* x, int ldx, int64_t m, int n_groups, int group_cols,
                       int ld_units, int il, const float* const* gamma, const float* save_mean,
                       const float* save_invstd, float* dx, int lddx, void* dx_bf16, int lddxb,
                       float* const* dgamma, float* const* dbeta, float* workspace, int64_t workspace_floats,
                       hcgs_stream stream) {
    HCGS_REQUIRE(dy && x && (dx || dx_bf16) && save_mean && save_invstd && workspace, "hcgs_bn_bwd: NULL argument");
    HCGS_REQUIRE(n_groups >= 1 && n_groups <= HCGS_MAX_GROUPS, "hcgs_bn_bwd: n_groups out of range");
    if (m <= 0) return 0;
    const int Ctot = n_groups * (il ? ld_units : group_cols);
    HCGS_REQUIRE(workspace_floats >= 2 * (int64_t)Ctot, "hcgs_bn_bwd: workspace too small");
    const int strips = (Ctot + 31) / 32;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t rows;
    const int splits = pick_splits(m, strips, workspace_floats, Ctot, &rows);
    bn_bwd_partial_kernel<<<dim3(strips, splits), 256, 0, st>>>(dy, lddy, x, ldx, m, Ctot, rows, save_mean,
                                                               save_invstd, workspace);
    HCGS_LAUNCH_CHECK();
    int rows_per_block = 256;
    int64_t yb = (m + rows_per_block - 1) / rows_per_block;
    while (yb > 65535) {
        rows_per_block *= 2;
        yb = (m + rows_per_block - 1) / rows_per_block;
    }
    bn_bwd_apply_kernel<<<dim3(strips, (unsigned)yb), 256, 0, st>>>(
        dy, lddy, x, ldx, m, Ctot, group_cols, n_groups, il, make_pack<const float>(gamma, n_groups), save_mean,
        save_invstd, workspace, splits, dx, lddx, (__nv_bfloat16*)dx_bf16, lddxb, make_pack<float>(dgamma, n_groups),
        make_pack<float>(dbeta, n_groups), rows_per_block);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_bn_bwd(const float* dy, int lddy, const float* x, int ldx, int64_t m, int n_groups,
                           int group_cols, const float* const* gamma, const float* save_mean,
                           const float* save_invstd, float* dx, int lddx, float* const* dgamma,
                           float* const* dbeta, float* workspace, int64_t workspace_floats, hcgs_stream stream) {
    return bn_bwd_impl(dy, lddy, x, ldx, m, n_groups, group_cols, group_cols, 0, gamma, save_mean, save_invstd, dx,
                       lddx, nullptr, 0, dgamma, dbeta, workspace, workspace_floats, stream);
}

extern "C" int hcgs_bn_bwd_il(const float* dy, const void* dy_bf16, const float* x, int64_t m, int n_groups,
                              int units, int ld_units, const float* const* gamma, const float* save_mean,
                              const float* save_invstd, float* dx, void* dx_bf16, float* const* dgamma,
                              float* const* dbeta, float* workspace, int64_t workspace_floats, hcgs_stream stream) {
    HCGS_REQUIRE((dy != nullptr) != (dy_bf16 != nullptr), "hcgs_bn_bwd_il: exactly one of dy / dy_bf16");
    HCGS_REQUIRE(x && (dx || dx_bf16) && save_mean && save_invstd && workspace, "hcgs_bn_bwd_il: NULL argument");
    HCGS_REQUIRE(n_groups >= 2 && n_groups <= HCGS_MAX_GROUPS && n_groups % 2 == 0,
                 "hcgs_bn_bwd_il: n_groups must be even and <= %d", HCGS_MAX_GROUPS);
    HCGS_REQUIRE(units >= 1 && ld_units >= units, "hcgs_bn_bwd_il: ld_units < units");
    if (m <= 0) return 0;
    HCGS_REQUIRE(workspace_floats >= 2 * (int64_t)n_groups * ld_units, "hcgs_bn_bwd_il: workspace too small");
    if (dy)
        return bn_bwd_il_launch<float>(dy, x, m, n_groups, units, ld_units, gamma, save_mean, save_invstd, dx, dx_bf16,
                                       dgamma, dbeta, workspace, workspace_floats, (cudaStream_t)stream);
    return bn_bwd_il_launch<__nv_bfloat16>((const __nv_bfloat16*)dy_bf16, x, m, n_groups, units, ld_units, gamma,
                                           save_mean, save_invstd, dx, dx_bf16, dgamma, dbeta, workspace,
                                           workspace_floats, (cudaStream_t)stream);
}

// Batch-norm statistics for the tensor-core build, without an apply pass (the recurrent kernel
// normalises on the fly): training = merge the per-tile {mean, M2} partials of hcgs_tc_gemm_nt
// and update the running statistics; eval = hand the running statistics over as mean / invstd.
extern "C" int hcgs_bn_stats_il(const float* partial, int64_t splits, int64_t rows_per_split, int64_t m, int n_groups,
                                int units, int ld_units, float* const* running_mean, float* const* running_var,
                                float eps, float momentum, int training, float* save_mean, float* save_invstd,
                                hcgs_stream stream) {
    HCGS_REQUIRE(save_mean && save_invstd, "hcgs_bn_stats_il: NULL output");
    HCGS_REQUIRE(n_groups >= 1 && n_groups <= HCGS_MAX_GROUPS && units >= 1 && ld_units >= units,
                 "hcgs_bn_stats_il: bad shape");
    if (training) {
        HCGS_REQUIRE(partial && splits >= 1 && rows_per_split >= 1 && m >= 1 &&
                         (splits - 1) * rows_per_split < m && splits * rows_per_split >= m,
                     "hcgs_bn_stats_il: partials do not cover m=%lld rows", (long long)m);
    } else {
        HCGS_REQUIRE(running_mean && running_var, "hcgs_bn_stats_il: eval mode needs running statistics");
        for (int g = 0; g < n_groups; ++g)
            HCGS_REQUIRE(running_mean[g] && running_var[g], "hcgs_bn_stats_il: eval mode needs running statistics");
    }
    const int Ctot = n_groups * ld_units;
    bn_stats_merge_kernel<<<(Ctot + 31) / 32, 256, 0, (cudaStream_t)stream>>>(
        partial, splits, m, rows_per_split, Ctot, units, n_groups, 1, eps, momentum, training,
        make_pack<float>(running_mean, n_groups), make_pack<float>(running_var, n_groups), save_mean, save_invstd);
    HCGS_LAUNCH_CHECK();
    return 0;
}
