// Output-layer epilogue: row log-softmax, prior normalisation, fused NLL / error / dlogits (fp32).
//
// One warp per frame (row); a row of P = 1944 .. 3480 log-probabilities stays in L1 between the
// passes, so HBM sees one read of the logits and one write of the result.
#include <cuda_bf16.h>

#include <algorithm>

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

// Backward of the output layer's log-softmax for the tensor-core build, fused with what follows it:
//   dz = dlogp - exp(logp) * rowsum(dlogp)   (autograd of nn.LogSoftmax, neural_networks.py:73-74)
//   db[p] += sum_m dz[m, p]                  (bias gradient, fp32, before rounding)
//   dz_bf16[m, :] = bf16(dz)                 (operand of the dW / dX tensor-core products; pad columns zero)
// One warp per row, persistent over rows; a lane keeps its 4*NV4 columns of the row in registers between the
// row sum and the write.  Its running column sums live in a per-warp private slice of shared memory (NV4 <= 16: one
// conflict-free 16-byte load + store per float4 of the row; in registers they cost 64 registers per thread and
// held the kernel at 16 warps per SM, latency bound at 2.9 TB/s) or in shared-memory atomics (wider rows).
template <int NV4>
__global__ void __launch_bounds__(256, NV4 <= 16 ? 3 : 2) logsoftmax_bwd_fused_kernel(const float* __restrict__ dlogp,
                                                                   const float* __restrict__ logp, int64_t M, int P,
                                                                   __nv_bfloat16* __restrict__ dz, int64_t ldo,
                                                                   float* __restrict__ db,
                                                                   const int64_t* __restrict__ hot_row,
                                                                   const int64_t* __restrict__ hot_col,
                                                                   const float* __restrict__ hot_val) {
    // hot_*: dlogp is NULL and the incoming gradient has ONE entry per row, entry m = (hot_row[m] == m, hot_col[m],
    // hot_val[m]) -- the gradient of an NLL loss kept in its sparse form: no dense [M, P] gradient is written or read.
    constexpr bool COLREG = NV4 <= 16;
    extern __shared__ float s_col[];  // [P] column sums of this CTA (+ [8][NV4*32] float4 private slices)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = threadIdx.x; p < P; p += 256) s_col[p] = 0.f;
    const int nv = (P + 127) / 128;  // float4 per lane actually used (<= NV4)
    float4* csum = reinterpret_cast<float4*>(s_col + ((P + 3) & ~3)) + warp * (NV4 * 32) + lane;  // [i*32]
    if (COLREG) {
#pragma unroll
        for (int i = 0; i < NV4; ++i) csum[i * 32] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    const int64_t wstride = (int64_t)gridDim.x * 8;
    for (int64_t m = (int64_t)blockIdx.x * 8 + warp; m < M; m += wstride) {
        const float4* dr = reinterpret_cast<const float4*>(dlogp + m * P);
        const float4* lr = reinterpret_cast<const float4*>(logp + m * P);
        // the row sum first; the second read of dlogp below hits L1 (a row is 4*P bytes, the warp just read it)
        float s = 0.f;
        int hot = -1;
        if (hot_val) {
            if (__ldg(hot_row + m) != m) __trap();   // not the one-entry-per-row form the caller promised: fail loudly
            const int64_t hc = __ldg(hot_col + m);
            if (hc >= 0 && hc < P) {
                hot = (int)hc;
                s = __ldg(hot_val + m);
            }
        } else {
#pragma unroll
            for (int i = 0; i < NV4; ++i) {
                const int c = (lane + 32 * i) * 4;
                if (i < nv && c < P) {
                    const float4 d = __ldg(dr + lane + 32 * i);
                    s += (d.x + d.y) + (d.z + d.w);
                }
            }
            s = warp_sum(s);
        }
        __nv_bfloat16* orow = dz + m * ldo;
#pragma unroll
        for (int i = 0; i < NV4; ++i) {
            const int c = (lane + 32 * i) * 4;
            if (i < nv && c < P) {
                const float4 l = __ldcs(lr + lane + 32 * i);
                float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!hot_val) {
                    d = __ldg(dr + lane + 32 * i);
                } else if ((unsigned)(hot - c) < 4u) {
                    const int k = hot - c;
                    d.x = k == 0 ? s : 0.f;
                    d.y = k == 1 ? s : 0.f;
                    d.z = k == 2 ? s : 0.f;
                    d.w = k == 3 ? s : 0.f;
                }
                float4 g;
                g.x = d.x - __expf(l.x) * s;
                g.y = d.y - __expf(l.y) * s;
                g.z = d.z - __expf(l.z) * s;
                g.w = d.w - __expf(l.w) * s;
                if (COLREG) {
                    float4 cs = csum[i * 32];
                    cs.x += g.x; cs.y += g.y; cs.z += g.z; cs.w += g.w;
                    csum[i * 32] = cs;
                } else {
                    atomicAdd(s_col + c, g.x); atomicAdd(s_col + c + 1, g.y);
                    atomicAdd(s_col + c + 2, g.z); atomicAdd(s_col + c + 3, g.w);
                }
                const __nv_bfloat162 lo = __floats2bfloat162_rn(g.x, g.y), hi = __floats2bfloat162_rn(g.z, g.w);
                uint2 pk;
                pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(orow + c) = pk;
            }
        }
        for (int c = P + lane; c < ldo; c += 32) orow[c] = __float2bfloat16(0.f);
    }
    if (db == nullptr) return;
    if (COLREG) {
#pragma unroll
        for (int i = 0; i < NV4; ++i) {
            const int c = (lane + 32 * i) * 4;
            if (i < nv && c < P) {
                const float4 cs = csum[i * 32];
                atomicAdd(s_col + c, cs.x); atomicAdd(s_col + c + 1, cs.y);
                atomicAdd(s_col + c + 2, cs.z); atomicAdd(s_col + c + 3, cs.w);
            }
        }
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += 256) {
        const float v = s_col[p];
        if (v != 0.f) atomicAdd(db + p, v);
    }
}

// cost_nll and cost_err of one output in ONE pass over the log-posteriors (utils.py:1931-1952): per row the NLL term
// -logp[m, label] and the error flag (first arg-max != label).  Per-row results (no atomics): the two means are taken by
// the caller over [M] values, deterministically.  One warp per row, 16-byte loads when P % 4 == 0.
__global__ void __launch_bounds__(256) nll_err_rows_kernel(const float* __restrict__ logp, int64_t ldl, int64_t M, int P,
                                                           const int64_t* __restrict__ labels,
                                                           float* __restrict__ loss_rows, float* __restrict__ err_rows) {
    const int lane = threadIdx.x & 31;
    const int64_t wstride = (int64_t)gridDim.x * 8;
    const bool vec = (P & 3) == 0 && (ldl & 3) == 0 && (((uintptr_t)logp) & 15) == 0;
    for (int64_t m = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); m < M; m += wstride) {
        const float* lr = logp + m * ldl;
        float best = -INFINITY;
        int besti = 0x7fffffff;
        if (vec) {
            for (int c = lane * 4; c < P; c += 128) {
                const float4 v = __ldg(reinterpret_cast<const float4*>(lr + c));
                if (v.x > best) { best = v.x; besti = c; }
                if (v.y > best) { best = v.y; besti = c + 1; }
                if (v.z > best) { best = v.z; besti = c + 2; }
                if (v.w > best) { best = v.w; besti = c + 3; }
            }
        } else {
            for (int c = lane; c < P; c += 32) {
                const float v = __ldg(lr + c);
                if (v > best) { best = v; besti = c; }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {   // first index of the maximum (torch.max tie rule)
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
        }
        if (lane == 0) {
            const int64_t label = __ldg(labels + m);
            loss_rows[m] = (label >= 0 && label < P) ? -__ldg(lr + label) : 0.f;
            err_rows[m] = ((int64_t)besti != label) ? 1.f : 0.f;
        }
    }
}

// Narrow outputs (the 48-class monophone head): a row is at most L * 4 columns, L = 8 or 16 lanes, so a warp takes 32 / L
// rows at a time instead of leaving most of its lanes idle (ncu: 10 % / 4 % of the DRAM peak with one warp per row).
template <int L>
__global__ void __launch_bounds__(256) nll_err_rows_small_kernel(const float* __restrict__ logp, int64_t ldl, int64_t M,
                                                                 int P, const int64_t* __restrict__ labels,
                                                                 float* __restrict__ loss_rows,
                                                                 float* __restrict__ err_rows) {
    constexpr int R = 32 / L;
    const int lane = threadIdx.x & 31, sub = lane / L, l = lane % L;
    const int64_t wstride = (int64_t)gridDim.x * 8 * R;
    for (int64_t m0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * R; m0 < M; m0 += wstride) {
        const int64_t m = m0 + sub;
        float best = -INFINITY;
        int besti = 0x7fffffff;
        const int c = l * 4;
        if (m < M && c < P) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(logp + m * ldl + c));
            best = v.x; besti = c;
            if (v.y > best) { best = v.y; besti = c + 1; }
            if (v.z > best) { best = v.z; besti = c + 2; }
            if (v.w > best) { best = v.w; besti = c + 3; }
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
        }
        if (l == 0 && m < M) {
            const int64_t label = __ldg(labels + m);
            loss_rows[m] = (label >= 0 && label < P) ? -__ldg(logp + m * ldl + label) : 0.f;
            err_rows[m] = ((int64_t)besti != label) ? 1.f : 0.f;
        }
    }
}

// log-softmax backward for narrow outputs, same contract as logsoftmax_bwd_fused_kernel (dense or one-hot gradient, bf16
// dz with zeroed pad columns, db += column sums): a lane owns four fixed columns, so the column sums live in registers.
template <int L>
__global__ void __launch_bounds__(256) logsoftmax_bwd_small_kernel(const float* __restrict__ dlogp,
                                                                   const float* __restrict__ logp, int64_t M, int P,
                                                                   __nv_bfloat16* __restrict__ dz, int64_t ldo,
                                                                   float* __restrict__ db,
                                                                   const int64_t* __restrict__ hot_row,
                                                                   const int64_t* __restrict__ hot_col,
                                                                   const float* __restrict__ hot_val) {
    constexpr int R = 32 / L;
    __shared__ float s_col[L * 4];
    for (int i = threadIdx.x; i < L * 4; i += 256) s_col[i] = 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31, sub = lane / L, l = lane % L;
    const int c = l * 4;
    float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
    const int64_t wstride = (int64_t)gridDim.x * 8 * R;
    for (int64_t m0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * R; m0 < M; m0 += wstride) {
        const int64_t m = m0 + sub;
        const bool live = m < M && c < P;
        float4 d = make_float4(0.f, 0.f, 0.f, 0.f), lg = make_float4(0.f, 0.f, 0.f, 0.f);
        float s = 0.f;
        if (live) lg = __ldcs(reinterpret_cast<const float4*>(logp + m * P + c));
        if (hot_val) {
            if (m < M) {
                if (__ldg(hot_row + m) != m) __trap();
                const int64_t hc = __ldg(hot_col + m);
                if (hc >= 0 && hc < P) {
                    s = __ldg(hot_val + m);
                    const int k = (int)hc - c;
                    if (live && (unsigned)k < 4u) {
                        d.x = k == 0 ? s : 0.f;
                        d.y = k == 1 ? s : 0.f;
                        d.z = k == 2 ? s : 0.f;
                        d.w = k == 3 ? s : 0.f;
                    }
                }
            }
        } else {
            if (live) d = __ldg(reinterpret_cast<const float4*>(dlogp + m * P + c));
            s = (d.x + d.y) + (d.z + d.w);
#pragma unroll
            for (int o = L / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        }
        if (live) {
            float4 g;
            g.x = d.x - __expf(lg.x) * s;
            g.y = d.y - __expf(lg.y) * s;
            g.z = d.z - __expf(lg.z) * s;
            g.w = d.w - __expf(lg.w) * s;
            cs.x += g.x; cs.y += g.y; cs.z += g.z; cs.w += g.w;
            const __nv_bfloat162 lo = __floats2bfloat162_rn(g.x, g.y), hi = __floats2bfloat162_rn(g.z, g.w);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t*>(&lo);
            pk.y = *reinterpret_cast<const uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(dz + m * ldo + c) = pk;
        }
        if (m < M)
            for (int cc = P + l; cc < ldo; cc += L) dz[m * ldo + cc] = __float2bfloat16(0.f);
    }
    if (db == nullptr) return;
    if (c < P) {
        atomicAdd(s_col + c, cs.x); atomicAdd(s_col + c + 1, cs.y);
        atomicAdd(s_col + c + 2, cs.z); atomicAdd(s_col + c + 3, cs.w);
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += 256) {
        const float v = s_col[p];
        if (v != 0.f) atomicAdd(db + p, v);
    }
}

}  // namespace hcgs

using namespace hcgs;

static int lsm_bwd_fused_launch(const float* dlogp, const int64_t* hot_row, const int64_t* hot_col, const float* hot_val,
                                const float* logp, int64_t m, int p, void* dz_bf16, int64_t ldo, float* db,
                                hcgs_stream stream) {
    HCGS_REQUIRE(p > 0 && p % 4 == 0 && p <= 4096, "hcgs_logsoftmax_bwd_fused: p = %d must be a multiple of 4, <= 4096", p);
    HCGS_REQUIRE(ldo >= p && ldo % 4 == 0, "hcgs_logsoftmax_bwd_fused: ldo = %lld must be a multiple of 4, >= p",
                 (long long)ldo);
    HCGS_REQUIRE((((uintptr_t)dlogp | (uintptr_t)logp) & 15) == 0 && ((uintptr_t)dz_bf16 & 7) == 0,
                 "hcgs_logsoftmax_bwd_fused: operands must be 16-byte aligned");
    if (m <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (p <= 64) {   // narrow heads: several rows per warp, column sums in registers
        const int blocks = (int)std::min<int64_t>((m + 15) / 16, 148 * 8);
        if (p <= 32)
            logsoftmax_bwd_small_kernel<8><<<blocks, 256, 0, st>>>(dlogp, logp, m, p, (__nv_bfloat16*)dz_bf16, ldo, db, hot_row,
                                                                   hot_col, hot_val);
        else
            logsoftmax_bwd_small_kernel<16><<<blocks, 256, 0, st>>>(dlogp, logp, m, p, (__nv_bfloat16*)dz_bf16, ldo, db,
                                                                    hot_row, hot_col, hot_val);
        HCGS_LAUNCH_CHECK();
        return 0;
    }
    if (p <= 2048) {
        const int blocks = (int)std::min<int64_t>((m + 7) / 8, 148 * 3);  // three resident CTAs per SM, one wave
        const size_t smem = (size_t)((p + 3) & ~3) * sizeof(float) + (size_t)8 * 16 * 32 * sizeof(float4);
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(logsoftmax_bwd_fused_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem));
        logsoftmax_bwd_fused_kernel<16><<<blocks, 256, smem, st>>>(dlogp, logp, m, p, (__nv_bfloat16*)dz_bf16, ldo, db,
                                                                   hot_row, hot_col, hot_val);
    } else {
        const int blocks = (int)std::min<int64_t>((m + 7) / 8, 148 * 2);
        const size_t smem = (size_t)p * sizeof(float);
        logsoftmax_bwd_fused_kernel<32><<<blocks, 256, smem, st>>>(dlogp, logp, m, p, (__nv_bfloat16*)dz_bf16, ldo, db,
                                                                   hot_row, hot_col, hot_val);
    }
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_logsoftmax_bwd_fused(const float* dlogp, const float* logp, int64_t m, int p, void* dz_bf16,
                                         int64_t ldo, float* db, hcgs_stream stream) {
    HCGS_REQUIRE(dlogp && logp && dz_bf16, "hcgs_logsoftmax_bwd_fused: NULL argument");
    return lsm_bwd_fused_launch(dlogp, nullptr, nullptr, nullptr, logp, m, p, dz_bf16, ldo, db, stream);
}

extern "C" int hcgs_logsoftmax_bwd_fused_onehot(const int64_t* rows, const int64_t* cols, const float* vals,
                                                const float* logp, int64_t m, int p, void* dz_bf16, int64_t ldo,
                                                float* db, hcgs_stream stream) {
    HCGS_REQUIRE(rows && cols && vals && logp && dz_bf16, "hcgs_logsoftmax_bwd_fused_onehot: NULL argument");
    return lsm_bwd_fused_launch(nullptr, rows, cols, vals, logp, m, p, dz_bf16, ldo, db, stream);
}

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

extern "C" int hcgs_nll_err_rows(const float* logp, int64_t ldl, int64_t m, int p, const int64_t* labels, float* loss_rows,
                                 float* err_rows, hcgs_stream stream) {
    HCGS_REQUIRE(logp && labels && loss_rows && err_rows, "hcgs_nll_err_rows: NULL argument");
    HCGS_REQUIRE(p >= 1 && ldl >= p, "hcgs_nll_err_rows: bad shape");
    if (m <= 0) return 0;
    const int blocks = (int)std::min<int64_t>((m + 7) / 8, 148 * 8);
    if (p <= 64 && p % 4 == 0 && ldl % 4 == 0 && ((uintptr_t)logp & 15) == 0) {
        if (p <= 32)
            nll_err_rows_small_kernel<8><<<blocks, 256, 0, (cudaStream_t)stream>>>(logp, ldl, m, p, labels, loss_rows, err_rows);
        else
            nll_err_rows_small_kernel<16><<<blocks, 256, 0, (cudaStream_t)stream>>>(logp, ldl, m, p, labels, loss_rows, err_rows);
        HCGS_LAUNCH_CHECK();
        return 0;
    }
    nll_err_rows_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(logp, ldl, m, p, labels, loss_rows, err_rows);
    HCGS_LAUNCH_CHECK();
    return 0;
}
