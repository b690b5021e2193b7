// Persistent recurrent LSTM kernels (fp32 parity build).
//
// One thread-block CLUSTER runs the whole time loop of one layer for NB = 8 sequences
// (reference loop: neural_networks.py:1063-1100).  CTA r of the cluster owns Hc hidden units
// (rows of the four recurrent matrices U_f,U_i,U_o,U_c, which share one HCGS mask) and keeps its
// slice of the compressed weights RESIDENT in shared memory for all T steps ("BELL-4": the
// non-empty 4x4 blocks of every 4-row group, 64 floats per block = 4 gates x 4 x 4).
//
//   forward step :  gate pre-activations = pre[t] + U h_{t-1}  (each CTA: its rows, all 8 sequences)
//                   -> fused sigmoid/act, cell update, dropout mask -> h_t slice
//                   -> slice pushed into every CTA's copy of h through distributed shared memory
//                   -> one cluster barrier per step.
//   backward step:  fused gate derivatives -> per-block partial products with the SAME resident
//                   blocks (transposed use) -> per-column sums pushed to the CTA that owns the
//                   column (reduce-scatter over DSMEM, summed in a fixed order => deterministic).
//
// Many clusters run side by side (one per group of 8 sequences); nothing but the cluster barrier
// synchronises, so independent sequence groups never wait for each other.
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace hcgs {

constexpr int NB = 8;   // sequences per cluster
constexpr int IPT = 2;  // pointwise items (unit, sequence) per thread
constexpr size_t kLnScratchBytes = 2048;   // 2 x NB floats + 16 x NB x 2 floats (clusters of up to 16 CTAs)

__host__ __device__ static inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

size_t lstm_fwd_smem_bytes(int H, int CL, int Hc, int HcP, int L, bool w_in_smem) {
    const int RG = Hc / 4;
    const int H4 = (H + 3) / 4 * 4;
    size_t b = 0;
    if (w_in_smem) b += align16((size_t)L * RG * 64 * 4);
    b += align16((size_t)L * RG * 2);
    b += align16((size_t)2 * H4 * NB * 4);
    b += align16((size_t)4 * NB * HcP * 4);
    b += align16((size_t)Hc * NB * 4);
    b += kLnScratchBytes;   // LayerNorm of h_t: per-sequence mean / denominator
    (void)CL;
    return b;
}

size_t lstm_bwd_smem_bytes(int H, int CL, int Hc, int max_blocks, int n_colgroups, bool w_in_smem) {
    size_t b = 0;
    if (w_in_smem) b += align16((size_t)max_blocks * 64 * 4);
    b += align16((size_t)max_blocks * 2);           // blk_a
    b += align16((size_t)(n_colgroups + 1) * 4);    // cg_ptr
    b += align16((size_t)max_blocks * 2);           // cg_list
    b += align16((size_t)4 * Hc * NB * 4);          // dpre_s
    b += align16((size_t)max_blocks * 32 * 4);      // contrib
    b += align16((size_t)2 * CL * Hc * NB * 4);     // slots
    b += kLnScratchBytes;                           // LayerNorm backward: statistics + cluster-wide sums [CL][NB][2]
    (void)H;
    return b;
}

struct LstmFwdArgs {
    const float* wpack;
    const uint16_t* bidx;
    const float* pre;
    const float* drop;
    float drop_scalar;
    float* h_out;
    float* c_out;
    float* gates_out;
    int T, B, H, CL, Hc, HcP, RG, L, NCG, act, w_in_smem;
    int act_bits;  // > 0: Quantize_inp of h_{k-1} before the four recurrent products of step k
    // act_bits > 0 with more than one cluster: the scale is the max |h| over the WHOLE batch, so the clusters exchange
    // their maxima once per time step through q_sync = {slot[3], arrivals} in global memory (all clusters co-resident,
    // checked by the launcher)
    unsigned int* q_sync;
    int n_groups;
    // LayerNorm of h_t inside the time loop (lstm_use_laynorm, neural_networks.py:1092-1095 with the class at :40-51):
    // h_t <- gamma * (h_t - mean) / (std + eps) + beta over the H units of every sequence, unbiased std; the normalised
    // h_t is what the next step's U h product and the layer output see.  ln_stats [T,B,2] = {mean, std + eps} for the
    // backward (may be NULL).
    const float* ln_gamma;
    const float* ln_beta;
    float ln_eps;
    float* ln_stats;
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// static scratch of the grid-wide scale exchange: one quantised recurrence runs on a device at a time (it needs every
// cluster co-resident, so two could not overlap anyway)
__device__ unsigned int g_q_sync[4];

__global__ void __launch_bounds__(512) lstm_fwd_kernel(const LstmFwdArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int group = blockIdx.x / a.CL;
    const int b0 = group * NB;
    const int nb = min(NB, a.B - b0);
    const int u0 = rank * a.Hc;
    const int nu = max(0, min(a.Hc, a.H - u0));
    const int H4 = a.NCG * 4;
    const int t = threadIdx.x, nthreads = blockDim.x;
    const int H = a.H, Hc = a.Hc, HcP = a.HcP, RG = a.RG, L = a.L;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    float* w_s = nullptr;
    if (a.w_in_smem) {
        w_s = reinterpret_cast<float*>(sp);
        sp += align16((size_t)L * RG * 64 * 4);
    }
    uint16_t* bidx_s = reinterpret_cast<uint16_t*>(sp);
    sp += align16((size_t)L * RG * 2);
    float* hbuf = reinterpret_cast<float*>(sp);
    sp += align16((size_t)2 * H4 * NB * 4);
    float* gbuf = reinterpret_cast<float*>(sp);
    sp += align16((size_t)4 * NB * HcP * 4);
    float* stage = reinterpret_cast<float*>(sp);
    sp += align16((size_t)Hc * NB * 4);
    float* ln_mu = reinterpret_cast<float*>(sp);   // [NB] mean, [NB] std + eps
    float* ln_den = ln_mu + NB;

    // ---- one-time load of the resident weights and index -----------------------------------
    const float* wsrc = a.wpack + (size_t)rank * L * RG * 64;
    if (a.w_in_smem) {
        const float4* s4 = reinterpret_cast<const float4*>(wsrc);
        float4* d4 = reinterpret_cast<float4*>(w_s);
        for (int i = t; i < L * RG * 16; i += nthreads) d4[i] = __ldg(s4 + i);
    }
    const float* wptr = a.w_in_smem ? w_s : wsrc;
    for (int i = t; i < L * RG; i += nthreads) {
        const uint16_t c = a.bidx[(size_t)rank * L * RG + i];
        bidx_s[i] = (c == 0xFFFF) ? (uint16_t)0 : c;  // padded blocks carry zero weights
    }
    for (int i = t; i < 2 * H4 * NB; i += nthreads) hbuf[i] = 0.f;
    __syncthreads();
    cluster.sync();  // every CTA's h buffers are zeroed before anybody pushes into them

    // ---- thread roles ------------------------------------------------------------------------
    const bool mv = t < RG * 16;  // mat-vec role: (row group, gate, sequence half, row half)
    const int a_l = t >> 4, g = (t >> 2) & 3, bh = (t >> 1) & 1, rh = t & 1;

    int it_b[IPT], it_ul[IPT];
    bool it_in[IPT], it_ok[IPT];
    float c_reg[IPT], dr[IPT], pf[IPT][4];
#pragma unroll
    for (int it = 0; it < IPT; ++it) {
        const int i = t + it * nthreads;
        it_in[it] = i < Hc * NB;
        it_b[it] = it_in[it] ? i / Hc : 0;
        it_ul[it] = it_in[it] ? i % Hc : 0;
        it_ok[it] = it_in[it] && it_ul[it] < nu && it_b[it] < nb;
        c_reg[it] = 0.f;
        dr[it] = a.drop_scalar;
        if (it_ok[it] && a.drop) dr[it] = __ldg(a.drop + (size_t)(b0 + it_b[it]) * H + u0 + it_ul[it]);
#pragma unroll
        for (int q = 0; q < 4; ++q) pf[it][q] = 0.f;
        if (it_ok[it] && a.T > 0) {
            const size_t base = ((size_t)0 * a.B + b0 + it_b[it]) * 4 * H + u0 + it_ul[it];
#pragma unroll
            for (int q = 0; q < 4; ++q) pf[it][q] = __ldg(a.pre + base + (size_t)q * H);
        }
    }
    const int chunks = Hc * NB / 4;  // float4 chunks of one h slice

    __shared__ float q_red[16];
    __shared__ float q_glob;
    const float q_scale = (float)(1 << (a.act_bits > 0 ? a.act_bits - 1 : 0)), q_inv = 1.0f / q_scale;

    // LayerNorm of the gathered h of time step `step`, in place on this CTA's copy hq [H4][NB]; every CTA of the cluster
    // normalises its own copy (same arithmetic, same result) and stores the rows of the units it owns.
    auto layer_norm_h = [&](float* hq, int step) {
        const int w = t >> 5, lane = t & 31;
        for (int b = w; b < NB; b += nthreads >> 5) {
            float mu = 0.f, den = 1.f;
            if (b < nb) {
                float sum = 0.f;
                for (int u = lane; u < H; u += 32) sum += hq[u * NB + b];
                mu = warp_sum(sum) / (float)H;
                float q = 0.f;
                for (int u = lane; u < H; u += 32) {
                    const float d = hq[u * NB + b] - mu;
                    q = fmaf(d, d, q);
                }
                den = sqrtf(warp_sum(q) / (float)(H - 1)) + a.ln_eps;
            }
            if (lane == 0) {
                ln_mu[b] = mu;
                ln_den[b] = den;
                if (rank == 0 && b < nb && a.ln_stats) {
                    float* st = a.ln_stats + ((size_t)step * a.B + b0 + b) * 2;
                    st[0] = mu;
                    st[1] = den;
                }
            }
        }
        __syncthreads();
        for (int i = t; i < H * NB; i += nthreads) {
            const int u = i / NB, b = i - u * NB;
            if (b >= nb) continue;
            const float v = __ldg(a.ln_gamma + u) * (hq[i] - ln_mu[b]) / ln_den[b] + __ldg(a.ln_beta + u);
            hq[i] = v;
            if (u >= u0 && u < u0 + nu) a.h_out[((size_t)step * a.B + b0 + b) * H + u] = v;
        }
        __syncthreads();
    };

    for (int k = 0; k < a.T; ++k) {
        const int cur = k & 1, nxt = cur ^ 1;
        const float* hc = hbuf + (size_t)cur * H4 * NB;
        if (a.ln_gamma && k > 0) layer_norm_h(hbuf + (size_t)cur * H4 * NB, k - 1);
        if (a.act_bits > 0 && k > 0) {
            // QuantizeLinear(if_inp_quant) on h_{k-1}: u{f,i,o,c}h each replace ht.data by Quantize_inp(ht.data)
            // (quantized_modules.py:212-219, neural_networks.py:1077-1088): scale = max |h| over the WHOLE [B,H]
            // tensor (this cluster holds the whole batch), four applications in a row.  The tensor object that was
            // appended to the output list is the one that gets swapped, so h_out[k-1] becomes the quantised copy.
            float* hq = hbuf + (size_t)cur * H4 * NB;
            float m = 0.f;
            for (int i = t; i < H4 * NB; i += nthreads) m = fmaxf(m, fabsf(hq[i]));
            m = warp_max(m);
            if ((t & 31) == 0) q_red[t >> 5] = m;
            __syncthreads();
            float var = 0.f;
            for (int w = 0; w < (nthreads + 31) / 32; ++w) var = fmaxf(var, q_red[w]);
            if (a.n_groups > 1) {
                // max over all clusters.  Step k uses slot k % 3; cluster 0 clears the slot of step k + 1 before it
                // arrives (every cluster read that slot at step k - 2, i.e. before it arrived at step k - 1, and
                // nobody writes it before this step's arrivals are complete).  |h| >= 0: the float bits order like
                // unsigned integers.
                if (t == 0) {
                    if (rank == 0) {
                        if (group == 0) *reinterpret_cast<volatile unsigned int*>(a.q_sync + (k + 1) % 3) = 0u;
                        atomicMax(a.q_sync + k % 3, __float_as_uint(var));
                        __threadfence();
                        atomicAdd(a.q_sync + 3, 1u);
                    }
                    const unsigned int target = (unsigned int)k * (unsigned int)a.n_groups;
                    unsigned int spins = 0;
                    while (ld_acquire_gpu(a.q_sync + 3) < target) {
                        if (++spins > (1u << 24)) __trap();   // a cluster that is not resident: fail, do not hang
                        __nanosleep(32);
                    }
                    q_glob = __uint_as_float(ld_acquire_gpu(a.q_sync + k % 3));
                }
                __syncthreads();
                var = q_glob;
            }
            if (var != 0.f) {
                for (int i = t; i < H4 * NB; i += nthreads) {
                    float v = hq[i];
#pragma unroll
                    for (int r = 0; r < 4; ++r) v = quant_act_once(v, var, q_scale, q_inv);
                    hq[i] = v;
                    const int u = i / NB, b = i - u * NB;
                    if (u >= u0 && u < u0 + nu && b < nb) a.h_out[((size_t)(k - 1) * a.B + b0 + b) * H + u] = v;
                }
            }
            __syncthreads();
        }
        // ---- U h_{k-1} for my rows ---------------------------------------------------------
        if (mv) {
            float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
            const float* wb = wptr + ((size_t)a_l * 4 + g) * 16 + rh * 8;
            const uint16_t* bi = bidx_s + a_l;
#pragma unroll 3
            for (int j = 0; j < L; ++j) {
                const int c = bi[(size_t)j * RG];
                const float4 w0 = *reinterpret_cast<const float4*>(wb + (size_t)j * RG * 64);
                const float4 w1 = *reinterpret_cast<const float4*>(wb + (size_t)j * RG * 64 + 4);
                const float* hp = hc + (size_t)c * 4 * NB + bh * 4;
                const float4 h0 = *reinterpret_cast<const float4*>(hp);
                const float4 h1 = *reinterpret_cast<const float4*>(hp + NB);
                const float4 h2 = *reinterpret_cast<const float4*>(hp + 2 * NB);
                const float4 h3 = *reinterpret_cast<const float4*>(hp + 3 * NB);
                acc0[0] = fmaf(w0.x, h0.x, fmaf(w0.y, h1.x, fmaf(w0.z, h2.x, fmaf(w0.w, h3.x, acc0[0]))));
                acc0[1] = fmaf(w0.x, h0.y, fmaf(w0.y, h1.y, fmaf(w0.z, h2.y, fmaf(w0.w, h3.y, acc0[1]))));
                acc0[2] = fmaf(w0.x, h0.z, fmaf(w0.y, h1.z, fmaf(w0.z, h2.z, fmaf(w0.w, h3.z, acc0[2]))));
                acc0[3] = fmaf(w0.x, h0.w, fmaf(w0.y, h1.w, fmaf(w0.z, h2.w, fmaf(w0.w, h3.w, acc0[3]))));
                acc1[0] = fmaf(w1.x, h0.x, fmaf(w1.y, h1.x, fmaf(w1.z, h2.x, fmaf(w1.w, h3.x, acc1[0]))));
                acc1[1] = fmaf(w1.x, h0.y, fmaf(w1.y, h1.y, fmaf(w1.z, h2.y, fmaf(w1.w, h3.y, acc1[1]))));
                acc1[2] = fmaf(w1.x, h0.z, fmaf(w1.y, h1.z, fmaf(w1.z, h2.z, fmaf(w1.w, h3.z, acc1[2]))));
                acc1[3] = fmaf(w1.x, h0.w, fmaf(w1.y, h1.w, fmaf(w1.z, h2.w, fmaf(w1.w, h3.w, acc1[3]))));
            }
            // gbuf[g][b][unit] , unit = a_l*4 + rh*2 + {0,1}
            float* gp = gbuf + ((size_t)g * NB + bh * 4) * HcP + a_l * 4 + rh * 2;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                gp[(size_t)b * HcP] = acc0[b];
                gp[(size_t)b * HcP + 1] = acc1[b];
            }
        }
        __syncthreads();
        // ---- gates, cell, hidden (one (unit, sequence) item at a time) ----------------------
#pragma unroll
        for (int it = 0; it < IPT; ++it) {
            if (!it_in[it]) continue;
            const int b = it_b[it], ul = it_ul[it];
            float hval = 0.f;
            if (it_ok[it]) {
                const float zf = gbuf[((size_t)0 * NB + b) * HcP + ul] + pf[it][0];
                const float zi = gbuf[((size_t)1 * NB + b) * HcP + ul] + pf[it][1];
                const float zo = gbuf[((size_t)2 * NB + b) * HcP + ul] + pf[it][2];
                const float zc = gbuf[((size_t)3 * NB + b) * HcP + ul] + pf[it][3];
                const float fg = sigmoidf_(zf), ig = sigmoidf_(zi), og = sigmoidf_(zo);
                const float gg = act_apply(a.act, zc);
                const float cn = ig * gg * dr[it] + fg * c_reg[it];
                c_reg[it] = cn;
                hval = og * act_apply(a.act, cn);
                const size_t row = (size_t)k * a.B + b0 + b;
                const int u = u0 + ul;
                if (!a.ln_gamma) a.h_out[row * H + u] = hval;   // with LayerNorm the normalised row is stored later
                if (a.c_out) a.c_out[row * H + u] = cn;
                if (a.gates_out) {
                    float* go = a.gates_out + row * 4 * H + u;
                    go[0] = fg;
                    go[(size_t)H] = ig;
                    go[(size_t)2 * H] = og;
                    go[(size_t)3 * H] = gg;
                }
                if (k + 1 < a.T) {
                    const size_t base = ((size_t)(k + 1) * a.B + b0 + b) * 4 * H + u;
#pragma unroll
                    for (int q = 0; q < 4; ++q) pf[it][q] = __ldg(a.pre + base + (size_t)q * H);
                }
            }
            stage[ul * NB + b] = hval;
        }
        __syncthreads();
        // ---- all-gather of h_k: push my slice into every CTA's next buffer (DSMEM) ----------
        for (int idx = t; idx < a.CL * chunks; idx += nthreads) {
            const int dst = idx / chunks, q = idx - dst * chunks;
            const int ul = q >> 1;
            if (u0 + ul < H4) {
                const float4 v = reinterpret_cast<const float4*>(stage)[q];
                float* remote = cluster.map_shared_rank(hbuf, dst);
                reinterpret_cast<float4*>(remote + (size_t)nxt * H4 * NB + (size_t)u0 * NB)[q] = v;
            }
        }
        cluster.sync();
    }
    // the last step's h is normalised too (it is only an output: nothing inside the loop consumes it)
    if (a.ln_gamma && a.T > 0) layer_norm_h(hbuf + (size_t)(a.T & 1) * H4 * NB, a.T - 1);
}

struct LstmBwdArgs {
    const float* wT;            // [n_blocks][ci][g][ri]
    const int* cta_blk_off;     // [CL+1]
    const uint16_t* blk_a;      // [n_blocks]
    const int* cg_ptr;          // [CL][NCG+1]
    const uint16_t* cg_list;    // [n_blocks]
    const float* dh_out;        // [T,B,H]
    const float* gates;         // [T,B,4H]
    const float* c;             // [T,B,H]
    const float* drop;
    float drop_scalar;
    float* dpre;                // [T,B,4H]
    int T, B, H, CL, Hc, RG, NCG, max_blocks, act, w_in_smem;
    // LayerNorm of h_t inside the loop (forward: LstmFwdArgs::ln_*).  dh arriving at step k -- from above and from step
    // k + 1 -- is the gradient w.r.t. the NORMALISED h_k; it is taken through the LayerNorm backward (two sums over the H
    // units of a sequence, i.e. over the whole cluster) before the gate derivatives.  ln_stats: {mean, std + eps} of the
    // forward; ln_dgamma / ln_dbeta: [n_groups][H] partial parameter gradients (one row per cluster, summed by the caller).
    const float* ln_gamma;
    const float* ln_stats;
    float ln_eps;
    float* ln_dgamma;
    float* ln_dbeta;
};

__global__ void __launch_bounds__(512) lstm_bwd_kernel(const LstmBwdArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int group = blockIdx.x / a.CL;
    const int b0 = group * NB;
    const int nb = min(NB, a.B - b0);
    const int u0 = rank * a.Hc;
    const int nu = max(0, min(a.Hc, a.H - u0));
    const int H4 = a.NCG * 4;
    const int t = threadIdx.x, nthreads = blockDim.x;
    const int H = a.H, Hc = a.Hc, CL = a.CL, NCG = a.NCG;
    const int blk0 = a.cta_blk_off[rank];
    const int nblk = a.cta_blk_off[rank + 1] - blk0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    float* w_s = nullptr;
    if (a.w_in_smem) {
        w_s = reinterpret_cast<float*>(sp);
        sp += align16((size_t)a.max_blocks * 64 * 4);
    }
    uint16_t* blk_a_s = reinterpret_cast<uint16_t*>(sp);
    sp += align16((size_t)a.max_blocks * 2);
    int* cg_ptr_s = reinterpret_cast<int*>(sp);
    sp += align16((size_t)(NCG + 1) * 4);
    uint16_t* cg_list_s = reinterpret_cast<uint16_t*>(sp);
    sp += align16((size_t)a.max_blocks * 2);
    float* dpre_s = reinterpret_cast<float*>(sp);
    sp += align16((size_t)4 * Hc * NB * 4);
    float* contrib = reinterpret_cast<float*>(sp);
    sp += align16((size_t)a.max_blocks * 32 * 4);
    float* slots = reinterpret_cast<float*>(sp);  // [2][CL][Hc][NB]
    sp += align16((size_t)2 * CL * Hc * NB * 4);
    float* ln_mu = reinterpret_cast<float*>(sp);   // [NB] mean, [NB] std + eps of the step being processed
    float* ln_den = ln_mu + NB;
    float* ln_red = ln_den + NB;                   // [CL][NB][2] per-CTA sums of g and g * (h - mean), pushed by every CTA

    const float* wsrc = a.wT + (size_t)blk0 * 64;
    if (a.w_in_smem) {
        const float4* s4 = reinterpret_cast<const float4*>(wsrc);
        float4* d4 = reinterpret_cast<float4*>(w_s);
        for (int i = t; i < nblk * 16; i += nthreads) d4[i] = __ldg(s4 + i);
    }
    const float* wptr = a.w_in_smem ? w_s : wsrc;
    for (int i = t; i < nblk; i += nthreads) {
        blk_a_s[i] = a.blk_a[blk0 + i];
        cg_list_s[i] = a.cg_list[blk0 + i];
    }
    for (int i = t; i <= NCG; i += nthreads) cg_ptr_s[i] = a.cg_ptr[(size_t)rank * (NCG + 1) + i];
    for (int i = t; i < 2 * CL * Hc * NB; i += nthreads) slots[i] = 0.f;
    __syncthreads();
    cluster.sync();

    int it_b[IPT], it_ul[IPT];
    bool it_in[IPT], it_ok[IPT];
    float dc_carry[IPT], dr[IPT];
    // prefetched operands of the step about to be processed
    float p_f[IPT], p_i[IPT], p_o[IPT], p_g[IPT], p_c[IPT], p_cm1[IPT], p_dh[IPT];
#pragma unroll
    for (int it = 0; it < IPT; ++it) {
        const int i = t + it * nthreads;
        it_in[it] = i < Hc * NB;
        it_b[it] = it_in[it] ? i / Hc : 0;
        it_ul[it] = it_in[it] ? i % Hc : 0;
        it_ok[it] = it_in[it] && it_ul[it] < nu && it_b[it] < nb;
        dc_carry[it] = 0.f;
        dr[it] = a.drop_scalar;
        if (it_ok[it] && a.drop) dr[it] = __ldg(a.drop + (size_t)(b0 + it_b[it]) * H + u0 + it_ul[it]);
        p_f[it] = p_i[it] = p_o[it] = p_g[it] = p_c[it] = p_cm1[it] = p_dh[it] = 0.f;
        if (it_ok[it] && a.T > 0) {
            const int k = a.T - 1;
            const size_t row = (size_t)k * a.B + b0 + it_b[it];
            const int u = u0 + it_ul[it];
            const float* gp = a.gates + row * 4 * H + u;
            p_f[it] = __ldg(gp);
            p_i[it] = __ldg(gp + (size_t)H);
            p_o[it] = __ldg(gp + (size_t)2 * H);
            p_g[it] = __ldg(gp + (size_t)3 * H);
            p_c[it] = __ldg(a.c + row * H + u);
            p_cm1[it] = (k > 0) ? __ldg(a.c + (row - a.B) * H + u) : 0.f;
            p_dh[it] = __ldg(a.dh_out + row * H + u);
        }
    }

    const bool ln = a.ln_gamma != nullptr;
    float ln_g[IPT], acc_dg[IPT], acc_db[IPT], dh_eff[IPT];
#pragma unroll
    for (int it = 0; it < IPT; ++it) {
        ln_g[it] = (ln && it_ok[it]) ? __ldg(a.ln_gamma + u0 + it_ul[it]) : 0.f;
        acc_dg[it] = acc_db[it] = dh_eff[it] = 0.f;
    }

    for (int k = a.T - 1; k >= 0; --k) {
        const int wr = k & 1, rd = wr ^ 1;
        if (ln) {
            // ---- LayerNorm backward of step k (y = gamma * xc / den + beta, xc = x - mean, den = std + eps, unbiased
            // std over N = H units):  dx_i = (g_i - S1 / N) / den - xc_i * S2 / ((N - 1) * std * den^2),
            // g = dy * gamma, S1 = sum g, S2 = sum g * xc over ALL units of the sequence (cluster-wide) ----
            if (t < NB) {
                const bool ok = t < nb;
                const float* st = a.ln_stats + ((size_t)k * a.B + b0 + (ok ? t : 0)) * 2;
                ln_mu[t] = ok ? __ldg(st) : 0.f;
                ln_den[t] = ok ? __ldg(st + 1) : 1.f;
            }
            __syncthreads();
            float xc_r[IPT], g_r[IPT];
#pragma unroll
            for (int it = 0; it < IPT; ++it) {
                if (!it_in[it]) continue;
                const int b = it_b[it], ul = it_ul[it];
                float g = 0.f, xc = 0.f;
                if (it_ok[it]) {
                    float dh = p_dh[it];
                    const float* sl = slots + ((size_t)rd * CL * Hc + ul) * NB + b;
                    for (int src = 0; src < CL; ++src) dh += sl[(size_t)src * Hc * NB];
                    xc = p_o[it] * act_apply(a.act, p_c[it]) - ln_mu[b];
                    g = dh * ln_g[it];
                    acc_db[it] += dh;
                    acc_dg[it] = fmaf(dh, xc / ln_den[b], acc_dg[it]);
                }
                xc_r[it] = xc;
                g_r[it] = g;
                dpre_s[((size_t)0 * Hc + ul) * NB + b] = g;        // planes 0 / 1 of the dpre tile are free until the
                dpre_s[((size_t)1 * Hc + ul) * NB + b] = g * xc;   // gate derivatives below
            }
            __syncthreads();
            for (int p = t >> 5; p < NB * 2; p += nthreads >> 5) {
                const int b = p >> 1, which = p & 1;
                float sum = 0.f;
                for (int ul = t & 31; ul < Hc; ul += 32) sum += dpre_s[((size_t)which * Hc + ul) * NB + b];
                sum = warp_sum(sum);
                if ((t & 31) == 0)
                    for (int dst = 0; dst < CL; ++dst) cluster.map_shared_rank(ln_red, dst)[(rank * NB + b) * 2 + which] = sum;
            }
            cluster.sync();
            const float n = (float)H;
#pragma unroll
            for (int it = 0; it < IPT; ++it) {
                if (!it_ok[it]) continue;
                const int b = it_b[it];
                float s1 = 0.f, s2 = 0.f;
                for (int src = 0; src < CL; ++src) {
                    s1 += ln_red[(src * NB + b) * 2];
                    s2 += ln_red[(src * NB + b) * 2 + 1];
                }
                const float den = ln_den[b], sd = den - a.ln_eps;
                dh_eff[it] = (g_r[it] - s1 / n) / den - xc_r[it] * s2 / ((n - 1.f) * sd * den * den);
            }
            __syncthreads();   // every thread is done with planes 0 / 1 and the statistics of this step
        }
        // ---- fused gate derivatives -------------------------------------------------------------
#pragma unroll
        for (int it = 0; it < IPT; ++it) {
            if (!it_in[it]) continue;
            const int b = it_b[it], ul = it_ul[it];
            float dpf = 0.f, dpi = 0.f, dpo = 0.f, dpg = 0.f;
            if (it_ok[it]) {
                float dh = dh_eff[it];
                if (!ln) {
                    dh = p_dh[it];
                    const float* sl = slots + ((size_t)rd * CL * Hc + ul) * NB + b;
                    for (int src = 0; src < CL; ++src) dh += sl[(size_t)src * Hc * NB];
                }
                const float fg = p_f[it], ig = p_i[it], og = p_o[it], gg = p_g[it];
                const float tc = act_apply(a.act, p_c[it]);
                const float dog = dh * tc;
                const float dct = dh * og * act_grad_from_out(a.act, tc) + dc_carry[it];
                const float dig = dct * gg * dr[it];
                const float dgg = dct * ig * dr[it];
                const float dfg = dct * p_cm1[it];
                dc_carry[it] = dct * fg;
                dpf = dfg * fg * (1.f - fg);
                dpi = dig * ig * (1.f - ig);
                dpo = dog * og * (1.f - og);
                dpg = dgg * act_grad_from_out(a.act, gg);
                const size_t row = (size_t)k * a.B + b0 + b;
                const int u = u0 + ul;
                float* dp = a.dpre + row * 4 * H + u;
                dp[0] = dpf;
                dp[(size_t)H] = dpi;
                dp[(size_t)2 * H] = dpo;
                dp[(size_t)3 * H] = dpg;
                if (k > 0) {
                    const size_t rowp = row - a.B;
                    const float* gp = a.gates + rowp * 4 * H + u;
                    p_f[it] = __ldg(gp);
                    p_i[it] = __ldg(gp + (size_t)H);
                    p_o[it] = __ldg(gp + (size_t)2 * H);
                    p_g[it] = __ldg(gp + (size_t)3 * H);
                    p_c[it] = p_cm1[it];
                    p_cm1[it] = (k > 1) ? __ldg(a.c + (rowp - a.B) * H + u) : 0.f;
                    p_dh[it] = __ldg(a.dh_out + rowp * H + u);
                }
            }
            dpre_s[((size_t)0 * Hc + ul) * NB + b] = dpf;
            dpre_s[((size_t)1 * Hc + ul) * NB + b] = dpi;
            dpre_s[((size_t)2 * Hc + ul) * NB + b] = dpo;
            dpre_s[((size_t)3 * Hc + ul) * NB + b] = dpg;
        }
        __syncthreads();
        // ---- per-block partial products: contrib[s][ci][b] = sum_{g,ri} U_g[ri][ci] * dpre_g[ri][b] ----
        for (int s = t >> 3; s < nblk; s += nthreads >> 3) {
            const int ci = (t >> 1) & 3, bh = t & 1;
            const int a_l = blk_a_s[s];
            const float* wb = wptr + ((size_t)s * 4 + ci) * 16;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const float4 w = *reinterpret_cast<const float4*>(wb + g * 4);
                const float* dp = dpre_s + ((size_t)g * Hc + a_l * 4) * NB + bh * 4;
                const float4 d0 = *reinterpret_cast<const float4*>(dp);
                const float4 d1 = *reinterpret_cast<const float4*>(dp + NB);
                const float4 d2 = *reinterpret_cast<const float4*>(dp + 2 * NB);
                const float4 d3 = *reinterpret_cast<const float4*>(dp + 3 * NB);
                acc[0] = fmaf(w.x, d0.x, fmaf(w.y, d1.x, fmaf(w.z, d2.x, fmaf(w.w, d3.x, acc[0]))));
                acc[1] = fmaf(w.x, d0.y, fmaf(w.y, d1.y, fmaf(w.z, d2.y, fmaf(w.w, d3.y, acc[1]))));
                acc[2] = fmaf(w.x, d0.z, fmaf(w.y, d1.z, fmaf(w.z, d2.z, fmaf(w.w, d3.z, acc[2]))));
                acc[3] = fmaf(w.x, d0.w, fmaf(w.y, d1.w, fmaf(w.z, d2.w, fmaf(w.w, d3.w, acc[3]))));
            }
            *reinterpret_cast<float4*>(contrib + ((size_t)s * 4 + ci) * NB + bh * 4) =
                make_float4(acc[0], acc[1], acc[2], acc[3]);
        }
        __syncthreads();
        // ---- column sums over my blocks, pushed to the CTA owning the column (reduce-scatter) ----
        for (int w = t; w < H4 * 2; w += nthreads) {
            const int col = w >> 1, bq = w & 1;
            const int c = col >> 2, ci = col & 3;
            float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int e = cg_ptr_s[c]; e < cg_ptr_s[c + 1]; ++e) {
                const int s = cg_list_s[e];
                const float4 v = *reinterpret_cast<const float4*>(contrib + ((size_t)s * 4 + ci) * NB + bq * 4);
                s4.x += v.x;
                s4.y += v.y;
                s4.z += v.z;
                s4.w += v.w;
            }
            const int owner = col / Hc, ul = col - owner * Hc;
            float* remote = cluster.map_shared_rank(slots, owner);
            *reinterpret_cast<float4*>(remote + (((size_t)wr * CL + rank) * Hc + ul) * NB + bq * 4) = s4;
        }
        cluster.sync();
    }
    if (ln) {
        // parameter gradients of the LayerNorm: sums over time sit in registers, the sequences of this cluster are added
        // in a fixed order; one row per cluster, the caller adds the rows
#pragma unroll
        for (int it = 0; it < IPT; ++it) {
            if (!it_in[it]) continue;
            dpre_s[((size_t)0 * Hc + it_ul[it]) * NB + it_b[it]] = acc_dg[it];
            dpre_s[((size_t)1 * Hc + it_ul[it]) * NB + it_b[it]] = acc_db[it];
        }
        __syncthreads();
        for (int i = t; i < 2 * nu; i += nthreads) {
            const int which = i / nu, ul = i - which * nu;
            float sum = 0.f;
            for (int b = 0; b < nb; ++b) sum += dpre_s[((size_t)which * Hc + ul) * NB + b];
            (which ? a.ln_dbeta : a.ln_dgamma)[(size_t)group * H + u0 + ul] = sum;
        }
    }
}

// ---- pack kernels: gather the masked dense U_g into the resident layouts -----------------------
__global__ void lstm_pack_fwd_kernel(const uint16_t* __restrict__ bidx, PtrPack<const float> U, int ldu, int H,
                                     int CL, int L, int RG, float* __restrict__ out) {
    const int64_t total = (int64_t)CL * L * RG * 64;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int ci = idx & 3, ri = (idx >> 2) & 3, g = (idx >> 4) & 3;
        const int64_t blk = idx >> 6;  // (r*L + j)*RG + a_l
        const int a_l = (int)(blk % RG);
        const int r = (int)(blk / ((int64_t)L * RG));
        const uint16_t c = bidx[blk];
        float v = 0.f;
        if (c != 0xFFFF) {
            const int u = (r * RG + a_l) * 4 + ri, col = c * 4 + ci;
            if (u < H && col < H) v = __ldg(U.p[g] + (int64_t)u * ldu + col);
        }
        out[idx] = v;
    }
}

__global__ void lstm_pack_bwd_kernel(const int* __restrict__ cta_blk_off, const uint16_t* __restrict__ blk_a,
                                     const uint16_t* __restrict__ blk_c, PtrPack<const float> U, int ldu, int H,
                                     int CL, int RG, int n_blocks, float* __restrict__ out) {
    const int64_t total = (int64_t)n_blocks * 64;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int ri = idx & 3, g = (idx >> 2) & 3, ci = (idx >> 4) & 3;
        const int s = (int)(idx >> 6);
        int r = 0;
        while (r + 1 < CL && cta_blk_off[r + 1] <= s) ++r;
        const int u = (r * RG + blk_a[s]) * 4 + ri, col = blk_c[s] * 4 + ci;
        float v = 0.f;
        if (u < H && col < H) v = __ldg(U.p[g] + (int64_t)u * ldu + col);
        out[idx] = v;
    }
}

template <typename Args>
static int launch_cluster(void (*kernel)(const Args), const Args& args, int n_clusters, int CL, int threads,
                          size_t smem, cudaStream_t st) {
    HCGS_CHECK_CUDA(cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (CL > 8)
        HCGS_CHECK_CUDA(cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(n_clusters * CL), 1, 1);
    cfg.blockDim = dim3((unsigned)threads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HCGS_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kernel, args));
    return 0;
}

template <typename Args>
static int max_resident_clusters(void (*kernel)(const Args), int CL, int threads, size_t smem, int* out) {
    HCGS_CHECK_CUDA(cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (CL > 8)
        HCGS_CHECK_CUDA(cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)CL, 1, 1);
    cfg.blockDim = dim3((unsigned)threads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HCGS_CHECK_CUDA(cudaOccupancyMaxActiveClusters(out, kernel, &cfg));
    return 0;
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_lstm_pack(const hcgs_plan* plan, const float* const* u, int ldu, float* fwd_pack,
                              float* bwd_pack, hcgs_stream stream) {
    HCGS_REQUIRE(plan && plan->recurrent && u, "hcgs_lstm_pack: needs a recurrent plan");
    const RecurrentPlan& rp = plan->rp;
    cudaStream_t st = (cudaStream_t)stream;
    PtrPack<const float> U = make_pack<const float>(u, 4);
    if (fwd_pack) {
        const int64_t total = rp.fwd_pack_floats;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        lstm_pack_fwd_kernel<<<blocks, 256, 0, st>>>(rp.d_bidx, U, ldu, plan->n_out, rp.CL, rp.L, rp.RG, fwd_pack);
        HCGS_LAUNCH_CHECK();
    }
    if (bwd_pack && rp.n_blocks > 0) {
        const int64_t total = rp.bwd_pack_floats;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        lstm_pack_bwd_kernel<<<blocks, 256, 0, st>>>(rp.d_cta_blk_off, rp.d_blk_a, rp.d_blk_c, U, ldu, plan->n_out,
                                                    rp.CL, rp.RG, rp.n_blocks, bwd_pack);
        HCGS_LAUNCH_CHECK();
    }
    return 0;
}

static int lstm_fwd_impl(const hcgs_plan* plan, const float* fwd_pack, const float* pre, const float* drop,
                         float drop_scalar, int act, int t_len, int batch, float* h_out, float* c_out, float* gates_out,
                         int act_bits, const float* ln_gamma, const float* ln_beta, float ln_eps, float* ln_stats,
                         hcgs_stream stream) {
    HCGS_REQUIRE(plan && plan->recurrent, "hcgs_lstm_fwd: needs a recurrent plan");
    HCGS_REQUIRE(fwd_pack && pre && h_out, "hcgs_lstm_fwd: NULL argument");
    HCGS_REQUIRE(act_bits >= 0 && act_bits <= 24, "hcgs_lstm_fwd: act_bits out of range");
    if (t_len <= 0 || batch <= 0) return 0;
    const RecurrentPlan& rp = plan->rp;
    HCGS_REQUIRE(rp.Hc * NB <= IPT * rp.threads, "hcgs_lstm_fwd: internal sizing error");
    LstmFwdArgs a;
    a.wpack = fwd_pack;
    a.bidx = rp.d_bidx;
    a.pre = pre;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.h_out = h_out;
    a.c_out = c_out;
    a.gates_out = gates_out;
    a.T = t_len;
    a.B = batch;
    a.H = plan->n_out;
    a.CL = rp.CL;
    a.Hc = rp.Hc;
    a.HcP = rp.HcP;
    a.RG = rp.RG;
    a.L = rp.L;
    a.NCG = rp.NCG;
    a.act = act;
    a.w_in_smem = rp.w_in_smem ? 1 : 0;
    a.act_bits = act_bits;
    a.ln_gamma = ln_gamma;
    a.ln_beta = ln_beta;
    a.ln_eps = ln_eps;
    a.ln_stats = ln_stats;
    const int n_clusters = (batch + NB - 1) / NB;
    a.n_groups = n_clusters;
    a.q_sync = nullptr;
    if (act_bits > 0 && n_clusters > 1) {
        // Quantize_inp of h_t scales by the max over the WHOLE batch (quantized_modules.py:104-108): the clusters meet
        // once per time step, which needs all of them on the GPU at the same time
        int resident = 0;
        if (int rc = max_resident_clusters(lstm_fwd_kernel, rp.CL, rp.threads, rp.smem_fwd, &resident)) return rc;
        HCGS_REQUIRE(n_clusters <= resident,
                     "hcgs_lstm_fwd: Quantize_inp of h_t needs the whole batch co-resident: %d sequences = %d clusters of %d "
                     "CTAs, this GPU holds %d clusters (<= %d sequences)", batch, n_clusters, rp.CL, resident, resident * NB);
        void* sym = nullptr;
        HCGS_CHECK_CUDA(cudaGetSymbolAddress(&sym, g_q_sync));
        a.q_sync = static_cast<unsigned int*>(sym);
        HCGS_CHECK_CUDA(cudaMemsetAsync(sym, 0, 4 * sizeof(unsigned int), (cudaStream_t)stream));
    }
    return launch_cluster(lstm_fwd_kernel, a, n_clusters, rp.CL, rp.threads, rp.smem_fwd, (cudaStream_t)stream);
}

extern "C" int hcgs_lstm_fwd(const hcgs_plan* plan, const float* fwd_pack, const float* pre, const float* drop,
                             float drop_scalar, int act, int t_len, int batch, float* h_out, float* c_out,
                             float* gates_out, int act_bits, hcgs_stream stream) {
    return lstm_fwd_impl(plan, fwd_pack, pre, drop, drop_scalar, act, t_len, batch, h_out, c_out, gates_out, act_bits,
                         nullptr, nullptr, 0.f, nullptr, stream);
}

extern "C" int hcgs_lstm_fwd_ln(const hcgs_plan* plan, const float* fwd_pack, const float* pre, const float* drop,
                                float drop_scalar, int act, int t_len, int batch, float* h_out, float* c_out,
                                float* gates_out, int act_bits, const float* ln_gamma, const float* ln_beta,
                                float ln_eps, float* ln_stats, hcgs_stream stream) {
    HCGS_REQUIRE(ln_gamma && ln_beta, "hcgs_lstm_fwd_ln: NULL LayerNorm parameters");
    HCGS_REQUIRE(plan && plan->n_out >= 2, "hcgs_lstm_fwd_ln: the unbiased std needs at least two hidden units");
    return lstm_fwd_impl(plan, fwd_pack, pre, drop, drop_scalar, act, t_len, batch, h_out, c_out, gates_out, act_bits,
                         ln_gamma, ln_beta, ln_eps, ln_stats, stream);
}

static int lstm_bwd_impl(const hcgs_plan* plan, const float* bwd_pack, const float* dh_out, const float* gates,
                         const float* c, const float* drop, float drop_scalar, int act, int t_len, int batch, float* dpre,
                         const float* ln_gamma, const float* ln_stats, float ln_eps, float* ln_dgamma, float* ln_dbeta,
                         hcgs_stream stream) {
    HCGS_REQUIRE(plan && plan->recurrent, "hcgs_lstm_bwd: needs a recurrent plan");
    HCGS_REQUIRE(bwd_pack && dh_out && gates && c && dpre, "hcgs_lstm_bwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    const RecurrentPlan& rp = plan->rp;
    LstmBwdArgs a;
    a.wT = bwd_pack;
    a.cta_blk_off = rp.d_cta_blk_off;
    a.blk_a = rp.d_blk_a;
    a.cg_ptr = rp.d_cg_ptr;
    a.cg_list = rp.d_cg_list;
    a.dh_out = dh_out;
    a.gates = gates;
    a.c = c;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.dpre = dpre;
    a.T = t_len;
    a.B = batch;
    a.H = plan->n_out;
    a.CL = rp.CL;
    a.Hc = rp.Hc;
    a.RG = rp.RG;
    a.NCG = rp.NCG;
    a.max_blocks = rp.max_cta_blocks;
    a.act = act;
    a.w_in_smem = rp.w_in_smem ? 1 : 0;
    a.ln_gamma = ln_gamma;
    a.ln_stats = ln_stats;
    a.ln_eps = ln_eps;
    a.ln_dgamma = ln_dgamma;
    a.ln_dbeta = ln_dbeta;
    const int n_clusters = (batch + NB - 1) / NB;
    return launch_cluster(lstm_bwd_kernel, a, n_clusters, rp.CL, rp.threads, rp.smem_bwd, (cudaStream_t)stream);
}

extern "C" int hcgs_lstm_bwd(const hcgs_plan* plan, const float* bwd_pack, const float* dh_out, const float* gates,
                             const float* c, const float* drop, float drop_scalar, int act, int t_len, int batch,
                             float* dpre, hcgs_stream stream) {
    return lstm_bwd_impl(plan, bwd_pack, dh_out, gates, c, drop, drop_scalar, act, t_len, batch, dpre, nullptr, nullptr,
                         0.f, nullptr, nullptr, stream);
}

extern "C" int hcgs_lstm_bwd_ln(const hcgs_plan* plan, const float* bwd_pack, const float* dh_out, const float* gates,
                                const float* c, const float* drop, float drop_scalar, int act, int t_len, int batch,
                                float* dpre, const float* ln_gamma, const float* ln_stats, float ln_eps,
                                float* ln_dgamma, float* ln_dbeta, hcgs_stream stream) {
    HCGS_REQUIRE(ln_gamma && ln_stats && ln_dgamma && ln_dbeta, "hcgs_lstm_bwd_ln: NULL LayerNorm argument");
    return lstm_bwd_impl(plan, bwd_pack, dh_out, gates, c, drop, drop_scalar, act, t_len, batch, dpre, ln_gamma, ln_stats,
                         ln_eps, ln_dgamma, ln_dbeta, stream);
}
