// Persistent recurrent kernels for the GRU family (fp32 build): GRU, minimalGRU, liGRU and the plain RNN.
//
// Reference loops (dense there; block-sparse here behind the same option names as the LSTM's HCGS keys):
//   GRU        neural_networks.py:1395-1408   z = s(wz + Uz h)  r = s(wr + Ur h)  a = wh + Uh (r*h)
//   minimalGRU neural_networks.py:1752-1763   z = s(wz + Uz h)                    a = wh + Uh (z*h)
//   liGRU      neural_networks.py:1574-1585   z = s(wz + Uz h)                    a = wh + Uh h
//   RNN        neural_networks.py:1907-1916                                       a = wh + Uh h
//   all:       hcand = act(a) * drop ;  h = z*h + (1-z)*hcand   (RNN: h = hcand)
//
// Same machine as lstm.cu: one thread-block cluster per NB = 8 sequences runs the whole time loop of a layer, CTA r
// owns Hc hidden units, the compressed recurrent weights stay resident in shared memory in the LSTM's "BELL-4" pack
// (four gate slots per kept 4x4 block: slot 0 = Uz, 1 = Ur, 2 = Uh, 3 unused), h_t is all-gathered through distributed
// shared memory.  GRU and minimalGRU are TWO-PHASE cells: the candidate's product runs on a gated copy m = r*h (z*h)
// that does not exist before the first product has finished, so a step has two exchanges (m, then h_t); liGRU and RNN
// need one.  Backward mirrors it: the gradient of m comes back through Uh^T (reduce-scatter 1), only then the reset /
// update gate derivatives exist and go through Ur^T / Uz^T (reduce-scatter 2).
//
// Buffers hold NG = 3 / 2 / 2 / 1 gate groups of H columns each, gate-major like the LSTM's fp32 build:
//   pre / gates / dpre [T, B, NG*H]; column offsets (in units of H) of z, r, h are derived from the cell kind.
#include <cooperative_groups.h>

#include <algorithm>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace hcgs {

namespace {

constexpr int GNB = 8;   // sequences per cluster
constexpr int GIPT = 2;  // pointwise items (unit, sequence) per thread

__host__ __device__ inline size_t g_align16(size_t x) { return (x + 15) & ~(size_t)15; }

struct CellInfo {
    int ng;          // gate groups in the time-batched buffers
    int oz, orr, oh; // column offset (units of H) of z / r / candidate, -1 = the cell has no such gate
    int two_phase;   // candidate product runs on a gated copy of h
};

__host__ __device__ inline CellInfo cell_info(int cell) {
    CellInfo c;
    switch (cell) {
        case HCGS_CELL_GRU: c.ng = 3, c.oz = 0, c.orr = 1, c.oh = 2, c.two_phase = 1; break;
        case HCGS_CELL_MINGRU: c.ng = 2, c.oz = 0, c.orr = -1, c.oh = 1, c.two_phase = 1; break;
        case HCGS_CELL_LIGRU: c.ng = 2, c.oz = 0, c.orr = -1, c.oh = 1, c.two_phase = 0; break;
        default: c.ng = 1, c.oz = -1, c.orr = -1, c.oh = 0, c.two_phase = 0; break;  // HCGS_CELL_RNN
    }
    return c;
}

size_t gru_fwd_smem(int H, int Hc, int HcP, int L, bool w_in_smem, bool two_phase) {
    const int RG = Hc / 4, H4 = (H + 3) / 4 * 4;
    size_t b = 0;
    if (w_in_smem) b += g_align16((size_t)L * RG * 64 * 4);
    b += g_align16((size_t)L * RG * 2);                        // block column index
    b += g_align16((size_t)2 * H4 * GNB * 4);                  // h double buffer
    if (two_phase) b += g_align16((size_t)H4 * GNB * 4);       // gated copy m
    b += g_align16((size_t)4 * GNB * HcP * 4);                 // recurrent products of my rows
    b += g_align16((size_t)Hc * GNB * 4);                      // stage of the slice being pushed
    return b;
}

size_t gru_bwd_smem(int CL, int Hc, int max_blocks, int n_colgroups, bool w_in_smem, bool two_phase) {
    size_t b = 0;
    if (w_in_smem) b += g_align16((size_t)max_blocks * 64 * 4);
    b += g_align16((size_t)max_blocks * 2);                    // row group of each block
    b += g_align16((size_t)(n_colgroups + 1) * 4);             // CSR over column groups
    b += g_align16((size_t)max_blocks * 2);                    // block slots by column group
    b += g_align16((size_t)4 * Hc * GNB * 4);                  // dpre of my rows
    b += g_align16((size_t)max_blocks * 32 * 4);               // per-block partial products
    b += g_align16((size_t)2 * CL * Hc * GNB * 4);             // received partial dh (double buffered)
    if (two_phase) b += g_align16((size_t)CL * Hc * GNB * 4);  // received partial dm
    return b;
}

struct GruFwdArgs {
    const float* wpack;
    const uint16_t* bidx;
    const float* pre;    // [T,B,NG*H]
    const float* drop;   // [B,H] or NULL
    float drop_scalar;
    float* h_out;        // [T,B,H]
    float* m_out;        // [T,B,H] gated copy fed to Uh (two-phase cells), or NULL
    float* gates_out;    // [T,B,NG*H] post-activation z, r, g = act(a), or NULL
    int T, B, H, CL, Hc, HcP, RG, L, NCG, act, w_in_smem, cell;
};

// push my [Hc][GNB] stage into slice u0 of `dst_local` in every CTA of the cluster
__device__ __forceinline__ void gru_push_slice(cg::cluster_group& cluster, const float* stage, float* dst_local,
                                               int CL, int chunks, int u0, int H4, int t, int nthreads) {
    for (int idx = t; idx < CL * chunks; idx += nthreads) {
        const int dst = idx / chunks, q = idx - dst * chunks;
        const int ul = q >> 1;
        if (u0 + ul < H4) {
            const float4 v = reinterpret_cast<const float4*>(stage)[q];
            float* remote = cluster.map_shared_rank(dst_local, dst);
            reinterpret_cast<float4*>(remote + (size_t)u0 * GNB)[q] = v;
        }
    }
}

__global__ void __launch_bounds__(512) gru_fwd_kernel(const GruFwdArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const CellInfo ci = cell_info(a.cell);
    const int rank = (int)cluster.block_rank();
    const int b0 = (blockIdx.x / a.CL) * GNB;
    const int nb = min(GNB, a.B - b0);
    const int u0 = rank * a.Hc;
    const int nu = max(0, min(a.Hc, a.H - u0));
    const int H4 = a.NCG * 4;
    const int t = threadIdx.x, nthreads = blockDim.x;
    const int H = a.H, Hc = a.Hc, HcP = a.HcP, RG = a.RG, L = a.L;
    const int ld = ci.ng * H;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    float* w_s = nullptr;
    if (a.w_in_smem) {
        w_s = reinterpret_cast<float*>(sp);
        sp += g_align16((size_t)L * RG * 64 * 4);
    }
    uint16_t* bidx_s = reinterpret_cast<uint16_t*>(sp);
    sp += g_align16((size_t)L * RG * 2);
    float* hbuf = reinterpret_cast<float*>(sp);
    sp += g_align16((size_t)2 * H4 * GNB * 4);
    float* mbuf = nullptr;
    if (ci.two_phase) {
        mbuf = reinterpret_cast<float*>(sp);
        sp += g_align16((size_t)H4 * GNB * 4);
    }
    float* gbuf = reinterpret_cast<float*>(sp);
    sp += g_align16((size_t)4 * GNB * HcP * 4);
    float* stage = reinterpret_cast<float*>(sp);

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
    for (int i = t; i < 2 * H4 * GNB; i += nthreads) hbuf[i] = 0.f;
    if (ci.two_phase)
        for (int i = t; i < H4 * GNB; i += nthreads) mbuf[i] = 0.f;
    __syncthreads();
    cluster.sync();

    // mat-vec role: (row group, gate slot, sequence half, row half); which slots work in which phase
    const bool mv = t < RG * 16;
    const int a_l = t >> 4, g = (t >> 2) & 3, bh = (t >> 1) & 1, rh = t & 1;
    bool gate_a, gate_b = false;
    if (a.cell == HCGS_CELL_GRU) gate_a = g < 2, gate_b = g == 2;
    else if (a.cell == HCGS_CELL_MINGRU) gate_a = g == 0, gate_b = g == 2;
    else if (a.cell == HCGS_CELL_LIGRU) gate_a = g == 0 || g == 2;
    else gate_a = g == 2;

    auto matvec = [&](const float* hv) {
        float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        const float* wb = wptr + ((size_t)a_l * 4 + g) * 16 + rh * 8;
        const uint16_t* bi = bidx_s + a_l;
#pragma unroll 3
        for (int j = 0; j < L; ++j) {
            const int c = bi[(size_t)j * RG];
            const float4 w0 = *reinterpret_cast<const float4*>(wb + (size_t)j * RG * 64);
            const float4 w1 = *reinterpret_cast<const float4*>(wb + (size_t)j * RG * 64 + 4);
            const float* hp = hv + (size_t)c * 4 * GNB + bh * 4;
            const float4 h0 = *reinterpret_cast<const float4*>(hp);
            const float4 h1 = *reinterpret_cast<const float4*>(hp + GNB);
            const float4 h2 = *reinterpret_cast<const float4*>(hp + 2 * GNB);
            const float4 h3 = *reinterpret_cast<const float4*>(hp + 3 * GNB);
            acc0[0] = fmaf(w0.x, h0.x, fmaf(w0.y, h1.x, fmaf(w0.z, h2.x, fmaf(w0.w, h3.x, acc0[0]))));
            acc0[1] = fmaf(w0.x, h0.y, fmaf(w0.y, h1.y, fmaf(w0.z, h2.y, fmaf(w0.w, h3.y, acc0[1]))));
            acc0[2] = fmaf(w0.x, h0.z, fmaf(w0.y, h1.z, fmaf(w0.z, h2.z, fmaf(w0.w, h3.z, acc0[2]))));
            acc0[3] = fmaf(w0.x, h0.w, fmaf(w0.y, h1.w, fmaf(w0.z, h2.w, fmaf(w0.w, h3.w, acc0[3]))));
            acc1[0] = fmaf(w1.x, h0.x, fmaf(w1.y, h1.x, fmaf(w1.z, h2.x, fmaf(w1.w, h3.x, acc1[0]))));
            acc1[1] = fmaf(w1.x, h0.y, fmaf(w1.y, h1.y, fmaf(w1.z, h2.y, fmaf(w1.w, h3.y, acc1[1]))));
            acc1[2] = fmaf(w1.x, h0.z, fmaf(w1.y, h1.z, fmaf(w1.z, h2.z, fmaf(w1.w, h3.z, acc1[2]))));
            acc1[3] = fmaf(w1.x, h0.w, fmaf(w1.y, h1.w, fmaf(w1.z, h2.w, fmaf(w1.w, h3.w, acc1[3]))));
        }
        float* gp = gbuf + ((size_t)g * GNB + bh * 4) * HcP + a_l * 4 + rh * 2;   // gbuf[slot][b][unit]
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            gp[(size_t)b * HcP] = acc0[b];
            gp[(size_t)b * HcP + 1] = acc1[b];
        }
    };

    int it_b[GIPT], it_ul[GIPT];
    bool it_in[GIPT], it_ok[GIPT];
    float hprev[GIPT], dr[GIPT], z_reg[GIPT], r_reg[GIPT];
#pragma unroll
    for (int it = 0; it < GIPT; ++it) {
        const int i = t + it * nthreads;
        it_in[it] = i < Hc * GNB;
        it_b[it] = it_in[it] ? i / Hc : 0;
        it_ul[it] = it_in[it] ? i % Hc : 0;
        it_ok[it] = it_in[it] && it_ul[it] < nu && it_b[it] < nb;
        hprev[it] = 0.f;
        z_reg[it] = r_reg[it] = 0.f;
        dr[it] = a.drop_scalar;
        if (it_ok[it] && a.drop) dr[it] = __ldg(a.drop + (size_t)(b0 + it_b[it]) * H + u0 + it_ul[it]);
    }
    const int chunks = Hc * GNB / 4;  // float4 chunks of one slice

    for (int k = 0; k < a.T; ++k) {
        const int cur = k & 1, nxt = cur ^ 1;
        const float* hc = hbuf + (size_t)cur * H4 * GNB;
        // ---- phase A: products on h_{k-1} -----------------------------------------------------------------------
        if (mv && gate_a) matvec(hc);
        __syncthreads();
        if (ci.oz >= 0) {
#pragma unroll
            for (int it = 0; it < GIPT; ++it) {
                if (!it_in[it]) continue;
                const int b = it_b[it], ul = it_ul[it];
                float m = 0.f;
                if (it_ok[it]) {
                    const float* pp = a.pre + ((size_t)k * a.B + b0 + b) * ld + u0 + ul;
                    z_reg[it] = sigmoidf_(gbuf[((size_t)0 * GNB + b) * HcP + ul] + __ldg(pp + (size_t)ci.oz * H));
                    if (ci.orr >= 0)
                        r_reg[it] = sigmoidf_(gbuf[((size_t)1 * GNB + b) * HcP + ul] + __ldg(pp + (size_t)ci.orr * H));
                    m = (ci.orr >= 0 ? r_reg[it] : z_reg[it]) * hprev[it];
                    if (ci.two_phase && a.m_out) a.m_out[((size_t)k * a.B + b0 + b) * H + u0 + ul] = m;
                }
                if (ci.two_phase) stage[ul * GNB + b] = m;
            }
        }
        if (ci.two_phase) {
            __syncthreads();
            gru_push_slice(cluster, stage, mbuf, a.CL, chunks, u0, H4, t, nthreads);
            cluster.sync();
            // ---- phase B: candidate product on the gated copy ---------------------------------------------------
            if (mv && gate_b) matvec(mbuf);
            __syncthreads();
        }
        // ---- candidate, new hidden state ----------------------------------------------------------------------------
#pragma unroll
        for (int it = 0; it < GIPT; ++it) {
            if (!it_in[it]) continue;
            const int b = it_b[it], ul = it_ul[it];
            float hval = 0.f;
            if (it_ok[it]) {
                const size_t row = (size_t)k * a.B + b0 + b;
                const int u = u0 + ul;
                const float av = gbuf[((size_t)2 * GNB + b) * HcP + ul] + __ldg(a.pre + row * ld + (size_t)ci.oh * H + u);
                const float gact = act_apply(a.act, av);
                const float hcand = gact * dr[it];
                hval = ci.oz >= 0 ? z_reg[it] * hprev[it] + (1.f - z_reg[it]) * hcand : hcand;
                hprev[it] = hval;
                a.h_out[row * H + u] = hval;
                if (a.gates_out) {
                    float* go = a.gates_out + row * ld + u;
                    if (ci.oz >= 0) go[(size_t)ci.oz * H] = z_reg[it];
                    if (ci.orr >= 0) go[(size_t)ci.orr * H] = r_reg[it];
                    go[(size_t)ci.oh * H] = gact;
                }
            }
            stage[ul * GNB + b] = hval;
        }
        __syncthreads();
        gru_push_slice(cluster, stage, hbuf + (size_t)nxt * H4 * GNB, a.CL, chunks, u0, H4, t, nthreads);
        cluster.sync();
    }
}

struct GruBwdArgs {
    const float* wT;          // [n_blocks][ci][g][ri]  (lstm_pack_bwd layout, slots z, r, h, -)
    const int* cta_blk_off;   // [CL+1]
    const uint16_t* blk_a;    // [n_blocks]
    const int* cg_ptr;        // [CL][NCG+1]
    const uint16_t* cg_list;  // [n_blocks]
    const float* dh_out;      // [T,B,H]
    const float* gates;       // [T,B,NG*H]  z, r, g = act(a)
    const float* h;           // [T,B,H]     forward output (h_{k-1} of every step)
    const float* drop;
    float drop_scalar;
    float* dpre;              // [T,B,NG*H]
    int T, B, H, CL, Hc, RG, NCG, max_blocks, act, w_in_smem, cell;
};

__global__ void __launch_bounds__(512) gru_bwd_kernel(const GruBwdArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    const CellInfo ci = cell_info(a.cell);
    const int rank = (int)cluster.block_rank();
    const int b0 = (blockIdx.x / a.CL) * GNB;
    const int nb = min(GNB, a.B - b0);
    const int u0 = rank * a.Hc;
    const int nu = max(0, min(a.Hc, a.H - u0));
    const int H4 = a.NCG * 4;
    const int t = threadIdx.x, nthreads = blockDim.x;
    const int H = a.H, Hc = a.Hc, CL = a.CL, NCG = a.NCG;
    const int ld = ci.ng * H;
    const int blk0 = a.cta_blk_off[rank];
    const int nblk = a.cta_blk_off[rank + 1] - blk0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    float* w_s = nullptr;
    if (a.w_in_smem) {
        w_s = reinterpret_cast<float*>(sp);
        sp += g_align16((size_t)a.max_blocks * 64 * 4);
    }
    uint16_t* blk_a_s = reinterpret_cast<uint16_t*>(sp);
    sp += g_align16((size_t)a.max_blocks * 2);
    int* cg_ptr_s = reinterpret_cast<int*>(sp);
    sp += g_align16((size_t)(NCG + 1) * 4);
    uint16_t* cg_list_s = reinterpret_cast<uint16_t*>(sp);
    sp += g_align16((size_t)a.max_blocks * 2);
    float* dpre_s = reinterpret_cast<float*>(sp);  // [4 slots][Hc][GNB]
    sp += g_align16((size_t)4 * Hc * GNB * 4);
    float* contrib = reinterpret_cast<float*>(sp);
    sp += g_align16((size_t)a.max_blocks * 32 * 4);
    float* slots = reinterpret_cast<float*>(sp);   // [2][CL][Hc][GNB]  partial dh_{k-1}
    sp += g_align16((size_t)2 * CL * Hc * GNB * 4);
    float* slots_m = reinterpret_cast<float*>(sp); // [CL][Hc][GNB]     partial dm (two-phase cells)

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
    for (int i = t; i < 2 * CL * Hc * GNB; i += nthreads) slots[i] = 0.f;
    if (ci.two_phase)
        for (int i = t; i < CL * Hc * GNB; i += nthreads) slots_m[i] = 0.f;
    for (int i = t; i < 4 * Hc * GNB; i += nthreads) dpre_s[i] = 0.f;
    __syncthreads();
    cluster.sync();

    // sum over my kept blocks of U_g^T dpre_g for the gate slots in `gmask`, column sums pushed to the CTA owning the
    // column (reduce-scatter over DSMEM, summed there in rank order => deterministic)
    auto transposed_products = [&](unsigned gmask, float* dst_local) {
        for (int s = t >> 3; s < nblk; s += nthreads >> 3) {
            const int cix = (t >> 1) & 3, bh = t & 1;
            const int a_l = blk_a_s[s];
            const float* wb = wptr + ((size_t)s * 4 + cix) * 16;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int g = 0; g < 3; ++g) {
                if (!(gmask >> g & 1)) continue;
                const float4 w = *reinterpret_cast<const float4*>(wb + g * 4);
                const float* dp = dpre_s + ((size_t)g * Hc + a_l * 4) * GNB + bh * 4;
                const float4 d0 = *reinterpret_cast<const float4*>(dp);
                const float4 d1 = *reinterpret_cast<const float4*>(dp + GNB);
                const float4 d2 = *reinterpret_cast<const float4*>(dp + 2 * GNB);
                const float4 d3 = *reinterpret_cast<const float4*>(dp + 3 * GNB);
                acc[0] = fmaf(w.x, d0.x, fmaf(w.y, d1.x, fmaf(w.z, d2.x, fmaf(w.w, d3.x, acc[0]))));
                acc[1] = fmaf(w.x, d0.y, fmaf(w.y, d1.y, fmaf(w.z, d2.y, fmaf(w.w, d3.y, acc[1]))));
                acc[2] = fmaf(w.x, d0.z, fmaf(w.y, d1.z, fmaf(w.z, d2.z, fmaf(w.w, d3.z, acc[2]))));
                acc[3] = fmaf(w.x, d0.w, fmaf(w.y, d1.w, fmaf(w.z, d2.w, fmaf(w.w, d3.w, acc[3]))));
            }
            *reinterpret_cast<float4*>(contrib + ((size_t)s * 4 + cix) * GNB + bh * 4) =
                make_float4(acc[0], acc[1], acc[2], acc[3]);
        }
        __syncthreads();
        for (int w = t; w < H4 * 2; w += nthreads) {
            const int col = w >> 1, bq = w & 1;
            const int c = col >> 2, cix = col & 3;
            float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int e = cg_ptr_s[c]; e < cg_ptr_s[c + 1]; ++e) {
                const int s = cg_list_s[e];
                const float4 v = *reinterpret_cast<const float4*>(contrib + ((size_t)s * 4 + cix) * GNB + bq * 4);
                s4.x += v.x;
                s4.y += v.y;
                s4.z += v.z;
                s4.w += v.w;
            }
            const int owner = col / Hc, ul = col - owner * Hc;
            float* remote = cluster.map_shared_rank(dst_local, owner);
            *reinterpret_cast<float4*>(remote + ((size_t)rank * Hc + ul) * GNB + bq * 4) = s4;
        }
        cluster.sync();
    };

    int it_b[GIPT] = {}, it_ul[GIPT] = {};
    bool it_in[GIPT], it_ok[GIPT];
    float carry[GIPT], dr[GIPT];
    float p_z[GIPT], p_r[GIPT], p_g[GIPT], p_hm1[GIPT], p_dh[GIPT];   // operands of the step about to be processed
    auto prefetch = [&](int it, int k) {
        const size_t row = (size_t)k * a.B + b0 + it_b[it];
        const int u = u0 + it_ul[it];
        const float* gp = a.gates + row * ld + u;
        p_z[it] = ci.oz >= 0 ? __ldg(gp + (size_t)ci.oz * H) : 0.f;
        p_r[it] = ci.orr >= 0 ? __ldg(gp + (size_t)ci.orr * H) : 0.f;
        p_g[it] = __ldg(gp + (size_t)ci.oh * H);
        p_hm1[it] = (k > 0) ? __ldg(a.h + (row - a.B) * H + u) : 0.f;   // h_{-1} = 0
        p_dh[it] = __ldg(a.dh_out + row * H + u);
    };
#pragma unroll
    for (int it = 0; it < GIPT; ++it) {
        const int i = t + it * nthreads;
        it_in[it] = i < Hc * GNB;
        it_b[it] = it_in[it] ? i / Hc : 0;
        it_ul[it] = it_in[it] ? i % Hc : 0;
        it_ok[it] = it_in[it] && it_ul[it] < nu && it_b[it] < nb;
        carry[it] = 0.f;
        dr[it] = a.drop_scalar;
        if (it_ok[it] && a.drop) dr[it] = __ldg(a.drop + (size_t)(b0 + it_b[it]) * H + u0 + it_ul[it]);
        p_z[it] = p_r[it] = p_g[it] = p_hm1[it] = p_dh[it] = 0.f;
        if (it_ok[it] && a.T > 0) prefetch(it, a.T - 1);
    }

    for (int k = a.T - 1; k >= 0; --k) {
        const int wr = k & 1, rd = wr ^ 1;
        float dz[GIPT], da[GIPT];
        // ---- candidate and update-gate derivatives from the total dh_k --------------------------------------------
#pragma unroll
        for (int it = 0; it < GIPT; ++it) {
            dz[it] = da[it] = 0.f;
            if (!it_in[it]) continue;
            const int b = it_b[it], ul = it_ul[it];
            if (it_ok[it]) {
                float dh = p_dh[it] + carry[it];
                const float* sl = slots + ((size_t)rd * CL * Hc + ul) * GNB + b;
                for (int src = 0; src < CL; ++src) dh += sl[(size_t)src * Hc * GNB];
                float dhc = dh;
                carry[it] = 0.f;
                if (ci.oz >= 0) {
                    const float z = p_z[it];
                    dz[it] = dh * (p_hm1[it] - p_g[it] * dr[it]);
                    dhc = dh * (1.f - z);
                    carry[it] = dh * z;
                }
                da[it] = dhc * dr[it] * act_grad_from_out(a.act, p_g[it]);
            }
            dpre_s[((size_t)2 * Hc + ul) * GNB + b] = da[it];
            if (!ci.two_phase && ci.oz >= 0) {
                const float z = p_z[it];
                dpre_s[((size_t)0 * Hc + ul) * GNB + b] = dz[it] * z * (1.f - z);
            }
        }
        __syncthreads();
        if (ci.two_phase) {
            // ---- reduce-scatter 1: dm = Uh^T da; then the reset / update gate derivatives ---------------------------
            transposed_products(1u << 2, slots_m);
#pragma unroll
            for (int it = 0; it < GIPT; ++it) {
                if (!it_in[it]) continue;
                const int b = it_b[it], ul = it_ul[it];
                float dpz = 0.f, dpr = 0.f;
                if (it_ok[it]) {
                    float dm = 0.f;
                    const float* sl = slots_m + (size_t)ul * GNB + b;
                    for (int src = 0; src < CL; ++src) dm += sl[(size_t)src * Hc * GNB];
                    const float z = p_z[it];
                    if (ci.orr >= 0) {
                        const float r = p_r[it];
                        dpr = dm * p_hm1[it] * r * (1.f - r);
                        carry[it] += dm * r;
                    } else {
                        dz[it] += dm * p_hm1[it];
                        carry[it] += dm * z;
                    }
                    dpz = dz[it] * z * (1.f - z);
                    dz[it] = dpz;      // keep for the global store below
                    if (ci.orr >= 0) p_r[it] = dpr;
                }
                dpre_s[((size_t)0 * Hc + ul) * GNB + b] = dpz;
                dpre_s[((size_t)1 * Hc + ul) * GNB + b] = dpr;
            }
            __syncthreads();
        }
        // ---- gradient w.r.t. the gate pre-activations of step k; operands of step k-1 ----------------------------------
#pragma unroll
        for (int it = 0; it < GIPT; ++it) {
            if (!it_ok[it]) continue;
            const size_t row = (size_t)k * a.B + b0 + it_b[it];
            float* dp = a.dpre + row * ld + u0 + it_ul[it];
            dp[(size_t)ci.oh * H] = da[it];
            if (ci.oz >= 0) {
                const float z = p_z[it];
                dp[(size_t)ci.oz * H] = ci.two_phase ? dz[it] : dz[it] * z * (1.f - z);
            }
            if (ci.orr >= 0) dp[(size_t)ci.orr * H] = p_r[it];
            if (k > 0) prefetch(it, k - 1);
        }
        // ---- reduce-scatter 2: dh_{k-1} += Uz^T dpre_z (+ Ur^T dpre_r) (+ Uh^T da for the one-phase cells) ---------
        unsigned gmask;
        if (a.cell == HCGS_CELL_GRU) gmask = 3u;
        else if (a.cell == HCGS_CELL_MINGRU) gmask = 1u;
        else if (a.cell == HCGS_CELL_LIGRU) gmask = 5u;
        else gmask = 4u;
        transposed_products(gmask, slots + (size_t)wr * CL * Hc * GNB);
    }
}

}  // namespace

// lstm.cu
template <typename Args>
static int gru_launch_cluster(void (*kernel)(const Args), const Args& args, int n_clusters, int CL, int threads,
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

static int gru_smem_limit() {
    int dev = 0, smem_max = 232448;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    return smem_max;
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_gru_fwd(const hcgs_plan* plan, int cell, const float* fwd_pack, const float* pre,
                            const float* drop, float drop_scalar, int act, int t_len, int batch, float* h_out,
                            float* m_out, float* gates_out, hcgs_stream stream) {
    HCGS_REQUIRE(plan && plan->recurrent, "hcgs_gru_fwd: needs a recurrent plan");
    HCGS_REQUIRE(cell >= HCGS_CELL_GRU && cell <= HCGS_CELL_RNN, "hcgs_gru_fwd: unknown cell kind %d", cell);
    HCGS_REQUIRE(fwd_pack && pre && h_out, "hcgs_gru_fwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    const RecurrentPlan& rp = plan->rp;
    const CellInfo ci = cell_info(cell);
    HCGS_REQUIRE(rp.Hc * GNB <= GIPT * rp.threads, "hcgs_gru_fwd: internal sizing error");
    const int limit = gru_smem_limit();
    bool w_in = rp.w_in_smem;
    size_t smem = gru_fwd_smem(plan->n_out, rp.Hc, rp.HcP, rp.L, w_in, ci.two_phase != 0);
    if (w_in && smem > (size_t)limit) {   // the extra gathered buffer pushed it over: weights stay in L2
        w_in = false;
        smem = gru_fwd_smem(plan->n_out, rp.Hc, rp.HcP, rp.L, false, ci.two_phase != 0);
    }
    HCGS_REQUIRE(smem <= (size_t)limit, "hcgs_gru_fwd: recurrent state does not fit shared memory (H=%d)", plan->n_out);
    GruFwdArgs a;
    a.wpack = fwd_pack;
    a.bidx = rp.d_bidx;
    a.pre = pre;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.h_out = h_out;
    a.m_out = ci.two_phase ? m_out : nullptr;
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
    a.w_in_smem = w_in ? 1 : 0;
    a.cell = cell;
    const int n_clusters = (batch + GNB - 1) / GNB;
    return gru_launch_cluster(gru_fwd_kernel, a, n_clusters, rp.CL, rp.threads, smem, (cudaStream_t)stream);
}

extern "C" int hcgs_gru_bwd(const hcgs_plan* plan, int cell, const float* bwd_pack, const float* dh_out,
                            const float* gates, const float* h, const float* drop, float drop_scalar, int act,
                            int t_len, int batch, float* dpre, hcgs_stream stream) {
    HCGS_REQUIRE(plan && plan->recurrent, "hcgs_gru_bwd: needs a recurrent plan");
    HCGS_REQUIRE(cell >= HCGS_CELL_GRU && cell <= HCGS_CELL_RNN, "hcgs_gru_bwd: unknown cell kind %d", cell);
    HCGS_REQUIRE(bwd_pack && dh_out && gates && h && dpre, "hcgs_gru_bwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    const RecurrentPlan& rp = plan->rp;
    const CellInfo ci = cell_info(cell);
    const int limit = gru_smem_limit();
    bool w_in = rp.w_in_smem;
    size_t smem = gru_bwd_smem(rp.CL, rp.Hc, rp.max_cta_blocks, rp.NCG, w_in, ci.two_phase != 0);
    if (w_in && smem > (size_t)limit) {
        w_in = false;
        smem = gru_bwd_smem(rp.CL, rp.Hc, rp.max_cta_blocks, rp.NCG, false, ci.two_phase != 0);
    }
    HCGS_REQUIRE(smem <= (size_t)limit, "hcgs_gru_bwd: recurrent state does not fit shared memory (H=%d)", plan->n_out);
    GruBwdArgs a;
    a.wT = bwd_pack;
    a.cta_blk_off = rp.d_cta_blk_off;
    a.blk_a = rp.d_blk_a;
    a.cg_ptr = rp.d_cg_ptr;
    a.cg_list = rp.d_cg_list;
    a.dh_out = dh_out;
    a.gates = gates;
    a.h = h;
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
    a.w_in_smem = w_in ? 1 : 0;
    a.cell = cell;
    const int n_clusters = (batch + GNB - 1) / GNB;
    return gru_launch_cluster(gru_bwd_kernel, a, n_clusters, rp.CL, rp.threads, smem, (cudaStream_t)stream);
}
