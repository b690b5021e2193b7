// Tensor-core persistent recurrent LSTM kernels (bf16 operands, fp32 accumulate / state).
//
// Reference loop: neural_networks.py:1063-1100 (forward) and its autograd.
//
// Why mma.sync and not tcgen05 here: the HCGS recurrent masks are 4x4 blocks inside 32x32 blocks
// (hcgsh_block = 32,4).  tcgen05.mma needs one K list shared by 128 accumulator rows, i.e. it can
// only skip at the 32x32 level (3x the weights, 3x the shared memory, twice the cluster size).
// With m16n8k16 the *batch* is the M dimension and one instruction multiplies 16 sequences by the
// four kept 4x4 blocks of one 4-row strip for two gates (n = 4 units x 2 gates, k = 4 blocks x 4
// columns), so the true 4x4 sparsity is exploited and the four gates of a (unit, sequence) item
// land in ONE thread's accumulators: gates, cell update and dropout are fused in registers with
// no shared-memory transpose.  The step is bound by the h_t exchange, not by the tensor pipe
// (measured: profiles/), which is what the design optimises.
//
// One thread-block cluster of CL CTAs runs NB = 16*NT sequences through all T steps of a layer:
//   forward : CTA r owns SPC 4-row strips of U (all four gates), resident in shared memory as
//             ready-made B fragments.  Per step every warp computes its strips for all NB
//             sequences, applies the fused pointwise LSTM update, and the CTA pushes its bf16
//             slice of h_t into every peer's shared memory (coalesced 16 B DSMEM stores);
//             arrival is signalled per warp on the peers' mbarriers (release/acquire at cluster
//             scope) -- no barrier.cluster and no fence over global stores inside the loop.
//   backward: the same row slice is stored column-strip-major.  Per step: fused gate derivatives
//             (pointwise, from the saved gates / cell) -> bf16 dpre in shared memory -> per
//             column strip an MMA chain over this CTA's kept row blocks (k = 4 gates x 4 rows)
//             -> fp32 partial dh for ALL columns -> pushed to the owner of each column slice
//             (reduce-scatter over DSMEM, summed by the owner in rank order => deterministic).
#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace hcgs {

constexpr int TC_MAX_WARPS = 16;
constexpr int TC_HPAD = 8;  // bf16 elements of padding per h row (bank spread, keeps 16 B alignment)
constexpr int TC_PPAD = 4;  // fp32 elements of padding per partial-dh row
constexpr int TC_DPAD = 8;  // bf16 elements of padding per dpre row

struct TcPlan {
    int CL = 0, H = 0, NS = 0;  // cluster size, hidden size, 4-row strips in total
    int SPC = 0, Hc = 0, HP = 0;  // strips per CTA (even), units per CTA, padded hidden = CL*Hc
    int SPW = 0, WARPS = 0, KS = 0;  // forward: strips per warp, warps per CTA, max k-steps per strip
    int n_blocks = 0, max_cta_blocks = 0;  // kept 4x4 blocks (total / largest CTA)
    // forward tables
    uint16_t* d_cg = nullptr;  // [CL][SPC][KS*4] column group of block j of the strip (0xFFFF = padding)
    uint8_t* d_nks = nullptr;  // [CL][SPC] k-steps actually used by the strip
    // backward tables (blocks of CTA r in column-strip-major order)
    int* d_blk_off = nullptr;     // [CL+1]
    int* d_cs_ptr = nullptr;      // [CL][NS+1] CSR over column strips into the CTA's block list
    uint16_t* d_blk_rs = nullptr;  // [n_blocks] local row strip of block e
    uint16_t* d_blk_cs = nullptr;  // [n_blocks] column strip of block e
    int64_t fwd_pack_u32 = 0, bwd_pack_u32 = 0;
};

void tc_plan_free(TcPlan* t) {
    if (!t) return;
    cudaFree(t->d_cg);
    cudaFree(t->d_nks);
    cudaFree(t->d_blk_off);
    cudaFree(t->d_cs_ptr);
    cudaFree(t->d_blk_rs);
    cudaFree(t->d_blk_cs);
    delete t;
}

template <typename T>
static int tc_upload(const std::vector<T>& v, T** dptr) {
    *dptr = nullptr;
    HCGS_CHECK_CUDA(cudaMalloc((void**)dptr, std::max<size_t>(v.size(), 1) * sizeof(T)));
    if (!v.empty()) HCGS_CHECK_CUDA(cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

static int tc_slot(int CL) { return CL == 4 ? 0 : (CL == 8 ? 1 : (CL == 16 ? 2 : -1)); }

// Build (or fetch) the tensor-core layouts of `plan` for cluster size CL.
static int tc_plan_get(const hcgs_plan* cplan, int CL, TcPlan** out) {
    hcgs_plan* plan = const_cast<hcgs_plan*>(cplan);
    HCGS_REQUIRE(plan && plan->recurrent && plan->h_blk44, "tensor-core recurrence needs a recurrent plan");
    const int slot = tc_slot(CL);
    HCGS_REQUIRE(slot >= 0, "tensor-core recurrence: cluster size must be 4, 8 or 16 (got %d)", CL);
    if (plan->tcp[slot]) {
        *out = plan->tcp[slot];
        return 0;
    }
    TcPlan* t = new TcPlan();
    t->CL = CL;
    t->H = plan->n_out;
    t->NS = plan->ns44;
    const int NS = t->NS;
    t->SPC = (NS + CL - 1) / CL;
    if (t->SPC & 1) ++t->SPC;  // Hc multiple of 8 -> 16-byte chunks of bf16
    t->Hc = t->SPC * 4;
    t->HP = CL * t->Hc;
    t->SPW = (t->SPC + TC_MAX_WARPS - 1) / TC_MAX_WARPS;
    t->WARPS = (t->SPC + t->SPW - 1) / t->SPW;
    const uint8_t* occ = plan->h_blk44;
    // forward: per strip the kept column groups
    std::vector<std::vector<uint16_t>> lists((size_t)CL * t->SPC);
    int L = 1;
    for (int s = 0; s < NS; ++s) {
        auto& l = lists[s];
        for (int c = 0; c < NS; ++c)
            if (occ[(size_t)s * NS + c]) l.push_back((uint16_t)c);
        L = std::max<int>(L, (int)l.size());
    }
    t->KS = (L + 3) / 4;
    std::vector<uint16_t> cgt((size_t)CL * t->SPC * t->KS * 4, 0xFFFF);
    std::vector<uint8_t> nks((size_t)CL * t->SPC, 0);
    for (int s = 0; s < CL * t->SPC; ++s) {
        const auto& l = lists[s];
        for (size_t j = 0; j < l.size(); ++j) cgt[(size_t)s * t->KS * 4 + j] = l[j];
        nks[s] = (uint8_t)((l.size() + 3) / 4);
    }
    // backward: per CTA, blocks grouped by column strip
    std::vector<int> blk_off(CL + 1, 0), cs_ptr((size_t)CL * (NS + 1), 0);
    std::vector<uint16_t> blk_rs, blk_cs;
    for (int r = 0; r < CL; ++r) {
        blk_off[r] = (int)blk_rs.size();
        int run = 0;
        for (int c = 0; c < NS; ++c) {
            cs_ptr[(size_t)r * (NS + 1) + c] = run;
            for (int sl = 0; sl < t->SPC; ++sl) {
                const int s = r * t->SPC + sl;
                if (s < NS && occ[(size_t)s * NS + c]) {
                    blk_rs.push_back((uint16_t)sl);
                    blk_cs.push_back((uint16_t)c);
                    ++run;
                }
            }
        }
        cs_ptr[(size_t)r * (NS + 1) + NS] = run;
        t->max_cta_blocks = std::max(t->max_cta_blocks, run);
    }
    blk_off[CL] = (int)blk_rs.size();
    t->n_blocks = (int)blk_rs.size();
    t->fwd_pack_u32 = (int64_t)CL * t->SPC * t->KS * 2 * 32 * 2;
    t->bwd_pack_u32 = (int64_t)std::max(t->n_blocks, 1) * 16 * 2;
    int rc = 0;
    rc |= tc_upload(cgt, &t->d_cg);
    rc |= tc_upload(nks, &t->d_nks);
    rc |= tc_upload(blk_off, &t->d_blk_off);
    rc |= tc_upload(cs_ptr, &t->d_cs_ptr);
    rc |= tc_upload(blk_rs, &t->d_blk_rs);
    rc |= tc_upload(blk_cs, &t->d_blk_cs);
    if (rc) {
        tc_plan_free(t);
        return 1;
    }
    plan->tcp[slot] = t;
    *out = t;
    return 0;
}

__host__ __device__ static inline size_t tc_align16(size_t x) { return (x + 15) & ~(size_t)15; }

static size_t tc_fwd_smem(const TcPlan& t, int NB) {
    size_t b = 0;
    b += tc_align16((size_t)t.SPC * t.KS * 2 * 32 * 8);               // B fragments
    b += tc_align16((size_t)t.SPC * t.KS * 4 * 2);                    // column groups
    b += tc_align16((size_t)t.SPC);                                   // k-steps per strip
    b += tc_align16((size_t)2 * NB * (t.HP + TC_HPAD) * 2);           // h double buffer (bf16)
    b += 64;                                                          // mbarriers
    return b;
}

static size_t tc_bwd_smem(const TcPlan& t, int NB) {
    size_t b = 0;
    b += tc_align16((size_t)t.max_cta_blocks * 16 * 8);               // B fragments (16 lanes x 8 B)
    b += tc_align16((size_t)t.max_cta_blocks * 2);                    // row strip of each block
    b += tc_align16((size_t)(t.NS + 1) * 4);                          // CSR over column strips
    b += tc_align16((size_t)NB * (t.SPC * 16 + TC_DPAD) * 2);         // dpre (bf16) [b][strip][g][i]
    b += tc_align16((size_t)NB * (t.HP + TC_PPAD) * 4);               // partial dh (fp32) [b][col]
    b += tc_align16((size_t)2 * t.CL * NB * t.Hc * 4);                // slots [2][src][b][ul]
    b += 64;
    return b;
}

// ---- PTX helpers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}

__device__ __forceinline__ void mbar_init(uint32_t addr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

// one release fence for a burst of relaxed arrives (a release per arrive serialises CL fences)
__device__ __forceinline__ void fence_release_cluster() { asm volatile("fence.acq_rel.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_cluster_relaxed(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    return ok;
}

// Bounded wait: a protocol bug must end in a trap (reported as a launch failure), never in a hang.
__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity) {
    for (int spin = 0; spin < (1 << 22); ++spin)
        if (mbar_try_wait(addr, parity)) return;
    __trap();
}

__device__ __forceinline__ void st_cluster_v4(uint32_t cluster_addr, const uint4& v) {
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

// ---- forward -----------------------------------------------------------------------------------
struct TcFwdArgs {
    const uint2* wfrag;   // [CL][SPC][KS][2][32]
    const uint16_t* cg;   // [CL][SPC][KS*4]
    const uint8_t* nks;   // [CL][SPC]
    const float* pre;     // element (row, g, u) at pre[row*ld + g*gs + u*us]
    int64_t ld;
    int gs, us;
    const float* drop;    // [B,H] or NULL
    float drop_scalar;
    float* h_out;         // [T,B,H] fp32
    __nv_bfloat16* hb_out;  // [T*B, hb_ld] bf16 copy for the tensor-core GEMMs, or NULL
    int hb_ld;
    float* c_out;         // [T,B,H] or NULL
    float* gates_out;     // same addressing as pre, or NULL
    int T, B, H, CL, SPC, SPW, KS, HP, act;
    int dbg;  // timing experiments only (HCGS_TC_DEBUG): 1 = no global stores, 2 = no prefetch loads, 4 = no MMA
};

// SPW = strips per warp (compile time so that the per-item state lives in registers)
template <int NT, int SPW>
__global__ void __launch_bounds__(TC_MAX_WARPS * 32, 1) lstm_tc_fwd_kernel(const TcFwdArgs a) {
    constexpr int NB = NT * 16;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int CL = a.CL, SPC = a.SPC, KS = a.KS, H = a.H;
    const int b0 = (blockIdx.x / CL) * NB;
    const int Hc = SPC * 4;
    const int u0 = rank * Hc;
    const int HPs = a.HP + TC_HPAD;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    uint2* wf_s = reinterpret_cast<uint2*>(sp);
    sp += tc_align16((size_t)SPC * KS * 2 * 32 * 8);
    uint16_t* cg_s = reinterpret_cast<uint16_t*>(sp);
    sp += tc_align16((size_t)SPC * KS * 4 * 2);
    uint8_t* nks_s = sp;
    sp += tc_align16((size_t)SPC);
    __nv_bfloat16* hbuf = reinterpret_cast<__nv_bfloat16*>(sp);
    sp += tc_align16((size_t)2 * NB * HPs * 2);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(sp);

    // ---- one-time loads ------------------------------------------------------------------------
    {
        const uint4* s4 = reinterpret_cast<const uint4*>(a.wfrag + (size_t)rank * SPC * KS * 64);
        uint4* d4 = reinterpret_cast<uint4*>(wf_s);
        for (int i = tid; i < SPC * KS * 32; i += nthreads) d4[i] = __ldg(s4 + i);
        for (int i = tid; i < SPC * KS * 4; i += nthreads) {
            const uint16_t c = a.cg[(size_t)rank * SPC * KS * 4 + i];
            cg_s[i] = (c == 0xFFFF) ? (uint16_t)0 : c;  // padded blocks carry zero weights
        }
        for (int i = tid; i < SPC; i += nthreads) nks_s[i] = a.nks[(size_t)rank * SPC + i];
        uint32_t* hz = reinterpret_cast<uint32_t*>(hbuf);
        for (int i = tid; i < NB * HPs; i += nthreads) hz[i] = 0u;  // both buffers (2*NB*HPs bf16)
        if (tid == 0) {
            mbar_init(smem_u32(&mbar[0]), (uint32_t)(CL * (nthreads >> 5)));
            mbar_init(smem_u32(&mbar[1]), (uint32_t)(CL * (nthreads >> 5)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    cluster.sync();

    // ---- per-thread items: strip i of this warp, m16 tile nt, row half hh ----------------------
    const int ui = lane & 3;   // unit inside the strip
    const int br = lane >> 2;  // batch row inside the m16 tile (and +8)
    float c_reg[SPW][NT][2], dr[SPW][NT][2], pf[SPW][NT][2][4];
    bool ok[SPW][NT][2];
#pragma unroll
    for (int i = 0; i < SPW; ++i) {
        const int sl = warp * SPW + i;
        const int u = u0 + sl * 4 + ui;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int b = b0 + nt * 16 + br + 8 * hh;
                ok[i][nt][hh] = (sl < SPC) && (u < H) && (b < a.B);
                c_reg[i][nt][hh] = 0.f;
                dr[i][nt][hh] = a.drop_scalar;
                if (ok[i][nt][hh] && a.drop) dr[i][nt][hh] = __ldg(a.drop + (size_t)b * H + u);
#pragma unroll
                for (int g = 0; g < 4; ++g) pf[i][nt][hh][g] = 0.f;
                if (ok[i][nt][hh] && a.T > 0) {
                    const float* p = a.pre + (size_t)b * a.ld + (size_t)u * a.us;
#pragma unroll
                    for (int g = 0; g < 4; ++g) pf[i][nt][hh][g] = __ldg(p + (size_t)g * a.gs);
                }
            }
    }
    const int chunks = Hc / 8;  // 16-byte chunks of one bf16 slice row
    const uint32_t hbuf_addr = smem_u32(hbuf);
    const uint32_t mbar_addr = smem_u32(mbar);

    for (int k = 0; k < a.T; ++k) {
        const int cur = k & 1, nxt = cur ^ 1;
        if (k > 0) mbar_wait(mbar_addr + cur * 8, (uint32_t)(((k - 1) >> 1) & 1));
        const __nv_bfloat16* hc = hbuf + (size_t)cur * NB * HPs;
        __nv_bfloat16* hn = hbuf + (size_t)nxt * NB * HPs;
        float o_h[SPW][NT][2], o_c[SPW][NT][2], o_g[SPW][NT][2][4];
#pragma unroll
        for (int i = 0; i < SPW; ++i) {
            const int sl = warp * SPW + i;
            if (sl >= SPC) continue;  // warp-uniform
            float acc[2][NT][4];
#pragma unroll
            for (int gp = 0; gp < 2; ++gp)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[gp][nt][q] = 0.f;
            const int nk = (a.dbg & 4) ? 0 : nks_s[sl];
            for (int ks = 0; ks < nk; ++ks) {
                const uint2 bA = wf_s[((size_t)(sl * KS + ks) * 2 + 0) * 32 + lane];
                const uint2 bB = wf_s[((size_t)(sl * KS + ks) * 2 + 1) * 32 + lane];
                const uint16_t* cgp = cg_s + (sl * KS + ks) * 4 + (ui >> 1);
                const int c0 = cgp[0] * 4 + (lane & 1) * 2, c2 = cgp[2] * 4 + (lane & 1) * 2;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const __nv_bfloat16* row = hc + (size_t)(nt * 16 + br) * HPs;
                    const uint32_t a0 = *reinterpret_cast<const uint32_t*>(row + c0);
                    const uint32_t a1 = *reinterpret_cast<const uint32_t*>(row + 8 * HPs + c0);
                    const uint32_t a2 = *reinterpret_cast<const uint32_t*>(row + c2);
                    const uint32_t a3 = *reinterpret_cast<const uint32_t*>(row + 8 * HPs + c2);
                    mma_bf16_16816(acc[0][nt], a0, a1, a2, a3, bA.x, bA.y);
                    mma_bf16_16816(acc[1][nt], a0, a1, a2, a3, bB.x, bB.y);
                }
            }
            // fused gates / cell / hidden for this thread's items of the strip
            const int ul = sl * 4 + ui;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    float hval = 0.f;
                    if (ok[i][nt][hh]) {
                        const float zf = acc[0][nt][hh * 2 + 0] + pf[i][nt][hh][0];
                        const float zi = acc[0][nt][hh * 2 + 1] + pf[i][nt][hh][1];
                        const float zo = acc[1][nt][hh * 2 + 0] + pf[i][nt][hh][2];
                        const float zc = acc[1][nt][hh * 2 + 1] + pf[i][nt][hh][3];
                        const float fg = sigmoidf_(zf), ig = sigmoidf_(zi), og = sigmoidf_(zo);
                        const float gg = act_apply(a.act, zc);
                        const float cn = ig * gg * dr[i][nt][hh] + fg * c_reg[i][nt][hh];
                        c_reg[i][nt][hh] = cn;
                        hval = og * act_apply(a.act, cn);
                        o_c[i][nt][hh] = cn;
                        o_g[i][nt][hh][0] = fg;
                        o_g[i][nt][hh][1] = ig;
                        o_g[i][nt][hh][2] = og;
                        o_g[i][nt][hh][3] = gg;
                    }
                    o_h[i][nt][hh] = hval;
                    hn[(size_t)(nt * 16 + br + 8 * hh) * HPs + u0 + ul] = __float2bfloat16_rn(hval);
                }
        }
        if (k + 1 < a.T) {
            __syncthreads();  // this CTA's slice of h_k is complete in the local buffer
            // push the slice into every peer (coalesced 16-byte DSMEM stores)
            for (int idx = tid; idx < NB * chunks; idx += nthreads) {
                const int b = idx / chunks, q = idx - b * chunks;
                const uint32_t off = (uint32_t)(((size_t)nxt * NB * HPs + (size_t)b * HPs + u0 + q * 8) * 2);
                const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(hbuf) + off);
                for (int d = 1; d < CL; ++d) st_cluster_v4(mapa_u32(hbuf_addr + off, (uint32_t)((rank + d) % CL)), v);
            }
            __syncwarp();
            if (lane == 0) {
                if (a.dbg & 8) {
                    for (int d = 0; d < CL; ++d)
                        mbar_arrive_cluster(mapa_u32(mbar_addr + nxt * 8, (uint32_t)((rank + d) % CL)));
                } else {
                    fence_release_cluster();
                    for (int d = 0; d < CL; ++d)
                        mbar_arrive_cluster_relaxed(mapa_u32(mbar_addr + nxt * 8, (uint32_t)((rank + d) % CL)));
                }
            }
        }
        // ---- global stores of step k (after the signal: the release never waits for them) and the
        //      prefetch of step k+1's pre-activations ---------------------------------------------
#pragma unroll
        for (int i = 0; i < SPW; ++i) {
            const int sl = warp * SPW + i;
            const int u = u0 + sl * 4 + ui;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    if (!ok[i][nt][hh]) continue;
                    const int b = b0 + nt * 16 + br + 8 * hh;
                    const size_t row = (size_t)k * a.B + b;
                    if (!(a.dbg & 1)) {
                        a.h_out[row * H + u] = o_h[i][nt][hh];
                        if (a.hb_out) a.hb_out[row * a.hb_ld + u] = __float2bfloat16_rn(o_h[i][nt][hh]);
                        if (a.c_out) a.c_out[row * H + u] = o_c[i][nt][hh];
                        if (a.gates_out) {
                            float* gp = a.gates_out + row * a.ld + (size_t)u * a.us;
#pragma unroll
                            for (int g = 0; g < 4; ++g) gp[(size_t)g * a.gs] = o_g[i][nt][hh][g];
                        }
                    }
                    if (k + 1 < a.T && !(a.dbg & 2)) {
                        const float* p = a.pre + (row + a.B) * a.ld + (size_t)u * a.us;
#pragma unroll
                        for (int g = 0; g < 4; ++g) pf[i][nt][hh][g] = __ldg(p + (size_t)g * a.gs);
                    }
                }
        }
    }
    cluster.sync();  // nobody exits while peers may still write into its shared memory
}

// ---- backward ----------------------------------------------------------------------------------
struct TcBwdArgs {
    const uint2* wfrag;      // [n_blocks][16]  (lanes 0..15 of the B fragment; lanes 16..31 are zero)
    const int* blk_off;      // [CL+1]
    const int* cs_ptr;       // [CL][NS+1]
    const uint16_t* blk_rs;  // [n_blocks]
    const float* dh_out;     // [T,B,H]
    const float* gates;      // element (row, g, u) at gates[row*ld + g*gs + u*us]
    int64_t ld;
    int gs, us;
    const float* c;          // [T,B,H]
    const float* drop;
    float drop_scalar;
    float* dpre;             // fp32, same addressing as gates
    __nv_bfloat16* dpreb;    // optional bf16 copy, same addressing with row stride ldb
    int64_t ldb;
    int T, B, H, CL, SPC, NS, HP, max_blocks, act;
};

template <int NT, int IT>
__global__ void __launch_bounds__(512, 1) lstm_tc_bwd_kernel(const TcBwdArgs a) {
    constexpr int NB = NT * 16;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int CL = a.CL, SPC = a.SPC, NS = a.NS, H = a.H;
    const int b0 = (blockIdx.x / CL) * NB;
    const int Hc = SPC * 4;
    const int u0 = rank * Hc;
    const int HPp = a.HP + TC_PPAD;
    const int DPS = SPC * 16 + TC_DPAD;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x, nwarps = nthreads >> 5;
    const int blk0 = a.blk_off[rank];
    const int nblk = a.blk_off[rank + 1] - blk0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw;
    uint2* wf_s = reinterpret_cast<uint2*>(sp);
    sp += tc_align16((size_t)a.max_blocks * 16 * 8);
    uint16_t* rs_s = reinterpret_cast<uint16_t*>(sp);
    sp += tc_align16((size_t)a.max_blocks * 2);
    int* csp_s = reinterpret_cast<int*>(sp);
    sp += tc_align16((size_t)(NS + 1) * 4);
    __nv_bfloat16* dpre_s = reinterpret_cast<__nv_bfloat16*>(sp);
    sp += tc_align16((size_t)NB * DPS * 2);
    float* partial = reinterpret_cast<float*>(sp);
    sp += tc_align16((size_t)NB * HPp * 4);
    float* slots = reinterpret_cast<float*>(sp);  // [2][CL][NB][Hc]
    sp += tc_align16((size_t)2 * CL * NB * Hc * 4);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(sp);

    {
        const uint4* s4 = reinterpret_cast<const uint4*>(a.wfrag + (size_t)blk0 * 16);
        uint4* d4 = reinterpret_cast<uint4*>(wf_s);
        for (int i = tid; i < nblk * 8; i += nthreads) d4[i] = __ldg(s4 + i);
        for (int i = tid; i < nblk; i += nthreads) rs_s[i] = a.blk_rs[blk0 + i];
        for (int i = tid; i <= NS; i += nthreads) csp_s[i] = a.cs_ptr[(size_t)rank * (NS + 1) + i];
        for (int i = tid; i < 2 * CL * NB * Hc; i += nthreads) slots[i] = 0.f;
        uint32_t* dz = reinterpret_cast<uint32_t*>(dpre_s);
        for (int i = tid; i < NB * DPS / 2; i += nthreads) dz[i] = 0u;
        if (tid == 0) {
            mbar_init(smem_u32(&mbar[0]), (uint32_t)(CL * nwarps));
            mbar_init(smem_u32(&mbar[1]), (uint32_t)(CL * nwarps));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    cluster.sync();

    // ---- pointwise items (unit fastest => coalesced global traffic) ----------------------------
    int it_b[IT], it_ul[IT];
    bool it_in[IT], it_ok[IT];
    float dc_carry[IT], dr[IT];
    float p_g[IT][4], p_c[IT], p_cm1[IT], p_dh[IT];
#pragma unroll
    for (int j = 0; j < IT; ++j) {
        const int idx = tid + j * nthreads;
        it_in[j] = idx < Hc * NB;
        it_b[j] = it_in[j] ? idx / Hc : 0;
        it_ul[j] = it_in[j] ? idx % Hc : 0;
        it_ok[j] = it_in[j] && (u0 + it_ul[j] < H) && (b0 + it_b[j] < a.B);
        dc_carry[j] = 0.f;
        dr[j] = a.drop_scalar;
        p_c[j] = p_cm1[j] = p_dh[j] = 0.f;
#pragma unroll
        for (int g = 0; g < 4; ++g) p_g[j][g] = 0.f;
        if (it_ok[j]) {
            const int u = u0 + it_ul[j];
            if (a.drop) dr[j] = __ldg(a.drop + (size_t)(b0 + it_b[j]) * H + u);
            if (a.T > 0) {
                const size_t row = (size_t)(a.T - 1) * a.B + b0 + it_b[j];
                const float* gp = a.gates + row * a.ld + (size_t)u * a.us;
#pragma unroll
                for (int g = 0; g < 4; ++g) p_g[j][g] = __ldg(gp + (size_t)g * a.gs);
                p_c[j] = __ldg(a.c + row * H + u);
                p_cm1[j] = (a.T > 1) ? __ldg(a.c + (row - a.B) * H + u) : 0.f;
                p_dh[j] = __ldg(a.dh_out + row * H + u);
            }
        }
    }
    const uint32_t slots_addr = smem_u32(slots);
    const uint32_t mbar_addr = smem_u32(mbar);
    const int br = lane >> 2;
    const int chunks = Hc / 4;  // 16-byte chunks of one fp32 slice row

    for (int k = a.T - 1; k >= 0; --k) {
        const int wr = k & 1, rd = wr ^ 1;
        if (k < a.T - 1) mbar_wait(mbar_addr + rd * 8, (uint32_t)(((a.T - 2 - k) >> 1) & 1));
        // ---- fused gate derivatives ------------------------------------------------------------
#pragma unroll
        for (int j = 0; j < IT; ++j) {
            if (!it_in[j]) continue;
            const int b = it_b[j], ul = it_ul[j];
            float dp[4] = {0.f, 0.f, 0.f, 0.f};
            if (it_ok[j]) {
                float dh = p_dh[j];
                if (k < a.T - 1) {
                    const float* sl = slots + ((size_t)rd * CL * NB + b) * Hc + ul;
                    for (int src = 0; src < CL; ++src) dh += sl[(size_t)src * NB * Hc];
                }
                const float fg = p_g[j][0], ig = p_g[j][1], og = p_g[j][2], gg = p_g[j][3];
                const float tc = act_apply(a.act, p_c[j]);
                const float dog = dh * tc;
                const float dct = dh * og * act_grad_from_out(a.act, tc) + dc_carry[j];
                const float dig = dct * gg * dr[j];
                const float dgg = dct * ig * dr[j];
                const float dfg = dct * p_cm1[j];
                dc_carry[j] = dct * fg;
                dp[0] = dfg * fg * (1.f - fg);
                dp[1] = dig * ig * (1.f - ig);
                dp[2] = dog * og * (1.f - og);
                dp[3] = dgg * act_grad_from_out(a.act, gg);
                const size_t row = (size_t)k * a.B + b0 + b;
                const int u = u0 + ul;
                float* o = a.dpre + row * a.ld + (size_t)u * a.us;
#pragma unroll
                for (int g = 0; g < 4; ++g) o[(size_t)g * a.gs] = dp[g];
                if (a.dpreb) {
                    __nv_bfloat16* ob = a.dpreb + row * a.ldb + (size_t)u * a.us;
#pragma unroll
                    for (int g = 0; g < 4; ++g) ob[(size_t)g * a.gs] = __float2bfloat16_rn(dp[g]);
                }
                if (k > 0) {
                    const size_t rowp = row - a.B;
                    const float* gp = a.gates + rowp * a.ld + (size_t)u * a.us;
#pragma unroll
                    for (int g = 0; g < 4; ++g) p_g[j][g] = __ldg(gp + (size_t)g * a.gs);
                    p_c[j] = p_cm1[j];
                    p_cm1[j] = (k > 1) ? __ldg(a.c + (rowp - a.B) * H + u) : 0.f;
                    p_dh[j] = __ldg(a.dh_out + rowp * H + u);
                }
            }
            __nv_bfloat16* d = dpre_s + (size_t)b * DPS + (ul >> 2) * 16 + (ul & 3);
#pragma unroll
            for (int g = 0; g < 4; ++g) d[g * 4] = __float2bfloat16_rn(dp[g]);
        }
        if (k == 0) break;  // dh_{-1} is not needed
        __syncthreads();
        // ---- partial dh for every column strip from this CTA's rows ---------------------------
        for (int cs = warp; cs < NS; cs += nwarps) {
            float acc[NT][4];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[nt][q] = 0.f;
            const int e1 = csp_s[cs + 1];
            const int koff = ((lane & 3) >> 1) * 4 + (lane & 1) * 2;  // (gate, row) pair inside the block
            for (int e = csp_s[cs]; e < e1; ++e) {
                const int rs = rs_s[e];
                uint2 bw = make_uint2(0u, 0u);
                if (lane < 16) bw = wf_s[(size_t)e * 16 + lane];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const __nv_bfloat16* base = dpre_s + (size_t)(nt * 16 + br) * DPS + rs * 16 + koff;
                    const uint32_t a0 = *reinterpret_cast<const uint32_t*>(base);
                    const uint32_t a1 = *reinterpret_cast<const uint32_t*>(base + 8 * DPS);
                    const uint32_t a2 = *reinterpret_cast<const uint32_t*>(base + 8);
                    const uint32_t a3 = *reinterpret_cast<const uint32_t*>(base + 8 * DPS + 8);
                    mma_bf16_16816(acc[nt], a0, a1, a2, a3, bw.x, bw.y);
                }
            }
            if ((lane & 3) < 2) {
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    float* p = partial + (size_t)(nt * 16 + br) * HPp + cs * 4 + (lane & 3) * 2;
                    *reinterpret_cast<float2*>(p) = make_float2(acc[nt][0], acc[nt][1]);
                    *reinterpret_cast<float2*>(p + 8 * HPp) = make_float2(acc[nt][2], acc[nt][3]);
                }
            }
        }
        __syncthreads();
        // ---- reduce-scatter: column slice d of the partial sums goes to CTA d -------------------
        for (int idx = tid; idx < CL * NB * chunks; idx += nthreads) {
            const int d = idx / (NB * chunks), rem = idx - d * (NB * chunks);
            const int b = rem / chunks, q = rem - b * chunks;
            const int col = d * Hc + q * 4;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (col < NS * 4) v = *reinterpret_cast<const uint4*>(partial + (size_t)b * HPp + col);
            const uint32_t off = (uint32_t)((((size_t)(wr * CL + rank) * NB + b) * Hc + q * 4) * 4);
            st_cluster_v4(mapa_u32(slots_addr + off, (uint32_t)d), v);
        }
        __syncwarp();
        if (lane == 0) {
            fence_release_cluster();
            for (int d = 0; d < CL; ++d)
                mbar_arrive_cluster_relaxed(mapa_u32(mbar_addr + wr * 8, (uint32_t)((rank + d) % CL)));
        }
    }
    cluster.sync();
}

// ---- pack kernels: masked dense fp32 U_g -> bf16 mma B fragments ------------------------------
__global__ void tc_pack_fwd_kernel(const uint16_t* __restrict__ cg, PtrPack<const float> U, int ldu, int H, int KS,
                                   int64_t total, uint32_t* __restrict__ out) {
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int j = idx & 1, lane = (idx >> 1) & 31, gp = (idx >> 6) & 1;
        const int64_t sk = idx >> 7;  // s*KS + ks
        const int64_t s = sk / KS;
        const int n = lane >> 2, unit_i = n >> 1, gate = gp * 2 + (n & 1);
        const int k0 = (lane & 3) * 2 + 8 * j, q = k0 >> 2, ci = k0 & 3;
        const uint16_t c = cg[sk * 4 + q];
        float lo = 0.f, hi = 0.f;
        if (c != 0xFFFF) {
            const int64_t row = s * 4 + unit_i;
            const int col = c * 4 + ci;
            if (row < H && col < H) lo = __ldg(U.p[gate] + row * ldu + col);
            if (row < H && col + 1 < H) hi = __ldg(U.p[gate] + row * ldu + col + 1);
        }
        out[idx] = pack_bf16x2(lo, hi);
    }
}

__global__ void tc_pack_bwd_kernel(const int* __restrict__ blk_off, const uint16_t* __restrict__ blk_rs,
                                   const uint16_t* __restrict__ blk_cs, PtrPack<const float> U, int ldu, int H,
                                   int CL, int SPC, int64_t total, uint32_t* __restrict__ out) {
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int j = idx & 1, lane = (idx >> 1) & 15;
        const int e = (int)(idx >> 5);
        int r = 0;
        while (r + 1 < CL && blk_off[r + 1] <= e) ++r;
        const int n = lane >> 2;
        const int k0 = (lane & 3) * 2 + 8 * j, g = k0 >> 2, i = k0 & 3;
        const int64_t row = (int64_t)(r * SPC + blk_rs[e]) * 4 + i;
        const int col = blk_cs[e] * 4 + n;
        float lo = 0.f, hi = 0.f;
        if (col < H) {
            if (row < H) lo = __ldg(U.p[g] + row * ldu + col);
            if (row + 1 < H) hi = __ldg(U.p[g] + (row + 1) * ldu + col);
        }
        out[idx] = pack_bf16x2(lo, hi);
    }
}

template <typename Args>
static int tc_launch(void (*kernel)(const Args), const Args& args, int n_clusters, int CL, int threads, size_t smem,
                     cudaStream_t st) {
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

static int tc_smem_limit() {
    int dev = 0, smem_max = 232448;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    return smem_max;
}

// Resolve (cluster size, sequences per cluster): 0 = choose.  Smallest cluster whose layouts fit.
static int tc_resolve(const hcgs_plan* plan, int cluster_size, int nb, TcPlan** out, int* nb_out) {
    HCGS_REQUIRE(nb == 0 || nb == 16 || nb == 32, "tensor-core recurrence: nb must be 0, 16 or 32");
    const int NB = nb ? nb : 16;
    const int limit = tc_smem_limit();
    const int cands[3] = {4, 8, 16};
    for (int ci = 0; ci < 3; ++ci) {
        if (cluster_size && cands[ci] != cluster_size) continue;
        TcPlan* t = nullptr;
        if (tc_plan_get(plan, cands[ci], &t)) return 1;
        const bool fits = tc_fwd_smem(*t, NB) <= (size_t)limit && tc_bwd_smem(*t, NB) <= (size_t)limit &&
                          t->SPW <= 4 && (NB == 16 || t->SPW <= 2) && (t->Hc * NB + 511) / 512 <= 10;
        if (fits) {
            *out = t;
            *nb_out = NB;
            return 0;
        }
    }
    set_error("tensor-core recurrence: H=%d does not fit (cluster %d, nb %d)", plan->n_out, cluster_size, NB);
    return 2;
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_lstm_tc_info(const hcgs_plan* plan, int cluster_size, int nb, int64_t* info) {
    HCGS_REQUIRE(plan && info, "hcgs_lstm_tc_info: NULL argument");
    TcPlan* t = nullptr;
    int NB = 0;
    if (int rc = tc_resolve(plan, cluster_size, nb, &t, &NB)) return rc;
    for (int i = 0; i < 16; ++i) info[i] = 0;
    info[0] = t->CL;
    info[1] = NB;
    info[2] = t->fwd_pack_u32;
    info[3] = t->bwd_pack_u32;
    info[4] = (int64_t)tc_fwd_smem(*t, NB);
    info[5] = (int64_t)tc_bwd_smem(*t, NB);
    info[6] = t->WARPS;
    info[7] = t->SPC;
    info[8] = t->KS;
    info[9] = t->n_blocks;
    info[10] = t->SPW;
    info[11] = t->HP;
    return 0;
}

extern "C" int hcgs_lstm_tc_pack(const hcgs_plan* plan, int cluster_size, const float* const* u, int ldu,
                                 uint32_t* fwd_pack, uint32_t* bwd_pack, hcgs_stream stream) {
    HCGS_REQUIRE(plan && u, "hcgs_lstm_tc_pack: NULL argument");
    TcPlan* t = nullptr;
    if (int rc = tc_plan_get(plan, cluster_size, &t)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    PtrPack<const float> U = make_pack<const float>(u, 4);
    if (fwd_pack) {
        const int64_t total = t->fwd_pack_u32;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        tc_pack_fwd_kernel<<<blocks, 256, 0, st>>>(t->d_cg, U, ldu, t->H, t->KS, total, fwd_pack);
        HCGS_LAUNCH_CHECK();
    }
    if (bwd_pack && t->n_blocks > 0) {
        const int64_t total = (int64_t)t->n_blocks * 32;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        tc_pack_bwd_kernel<<<blocks, 256, 0, st>>>(t->d_blk_off, t->d_blk_rs, t->d_blk_cs, U, ldu, t->H, t->CL,
                                                  t->SPC, total, bwd_pack);
        HCGS_LAUNCH_CHECK();
    }
    return 0;
}

namespace hcgs {

typedef void (*TcFwdKernel)(const TcFwdArgs);
typedef void (*TcBwdKernel)(const TcBwdArgs);

static TcFwdKernel tc_pick_fwd(int NT, int SPW) {
    if (NT == 1) {
        switch (SPW) {
            case 1: return lstm_tc_fwd_kernel<1, 1>;
            case 2: return lstm_tc_fwd_kernel<1, 2>;
            case 3: return lstm_tc_fwd_kernel<1, 3>;
            case 4: return lstm_tc_fwd_kernel<1, 4>;
        }
    } else if (NT == 2) {
        switch (SPW) {
            case 1: return lstm_tc_fwd_kernel<2, 1>;
            case 2: return lstm_tc_fwd_kernel<2, 2>;
        }
    }
    return nullptr;
}

static TcBwdKernel tc_pick_bwd(int NT, int items) {
    if (NT == 1) {
        if (items <= 2) return lstm_tc_bwd_kernel<1, 2>;
        if (items <= 4) return lstm_tc_bwd_kernel<1, 4>;
        if (items <= 6) return lstm_tc_bwd_kernel<1, 6>;
        if (items <= 8) return lstm_tc_bwd_kernel<1, 8>;
        if (items <= 10) return lstm_tc_bwd_kernel<1, 10>;
    } else if (NT == 2) {
        if (items <= 2) return lstm_tc_bwd_kernel<2, 2>;
        if (items <= 4) return lstm_tc_bwd_kernel<2, 4>;
        if (items <= 6) return lstm_tc_bwd_kernel<2, 6>;
        if (items <= 8) return lstm_tc_bwd_kernel<2, 8>;
        if (items <= 10) return lstm_tc_bwd_kernel<2, 10>;
    }
    return nullptr;
}

}  // namespace hcgs

extern "C" int hcgs_lstm_tc_fwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                                const float* pre, int64_t ld, int gate_stride, int unit_stride, const float* drop,
                                float drop_scalar, int act, int t_len, int batch, float* h_out, void* hb_out,
                                int hb_ld, float* c_out, float* gates_out, hcgs_stream stream) {
    HCGS_REQUIRE(plan && fwd_pack && pre && h_out, "hcgs_lstm_tc_fwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    TcPlan* t = nullptr;
    int NB = 0;
    if (int rc = tc_resolve(plan, cluster_size, nb, &t, &NB)) return rc;
    TcFwdArgs a;
    a.wfrag = reinterpret_cast<const uint2*>(fwd_pack);
    a.cg = t->d_cg;
    a.nks = t->d_nks;
    a.pre = pre;
    a.ld = ld;
    a.gs = gate_stride;
    a.us = unit_stride;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.h_out = h_out;
    a.hb_out = reinterpret_cast<__nv_bfloat16*>(hb_out);
    a.hb_ld = hb_ld;
    a.c_out = c_out;
    a.gates_out = gates_out;
    a.T = t_len;
    a.B = batch;
    a.H = t->H;
    a.CL = t->CL;
    a.SPC = t->SPC;
    a.SPW = t->SPW;
    a.KS = t->KS;
    a.HP = t->HP;
    a.act = act;
    {
        const char* e = getenv("HCGS_TC_DEBUG");
        a.dbg = e ? atoi(e) : 0;
    }
    TcFwdKernel k = tc_pick_fwd(NB / 16, t->SPW);
    HCGS_REQUIRE(k != nullptr, "hcgs_lstm_tc_fwd: no kernel for nb=%d strips/warp=%d", NB, t->SPW);
    const int n_clusters = (batch + NB - 1) / NB;
    return tc_launch(k, a, n_clusters, t->CL, t->WARPS * 32, tc_fwd_smem(*t, NB), (cudaStream_t)stream);
}

extern "C" int hcgs_lstm_tc_bwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* bwd_pack,
                                const float* dh_out, const float* gates, int64_t ld, int gate_stride,
                                int unit_stride, const float* c, const float* drop, float drop_scalar, int act,
                                int t_len, int batch, float* dpre, void* dpre_bf16, int64_t ld_bf16,
                                hcgs_stream stream) {
    HCGS_REQUIRE(plan && bwd_pack && dh_out && gates && c && dpre, "hcgs_lstm_tc_bwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    TcPlan* t = nullptr;
    int NB = 0;
    if (int rc = tc_resolve(plan, cluster_size, nb, &t, &NB)) return rc;
    TcBwdArgs a;
    a.wfrag = reinterpret_cast<const uint2*>(bwd_pack);
    a.blk_off = t->d_blk_off;
    a.cs_ptr = t->d_cs_ptr;
    a.blk_rs = t->d_blk_rs;
    a.dh_out = dh_out;
    a.gates = gates;
    a.ld = ld;
    a.gs = gate_stride;
    a.us = unit_stride;
    a.c = c;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.dpre = dpre;
    a.dpreb = reinterpret_cast<__nv_bfloat16*>(dpre_bf16);
    a.ldb = ld_bf16;
    a.T = t_len;
    a.B = batch;
    a.H = t->H;
    a.CL = t->CL;
    a.SPC = t->SPC;
    a.NS = t->NS;
    a.HP = t->HP;
    a.max_blocks = std::max(t->max_cta_blocks, 1);
    a.act = act;
    const int items = (t->Hc * NB + 511) / 512;
    TcBwdKernel k = tc_pick_bwd(NB / 16, items);
    HCGS_REQUIRE(k != nullptr, "hcgs_lstm_tc_bwd: no kernel for nb=%d items=%d", NB, items);
    const int n_clusters = (batch + NB - 1) / NB;
    return tc_launch(k, a, n_clusters, t->CL, 512, tc_bwd_smem(*t, NB), (cudaStream_t)stream);
}
