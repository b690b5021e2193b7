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
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace hcgs {

constexpr int TC_MAX_WARPS = 16;
constexpr int TC_MAX_WARPS2 = 20;  // forward kernels with <= 2 strips per warp (register budget 65536 / 640)
constexpr int TCB_CW = 18;         // warps of the backward kernel

struct TcPlan {
    int CL = 0, H = 0, NS = 0;  // cluster size, hidden size, 4-row strips in total
    int SPC = 0, Hc = 0, HP = 0;  // strips per CTA (even), units per CTA, padded hidden = CL*Hc
    int SPW = 0, WARPS = 0, KS = 0;  // forward: strips per warp, warps per CTA, max k-steps per strip
    int n_blocks = 0, max_cta_blocks = 0;  // kept 4x4 blocks (total / largest CTA)
    // forward tables
    uint16_t* d_cg = nullptr;  // [CL][SPC][KS*4] column group of block j of the strip (0xFFFF = padding)
    uint8_t* d_nks = nullptr;  // [CL][SPC] k-steps actually used by the strip
    // backward tables: the kept blocks of CTA r in task order (warp, task, row-strip group, column strip)
    int* d_blk_off = nullptr;      // [CL+1]
    uint16_t* d_blk_rs = nullptr;  // [n_blocks] local row strip of block e (pack kernel)
    uint16_t* d_blk_cs = nullptr;  // [n_blocks] column strip of block e (pack kernel)
    int* d_grp_off = nullptr;      // [CL+1] row-strip groups per CTA
    uint32_t* d_grp = nullptr;     // [n_groups] (local row strip * 32) | (kept column strips of the task's 8) << 16
    int* d_task_off = nullptr;     // [CL+1]
    int* d_task = nullptr;         // [n_tasks][8] {g0, g1, blk0, jlo, nst, dest, sl0, 0} (CTA-relative indices)
    int* d_wtp = nullptr;          // [CL][TCB_CW+1] tasks of each compute warp
    int* d_rx = nullptr;           // [CL][3] {bytes of partial dh the CTA receives from its peers per time step,
                                   //          peers that send it nothing, bit mask of the peers it sends nothing to}
    int max_cta_groups = 0, max_cta_tasks = 0;
    int64_t fwd_pack_u32 = 0, bwd_pack_u32 = 0;
};

void tc_plan_free(TcPlan* t) {
    if (!t) return;
    cudaFree(t->d_cg);
    cudaFree(t->d_nks);
    cudaFree(t->d_blk_off);
    cudaFree(t->d_grp_off);
    cudaFree(t->d_grp);
    cudaFree(t->d_task_off);
    cudaFree(t->d_task);
    cudaFree(t->d_wtp);
    cudaFree(t->d_rx);
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
    // forward compute warps (two more warps drive the copy engine): up to 18 warps of <= 2 strips (short
    // per-warp dependency chains, 640 threads x 102 registers), else up to 14 warps of 3..4 strips (512 x 128)
    t->SPW = (t->SPC + (TC_MAX_WARPS2 - 2) - 1) / (TC_MAX_WARPS2 - 2);
    if (t->SPW > 2) t->SPW = (t->SPC + (TC_MAX_WARPS - 2) - 1) / (TC_MAX_WARPS - 2);
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
    // backward: the 8-strip (32-column) column blocks of U are the tasks of the warps.  A task keeps the partial dh of
    // its <= 8 column strips in accumulators and walks the CTA's row strips: one A tile (the bf16 dpre of a row strip,
    // 16 sequences x 4 rows x 4 gates) serves every kept block of that row strip inside the column block ("group").
    // Tasks are cut at destination-CTA boundaries, heavy ones at the middle of the column block, and spread over the
    // warps longest-first.  Column blocks without a kept block in the CTA's rows have no task: their partial sums are
    // zero and the owner's receive slots keep the zeros they were initialised with.
    struct HTask {
        int base, jlo, nst, ngroups, nblocks, cost;
    };
    const int CS = CL * t->SPC;
    // issue slots of a task (the group loop is issue bound: profiles/r02_ncu_lstm_tc_bwd3.md): ~20 per group plus 3 per
    // column strip the loop variant tests (8, or 4 for tasks inside one half of the column block), ~30 per pushed strip
    auto measure = [&](int r, HTask& k) {
        k.ngroups = k.nblocks = 0;
        for (int sl = 0; sl < t->SPC; ++sl) {
            const int s = r * t->SPC + sl;
            if (s >= NS) break;
            int n = 0;
            for (int j = k.jlo; j < k.jlo + k.nst; ++j)
                if (k.base + j < NS && occ[(size_t)s * NS + k.base + j]) ++n;
            if (n) ++k.ngroups, k.nblocks += n;
        }
        const bool half = k.jlo + k.nst <= 4 || k.jlo >= 4;
        k.cost = k.ngroups ? 150 + k.ngroups * (20 + 3 * (half ? 4 : 8)) + 30 * k.nst : 0;
    };
    std::vector<int> blk_off(CL + 1, 0), grp_off(CL + 1, 0), task_off(CL + 1, 0), wtp((size_t)CL * (TCB_CW + 1), 0),
        rx((size_t)CL * 3, 0), pair((size_t)CL * CL, 0), tasks;
    std::vector<uint16_t> blk_rs, blk_cs;
    std::vector<uint32_t> grp;
    // warp 0 leads (mbarrier waits, TMA, handshake: ~3 k cycles of a step) and takes no tasks unless there are more
    // tasks than the other warps can hold
    const int TW = TCB_CW - 1;
    for (int r = 0; r < CL; ++r) {
        std::vector<HTask> work;
        for (int base = 0; base < NS; base += 8) {
            int j = 0;
            while (j < 8 && base + j < NS) {
                const int d = (base + j) / t->SPC;
                int je = j;
                while (je < 8 && base + je < NS && (base + je) / t->SPC == d) ++je;
                HTask k{base, j, je - j, 0, 0, 0};
                measure(r, k);
                if (k.cost) work.push_back(k);
                j = je;
            }
        }
        // a task that spans both halves of its column block and is heavier than a warp's fair share is cut at strip 4:
        // both halves still walk nearly every row strip, but their loop tests 4 strips per group instead of 8
        int64_t total = 0;
        for (const HTask& k : work) total += k.cost;
        static const int split_pct = getenv("HCGS_TC_SPLIT") ? atoi(getenv("HCGS_TC_SPLIT")) : 150;  // tuning only
        static const int lead_cost = getenv("HCGS_TC_LEAD") ? atoi(getenv("HCGS_TC_LEAD")) : 300;     // tuning only
        const int64_t fair = std::max<int64_t>(300, (total * split_pct / 100 + TW - 1) / TW);
        for (size_t i = 0; i < work.size(); ++i) {
            const HTask w = work[i];
            if (w.cost <= fair || w.jlo >= 4 || w.jlo + w.nst <= 4) continue;
            HTask lo = w, hi = w;
            lo.nst = 4 - w.jlo;
            hi.jlo = 4;
            hi.nst = w.jlo + w.nst - 4;
            measure(r, lo);
            measure(r, hi);
            if (!lo.cost || !hi.cost) continue;
            work[i] = lo;
            work.insert(work.begin() + i + 1, hi);
            ++i;
        }
        std::stable_sort(work.begin(), work.end(), [](const HTask& x, const HTask& y) { return x.cost > y.cost; });
        std::vector<std::vector<int>> per_warp(TCB_CW);
        std::vector<int64_t> load(TCB_CW, 0);
        load[0] = lead_cost;
        for (size_t i = 0; i < work.size(); ++i) {
            int w = 0;
            for (int x = 1; x < TCB_CW; ++x)
                if (load[x] < load[w]) w = x;
            per_warp[w].push_back((int)i);
            load[w] += work[i].cost;
        }
        blk_off[r] = (int)blk_rs.size();
        grp_off[r] = (int)grp.size();
        task_off[r] = (int)tasks.size() / 8;
        int nt = 0;
        for (int w = 0; w < TCB_CW; ++w) {
            wtp[(size_t)r * (TCB_CW + 1) + w] = nt;
            for (int i : per_warp[w]) {
                const HTask& k = work[i];
                const int g0 = (int)grp.size() - grp_off[r], b0 = (int)blk_rs.size() - blk_off[r];
                for (int sl = 0; sl < t->SPC; ++sl) {
                    const int s = r * t->SPC + sl;
                    if (s >= NS) break;
                    uint32_t m = 0;
                    for (int j = k.jlo; j < k.jlo + k.nst; ++j)
                        if (k.base + j < NS && occ[(size_t)s * NS + k.base + j]) {
                            m |= 1u << j;
                            blk_rs.push_back((uint16_t)sl);
                            blk_cs.push_back((uint16_t)(k.base + j));
                        }
                    if (m) grp.push_back((uint32_t)(sl * 32) | (m << 16));
                }
                const int first = k.base + k.jlo, dest = first / t->SPC;
                const int rec[8] = {g0, (int)grp.size() - grp_off[r], b0, k.jlo, k.nst, dest, first - dest * t->SPC, 0};
                tasks.insert(tasks.end(), rec, rec + 8);
                if (dest != r) rx[(size_t)dest * 3] += k.nst * 256, pair[(size_t)r * CL + dest] += k.nst * 256;
                ++nt;
            }
        }
        wtp[(size_t)r * (TCB_CW + 1) + TCB_CW] = nt;
        t->max_cta_blocks = std::max(t->max_cta_blocks, (int)blk_rs.size() - blk_off[r]);
        t->max_cta_groups = std::max(t->max_cta_groups, (int)grp.size() - grp_off[r]);
        t->max_cta_tasks = std::max(t->max_cta_tasks, nt);
        (void)CS;
    }
    // a pair of CTAs without traffic still exchanges one empty arrival per step: the reduce-scatter is what keeps the
    // cluster in lock step (nobody runs more than one step ahead), and the handshake barriers count on that
    for (int src = 0; src < CL; ++src)
        for (int dst = 0; dst < CL; ++dst)
            if (src != dst && pair[(size_t)src * CL + dst] == 0) {
                ++rx[(size_t)dst * 3 + 1];
                rx[(size_t)src * 3 + 2] |= 1 << dst;
            }
    blk_off[CL] = (int)blk_rs.size();
    grp_off[CL] = (int)grp.size();
    task_off[CL] = (int)tasks.size() / 8;
    t->n_blocks = (int)blk_rs.size();
    t->fwd_pack_u32 = (int64_t)CL * t->SPC * t->KS * 2 * 32 * 2;
    t->bwd_pack_u32 = (int64_t)std::max(t->n_blocks, 1) * 16 * 2;
    int rc = 0;
    rc |= tc_upload(cgt, &t->d_cg);
    rc |= tc_upload(nks, &t->d_nks);
    rc |= tc_upload(blk_off, &t->d_blk_off);
    rc |= tc_upload(grp_off, &t->d_grp_off);
    rc |= tc_upload(grp, &t->d_grp);
    rc |= tc_upload(task_off, &t->d_task_off);
    rc |= tc_upload(tasks, &t->d_task);
    rc |= tc_upload(wtp, &t->d_wtp);
    rc |= tc_upload(rx, &t->d_rx);
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

static size_t tc_fwd2_smem(const TcPlan& t, int NB) {
    const int HcS = (t.Hc % 16 == 8) ? t.Hc : t.Hc + 8;
    size_t b = 128;                                                   // alignment slack (TMA boxes: 128 B)
    b += (size_t)2 * NB * t.Hc * 16;                                  // pre-activation / gate stages (dense boxes)
    b += 2 * (((size_t)NB * t.Hc * 4 + 127) & ~(size_t)127);          // c and h output stages
    b += 2 * (((size_t)NB * t.Hc * 2 + 127) & ~(size_t)127);          // dense bf16 copy of the slice + its low part
    b += ((size_t)2 * t.CL * NB * HcS * 2 + 127) & ~(size_t)127;      // slice-major h double buffer (bf16)
    b += tc_align16((size_t)t.SPC * t.KS * 2 * 32 * 8);               // B fragments
    b += 3 * tc_align16((size_t)t.Hc * 16);                           // fused batch-norm coefficients
    b += 64;                                                          // mbarriers
    return b;
}
static size_t tc_fwd_smem_v(const TcPlan& t, int NB) { return tc_fwd2_smem(t, NB); }

static size_t tc_bwd3_smem(const TcPlan& t, int NB) {
    size_t b = 128;                                                   // alignment slack (TMA boxes: 128 B)
    b += (size_t)2 * NB * t.Hc * 16;                                  // gate stages (dense boxes)
    b += ((size_t)2 * NB * t.Hc * 4 + 127) & ~(size_t)127;            // c ring
    b += ((size_t)NB * t.Hc * 4 + 127) & ~(size_t)127;                // dh
    b += ((size_t)t.CL * t.SPC * 256 + 127) & ~(size_t)127;           // partial-dh slots [source][strip][16 x 4]
    b += tc_align16((size_t)NB * (t.Hc * 4 + 8) * 2);                 // bf16 dpre tile (MMA A operand)
    b += tc_align16((size_t)std::max(t.max_cta_blocks, 1) * 16 * 8);  // B fragments (16 lanes x 8 B)
    b += tc_align16((size_t)(std::max(t.max_cta_groups, 1) + 2) * 4); // row-strip groups (+2 spare)
    b += tc_align16((size_t)std::max(t.max_cta_tasks, 1) * 32);       // tasks
    b += tc_align16((size_t)(TCB_CW + 1) * 4);                        // tasks of each warp
    b += 64;
    return b;
}
static size_t tc_bwd_smem_v(const TcPlan& t, int NB) { return tc_bwd3_smem(t, NB); }

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

// Waits of the asynchronous kernels: the default (.acquire.cta) form with a suspend-time hint parks the thread
// in hardware instead of polling -- in the profile of the polling form 37 % of all executed instructions of the
// backward kernel were this loop, stealing issue slots from the working warps.  Transaction-count completion
// (TMA, st.async) makes the written data visible to the waiter without a cluster-scope acquire.
__device__ __forceinline__ uint32_t mbar_try_wait_park(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity), "r"(20000u)
        : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait_park(uint32_t addr, uint32_t parity) {
    for (int spin = 0; spin < (1 << 20); ++spin)
        if (mbar_try_wait_park(addr, parity)) return;
    __trap();
}

// Bounded wait: a protocol bug must end in a trap (reported as a launch failure), never in a hang.
__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity) {
    for (int spin = 0; spin < (1 << 22); ++spin)
        if (mbar_try_wait(addr, parity)) return;
    __trap();
}
// Development variant (HCGS_TC_DEBUG bit 512): the first wait that times out records {tag, block, warp, parity,
// step} and every later wait returns at once, so the kernel ends normally and hcgs_tc_trace_read shows who hung.
__device__ int g_tc_fail[8];
__device__ __forceinline__ void mbar_wait_dbg(uint32_t addr, uint32_t parity, int dbg, int tag, int step) {
    if (!(dbg & 512)) {
        mbar_wait_park(addr, parity);
        return;
    }
    for (int spin = 0; spin < (1 << 18); ++spin) {
        if (mbar_try_wait(addr, parity)) return;
        if ((spin & 255) == 255 && *reinterpret_cast<volatile int*>(&g_tc_fail[0])) return;
    }
    if (atomicCAS(&g_tc_fail[0], 0, 1) == 0) {
        g_tc_fail[1] = tag;
        g_tc_fail[2] = (int)blockIdx.x;
        g_tc_fail[3] = (int)(threadIdx.x >> 5);
        g_tc_fail[4] = (int)parity;
        g_tc_fail[5] = step;
        __threadfence();
    }
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
// Fast transcendental forms for the bf16 build (MUFU.EX2 / MUFU.RCP; abs. error ~1e-6, far below
// the bf16 operand rounding): the libm tanhf/expf of the fp32 parity kernels cost ~200 issue slots
// per (unit, sequence) item and were the largest share of a step.
__device__ __forceinline__ float fast_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float fast_tanh(float x) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * x)); }
__device__ __forceinline__ float act_apply_fast(int act, float x) {
    if (act == HCGS_ACT_TANH) return fast_tanh(x);
    if (act == HCGS_ACT_SIGMOID) return fast_sigmoid(x);
    return act_apply(act, x);
}

// phase timing for kernel tuning (HCGS_TC_DEBUG bit 64): thread 0 of every CTA accumulates clock64 deltas per
// phase in shared memory; CTA 0 publishes them (hcgs_tc_trace_read).  Costs two predicated branches per phase.
__device__ long long g_tc_trace[16];
__device__ long long g_tc_tl[20 * 16];  // HCGS_TC_DEBUG bit 128: clock stamps of every warp of CTA 0 at one time step
#define TC_TL(i)                                                             \
    if ((a.dbg & 128) && lane == 0 && blockIdx.x == 0 && s == a.T / 2) g_tc_tl[warp * 16 + (i)] = clock64();
#define TC_TRACE(p)                                   \
    if (a.dbg & 64) {                                 \
        if (threadIdx.x == 0) {                       \
            const long long now_ = clock64();         \
            trace_s[p] += now_ - trace_s[15];         \
            trace_s[15] = now_;                       \
        }                                             \
    }
#define TC_TRACE_INIT()                                        \
    __shared__ long long trace_s[16];                          \
    if (threadIdx.x == 0) {                                    \
        for (int i_ = 0; i_ < 15; ++i_) trace_s[i_] = 0;       \
        trace_s[15] = clock64();                               \
    }
#define TC_TRACE_FLUSH()                                                          \
    if ((a.dbg & 64) && threadIdx.x == 0 && blockIdx.x == 0)                      \
        for (int i_ = 0; i_ < 15; ++i_) g_tc_trace[i_] = trace_s[i_];

__host__ __device__ inline size_t tc_align128(size_t b) { return (b + 127) & ~(size_t)127; }

struct TcFwdArgs {
    const uint2* wfrag;   // [CL][SPC][KS][2][32]
    const uint16_t* cg;   // [CL][SPC][KS*4]
    const uint8_t* nks;   // [CL][SPC]
    const float* pre;     // [T*B][4*ldH] gate-interleaved: (row, u, g) at pre[row*4*ldH + u*4 + g]
    const float* drop;    // [B,H] or NULL
    float drop_scalar;
    float* h_out;           // [T*B][ldH] fp32 (pad columns zero) or NULL
    __nv_bfloat16* hb_out;  // [T*B][ldH] bf16 copy for the tensor-core GEMMs, or NULL
    __nv_bfloat16* hl_out;  // [T*B][ldH] bf16(h - float(bf16(h))): second term of the next layer's two-term projection, or NULL
    float* c_out;           // [T*B][ldH] or NULL
    float* gates_out;       // [T*B][4*ldH] gate-interleaved, or NULL
    // fused batch norm of the pre-activations (all NULL = pre is used as is): z = (pre - mean) * invstd * gamma + beta
    const float* bn_mean;    // [4*ldH] gate-interleaved (column u*4+g)
    const float* bn_invstd;  // [4*ldH]
    PtrPack<const float> bn_gamma, bn_beta;  // per gate [H] (NULL = 1 / 0)
    int T, B, H, ldH, CL, SPC, KS, HP, act;
    int dbg;  // timing experiments only (HCGS_TC_DEBUG): 1 = no global stores, 2 = no prefetch, 4 = no MMA
    // TMA build: [T][B][cols] tensor maps with boxes {BW | Hc, 16, 1}; the pre / gates slice of a CTA is NBX boxes
    int BW, NBX;
    alignas(64) CUtensorMap tm_pre, tm_gates, tm_c, tm_h, tm_hb, tm_hl;
};

// NT m16 tiles of sequences per cluster, SPW strips per warp, KSM >= k-steps per strip (compile time
// so that the gather offsets and the per-item state live in registers).
template <int ACT>
__device__ __forceinline__ float act_fwd_t(int act, float x) {
    if (ACT == HCGS_ACT_TANH) return fast_tanh(x);
    return act_apply_fast(act, x);
}

// ---- forward, bulk-copy build --------------------------------------------------------------------
// All data movement is done by the bulk-copy engine instead of the LSU (measured with HCGS_TC_DEBUG=64 on a first
// LSU-based version: pushes, release fence + arrives, global stores and prefetch issue were >55 % of a step):
//   * h exchange   : every CTA owns slice `rank` of the slice-major h buffer [2][CL][NB][HcS]; after
//                    the gate math ONE cp.async.bulk (shared::cta -> shared::cluster) per peer moves the
//                    whole slice and completes transaction bytes on the peer's mbarrier -- no release
//                    fence, no arrive messages, no third block barrier;
//   * pre stream   : one cp.async.bulk global -> shared per sequence row and step (mbarrier tx);
//   * outputs      : gates (in place in the pre stage), c, h (fp32) and the bf16 copy of h leave by
//                    one cp.async.bulk shared -> global per row and tensor.
__device__ __forceinline__ void mbar_expect_tx(uint32_t addr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2c(uint32_t dst_cluster, uint32_t src, uint32_t bytes, uint32_t mbar_cluster) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     dst_cluster),
                 "r"(src), "r"(bytes), "r"(mbar_cluster)
                 : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* m, int c0, int c1, int c2, uint32_t mbar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(m), "r"(c0), "r"(c1), "r"(c2), "r"(mbar)
        : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, int c0, int c1, int c2, uint32_t src) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(m),
                 "r"(src), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void st_async_v4(uint32_t cluster_addr, const uint4& v, uint32_t mbar_cluster) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(
                     cluster_addr),
                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(mbar_cluster)
                 : "memory");
}
__device__ __forceinline__ void st_async_b64(uint32_t cluster_addr, uint64_t v, uint32_t mbar_cluster) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(cluster_addr),
                 "l"(v), "r"(mbar_cluster)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_local(uint32_t addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float clamp17(float x) { return fminf(fmaxf(x, -17.f), 17.f); }

// bf16 elements per row of one h slice: a multiple of 8 (16-byte rows for the bulk copies) that is 8 mod 16,
// so the eight rows of an A-fragment gather fall into different bank groups
__host__ __device__ inline int tc_hcs(int Hc) { return (Hc % 16 == 8) ? Hc : Hc + 8; }

template <int NT, int SPW, int KSM, int ACT>
__global__ void __launch_bounds__((SPW <= 2 ? TC_MAX_WARPS2 : TC_MAX_WARPS) * 32, 1)
    lstm_tc_fwd2_kernel(const __grid_constant__ TcFwdArgs a) {
    constexpr int NB = NT * 16;
    static_assert(NT == 1, "one m16 tile of sequences per cluster");
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int CL = a.CL, SPC = a.SPC, KS = a.KS, H = a.H, ldH = a.ldH;
    const int b0 = (blockIdx.x / CL) * NB;
    const int Hc = SPC * 4;
    const int u0 = rank * Hc;
    const int nv = max(0, min(Hc, ldH - u0));  // units of this CTA inside the padded tensors (multiple of 8)
    const int HcS = tc_hcs(Hc);
    const int BW = a.BW, NBX = a.NBX;  // pre / gate stage = NBX boxes of [16][BW] floats
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x;
    // the last two warps only drive the copy engine (a TMA issue costs its thread 150..600 cycles): DMA warp 0
    // loads the pre boxes and stores the gate boxes, DMA warp 1 stores c / h / bf16 h and arms the h barrier
    const int CW = (nthreads >> 5) - 2;  // compute warps
    const bool is_dma = warp >= CW;
    const int cthreads = CW * 32;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);  // TMA boxes: 128-byte aligned
    float* pre_s = reinterpret_cast<float*>(sp);  // [2][NBX][16][BW]
    sp += (size_t)2 * NB * Hc * 16;
    float* c_s = reinterpret_cast<float*>(sp);  // [16][Hc]
    sp += tc_align128((size_t)NB * Hc * 4);
    float* h_s = reinterpret_cast<float*>(sp);
    sp += tc_align128((size_t)NB * Hc * 4);
    __nv_bfloat16* hb_s = reinterpret_cast<__nv_bfloat16*>(sp);  // [16][Hc] dense bf16 copy for the store
    sp += tc_align128((size_t)NB * Hc * 2);
    __nv_bfloat16* hl_s = reinterpret_cast<__nv_bfloat16*>(sp);  // [16][Hc] low part bf16(h - hb) for the store
    sp += tc_align128((size_t)NB * Hc * 2);
    __nv_bfloat16* hbuf = reinterpret_cast<__nv_bfloat16*>(sp);  // [2][CL][NB][HcS]
    sp += tc_align128((size_t)2 * CL * NB * HcS * 2);
    uint2* wf_s = reinterpret_cast<uint2*>(sp);
    sp += tc_align16((size_t)SPC * KS * 2 * 32 * 8);
    float4* bn_m = reinterpret_cast<float4*>(sp);
    sp += tc_align16((size_t)Hc * 16);
    float4* bn_a = reinterpret_cast<float4*>(sp);
    sp += tc_align16((size_t)Hc * 16);
    float4* bn_b = reinterpret_cast<float4*>(sp);
    sp += tc_align16((size_t)Hc * 16);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(sp);
    const bool use_bn = a.bn_mean != nullptr;
    const size_t slice = (size_t)NB * HcS;  // bf16 elements of one slice
    const uint32_t push_bytes = (uint32_t)NB * Hc * 2;  // bytes one peer receives from this CTA per step
    const uint32_t mbar_addr = smem_u32(mbar);
    const uint32_t hbuf_addr = smem_u32(hbuf);
    const uint32_t stage_bytes = (uint32_t)NB * Hc * 16;
    const int stage_floats = NB * Hc * 4;

    TC_TRACE_INIT();
    if (use_bn) {
        for (int i = tid; i < Hc * 4; i += nthreads) {
            const int ul = i >> 2, g = i & 3, u = u0 + ul;
            float m = 0.f, sc = 0.f, sh = 0.f;
            if (u < H) {
                m = __ldg(a.bn_mean + (size_t)u * 4 + g);
                sc = __ldg(a.bn_invstd + (size_t)u * 4 + g) * (a.bn_gamma.p[g] ? __ldg(a.bn_gamma.p[g] + u) : 1.f);
                sh = a.bn_beta.p[g] ? __ldg(a.bn_beta.p[g] + u) : 0.f;
            }
            reinterpret_cast<float*>(bn_m)[i] = m;
            reinterpret_cast<float*>(bn_a)[i] = sc;
            reinterpret_cast<float*>(bn_b)[i] = sh;
        }
    }
    {
        const uint4* s4 = reinterpret_cast<const uint4*>(a.wfrag + (size_t)rank * SPC * KS * 64);
        uint4* d4 = reinterpret_cast<uint4*>(wf_s);
        for (int i = tid; i < SPC * KS * 32; i += nthreads) d4[i] = __ldg(s4 + i);
        uint32_t* hz = reinterpret_cast<uint32_t*>(hbuf);
        for (int i = tid; i < (int)(CL * slice); i += nthreads) hz[i] = 0u;  // both buffers (2*CL*slice bf16)
        for (int i = tid; i < NB * Hc; i += nthreads) c_s[i] = h_s[i] = 0.f;
        uint32_t* hbz = reinterpret_cast<uint32_t*>(hb_s);
        uint32_t* hlz = reinterpret_cast<uint32_t*>(hl_s);
        for (int i = tid; i < NB * Hc / 2; i += nthreads) hbz[i] = hlz[i] = 0u;
        if (tid == 0) {
            // [0..1] h_k complete: CW compute-warp arrivals (own slice) + the arm + the peers' transaction bytes
            // [2..3] pre boxes landed   [4..5] outputs of a step complete (CW arrivals)   [6] c/h/hb stages drained
            mbar_init(mbar_addr + 0, (uint32_t)CW + 1u);
            mbar_init(mbar_addr + 8, (uint32_t)CW + 1u);
            mbar_init(mbar_addr + 16, 1u);
            mbar_init(mbar_addr + 24, 1u);
            mbar_init(mbar_addr + 32, (uint32_t)CW);
            mbar_init(mbar_addr + 40, (uint32_t)CW);
            mbar_init(mbar_addr + 48, 1u);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    if (tid == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_pre) : "memory");
        if (a.gates_out) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_gates) : "memory");
        if (a.c_out) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_c) : "memory");
        if (a.h_out) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_h) : "memory");
        if (a.hb_out) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_hb) : "memory");
        if (a.hl_out) asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_hl) : "memory");
    }
    fence_proxy_async();
    __syncthreads();

    // pre-activations of step `step` -> stage (step & 1): NBX boxes, issued by lane 0 of warps 0..NBX-1
    auto load_pre = [&](int step) {  // DMA warp 0, lane 0
        if (step >= a.T || nv == 0 || warp != CW || lane != 0) return;
        const uint32_t bar = mbar_addr + 8 * (2 + (step & 1));
        mbar_expect_tx(bar, stage_bytes);
        for (int x = 0; x < NBX; ++x)
            tma_load_3d(smem_u32(pre_s + (size_t)(step & 1) * stage_floats + (size_t)x * NB * BW), &a.tm_pre,
                        u0 * 4 + x * BW, b0, step, bar);
    };
    load_pre(0);
    load_pre(1);

    // ---- per-thread constants: gather offsets of this warp's strips, item state ----------------
    const int ui = lane & 3;   // unit inside the strip
    const int br = lane >> 2;  // batch row inside the m16 tile (and +8)
    int off0[SPW][KSM], off2[SPW][KSM], pzo[SPW];
    float c_reg[SPW][2], dr[SPW][2];
    bool ok[SPW][2];
#pragma unroll
    for (int i = 0; i < SPW; ++i) {
        const int sl = warp * SPW + i;
#pragma unroll
        for (int ks = 0; ks < KSM; ++ks) {
            off0[i][ks] = off2[i][ks] = 0;
            if (sl < SPC && ks < KS) {
                const uint16_t* cgp = a.cg + ((size_t)(rank * SPC + sl) * KS + ks) * 4 + (ui >> 1);
                const uint16_t c0 = cgp[0], c2 = cgp[2];
                const int col0 = (c0 == 0xFFFF ? 0 : c0 * 4) + (lane & 1) * 2;  // padded blocks carry zero weights
                const int col2 = (c2 == 0xFFFF ? 0 : c2 * 4) + (lane & 1) * 2;
                off0[i][ks] = (col0 / Hc) * (int)slice + col0 % Hc;
                off2[i][ks] = (col2 / Hc) * (int)slice + col2 % Hc;
            }
        }
        const int ul = sl * 4 + ui;
        const int u = u0 + ul;
        pzo[i] = ((ul * 4) / BW) * (NB * BW) + (ul * 4) % BW;  // box, column inside the box (row term added later)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int b = b0 + br + 8 * hh;
            ok[i][hh] = (sl < SPC) && !is_dma && (u < H) && (b < a.B);
            c_reg[i][hh] = 0.f;
            dr[i][hh] = a.drop_scalar;
            if (ok[i][hh] && a.drop) dr[i][hh] = __ldg(a.drop + (size_t)b * H + u);
        }
    }
    cluster.sync();  // every CTA's barriers are initialised and its buffers zeroed before any peer writes

    // strips of this warp and its share of the slice (consecutive units => one 8- or 16-byte chunk per row)
    const int s0w = warp * SPW;
    const int nstr = max(0, min(SPW, SPC - s0w));
    const bool push16 = (nstr & 1) == 0 && (SPW & 1) == 0;
    const int pch = push16 ? nstr / 2 : nstr;  // chunks per row
    const bool coop_push = CL > 4;
    const uint32_t mb_h = mbar_addr, mb_pre = mbar_addr + 16, mb_out = mbar_addr + 32, mb_free = mbar_addr + 48;

    for (int k = 0; k < a.T; ++k) {
        const int cur = k & 1, nxt = cur ^ 1;
        const __nv_bfloat16* hc = hbuf + (size_t)cur * CL * slice;
        __nv_bfloat16* hn = hbuf + (size_t)nxt * CL * slice + (size_t)rank * slice;  // own slice of h_k
        float* stage = pre_s + (size_t)cur * stage_floats;
        if (!is_dma) {
            // ================= compute warps: no block-wide barrier, only mbarrier waits ====================
            TC_TRACE(0);
            if (k > 0) mbar_wait_park(mb_h + cur * 8, (uint32_t)(((k - 1) >> 1) & 1));  // h_{k-1}: own warps + peers
            TC_TRACE(2);
            // ---- U h_{k-1}: gather A fragments from the h buffer, B fragments resident ---------------
            float acc[SPW][2][4];
    #pragma unroll
            for (int i = 0; i < SPW; ++i) {
    #pragma unroll
                for (int gp = 0; gp < 2; ++gp)
    #pragma unroll
                    for (int q = 0; q < 4; ++q) acc[i][gp][q] = 0.f;
            }
            if (k > 0) {
    #pragma unroll
                for (int ks = 0; ks < KSM; ++ks) {
                    if (ks < KS) {  // uniform
    #pragma unroll
                        for (int i = 0; i < SPW; ++i) {
                            const int sl = warp * SPW + i;
                            if (sl < SPC) {  // warp-uniform
                                const uint2 bA = wf_s[((size_t)(sl * KS + ks) * 2 + 0) * 32 + lane];
                                const uint2 bB = wf_s[((size_t)(sl * KS + ks) * 2 + 1) * 32 + lane];
                                const __nv_bfloat16* row = hc + (size_t)br * HcS;
                                const uint32_t a0 = *reinterpret_cast<const uint32_t*>(row + off0[i][ks]);
                                const uint32_t a1 = *reinterpret_cast<const uint32_t*>(row + 8 * HcS + off0[i][ks]);
                                const uint32_t a2 = *reinterpret_cast<const uint32_t*>(row + off2[i][ks]);
                                const uint32_t a3 = *reinterpret_cast<const uint32_t*>(row + 8 * HcS + off2[i][ks]);
                                if (a.dbg & 4) {  // timing experiment: operand loads without the tensor-core instructions
                                    acc[i][0][0] += __uint_as_float((a0 ^ a1 ^ a2 ^ a3 ^ bA.x ^ bA.y ^ bB.x ^ bB.y) & 0x3f800000u);
                                    continue;
                                }
                                mma_bf16_16816(acc[i][0], a0, a1, a2, a3, bA.x, bA.y);
                                mma_bf16_16816(acc[i][1], a0, a1, a2, a3, bB.x, bB.y);
                            }
                        }
                    }
                }
            }
            TC_TRACE(3);
            if (nv > 0) mbar_wait_park(mb_pre + cur * 8, (uint32_t)((k >> 1) & 1));  // pre boxes of step k landed
            if (k > 0) mbar_wait_park(mb_free, (uint32_t)((k - 1) & 1));  // c / h / hb stages of step k-1 were stored out
            TC_TRACE(1);
            // ---- fused gates / cell / hidden; gates overwrite the staged pre-activations in place ---
    #pragma unroll
            for (int i = 0; i < SPW; ++i) {
                const int sl = warp * SPW + i;
                if (sl >= SPC) continue;  // warp-uniform
                const int ul = sl * 4 + ui;
                float4 bm = make_float4(0.f, 0.f, 0.f, 0.f), ba = make_float4(1.f, 1.f, 1.f, 1.f), bb = bm;
                if (use_bn) {
                    bm = bn_m[ul];
                    ba = bn_a[ul];
                    bb = bn_b[ul];
                }
    #pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    // branch free: absent sequences / pad units run on zeros and are masked at the end, so the
                    // two items and the strips of a thread overlap their MUFU / FMA chains
                    const int bl = br + 8 * hh;
                    const float okf = ok[i][hh] ? 1.f : 0.f;
                    float4* pz = reinterpret_cast<float4*>(stage + pzo[i] + bl * BW);
                    float4 p = *pz;
                    if (use_bn) {
                        p.x = (p.x - bm.x) * ba.x + bb.x;
                        p.y = (p.y - bm.y) * ba.y + bb.y;
                        p.z = (p.z - bm.z) * ba.z + bb.z;
                        p.w = (p.w - bm.w) * ba.w + bb.w;
                    }
                    const float zf = acc[i][0][hh * 2 + 0] + p.x;
                    const float zi = acc[i][0][hh * 2 + 1] + p.y;
                    const float zo = acc[i][1][hh * 2 + 0] + p.z;
                    const float zc = acc[i][1][hh * 2 + 1] + p.w;
                    float fg, ig, og, gg, cn, hval;
                    if (ACT == HCGS_ACT_TANH) {
                        // four activations, one reciprocal: 1/(1+e_x) for x in {f, i, o} and 1 - 2/(1+e^{2 z_c});
                        // arguments clamped to +-17 (sigmoid(-17) = 4e-8) so the product of the four denominators
                        // stays far inside the fp32 range
                        const float da = 1.f + __expf(-clamp17(zf)), db = 1.f + __expf(-clamp17(zi));
                        const float dc = 1.f + __expf(-clamp17(zo)), dd = 1.f + __expf(clamp17(2.f * zc));
                        const float ab = da * db, cd = dc * dd;
                        const float r = rcp_approx(ab * cd);
                        const float rab = r * cd, rcd = r * ab;
                        fg = rab * db;
                        ig = rab * da;
                        og = rcd * dd;
                        gg = 1.f - 2.f * (rcd * dc);
                        cn = ig * gg * dr[i][hh] + fg * c_reg[i][hh];
                        hval = og * (1.f - 2.f * rcp_approx(1.f + __expf(clamp17(2.f * cn))));
                    } else {
                        fg = fast_sigmoid(zf), ig = fast_sigmoid(zi), og = fast_sigmoid(zo);
                        gg = act_apply_fast(a.act, zc);
                        cn = ig * gg * dr[i][hh] + fg * c_reg[i][hh];
                        hval = og * act_apply_fast(a.act, cn);
                    }
                    cn *= okf;
                    hval *= okf;
                    c_reg[i][hh] = cn;
                    *pz = make_float4(fg * okf, ig * okf, og * okf, gg * okf);
                    c_s[bl * Hc + ul] = cn;
                    h_s[bl * Hc + ul] = hval;
                    const __nv_bfloat16 hbv = __float2bfloat16_rn(hval);
                    hn[(size_t)bl * HcS + ul] = hbv;
                    hb_s[bl * Hc + ul] = hbv;
                    hl_s[bl * Hc + ul] = __float2bfloat16_rn(hval - __bfloat162float(hbv));
                }
            }
            fence_proxy_async();  // generic writes above -> visible to the TMA stores of the DMA warps
            __syncwarp();
            TC_TRACE(6);
            // (the DMA warps are released AFTER the per-warp push: their TMA stores compete with the remote stores
            //  for the shared-memory port and the push is on the critical path -- measured 0.9 us/step)
            if (lane == 0 && (coop_push || k + 1 >= a.T)) mbar_arrive_local(mb_out + cur * 8);
            if (k + 1 < a.T) {
                // h_k -> every peer with st.async: remote stores that complete transaction bytes on the peer's
                // barrier (no release fence, no arrive message).
                if (!coop_push) {
                    // per warp, right after its own gate math, so that the DSMEM traffic overlaps the other warps
                    const uint32_t base = (uint32_t)((((size_t)nxt * CL + rank) * slice + (size_t)s0w * 4) * 2);
                    const int total = NB * pch * 3;  // this path is the cluster of 4 (coop_push = CL > 4)
                    for (int q = lane; q < total; q += 32) {
                        const int row = q & 15, rest = q >> 4;
                        const int d = rest % 3, ch = rest / 3;  // constants: no integer division in the time loop
                        const uint32_t peer = (uint32_t)((rank + 1 + d) & 3);
                        const uint32_t off = base + (uint32_t)(row * HcS * 2 + ch * (push16 ? 16 : 8));
                        const unsigned char* src = reinterpret_cast<const unsigned char*>(hbuf) + off;
                        if (push16)
                            st_async_v4(mapa_u32(hbuf_addr + off, peer), *reinterpret_cast<const uint4*>(src),
                                        mapa_u32(mb_h + nxt * 8, peer));
                        else
                            st_async_b64(mapa_u32(hbuf_addr + off, peer), *reinterpret_cast<const uint64_t*>(src),
                                         mapa_u32(mb_h + nxt * 8, peer));
                    }
                } else {
                    // large clusters: many small per-warp bursts cost more than one cooperative sweep over the
                    // finished slice (measured at cluster 8), so wait for all compute warps first
                    mbar_wait_park(mb_out + cur * 8, (uint32_t)((k >> 1) & 1));
                    const uint32_t base = (uint32_t)(((size_t)nxt * CL + rank) * slice * 2);
                    const int rchunks = Hc / 8;  // 16-byte chunks per row
                    for (int q = tid; q < NB * rchunks * (CL - 1); q += cthreads) {
                        const int d = q / (NB * rchunks), rc = q - d * (NB * rchunks);
                        const int row = rc / rchunks, ch = rc - row * rchunks;
                        const uint32_t peer = (uint32_t)((rank + 1 + d) % CL);
                        const uint32_t off = base + (uint32_t)(row * HcS * 2 + ch * 16);
                        const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(hbuf) + off);
                        st_async_v4(mapa_u32(hbuf_addr + off, peer), v, mapa_u32(mb_h + nxt * 8, peer));
                    }
                    __syncwarp();
                }
                if (lane == 0) {
                    if (!coop_push) mbar_arrive_local(mb_out + cur * 8);
                    mbar_arrive_local(mb_h + nxt * 8);
                }
            }
            TC_TRACE(8);
        } else if (lane == 0) {
            // ================= DMA warps: drive the copy engine ==============================================
            if (warp == CW) {
                mbar_wait_park(mb_out + cur * 8, (uint32_t)((k >> 1) & 1));  // gate math of step k is complete
                if (nv > 0 && a.gates_out) {
                    for (int x = 0; x < NBX; ++x)
                        tma_store_3d(&a.tm_gates, u0 * 4 + x * BW, b0, k, smem_u32(stage + (size_t)x * NB * BW));
                    bulk_commit();
                    bulk_wait_read<0>();
                }
                load_pre(k + 2);  // into the stage just drained
            } else {
                // arm this CTA's barrier for the peers' slices of h_k
                if (k + 1 < a.T) mbar_expect_tx(mb_h + nxt * 8, (uint32_t)(CL - 1) * push_bytes);
                mbar_wait_park(mb_out + cur * 8, (uint32_t)((k >> 1) & 1));
                if (nv > 0) {
                    if (a.hb_out) tma_store_3d(&a.tm_hb, u0, b0, k, smem_u32(hb_s));
                    if (a.hl_out) tma_store_3d(&a.tm_hl, u0, b0, k, smem_u32(hl_s));
                    if (a.c_out) tma_store_3d(&a.tm_c, u0, b0, k, smem_u32(c_s));
                    if (a.h_out) tma_store_3d(&a.tm_h, u0, b0, k, smem_u32(h_s));
                    bulk_commit();
                    bulk_wait_read<0>();
                }
                mbar_arrive_local(mb_free);
            }
        }
    }
    if (is_dma && lane == 0) bulk_wait_all<0>();
    TC_TRACE_FLUSH();
    cluster.sync();  // nobody exits while peers may still write into its shared memory
}

// ---- backward ----------------------------------------------------------------------------------
struct TcBwdArgs {
    const uint2* wfrag;       // [n_blocks][16]  (lanes 0..15 of the B fragment; n columns 4..7 are not used)
    const int* blk_off;       // [CL+1]
    const int* grp_off;       // [CL+1]
    const uint32_t* grp;      // [n_groups]
    const int* task_off;      // [CL+1]
    const int* task;          // [n_tasks][8]
    const int* wtp;           // [CL][TCB_CW+1]
    const int* rx;            // [CL][3] {bytes received per step, silent sources, mask of silent destinations}
    const float* dh_out;      // [T*B][ldH]
    const float* gates;       // [T*B][4*ldH] gate-interleaved
    const float* c;           // [T*B][ldH]
    const float* drop;
    float drop_scalar;
    float* dpre;             // [T*B][4*ldH] gate-interleaved fp32
    __nv_bfloat16* dpreb;    // optional bf16 copy, same layout
    int T, B, H, ldH, CL, SPC, NS, HP, max_blocks, max_groups, max_tasks, act;
    int dbg;  // timing experiments only (HCGS_TC_DEBUG): 1 = no global stores, 4 = no MMA
    // [T][B][cols] tensor maps with boxes {BW | Hc, 16, 1}
    int BW, NBX;
    alignas(64) CUtensorMap tm_gates, tm_c, tm_dh;
};

// the m16 x k16 bf16 A tile of one mma in ONE shared-memory instruction: lane i passes the address of row (i & 15),
// k-half (i >> 4); registers come back in a0..a3 fragment order
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
// register-only form the compiler may schedule freely
__device__ __forceinline__ void mma_bf16_16816_s(float (&d)[4], const uint32_t (&af)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(af[0]), "r"(af[1]), "r"(af[2]), "r"(af[3]), "r"(b0), "r"(b1));
}
// B fragments are written once before the time loop: a plain (schedulable) shared load
__device__ __forceinline__ uint2 lds_u2(uint32_t addr) {
    uint2 v;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

// The row-strip groups [g0, g1) of one task: per group one A tile (ldmatrix.x4) and one mma per kept block into the
// accumulator of its column strip J0 <= j < J0 + JN (the loop is issue bound, so tasks inside one half of a column block
// use the variant that tests four strips).  Lane L votes bit (L & 7) of the group's mask: the ballot is the mask again,
// but in a form the compiler knows to be warp-uniform -- uniform predicates, no re-convergence instructions around the
// predicated mma (redux.sync would do the same and is far slower: one per ~30 cycles per SM).  The word and the A tile
// of group g + 1 are fetched before the mma of group g (the table has one spare entry behind the last group).
template <int J0, int JN>
__device__ __forceinline__ void tc_run_groups(float (&acc)[8][4], const uint32_t* grp_s, int g0, int g1,
                                              uint32_t lm_base, uint32_t bp, uint32_t lane_bit) {
    uint32_t meta = grp_s[g0];
    uint32_t af[4];
    ldmatrix_x4(af, lm_base + (meta & 0xFFFFu));
    for (int g = g0; g < g1; ++g) {
        const uint32_t meta_n = grp_s[g + 1];
        uint32_t an[4];
        ldmatrix_x4(an, lm_base + (meta_n & 0xFFFFu));
        const uint32_t kept = __ballot_sync(0xffffffffu, (meta & lane_bit) != 0u);
#pragma unroll
        for (int j = J0; j < J0 + JN; ++j)
            if (kept & (1u << j)) {
                const uint2 bw = lds_u2(bp);
                bp += 128u;
                mma_bf16_16816_s(acc[j], af, bw.x, bw.y);
            }
        meta = meta_n;
#pragma unroll
        for (int x = 0; x < 4; ++x) af[x] = an[x];
    }
}

// ---- backward, third build ----------------------------------------------------------------------
// 18 warps, all of them compute.  Per time step k (descending):
//   item phase : every thread owns one unit of the CTA's slice and up to IT sequences: dh_k = dh_out[k] + the partial
//                sums the CLUSTER pushed into this CTA's slots during step k+1; gate derivatives from the TMA-staged
//                gates / c; dpre leaves as fp32 + bf16 global stores and as the bf16 A tile (unit-major,
//                k = row-in-strip * 4 + gate) in shared memory;
//   MMA phase  : a warp walks its tasks (plan: column block of <= 8 strips x this CTA's rows).  Per row-strip group
//                ONE ldmatrix.x4 brings the 16 x 16 A tile, then one mma per kept 4x4 block of the group into the
//                accumulator of its column strip (uniform predicates from the group's bit mask); round 1 loaded the
//                A tile once per block and that shared-memory traffic bounded the phase;
//   push       : a finished task leaves straight from the accumulators: two shuffles turn a strip's accumulator
//                layout into 16-byte row pieces and ONE st.async.v4 (16 lanes, 256 contiguous bytes) per strip writes
//                them into the owner's slot [source][strip][16 rows x 4 columns] and completes transaction bytes on
//                its barrier (tools/probe/dsmem_push.cu: 2.4 k cycles per exchange against 3.9 k for the 8-byte form
//                of round 1; staged bulk copies are slower than either).
// Synchronisation: inside the CTA two NAMED barriers per step (inputs ready / A tile complete); only thread 0 waits on
// mbarriers (TMA inputs, partial sums of the peers) and issues the TMA prefetches, the arm of the receive barrier and
// the consumed-handshake.  Round 1 and the first version of this build let all 18 warps + 2 DMA warps wait on
// mbarriers: every transaction byte that lands wakes every sleeping waiter, and the wake-up / re-check loop was 54 % of
// all executed warp instructions (ncu, profiles/r02_ncu_lstm_tc_bwd3.md) -- issue slots taken from the working warps.
// The rows of a strip's slot are XOR-swizzled with the strip index so that the item phase reads them without bank
// conflicts; the slots are single-buffered, the consumed-handshake (one remote arrive per peer and step) guards
// their reuse; the owner sums the sources in rank order (deterministic).
template <int ID, int THREADS>
__device__ __forceinline__ void named_bar_sync() {
    asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(THREADS) : "memory");
}

template <int IT, int ACT>
__global__ void __launch_bounds__(TCB_CW * 32, 1) lstm_tc_bwd3_kernel(const __grid_constant__ TcBwdArgs a) {
    constexpr int NB = 16;
    constexpr int CW = TCB_CW;
    constexpr int cthreads = CW * 32;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int CL = a.CL, SPC = a.SPC, H = a.H, ldH = a.ldH;
    const int b0 = (blockIdx.x / CL) * NB;
    const int Hc = SPC * 4;
    const int u0 = rank * Hc;
    const int nv = max(0, min(Hc, ldH - u0));
    const int DPS = Hc * 4 + 8;  // bf16 elements per dpre row: pitch = 16 mod 128 bytes (Hc multiple of 8)
    const int BW = a.BW, NBX = a.NBX;
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int blk0 = a.blk_off[rank];
    const int nblk = a.blk_off[rank + 1] - blk0;
    const int grp0 = a.grp_off[rank], ngrp = a.grp_off[rank + 1] - grp0;
    const int task0 = a.task_off[rank], ntask = a.task_off[rank + 1] - task0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* sp = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    float* gates_s = reinterpret_cast<float*>(sp);  // [2][NBX][16][BW]
    sp += (size_t)2 * NB * Hc * 16;
    float* c_s = reinterpret_cast<float*>(sp);  // [2][16][Hc]
    sp += tc_align128((size_t)2 * NB * Hc * 4);
    float* dh_s = reinterpret_cast<float*>(sp);  // [16][Hc]
    sp += tc_align128((size_t)NB * Hc * 4);
    float* slots = reinterpret_cast<float*>(sp);  // [CL][SPC][16 x 4]: partial dh of the next step, one slot per source
    sp += tc_align128((size_t)CL * SPC * 256);
    __nv_bfloat16* dpre_s = reinterpret_cast<__nv_bfloat16*>(sp);  // [16][DPS], unit-major (u*4+g)
    sp += tc_align16((size_t)NB * DPS * 2);
    uint2* wf_s = reinterpret_cast<uint2*>(sp);
    sp += tc_align16((size_t)a.max_blocks * 16 * 8);
    uint32_t* grp_s = reinterpret_cast<uint32_t*>(sp);
    sp += tc_align16((size_t)(a.max_groups + 2) * 4);
    int4* task_s = reinterpret_cast<int4*>(sp);
    sp += tc_align16((size_t)a.max_tasks * 32);
    int* wtp_s = reinterpret_cast<int*>(sp);
    sp += tc_align16((size_t)(CW + 1) * 4);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(sp);
    const uint32_t mbar_addr = smem_u32(mbar);
    // [0..1] inputs of a step landed (two arms: gates two steps ahead, c / dh one step ahead)
    // [2] slots full (arm + silent peers' arrivals + peers' bytes)  [3] peers have consumed their slots (CL-1 arrivals)
    const uint32_t mb_in = mbar_addr, mb_full = mbar_addr + 16, mb_free = mbar_addr + 24;
    const uint32_t slots_addr = smem_u32(slots);
    const int stage_floats = NB * Hc * 4;
    const uint32_t stage_bytes = (uint32_t)stage_floats * 4, box_bytes = (uint32_t)NB * Hc * 4;
    const uint32_t rx_bytes = (uint32_t)a.rx[rank * 3];
    const bool staged = (CL - 1) * 256 <= 1024;  // the foreign strips fit a gate stage (SPC * 1024 bytes): CL <= 5
    const uint32_t stage_addr0 = smem_u32(gates_s);
    const int silent_in = a.rx[rank * 3 + 1], silent_out = a.rx[rank * 3 + 2];

    TC_TRACE_INIT();
    {
        const uint4* s4 = reinterpret_cast<const uint4*>(a.wfrag + (size_t)blk0 * 16);
        uint4* d4 = reinterpret_cast<uint4*>(wf_s);
        for (int i = tid; i < nblk * 8; i += blockDim.x) d4[i] = __ldg(s4 + i);
        for (int i = tid; i < ngrp + 2; i += blockDim.x) grp_s[i] = i < ngrp ? a.grp[grp0 + i] : 0u;
        const int4* t4 = reinterpret_cast<const int4*>(a.task) + (size_t)task0 * 2;
        for (int i = tid; i < ntask * 2; i += blockDim.x) task_s[i] = __ldg(t4 + i);
        for (int i = tid; i <= CW; i += blockDim.x) wtp_s[i] = a.wtp[(size_t)rank * (CW + 1) + i];
        for (int i = tid; i < CL * SPC * 64; i += blockDim.x) slots[i] = 0.f;
        uint32_t* dz = reinterpret_cast<uint32_t*>(dpre_s);
        for (int i = tid; i < NB * DPS / 2; i += blockDim.x) dz[i] = 0u;
        if (tid == 0) {
            mbar_init(mb_in, 2u);
            mbar_init(mb_in + 8, 2u);
            mbar_init(mb_full, (uint32_t)(1 + silent_in));
            mbar_init(mb_free, (uint32_t)(CL - 1));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_gates) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_c) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&a.tm_dh) : "memory");
        }
    }
    fence_proxy_async();
    __syncthreads();

    // thread 0: inputs of time step kk.  Gate boxes go to stage kk&1 two steps ahead (first arm of barrier kk&1);
    // dh_{kk} (single buffer) and c_{kk-1} (c buffer (kk-1)&1) one step ahead (second arm)
    auto load_gates = [&](int kk) {
        const uint32_t bar = mb_in + 8 * (kk & 1);
        mbar_expect_tx(bar, stage_bytes);
        for (int x = 0; x < NBX; ++x)
            tma_load_3d(smem_u32(gates_s + (size_t)(kk & 1) * stage_floats + (size_t)x * NB * BW), &a.tm_gates,
                        u0 * 4 + x * BW, b0, kk, bar);
    };
    auto load_small = [&](int kk, bool with_c_kk) {
        const uint32_t bar = mb_in + 8 * (kk & 1);
        mbar_expect_tx(bar, box_bytes * (1u + (kk >= 1 ? 1u : 0u) + (with_c_kk ? 1u : 0u)));
        tma_load_3d(smem_u32(dh_s), &a.tm_dh, u0, b0, kk, bar);
        if (kk >= 1) tma_load_3d(smem_u32(c_s + (size_t)((kk - 1) & 1) * NB * Hc), &a.tm_c, u0, b0, kk - 1, bar);
        if (with_c_kk) tma_load_3d(smem_u32(c_s + (size_t)(kk & 1) * NB * Hc), &a.tm_c, u0, b0, kk, bar);
    };
    if (tid == 0 && nv > 0) {
        load_gates(a.T - 1);
        load_small(a.T - 1, true);
        if (a.T > 1) load_gates(a.T - 2);
    }

    // ---- pointwise items: thread -> unit `ul` of the slice (fastest over the lanes => conflict-free shared loads,
    // coalesced 16-byte global stores) and the sequences bq, bq + BPP, ...; every offset is affine in the slot j
    const int BPP = min(NB, cthreads / Hc);  // sequences per pass
    const bool active = tid < BPP * Hc;
    const int bq = active ? tid / Hc : 0, ul = active ? tid - (tid / Hc) * Hc : 0;
    const int u = u0 + ul;
    const bool unit_ok = active && u < H, unit_pad = active && u >= H && u < ldH;
    const int swz = (ul >> 2) & 7;
    const int sm0 = bq * Hc + ul;                                                          // dh_s / c_s
    const int gs0 = ((ul * 4) / BW) * (NB * BW) + bq * BW + (ul * 4) % BW;                 // inside a gate stage
    const int ds0 = bq * DPS + ul * 4;                                                     // bf16 dpre tile
    const int so0 = (ul >> 2) * 64 + (ul & 3);                                             // inside one source's slots
    const uint32_t go0 = (uint32_t)((size_t)(b0 + bq) * 4 * ldH + (size_t)u * 4);          // inside one step of dpre
    float dc_carry[IT], dr[IT];
#pragma unroll
    for (int j = 0; j < IT; ++j) {
        const int b = bq + j * BPP;
        dc_carry[j] = 0.f;
        dr[j] = a.drop_scalar;
        if (unit_ok && b < NB && b0 + b < a.B && a.drop) dr[j] = __ldg(a.drop + (size_t)(b0 + b) * H + u);
    }
    const int br = lane >> 2, q = lane & 3;
    const int wt0 = wtp_s[warp], wt1 = wtp_s[warp + 1];
    const uint32_t lm_base = smem_u32(dpre_s) + (uint32_t)(((lane & 15) * DPS + (lane >> 4) * 8) * 2);
    const uint32_t wf_lane = smem_u32(wf_s) + (uint32_t)(lane & 15) * 8u;
    const uint32_t lane_bit = 0x10000u << (lane & 7);
    cluster.sync();

    for (int s = 0; s < a.T; ++s) {
        const int k = a.T - 1 - s;
        TC_TRACE(0);
        if (tid == 0) {
            if (nv > 0) mbar_wait_dbg(mb_in + 8 * (k & 1), (uint32_t)((s >> 1) & 1), a.dbg, 1, s);
            if (s > 0) mbar_wait_dbg(mb_full, (uint32_t)((s - 1) & 1), a.dbg, 3, s);  // partial dh_k from every CTA has landed
        }
        TC_TL(0);
        named_bar_sync<1, cthreads>();  // inputs and partial sums are in place; everybody's local slot writes are done
        TC_TRACE(1);
        TC_TL(1);
        if (active) {
            const float* gst = gates_s + (size_t)(k & 1) * stage_floats;
            const float* ck_s = c_s + (size_t)(k & 1) * NB * Hc;
            const float* cm_s = c_s + (size_t)((k & 1) ^ 1) * NB * Hc;
            float* dpre_k = a.dpre + (size_t)k * a.B * 4 * ldH;
            __nv_bfloat16* dpreb_k = a.dpreb + (size_t)k * a.B * 4 * ldH;
            const bool st32 = a.dpre != nullptr && !(a.dbg & 1), st16 = a.dpreb != nullptr && !(a.dbg & 1);
            const bool has_prev = k > 0;  // c_{-1} = 0 (the other c buffer then holds an older step or nothing)
#pragma unroll
            for (int j = 0; j < IT; ++j) {
                const int b = bq + j * BPP;
                if (b >= NB) break;
                const bool ok = unit_ok && b0 + b < a.B;
                const int sm = sm0 + j * BPP * Hc;
                float dh = dh_s[sm];
                if (s > 0) {
                    const float* sl_p = slots + so0 + ((b ^ swz) << 2);
                    if (CL == 4) {  // the common cluster: four independent loads, same summation order
                        const float p0 = sl_p[0], p1 = sl_p[(size_t)SPC * 64], p2 = sl_p[(size_t)2 * SPC * 64],
                                    p3 = sl_p[(size_t)3 * SPC * 64];
                        dh = (((dh + p0) + p1) + p2) + p3;
                    } else {
                        for (int src = 0; src < CL; ++src) dh += sl_p[(size_t)src * SPC * 64];
                    }
                }
                const float4 g4 = *reinterpret_cast<const float4*>(gst + gs0 + j * BPP * BW);
                const float cval = ck_s[sm];
                const float cmv = cm_s[sm];
                const float cm1 = has_prev ? cmv : 0.f;
                const float fg = g4.x, ig = g4.y, og = g4.z, gg = g4.w;
                const float tc = (ACT == HCGS_ACT_TANH) ? fast_tanh(cval) : act_apply_fast(a.act, cval);
                const float dog = dh * tc;
                const float dct = dh * og * ((ACT == HCGS_ACT_TANH) ? (1.f - tc * tc) : act_grad_from_out(a.act, tc)) +
                                  dc_carry[j];
                const float dig = dct * gg * dr[j];
                const float dgg = dct * ig * dr[j];
                const float dfg = dct * cm1;
                float4 dp;
                dp.x = dfg * fg * (1.f - fg);
                dp.y = dig * ig * (1.f - ig);
                dp.z = dog * og * (1.f - og);
                dp.w = dgg * ((ACT == HCGS_ACT_TANH) ? (1.f - gg * gg) : act_grad_from_out(a.act, gg));
                if (!ok) dp = make_float4(0.f, 0.f, 0.f, 0.f);
                dc_carry[j] = ok ? dct * fg : 0.f;
                const uint2 pb = make_uint2(pack_bf16x2(dp.x, dp.y), pack_bf16x2(dp.z, dp.w));
                if (ok || (unit_pad && b0 + b < a.B)) {
                    const uint32_t go = go0 + (uint32_t)(j * BPP) * 4u * (uint32_t)ldH;
                    if (st32) *reinterpret_cast<float4*>(dpre_k + go) = dp;
                    if (st16) *reinterpret_cast<uint2*>(dpreb_k + go) = pb;
                }
                *reinterpret_cast<uint2*>(dpre_s + ds0 + j * BPP * DPS) = pb;
            }
        }
        TC_TRACE(2);
        TC_TL(2);
        if (k == 0) break;
        named_bar_sync<2, cthreads>();  // the bf16 dpre tile is complete; gates / c / dh / slots of step k are consumed
        TC_TRACE(3);
        TC_TL(3);
        bool freed = false;
        if (tid == 0) {
            // arm for the partial dh_{k-1} (the previous phase of this barrier is complete: waited for above), tell the
            // peers that this CTA's slots may be overwritten (relaxed: nothing is published, the slot reads are
            // complete; a release per arrive costs ~1000 cycles), prefetch the inputs of the next steps
            mbar_expect_tx(mb_full, rx_bytes);
            for (int d = 1; d < CL; ++d) mbar_arrive_cluster_relaxed(mapa_u32(mb_free, (uint32_t)((rank + d) % CL)));
            if (nv > 0) {
                if (!staged && k >= 2) load_gates(k - 2);
                load_small(k - 1, false);
            }
            if (staged) {
                // the gate stage consumed at step s-1 was the staging area of that step's pushes.  Once every peer has
                // consumed ITS slots of this step, those copies have landed (they were part of what the peers waited
                // for): only now may the TMA overwrite the stage with the gates of step k-1
                mbar_wait_dbg(mb_free, (uint32_t)(s & 1), a.dbg, 6, s);
                freed = true;
                if (nv > 0 && s > 0) load_gates(k - 1);
            }
        }
        TC_TL(4);
        // ---- partial dh_{k-1}: this warp's tasks
        for (int ti = wt0; ti < wt1; ++ti) {
            const int4 ta = task_s[2 * ti], tb = task_s[2 * ti + 1];  // {g0, g1, blk0, jlo}, {nst, dest, sl0, -}
            float acc[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            const uint32_t bp = wf_lane + (uint32_t)ta.z * 128u;
            const int g1 = (a.dbg & 4) ? ta.x : ta.y;
            if (ta.w + tb.x <= 4) tc_run_groups<0, 4>(acc, grp_s, ta.x, g1, lm_base, bp, lane_bit);
            else if (ta.w >= 4) tc_run_groups<4, 4>(acc, grp_s, ta.x, g1, lm_base, bp, lane_bit);
            else tc_run_groups<0, 8>(acc, grp_s, ta.x, g1, lm_base, bp, lane_bit);
            if (ti == wt0) { TC_TL(5); }
            // ---- the task's strips leave from the accumulators (lane (br, q) holds columns 2q, 2q+1 of rows br and
            // br+8 for q < 2; n columns 4..7 are unused)
            const int d = tb.y;
            const bool remote = d != rank;
            if (remote && !freed) {  // the peers have consumed the slots of this step
                TC_TRACE(4);
                mbar_wait_dbg(mb_free, (uint32_t)(s & 1), a.dbg, 5, s);
                freed = true;
                TC_TRACE(5);
            }
            if (ti == wt0) { TC_TL(6); }
            const uint32_t obase = slots_addr + (uint32_t)((rank * SPC + tb.z) * 256);
            // staged: remote strips are written to the gate stage consumed by this step's item phase (compact index over
            // the CL-1 foreign slices) and ONE bulk copy per task moves them: the copy engine does the DSMEM transfer,
            // the LSU queue stays free for the other warps' mma operands (direct st.async pushes back-pressure it)
            const int rel = d - rank - 1 + (d <= rank ? CL : 0);  // compact index of the destination among the peers
            const uint32_t sbase = stage_addr0 + (uint32_t)(k & 1) * stage_bytes + (uint32_t)((rel * SPC + tb.z) * 256);
            const bool direct = remote && !staged;
            const uint32_t rbase = direct ? mapa_u32(obase, (uint32_t)d) : (remote ? sbase : obase);
            const uint32_t rbar = remote ? mapa_u32(mb_full, (uint32_t)d) : 0u;
            if (!direct) {
                // local slots or staging: the accumulator layout goes out as it is, two 8-byte stores per strip
                // (lanes q < 2 hold columns 2q, 2q+1 of rows br and br + 8; row r of a strip sits at (r ^ swizzle) * 16)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int tj = j - ta.w;
                    if (tj >= 0 && tj < tb.x && q < 2) {
                        const uint32_t ad = rbase + (uint32_t)(tj * 256 + ((br ^ ((tb.z + tj) & 7)) << 4) + q * 8);
                        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(ad), "f"(acc[j][0]), "f"(acc[j][1]) : "memory");
                        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(ad + 128u), "f"(acc[j][2]), "f"(acc[j][3])
                                     : "memory");
                    }
                }
            } else {
                // direct push (clusters of 8 and 16: the foreign strips do not fit a gate stage): after one exchange with
                // the neighbour lane q = 0 owns the 16 bytes of row br and q = 1 those of row br + 8; one st.async.v4 per
                // strip = 256 contiguous bytes
                const int row = br + 8 * (q & 1);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int tj = j - ta.w;
                    if (tj >= 0 && tj < tb.x) {  // warp-uniform
                        const float sx = (q & 1) ? acc[j][0] : acc[j][2];
                        const float sy = (q & 1) ? acc[j][1] : acc[j][3];
                        const float rx = __shfl_xor_sync(0xffffffffu, sx, 1);
                        const float ry = __shfl_xor_sync(0xffffffffu, sy, 1);
                        uint4 v;
                        v.x = __float_as_uint((q & 1) ? rx : acc[j][0]);
                        v.y = __float_as_uint((q & 1) ? ry : acc[j][1]);
                        v.z = __float_as_uint((q & 1) ? acc[j][2] : rx);
                        v.w = __float_as_uint((q & 1) ? acc[j][3] : ry);
                        const uint32_t off = (uint32_t)(tj * 256 + ((row ^ ((tb.z + tj) & 7)) << 4));
                        if (q < 2) st_async_v4(rbase + off, v, rbar);
                    }
                }
            }
            if (remote && staged) {
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) bulk_s2c(mapa_u32(obase, (uint32_t)d), sbase, (uint32_t)tb.x * 256u, rbar);
            }
        }
        TC_TL(7);
        if (tid == 0 && silent_out) {  // peers this CTA has no partial sums for: an empty arrival keeps them in step
            if (!freed) mbar_wait_dbg(mb_free, (uint32_t)(s & 1), a.dbg, 9, s);
            for (int d = 0; d < CL; ++d)
                if (silent_out & (1 << d)) mbar_arrive_cluster_relaxed(mapa_u32(mb_full, (uint32_t)d));
        }
        TC_TRACE(4);
    }
    TC_TRACE_FLUSH();
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
                                   int CL, int SPC, int64_t total, int unit_major, uint32_t* __restrict__ out) {
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int j = idx & 1, lane = (idx >> 1) & 15;
        const int e = (int)(idx >> 5);
        int r = 0;
        while (r + 1 < CL && blk_off[r + 1] <= e) ++r;
        const int n = lane >> 2;
        const int k0 = (lane & 3) * 2 + 8 * j;
        const int col = blk_cs[e] * 4 + n;
        float lo = 0.f, hi = 0.f;
        if (unit_major) {  // k = row-in-block * 4 + gate: the pair (k0, k0+1) is two gates of one row
            const int i = k0 >> 2, g = k0 & 3;
            const int64_t row = (int64_t)(r * SPC + blk_rs[e]) * 4 + i;
            if (col < H && row < H) {
                lo = __ldg(U.p[g] + row * ldu + col);
                hi = __ldg(U.p[g + 1] + row * ldu + col);
            }
        } else {  // k = gate * 4 + row-in-block: two rows of one gate
            const int g = k0 >> 2, i = k0 & 3;
            const int64_t row = (int64_t)(r * SPC + blk_rs[e]) * 4 + i;
            if (col < H) {
                if (row < H) lo = __ldg(U.p[g] + row * ldu + col);
                if (row + 1 < H) hi = __ldg(U.p[g] + (row + 1) * ldu + col);
            }
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
    HCGS_REQUIRE(nb == 0 || nb == 16, "tensor-core recurrence: nb must be 0 or 16 sequences per cluster");
    const int NB = nb ? nb : 16;
    const int limit = tc_smem_limit();
    const int cands[3] = {4, 8, 16};
    for (int ci = 0; ci < 3; ++ci) {
        if (cluster_size && cands[ci] != cluster_size) continue;
        TcPlan* t = nullptr;
        if (tc_plan_get(plan, cands[ci], &t)) return 1;
        const bool fits = tc_fwd_smem_v(*t, NB) <= (size_t)limit && tc_bwd_smem_v(*t, NB) <= (size_t)limit &&
                          t->SPW <= 4 && t->KS <= 8 && t->Hc <= 256;
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

namespace hcgs {
static int tc_max_active_clusters(const TcPlan& t, int NB, int* fwd_clusters, int* bwd_clusters);
}

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
    info[4] = (int64_t)tc_fwd_smem_v(*t, NB);
    info[5] = (int64_t)tc_bwd_smem_v(*t, NB);
    info[6] = t->WARPS;
    info[7] = t->SPC;
    info[8] = t->KS;
    info[9] = t->n_blocks;
    info[10] = t->SPW;
    info[11] = t->HP;
    int fc = 0, bc = 0;
    if (tc_max_active_clusters(*t, NB, &fc, &bc) == 0) {
        info[12] = fc;
        info[13] = bc;
    }
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
                                                  t->SPC, total, 1, bwd_pack);
        HCGS_LAUNCH_CHECK();
    }
    return 0;
}

namespace hcgs {

// boxes of the TMA builds: a CTA's [16][Hc*4] fp32 slice as NBX boxes of BW floats; BW % 32 == 16 makes the
// float4 accesses of the gate math bank-conflict free
static int tc_pick_box(int Hc, int* BW, int* NBX) {
    *BW = *NBX = 0;
    const int W = Hc * 4;
    for (int nbx = 1; nbx <= Hc; ++nbx) {
        if (W % nbx) continue;
        const int bw = W / nbx;
        if (bw > 256 || bw % 4) continue;
        if (*NBX == 0 || bw % 32 == 16) *BW = bw, *NBX = nbx;
        if (bw % 32 == 16 || bw < 64) break;
    }
    HCGS_REQUIRE(*NBX > 0 && Hc <= 256, "tensor-core recurrence: a slice of %d units has no TMA box", Hc);
    return 0;
}

typedef void (*TcFwdKernel)(const TcFwdArgs);
typedef void (*TcBwdKernel)(const TcBwdArgs);

template <int SPW, int ACT>
static TcFwdKernel tc_pick_fwd_ks(int KS) {
    if (KS <= 4) return lstm_tc_fwd2_kernel<1, SPW, 4, ACT>;
    if (KS <= 8) return lstm_tc_fwd2_kernel<1, SPW, 8, ACT>;
    return nullptr;
}

template <int ACT>
static TcFwdKernel tc_pick_fwd_act(int SPW, int KS) {
    switch (SPW) {
        case 1: return tc_pick_fwd_ks<1, ACT>(KS);
        case 2: return tc_pick_fwd_ks<2, ACT>(KS);
        case 3: return tc_pick_fwd_ks<3, ACT>(KS);
        case 4: return tc_pick_fwd_ks<4, ACT>(KS);
    }
    return nullptr;
}

static TcFwdKernel tc_pick_fwd(int NT, int SPW, int KS, int act) {
    if (NT != 1) return nullptr;
    return act == HCGS_ACT_TANH ? tc_pick_fwd_act<HCGS_ACT_TANH>(SPW, KS) : tc_pick_fwd_act<-1>(SPW, KS);
}

// slots of sequences per compute thread of the backward item phase: 16 sequences / (576 threads / Hc units)
static int tc_bwd_items(int Hc) {
    const int bpp = std::min(16, (TCB_CW * 32) / std::max(Hc, 1));
    return bpp > 0 ? (16 + bpp - 1) / bpp : 99;
}
template <int ACT>
static TcBwdKernel tc_pick_bwd3(int items) {
    if (items <= 1) return lstm_tc_bwd3_kernel<1, ACT>;
    if (items <= 2) return lstm_tc_bwd3_kernel<2, ACT>;
    if (items <= 4) return lstm_tc_bwd3_kernel<4, ACT>;
    if (items <= 8) return lstm_tc_bwd3_kernel<8, ACT>;
    return nullptr;
}
static TcBwdKernel tc_pick_bwd(int NT, int items, int act = HCGS_ACT_TANH) {
    if (NT != 1) return nullptr;
    return act == HCGS_ACT_TANH ? tc_pick_bwd3<HCGS_ACT_TANH>(items) : tc_pick_bwd3<-1>(items);
}

// co-resident clusters of the two kernels on the current device (0 when the query is unavailable)
static int tc_max_active_clusters(const TcPlan& t, int NB, int* fwd_clusters, int* bwd_clusters) {
    *fwd_clusters = *bwd_clusters = 0;
    const void* ks[2] = {(const void*)tc_pick_fwd(NB / 16, t.SPW, t.KS, HCGS_ACT_TANH),
                         (const void*)tc_pick_bwd(NB / 16, tc_bwd_items(t.Hc))};
    const size_t smem[2] = {tc_fwd_smem_v(t, NB), tc_bwd_smem_v(t, NB)};
    const int threads[2] = {(t.WARPS + 2) * 32, TCB_CW * 32};
    int* outs[2] = {fwd_clusters, bwd_clusters};
    for (int i = 0; i < 2; ++i) {
        if (!ks[i]) continue;
        if (cudaFuncSetAttribute(ks[i], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem[i]) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        if (t.CL > 8) cudaFuncSetAttribute(ks[i], cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)(t.CL * 64), 1, 1);
        cfg.blockDim = dim3((unsigned)threads[i], 1, 1);
        cfg.dynamicSmemBytes = smem[i];
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)t.CL;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, ks[i], &cfg) == cudaSuccess) *outs[i] = n;
        else cudaGetLastError();
    }
    return 0;
}

}  // namespace hcgs

static int lstm_tc_fwd_impl(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                            const float* pre, int ld_h, const float* drop, float drop_scalar, int act,
                            int t_len, int batch, float* h_out, void* hb_out, void* hl_out, float* c_out,
                            float* gates_out, const float* bn_mean, const float* bn_invstd,
                            const float* const* bn_gamma, const float* const* bn_beta, hcgs_stream stream) {
    HCGS_REQUIRE(plan && fwd_pack && pre, "hcgs_lstm_tc_fwd: NULL argument");
    HCGS_REQUIRE((bn_mean == nullptr) == (bn_invstd == nullptr), "hcgs_lstm_tc_fwd: bn_mean and bn_invstd go together");
    HCGS_REQUIRE(h_out || hb_out, "hcgs_lstm_tc_fwd: no output requested");
    if (t_len <= 0 || batch <= 0) return 0;
    TcPlan* t = nullptr;
    int NB = 0;
    if (int rc = tc_resolve(plan, cluster_size, nb, &t, &NB)) return rc;
    HCGS_REQUIRE(ld_h >= t->H && ld_h % 8 == 0, "hcgs_lstm_tc_fwd: ld_h=%d must be a multiple of 8 and >= H=%d", ld_h,
                 t->H);
    TcFwdArgs a;
    a.wfrag = reinterpret_cast<const uint2*>(fwd_pack);
    a.cg = t->d_cg;
    a.nks = t->d_nks;
    a.pre = pre;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.h_out = h_out;
    a.hb_out = reinterpret_cast<__nv_bfloat16*>(hb_out);
    a.hl_out = reinterpret_cast<__nv_bfloat16*>(hl_out);
    a.c_out = c_out;
    a.gates_out = gates_out;
    a.bn_mean = bn_mean;
    a.bn_invstd = bn_invstd;
    a.bn_gamma = make_pack<const float>(bn_mean ? bn_gamma : nullptr, 4);
    a.bn_beta = make_pack<const float>(bn_mean ? bn_beta : nullptr, 4);
    a.T = t_len;
    a.B = batch;
    a.H = t->H;
    a.ldH = ld_h;
    a.CL = t->CL;
    a.SPC = t->SPC;
    a.KS = t->KS;
    a.HP = t->HP;
    a.act = act;
    {
        const char* e = getenv("HCGS_TC_DEBUG");
        a.dbg = e ? atoi(e) : 0;
    }
    a.BW = a.NBX = 0;
    {
        if (int rc = tc_pick_box(t->Hc, &a.BW, &a.NBX)) return rc;
        const uint64_t B64 = (uint64_t)batch, T64 = (uint64_t)t_len, ld = (uint64_t)ld_h;
        if (int rc = make_tmap_3d(&a.tm_pre, 4, pre, 4 * ld, B64, T64, 16 * ld, B64 * 16 * ld, a.BW, 16)) return rc;
        if (gates_out)
            if (int rc = make_tmap_3d(&a.tm_gates, 4, gates_out, 4 * ld, B64, T64, 16 * ld, B64 * 16 * ld, a.BW, 16))
                return rc;
        if (c_out)
            if (int rc = make_tmap_3d(&a.tm_c, 4, c_out, ld, B64, T64, 4 * ld, B64 * 4 * ld, t->Hc, 16)) return rc;
        if (h_out)
            if (int rc = make_tmap_3d(&a.tm_h, 4, h_out, ld, B64, T64, 4 * ld, B64 * 4 * ld, t->Hc, 16)) return rc;
        if (hb_out)
            if (int rc = make_tmap_3d(&a.tm_hb, 2, hb_out, ld, B64, T64, 2 * ld, B64 * 2 * ld, t->Hc, 16)) return rc;
        if (hl_out)
            if (int rc = make_tmap_3d(&a.tm_hl, 2, hl_out, ld, B64, T64, 2 * ld, B64 * 2 * ld, t->Hc, 16)) return rc;
    }
    TcFwdKernel k = tc_pick_fwd(NB / 16, t->SPW, t->KS, act);
    HCGS_REQUIRE(k != nullptr, "hcgs_lstm_tc_fwd: no kernel for nb=%d strips/warp=%d k-steps=%d", NB, t->SPW, t->KS);
    const int n_clusters = (batch + NB - 1) / NB;
    return tc_launch(k, a, n_clusters, t->CL, (t->WARPS + 2) * 32, tc_fwd_smem_v(*t, NB),
                     (cudaStream_t)stream);
}

extern "C" int hcgs_lstm_tc_fwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                                const float* pre, int ld_h, const float* drop, float drop_scalar, int act,
                                int t_len, int batch, float* h_out, void* hb_out, float* c_out, float* gates_out,
                                const float* bn_mean, const float* bn_invstd, const float* const* bn_gamma,
                                const float* const* bn_beta, hcgs_stream stream) {
    return lstm_tc_fwd_impl(plan, cluster_size, nb, fwd_pack, pre, ld_h, drop, drop_scalar, act, t_len, batch, h_out, hb_out,
                            nullptr, c_out, gates_out, bn_mean, bn_invstd, bn_gamma, bn_beta, stream);
}

extern "C" int hcgs_lstm_tc_fwd_split(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* fwd_pack,
                                      const float* pre, int ld_h, const float* drop, float drop_scalar, int act,
                                      int t_len, int batch, float* h_out, void* hb_out, void* hl_out, float* c_out,
                                      float* gates_out, const float* bn_mean, const float* bn_invstd,
                                      const float* const* bn_gamma, const float* const* bn_beta, hcgs_stream stream) {
    HCGS_REQUIRE(hb_out && hl_out, "hcgs_lstm_tc_fwd_split: the high and the low part go together");
    return lstm_tc_fwd_impl(plan, cluster_size, nb, fwd_pack, pre, ld_h, drop, drop_scalar, act, t_len, batch, h_out, hb_out,
                            hl_out, c_out, gates_out, bn_mean, bn_invstd, bn_gamma, bn_beta, stream);
}

extern "C" int hcgs_lstm_tc_bwd(const hcgs_plan* plan, int cluster_size, int nb, const uint32_t* bwd_pack,
                                const float* dh_out, const float* gates, const float* c, int ld_h,
                                const float* drop, float drop_scalar, int act, int t_len, int batch, float* dpre,
                                void* dpre_bf16, hcgs_stream stream) {
    HCGS_REQUIRE(plan && bwd_pack && dh_out && gates && c && (dpre || dpre_bf16), "hcgs_lstm_tc_bwd: NULL argument");
    if (t_len <= 0 || batch <= 0) return 0;
    TcPlan* t = nullptr;
    int NB = 0;
    if (int rc = tc_resolve(plan, cluster_size, nb, &t, &NB)) return rc;
    HCGS_REQUIRE(ld_h >= t->H && ld_h % 8 == 0, "hcgs_lstm_tc_bwd: ld_h=%d must be a multiple of 8 and >= H=%d", ld_h,
                 t->H);
    TcBwdArgs a;
    a.wfrag = reinterpret_cast<const uint2*>(bwd_pack);
    a.blk_off = t->d_blk_off;
    a.grp_off = t->d_grp_off;
    a.grp = t->d_grp;
    a.task_off = t->d_task_off;
    a.task = t->d_task;
    a.wtp = t->d_wtp;
    a.rx = t->d_rx;
    {
        const char* e = getenv("HCGS_TC_DEBUG");
        a.dbg = e ? atoi(e) : 0;
    }
    a.dh_out = dh_out;
    a.gates = gates;
    a.c = c;
    a.drop = drop;
    a.drop_scalar = drop_scalar;
    a.dpre = dpre;
    a.dpreb = reinterpret_cast<__nv_bfloat16*>(dpre_bf16);
    a.T = t_len;
    a.B = batch;
    a.H = t->H;
    a.ldH = ld_h;
    a.CL = t->CL;
    a.SPC = t->SPC;
    a.NS = t->NS;
    a.HP = t->HP;
    a.max_blocks = std::max(t->max_cta_blocks, 1);
    a.max_groups = std::max(t->max_cta_groups, 1);
    a.max_tasks = std::max(t->max_cta_tasks, 1);
    a.act = act;
    const int items = tc_bwd_items(t->Hc);
    a.BW = a.NBX = 0;
    {
        if (int rc = tc_pick_box(t->Hc, &a.BW, &a.NBX)) return rc;
        const uint64_t B64 = (uint64_t)batch, T64 = (uint64_t)t_len, ld = (uint64_t)ld_h;
        if (int rc = make_tmap_3d(&a.tm_gates, 4, gates, 4 * ld, B64, T64, 16 * ld, B64 * 16 * ld, a.BW, 16)) return rc;
        if (int rc = make_tmap_3d(&a.tm_c, 4, c, ld, B64, T64, 4 * ld, B64 * 4 * ld, t->Hc, 16)) return rc;
        if (int rc = make_tmap_3d(&a.tm_dh, 4, dh_out, ld, B64, T64, 4 * ld, B64 * 4 * ld, t->Hc, 16)) return rc;
    }
    TcBwdKernel k = tc_pick_bwd(NB / 16, items, act);
    HCGS_REQUIRE(k != nullptr, "hcgs_lstm_tc_bwd: no kernel for nb=%d items=%d", NB, items);
    const int n_clusters = (batch + NB - 1) / NB;
    return tc_launch(k, a, n_clusters, t->CL, TCB_CW * 32, tc_bwd_smem_v(*t, NB),
                     (cudaStream_t)stream);
}

// tuning aid (HCGS_TC_DEBUG bit 128): clock stamps of the warps of CTA 0 at the middle time step, 20 x 16 values
extern "C" int hcgs_tc_timeline_read(int64_t* out320) {
    HCGS_REQUIRE(out320, "hcgs_tc_timeline_read: NULL argument");
    long long h[320];
    HCGS_CHECK_CUDA(cudaMemcpyFromSymbol(h, hcgs::g_tc_tl, sizeof(h)));
    for (int i = 0; i < 320; ++i) out320[i] = (int64_t)h[i];
    return 0;
}

// tuning aid: per-phase clock totals of CTA 0 / thread 0 of the last traced launch (HCGS_TC_DEBUG bit 64)
extern "C" int hcgs_tc_trace_read(int64_t* out16) {
    HCGS_REQUIRE(out16, "hcgs_tc_trace_read: NULL argument");
    long long h[16];
    HCGS_CHECK_CUDA(cudaMemcpyFromSymbol(h, hcgs::g_tc_trace, sizeof(h)));
    for (int i = 0; i < 16; ++i) out16[i] = (int64_t)h[i];
    int f[8];
    HCGS_CHECK_CUDA(cudaMemcpyFromSymbol(f, hcgs::g_tc_fail, sizeof(f)));
    if (f[0]) {  // a wait timed out in debug mode: {tag, block, warp, parity, step} packed into out16[15]
        out16[15] = -((int64_t)f[1] * 100000000 + (int64_t)f[2] * 1000000 + (int64_t)f[3] * 10000 + (int64_t)f[4] * 1000 + f[5]);
        int z[8] = {0};
        HCGS_CHECK_CUDA(cudaMemcpyToSymbol(hcgs::g_tc_fail, z, sizeof(z)));
    }
    return 0;
}
