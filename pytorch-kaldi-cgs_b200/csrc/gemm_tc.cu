// Block-sparse bf16 GEMM family on the 5th-generation tensor cores (tcgen05 + TMEM), fed by TMA.
//
//   nt   : Y[m, (r,u,g)]  = sum_k X[m,k] W_g[r*32+u, k]          time-batched input projections
//          (reference: the four F.linear calls neural_networks.py:951-954, MLP :306)
//   nn   : dX[m, k]       = sum_{r,u,g} dY[m,(r,u,g)] W_g[r*32+u,k]   data gradient
//   sddmm: dW_g[r*32+u,k] = mask * sum_m dY[m,(r,u,g)] X[m,k]      weight gradient, kept tiles only
//
// Sparsity: the plan's 32x32 tile index (HCGS level-1 blocks); a masked tile is never loaded,
// multiplied or written.  The NG gates of an LSTM layer share one mask, so one accumulator tile is
// 128 = 32 units x 4 gates wide with the gate index fastest ("gate-interleaved", the layout the
// recurrent kernels stream as one float4 per (sequence, unit) item).
//
// Operands: activations are bf16 row-major with a padded leading dimension and reach shared
// memory through 2-D TMA tensor maps (SWIZZLE_64B / SWIZZLE_128B boxes); weights are re-packed
// every step (they change every optimizer step) into pre-swizzled bf16 tiles that one
// cp.async.bulk brings in.  Accumulators live in TMEM; one elected thread issues tcgen05.mma;
// smem stages are recycled through tcgen05.commit -> mbarrier; the epilogue reads TMEM with
// tcgen05.ld.  Every mbarrier wait is bounded (trap, never hang).
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace hcgs {

// ---- PTX helpers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t g_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void g_mbar_init(uint32_t addr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}
__device__ __forceinline__ void g_mbar_expect_tx(uint32_t addr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t g_mbar_try_wait(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    return ok;
}
__device__ __forceinline__ void g_mbar_wait(uint32_t addr, uint32_t parity) {
    for (int spin = 0; spin < (1 << 24); ++spin)
        if (g_mbar_try_wait(addr, parity)) return;
    __trap();  // a pipeline bug must surface as a launch error, never as a hang
}

__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* map, int c0, int c1, uint32_t mbar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_dst), "l"(map), "r"(c0), "r"(c1), "r"(mbar)
        : "memory");
}
__device__ __forceinline__ void bulk_load_1d(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_dst),
                 "l"(gsrc), "r"(bytes), "r"(mbar)
                 : "memory");
}

// one lane of a CONVERGENT warp.  The MMA-issuing warps below run their loops on warp-uniform values (warp index and
// loop bounds passed through a shuffle) and let the elected lane issue: the tcgen05 operands then sit in uniform
// registers.  Issued from a `lane == 0` branch, every tcgen05.mma is wrapped in an ELECT / R2UR.BROADCAST loop.
__device__ __forceinline__ bool out_elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// whole warp; NCOLS power of two >= 32
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result_addr) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result_addr),
                 "n"(NCOLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem], bf16 operands, fp32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all previously issued MMAs of this thread arrive on the mbarrier when they have completed
__device__ __forceinline__ void umma_commit(uint32_t mbar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar)
                 : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives TMEM lane (base lane + i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- UMMA descriptors (cute/arch/mma_sm100_desc.hpp bit layout) --------------------------------
// shared-memory matrix descriptor: start address, leading / stride byte offsets (all >> 4),
// version 1 (Blackwell), swizzle mode in bits [61,64): 2 = 128B, 4 = 64B, 6 = 32B, 0 = none.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(layout_type & 7) << 61;
    return d;
}
constexpr uint32_t UMMA_SW128 = 2, UMMA_SW64 = 4;

// instruction descriptor, kind::f16: fp32 accumulate, bf16 A and B; a_mn / b_mn = 1 for MN-major
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- host: TMA tensor maps --------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = (EncodeTiledFn)p;
    return fn;
}

// bf16 row-major [rows, cols] with leading dimension ld (elements); box = box_cols x box_rows
static int make_tmap_bf16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_cols,
                          int box_rows, CUtensorMapSwizzle swz) {
    EncodeTiledFn fn = get_encode_fn();
    HCGS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
    HCGS_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0,
                 "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (ld=%lld)", (long long)ld);
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    HCGS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld", (int)r,
                 (long long)rows, (long long)cols, (long long)ld);
    return 0;
}

}  // namespace hcgs (reopened below)

// [d2][d1][d0] tensor (d0 contiguous) of 2-byte (bf16) or 4-byte (fp32) elements, no swizzle; used by the
// recurrent kernels for their per-step [1][16 sequences][slice] boxes.  `map` points to a CUtensorMap.
int hcgs::make_tmap_3d(void* map, int elem_bytes, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                       uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0, uint32_t box1) {
    EncodeTiledFn fn = get_encode_fn();
    HCGS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
    HCGS_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "make_tmap_3d: element size %d", elem_bytes);
    HCGS_REQUIRE(((uintptr_t)base & 15) == 0 && stride1_bytes % 16 == 0 && stride2_bytes % 16 == 0,
                 "TMA operand must be 16-byte aligned with 16-byte multiple pitches");
    HCGS_REQUIRE(box0 >= 1 && box0 <= 256 && box1 >= 1 && box1 <= 256 && (box0 * (uint32_t)elem_bytes) % 16 == 0,
                 "make_tmap_3d: bad box %u x %u", box0, box1);
    cuuint64_t gdim[3] = {d0, d1, d2};
    cuuint64_t gstr[2] = {stride1_bytes, stride2_bytes};
    cuuint32_t box[3] = {box0, box1, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn((CUtensorMap*)map, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                    3, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    HCGS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (3d) failed (%d)", (int)r);
    return 0;
}

namespace hcgs {

// ---- weight packing: masked dense fp32 -> pre-swizzled bf16 tiles ------------------------------
// nt tiles, CSR order (tile t = (r, c) from the plan's flat list): [32*NG rows n = u*NG+g][32 k],
// 64-byte rows in the SWIZZLE_64B pattern (16-byte chunk index ^= (row >> 1) & 3).
// split = 1 (two-term operands, hcgs_tc_pack_nt_split): tile t becomes the pair [hi tile | lo tile] with
// hi = bf16(w), lo = bf16(w - hi): hi + lo carries 16 mantissa bits of the fp32 master.
template <int NG>
__global__ void tc_pack_nt_kernel(const int* __restrict__ tile_rc, int nz, PtrPack<const float> W, int ldw,
                                  int n_out, int n_in, __nv_bfloat16* __restrict__ out, int split) {
    const int64_t total = (int64_t)nz * 32 * NG * 4;  // 16-byte chunks
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int c16 = idx & 3;
        const int n = (int)((idx >> 2) % (32 * NG));
        const int t = (int)(idx / (4 * 32 * NG));
        const int r = tile_rc[2 * t], c = tile_rc[2 * t + 1];
        const int u = n / NG, g = n % NG;
        const int row = r * 32 + u, col0 = c * 32 + c16 * 8;
        uint32_t w[4], w2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo = 0.f, hi = 0.f;
            if (row < n_out) {
                if (col0 + 2 * j < n_in) lo = __ldg(W.p[g] + (int64_t)row * ldw + col0 + 2 * j);
                if (col0 + 2 * j + 1 < n_in) hi = __ldg(W.p[g] + (int64_t)row * ldw + col0 + 2 * j + 1);
            }
            const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
            w[j] = *reinterpret_cast<const uint32_t*>(&v);
            const __nv_bfloat162 r2 = __floats2bfloat162_rn(lo - __low2float(v), hi - __high2float(v));
            w2[j] = *reinterpret_cast<const uint32_t*>(&r2);
        }
        const size_t tile_bytes = 32 * NG * 64;
        unsigned char* dst = reinterpret_cast<unsigned char*>(out) + (size_t)t * tile_bytes * (split ? 2 : 1) + n * 64 +
                             ((c16 ^ ((n >> 1) & 3)) << 4);
        *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
        if (split) *reinterpret_cast<uint4*>(dst + tile_bytes) = make_uint4(w2[0], w2[1], w2[2], w2[3]);
    }
}

// ---- nt: Y = X W^T ------------------------------------------------------------------------------
struct TcNtArgs {
    const __nv_bfloat16* wpack;  // nt tiles, CSR order
    const int* row_ptr;          // [tr+1]
    const int* col_idx;          // [nz]
    float* y;                    // [M, ldy] fp32; column of (r, u, g) = (r*32+u)*NG + g
    int64_t ldy;
    int64_t M;
    int n_out;                   // output units (rows of one W_g)
    PtrPack<const float> bias;   // per gate [n_out] or NULL
    float* stats;                // optional [m-tiles][ldy][2]: per 128-row tile and column {mean, M2} (batch norm)
    int split;                   // 1: two-term operands, three products per k tile (x_hi w_hi + x_lo w_hi + x_hi w_lo)
};

constexpr int NT_STAGES = 4;
constexpr int NT_A_BYTES = 128 * 64;  // 128 rows x 32 bf16, SWIZZLE_64B
constexpr int NT_EST = 132;           // floats per staged output row (NG = 4): 16-byte aligned, conflict free

template <int NG>
__host__ __device__ constexpr int nt_smem_bytes() {
    // NG = 4: the drained pipeline stages are reused as the [128][NT_EST] fp32 staging tile of the epilogue
    return 1024 + (NG == 4 ? (NT_STAGES * (NT_A_BYTES + 32 * NG * 64) > 128 * NT_EST * 4
                                  ? NT_STAGES * (NT_A_BYTES + 32 * NG * 64) : 128 * NT_EST * 4) + 128
                           : NT_STAGES * (NT_A_BYTES + 32 * NG * 64)) + 256;
}

template <int NG>
__global__ void __launch_bounds__(128) tc_gemm_nt_kernel(const __grid_constant__ CUtensorMap tmX,
                                                         const __grid_constant__ CUtensorMap tmXlo, const TcNtArgs a) {
    constexpr int BN = 32 * NG;
    constexpr int B_BYTES = BN * 64;
    constexpr int TCOLS = BN < 32 ? 32 : BN;
    constexpr uint32_t IDESC = umma_idesc_bf16(128, BN, 0, 0);
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    unsigned char* sA = smem;
    unsigned char* sB = smem + NT_STAGES * NT_A_BYTES;
    constexpr int STAGE_BYTES = NT_STAGES * (NT_A_BYTES + B_BYTES);
    constexpr int BAR_OFF = (NG == 4 && 128 * NT_EST * 4 > STAGE_BYTES) ? ((128 * NT_EST * 4 + 127) & ~127) : STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * NT_STAGES + 1);
    const uint32_t full0 = g_smem_u32(bars), empty0 = g_smem_u32(bars + NT_STAGES),
                   done = g_smem_u32(bars + 2 * NT_STAGES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);  // provably warp-uniform copy (see out_elect_one)
    const int r = blockIdx.x;
    const int64_t m0 = (int64_t)blockIdx.y * 128;
    const int k0 = a.row_ptr[r];
    // pipeline items: one per kept k tile, or three in split mode (the lo x lo term, 2^-16 relative, is dropped)
    const int nk = (a.row_ptr[r + 1] - k0) * (a.split ? 3 : 1);

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < NT_STAGES; ++s) {
            g_mbar_init(full0 + 8 * s, 1);
            g_mbar_init(empty0 + 8 * s, 1);
        }
        g_mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc<TCOLS>(g_smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ---- TMA producer ----------------------------------------------------------------------
        for (int i = 0; i < nk; ++i) {
            const int s = i % NT_STAGES;
            if (i >= NT_STAGES) g_mbar_wait(empty0 + 8 * s, (uint32_t)(((i / NT_STAGES) - 1) & 1));
            g_mbar_expect_tx(full0 + 8 * s, NT_A_BYTES + B_BYTES);
            int kt = i, piece = 0;
            if (a.split) {
                kt = i / 3;
                piece = i - 3 * kt;
            }
            tma_load_2d(g_smem_u32(sA + s * NT_A_BYTES), piece == 1 ? &tmXlo : &tmX, a.col_idx[k0 + kt] * 32, (int)m0,
                        full0 + 8 * s);
            const size_t wtile = a.split ? (size_t)(k0 + kt) * 2 + (piece == 2 ? 1 : 0) : (size_t)(k0 + kt);
            bulk_load_1d(g_smem_u32(sB + s * B_BYTES), reinterpret_cast<const unsigned char*>(a.wpack) + wtile * B_BYTES,
                         B_BYTES, full0 + 8 * s);
        }
    } else if (warp_u == 1) {
        // ---- MMA issuer (convergent warp, elected lane) -----------------------------------------
        const int nk_u = __shfl_sync(0xffffffffu, nk, 0);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        for (int i = 0; i < nk_u; ++i) {
            const int s = i % NT_STAGES;
            g_mbar_wait(full0 + 8 * s, (uint32_t)((i / NT_STAGES) & 1));
            tc_fence_after();
            const uint32_t aaddr = g_smem_u32(sA + s * NT_A_BYTES), baddr = g_smem_u32(sB + s * B_BYTES);
            if (out_elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {  // 32 k per stage = 2 x (K = 16); +32 bytes inside the swizzle atom
                    const uint64_t ad = umma_smem_desc(aaddr + kk * 32, 16, 512, UMMA_SW64);
                    const uint64_t bd = umma_smem_desc(baddr + kk * 32, 16, 512, UMMA_SW64);
                    umma_bf16(tmem_u, ad, bd, IDESC, (i > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(empty0 + 8 * s);  // stage reusable once these MMAs have read it
            }
            __syncwarp();
        }
        if (out_elect_one()) umma_commit(done);
        __syncwarp();
    }
    // ---- epilogue: TMEM -> registers -> global ------------------------------------------------------
    __syncwarp();
    if (nk > 0) {
        g_mbar_wait(done, 0);
        tc_fence_after();
    }
    if constexpr (NG == 4) {
        // ---- staged epilogue: TMEM -> shared tile -> (column statistics) -> coalesced 512-byte row stores ----
        float* st = reinterpret_cast<float*>(smem);
        const int rows = (int)((a.M - m0 < 128) ? (a.M - m0) : 128);
        const int row = warp * 32 + lane;
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
            float v[32];
            if (nk > 0) {
                tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0.f;
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int n = cb * 32 + i, u = r * 32 + (n >> 2), g = n & 3;
                if (u >= a.n_out) v[i] = 0.f;  // pad units stay zero
                else if (a.bias.p[g]) v[i] += __ldg(a.bias.p[g] + u);
            }
            float4* d4 = reinterpret_cast<float4*>(st + (size_t)row * NT_EST + cb * 32);
#pragma unroll
            for (int i = 0; i < 8; ++i) d4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        __syncthreads();
        const int col_t = r * 128 + (int)threadIdx.x;  // this thread's output column for the statistics
        if (a.stats && col_t < a.ldy) {
            // fixed summation order (four interleaved chains) => deterministic
            const float* colp = st + threadIdx.x;
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
            int i = 0;
            for (; i + 4 <= rows; i += 4) {
#pragma unroll
                for (int q = 0; q < 4; ++q) s4[q] += colp[(size_t)(i + q) * NT_EST];
            }
            for (; i < rows; ++i) s4[0] += colp[(size_t)i * NT_EST];
            const float mean = ((s4[0] + s4[1]) + (s4[2] + s4[3])) / (float)rows;
            float q4[4] = {0.f, 0.f, 0.f, 0.f};
            for (i = 0; i + 4 <= rows; i += 4) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float d = colp[(size_t)(i + q) * NT_EST] - mean;
                    q4[q] = fmaf(d, d, q4[q]);
                }
            }
            for (; i < rows; ++i) {
                const float d = colp[(size_t)i * NT_EST] - mean;
                q4[0] = fmaf(d, d, q4[0]);
            }
            const float m2 = (q4[0] + q4[1]) + (q4[2] + q4[3]);
            *reinterpret_cast<float2*>(a.stats + ((size_t)blockIdx.y * a.ldy + col_t) * 2) = make_float2(mean, m2);
        }
        const int c4 = r * 128 + lane * 4;  // first of this lane's four output columns
        const bool vec = ((a.ldy & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.y) & 15) == 0);
        for (int i = warp; i < rows; i += 4) {
            const float4 v = *reinterpret_cast<const float4*>(st + (size_t)i * NT_EST + lane * 4);
            float* dst = a.y + (m0 + i) * a.ldy + c4;
            if (vec && c4 + 4 <= a.ldy) {
                __stcs(reinterpret_cast<float4*>(dst), v);
            } else {
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (c4 + q < a.ldy) dst[q] = e[q];
            }
        }
        tc_fence_before();
        __syncthreads();
        if (warp == 1) tmem_dealloc<TCOLS>(tmem);
        return;
    }
    const int64_t m = m0 + warp * 32 + lane;
#pragma unroll 1
    for (int cb = 0; cb < BN / 32; ++cb) {
        float v[32];
        if (nk > 0) {
            tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
        if (m < a.M) {
            const int col0 = r * 32 * NG + cb * 32;  // output column of v[0]
            float* dst = a.y + m * a.ldy + col0;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int n = cb * 32 + i, u = r * 32 + n / NG, g = n % NG;
                if (u < a.n_out) {
                    float o = v[i];
                    if (a.bias.p[g]) o += __ldg(a.bias.p[g] + u);
                    v[i] = o;
                }
            }
            const bool vec = ((a.ldy & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.y) & 15) == 0) &&
                             (r * 32 + 32 <= a.n_out);
            if (vec) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    *reinterpret_cast<float4*>(dst + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int u = r * 32 + (cb * 32 + i) / NG;
                    if (u < a.n_out) dst[i] = v[i];
                    else if (col0 + i < a.ldy) dst[i] = 0.f;  // pad columns of a padded output stay zero
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TCOLS>(tmem);
}

}  // namespace hcgs

using namespace hcgs;

static int tc_pack_nt_impl(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack, int split,
                           hcgs_stream stream) {
    HCGS_REQUIRE(plan && w && wpack, "hcgs_tc_pack_nt: NULL argument");
    HCGS_REQUIRE(n_w == 1 || n_w == 4, "hcgs_tc_pack_nt: n_w must be 1 or 4");
    if (plan->nz_tiles == 0) return 0;
    const int64_t total = (int64_t)plan->nz_tiles * 32 * n_w * 4;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_w == 4)
        tc_pack_nt_kernel<4><<<blocks, 256, 0, st>>>(plan->d_tile_rc, plan->nz_tiles, make_pack<const float>(w, 4), ldw,
                                                     plan->n_out, plan->n_in, (__nv_bfloat16*)wpack, split);
    else
        tc_pack_nt_kernel<1><<<blocks, 256, 0, st>>>(plan->d_tile_rc, plan->nz_tiles, make_pack<const float>(w, 1), ldw,
                                                     plan->n_out, plan->n_in, (__nv_bfloat16*)wpack, split);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_tc_pack_nt(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack,
                               hcgs_stream stream) {
    return tc_pack_nt_impl(plan, n_w, w, ldw, wpack, 0, stream);
}

extern "C" int hcgs_tc_pack_nt_split(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack2,
                                     hcgs_stream stream) {
    return tc_pack_nt_impl(plan, n_w, w, ldw, wpack2, 1, stream);
}

static int tc_gemm_nt_impl(const hcgs_plan* plan, int n_w, const void* wpack, const void* x_bf16, int64_t ldx,
                           const void* x_lo, int64_t ldx_lo, int64_t m, const float* const* bias, float* y,
                           int64_t ldy, float* stats, hcgs_stream stream) {
    HCGS_REQUIRE(plan && wpack && x_bf16 && y, "hcgs_tc_gemm_nt: NULL argument");
    HCGS_REQUIRE(stats == nullptr || n_w == 4, "hcgs_tc_gemm_nt: fused column statistics need n_w = 4");
    HCGS_REQUIRE(n_w == 1 || n_w == 4, "hcgs_tc_gemm_nt: n_w must be 1 or 4");
    if (m <= 0) return 0;
    HCGS_REQUIRE(ldy >= (int64_t)plan->tr * 32 * n_w || ldy >= (int64_t)plan->n_out * n_w,
                 "hcgs_tc_gemm_nt: ldy too small");
    CUtensorMap tm, tml;
    if (int rc = make_tmap_bf16(&tm, x_bf16, m, plan->n_in, ldx, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    if (x_lo) {
        if (int rc = make_tmap_bf16(&tml, x_lo, m, plan->n_in, ldx_lo, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    } else {
        tml = tm;
    }
    TcNtArgs a;
    a.split = x_lo ? 1 : 0;
    a.wpack = (const __nv_bfloat16*)wpack;
    a.row_ptr = plan->d_row_ptr;
    a.col_idx = plan->d_col_idx;
    a.y = y;
    a.ldy = ldy;
    a.M = m;
    a.n_out = plan->n_out;
    a.bias = make_pack<const float>(bias, n_w);
    a.stats = stats;
    const int64_t mt = (m + 127) / 128;
    HCGS_REQUIRE(mt <= 65535, "hcgs_tc_gemm_nt: too many rows (%lld)", (long long)m);
    dim3 grid(plan->tr, (unsigned)mt, 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_w == 4) {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_nt_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             nt_smem_bytes<4>()));
        tc_gemm_nt_kernel<4><<<grid, 128, nt_smem_bytes<4>(), st>>>(tm, tml, a);
    } else {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_nt_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             nt_smem_bytes<1>()));
        tc_gemm_nt_kernel<1><<<grid, 128, nt_smem_bytes<1>(), st>>>(tm, tml, a);
    }
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_tc_gemm_nt(const hcgs_plan* plan, int n_w, const void* wpack, const void* x_bf16, int64_t ldx,
                               int64_t m, const float* const* bias, float* y, int64_t ldy, float* stats,
                               hcgs_stream stream) {
    return tc_gemm_nt_impl(plan, n_w, wpack, x_bf16, ldx, nullptr, 0, m, bias, y, ldy, stats, stream);
}

extern "C" int hcgs_tc_gemm_nt_split(const hcgs_plan* plan, int n_w, const void* wpack2, const void* x_hi,
                                     int64_t ldx_hi, const void* x_lo, int64_t ldx_lo, int64_t m,
                                     const float* const* bias, float* y, int64_t ldy, float* stats,
                                     hcgs_stream stream) {
    HCGS_REQUIRE(x_lo != nullptr, "hcgs_tc_gemm_nt_split: x_lo is NULL");
    return tc_gemm_nt_impl(plan, n_w, wpack2, x_hi, ldx_hi, x_lo, ldx_lo, m, bias, y, ldy, stats, stream);
}

// fp32 [rows, cols] (ld = ldx) -> bf16 [rows, ldo] with zero pad columns (cols <= ldo)
namespace hcgs {
__global__ void cast_bf16_kernel(const float* __restrict__ x, int64_t ldx, int64_t rows, int cols,
                                 __nv_bfloat16* __restrict__ out, int64_t ldo) {
    const int64_t total = rows * (ldo / 2);
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = idx / (ldo / 2);
        const int c = (int)(idx - r * (ldo / 2)) * 2;
        const float lo = c < cols ? __ldg(x + r * ldx + c) : 0.f;
        const float hi = c + 1 < cols ? __ldg(x + r * ldx + c + 1) : 0.f;
        *reinterpret_cast<__nv_bfloat162*>(out + r * ldo + c) = __floats2bfloat162_rn(lo, hi);
    }
}
}  // namespace hcgs

namespace hcgs {
// two-term split of fp32 rows: hi = bf16(x) (optional output), lo = bf16(x - hi); pad columns zero
__global__ void split_bf16_kernel(const float* __restrict__ x, int64_t ldx, int64_t rows, int cols,
                                  __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo, int64_t ldo) {
    const int64_t total = rows * (ldo / 2);
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = idx / (ldo / 2);
        const int c = (int)(idx - r * (ldo / 2)) * 2;
        const float a = c < cols ? __ldg(x + r * ldx + c) : 0.f;
        const float b = c + 1 < cols ? __ldg(x + r * ldx + c + 1) : 0.f;
        const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        if (hi) *reinterpret_cast<__nv_bfloat162*>(hi + r * ldo + c) = h;
        *reinterpret_cast<__nv_bfloat162*>(lo + r * ldo + c) =
            __floats2bfloat162_rn(a - __low2float(h), b - __high2float(h));
    }
}
}  // namespace hcgs

extern "C" int hcgs_split_bf16(const float* x, int64_t ldx, int64_t rows, int cols, void* hi, void* lo, int64_t ldo,
                               hcgs_stream stream) {
    HCGS_REQUIRE(x && lo, "hcgs_split_bf16: NULL argument");
    HCGS_REQUIRE(ldo >= cols && (ldo & 1) == 0, "hcgs_split_bf16: ldo must be even and >= cols");
    if (rows <= 0) return 0;
    const int64_t total = rows * (ldo / 2);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    split_bf16_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, rows, cols, (__nv_bfloat16*)hi,
                                                                (__nv_bfloat16*)lo, ldo);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_cast_bf16(const float* x, int64_t ldx, int64_t rows, int cols, void* out, int64_t ldo,
                              hcgs_stream stream) {
    HCGS_REQUIRE(x && out, "hcgs_cast_bf16: NULL argument");
    HCGS_REQUIRE(ldo >= cols && (ldo & 1) == 0, "hcgs_cast_bf16: ldo must be even and >= cols");
    if (rows <= 0) return 0;
    const int64_t total = rows * (ldo / 2);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    cast_bf16_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, rows, cols, (__nv_bfloat16*)out, ldo);
    HCGS_LAUNCH_CHECK();
    return 0;
}

namespace hcgs {

// ---- nn: dX = dY W ------------------------------------------------------------------------------
// Work item = (column block c of W, 128-row tile of dY).  The reduction runs over the kept tile
// rows r of that column (CSC) and the 32*NG (unit, gate) columns of dY that belong to r.
//   A = dY box [128 rows x KCH k] K-major (TMA), B = W^T tile [32 kcols x KCH k] K-major (pre-swizzled)
//   NG = 4: KCH = 64 (SWIZZLE_128B), two sub-tiles per r;  NG = 1: KCH = 32 (SWIZZLE_64B).
template <int NG>
struct NnCfg {
    static constexpr int KCH = NG == 4 ? 64 : 32;
    static constexpr int NSUB = NG == 4 ? 2 : 1;
    static constexpr int A_SUB = 128 * KCH * 2;
    static constexpr int B_SUB = 32 * KCH * 2;
    static constexpr int STAGE = A_SUB + B_SUB;     // one k sub-tile per stage: 20 KB (NG = 4) / 10 KB (NG = 1)
    static constexpr int STAGES = NG == 4 ? 2 : 4;  // ~41 KB per CTA: five CTAs per SM (the CTAs are short and latency
                                                    // bound; with both sub-tiles in one 40 KB stage only two fitted)
    static constexpr uint32_t SW = NG == 4 ? UMMA_SW128 : UMMA_SW64;
    static constexpr uint32_t SBO = 8 * KCH * 2;
    static constexpr int SMEM = 1024 + STAGES * STAGE + 256;
};

template <int NG>
__global__ void tc_pack_nn_kernel(const int* __restrict__ csc_rc, int nz, PtrPack<const float> W, int ldw, int n_out,
                                  int n_in, __nv_bfloat16* __restrict__ out) {
    using C = NnCfg<NG>;
    constexpr int CPR = C::KCH / 8;  // 16-byte chunks per row
    const int64_t total = (int64_t)nz * C::NSUB * 32 * CPR;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int c16 = (int)(idx % CPR);
        const int kc = (int)((idx / CPR) % 32);
        const int h = (int)((idx / (CPR * 32)) % C::NSUB);
        const int e = (int)(idx / (CPR * 32 * C::NSUB));
        const int r = csc_rc[2 * e], c = csc_rc[2 * e + 1];
        const int col = c * 32 + kc;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = h * C::KCH + c16 * 8 + j;
            const int u = n / NG, g = n % NG;
            const int row = r * 32 + u;
            v[j] = (row < n_out && col < n_in) ? __ldg(W.p[g] + (int64_t)row * ldw + col) : 0.f;
        }
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 p = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            w[j] = *reinterpret_cast<const uint32_t*>(&p);
        }
        const int sw = (NG == 4) ? (kc & 7) : ((kc >> 1) & 3);
        unsigned char* dst = reinterpret_cast<unsigned char*>(out) + (size_t)e * (C::NSUB * C::B_SUB) + h * C::B_SUB +
                             kc * (C::KCH * 2) + ((c16 ^ sw) << 4);
        *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

struct TcNnArgs {
    const __nv_bfloat16* wpack;  // nn tiles, CSC order
    const int* col_ptr;          // [tc+1]
    const int* row_idx;          // [nz]
    float* dx;                   // [M, lddx]
    int64_t lddx;
    int64_t M;
    int n_in;
};

template <int NG>
__global__ void __launch_bounds__(128) tc_gemm_nn_kernel(const __grid_constant__ CUtensorMap tmDY, const TcNnArgs a) {
    using C = NnCfg<NG>;
    constexpr uint32_t IDESC = umma_idesc_bf16(128, 32, 0, 0);
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::STAGES + 1);
    const uint32_t full0 = g_smem_u32(bars), empty0 = g_smem_u32(bars + C::STAGES),
                   done = g_smem_u32(bars + 2 * C::STAGES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);  // provably warp-uniform copy (see out_elect_one)
    const int c = blockIdx.x;
    const int64_t m0 = (int64_t)blockIdx.y * 128;
    const int k0 = a.col_ptr[c], nk = a.col_ptr[c + 1] - k0;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            g_mbar_init(full0 + 8 * s, 1);
            g_mbar_init(empty0 + 8 * s, 1);
        }
        g_mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc<32>(g_smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < nk * C::NSUB; ++i) {
            const int s = i % C::STAGES;
            if (i >= C::STAGES) g_mbar_wait(empty0 + 8 * s, (uint32_t)(((i / C::STAGES) - 1) & 1));
            g_mbar_expect_tx(full0 + 8 * s, C::STAGE);
            const int e = k0 + i / C::NSUB, h = i % C::NSUB;
            const int r = a.row_idx[e];
            unsigned char* st = smem + s * C::STAGE;
            tma_load_2d(g_smem_u32(st), &tmDY, r * 32 * NG + h * C::KCH, (int)m0, full0 + 8 * s);
            bulk_load_1d(g_smem_u32(st + C::A_SUB),
                         reinterpret_cast<const unsigned char*>(a.wpack) + ((size_t)e * C::NSUB + h) * C::B_SUB, C::B_SUB,
                         full0 + 8 * s);
        }
    } else if (warp_u == 1) {
        const int nk_u = __shfl_sync(0xffffffffu, nk, 0) * C::NSUB;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        for (int i = 0; i < nk_u; ++i) {
            const int s = i % C::STAGES;
            g_mbar_wait(full0 + 8 * s, (uint32_t)((i / C::STAGES) & 1));
            tc_fence_after();
            const uint32_t base = g_smem_u32(smem + s * C::STAGE);
            if (out_elect_one()) {
#pragma unroll
                for (int kk = 0; kk < C::KCH / 16; ++kk) {
                    const uint64_t ad = umma_smem_desc(base + kk * 32, 16, C::SBO, C::SW);
                    const uint64_t bd = umma_smem_desc(base + C::A_SUB + kk * 32, 16, C::SBO, C::SW);
                    umma_bf16(tmem_u, ad, bd, IDESC, (i > 0 || kk > 0) ? 1u : 0u);
                }
                umma_commit(empty0 + 8 * s);
            }
            __syncwarp();
        }
        if (out_elect_one()) umma_commit(done);
        __syncwarp();
    }
    __syncwarp();
    float v[32];
    if (nk > 0) {
        g_mbar_wait(done, 0);
        tc_fence_after();
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
    } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
    }
    const int64_t m = m0 + warp * 32 + lane;
    if (m < a.M) {
        float* dst = a.dx + m * a.lddx + c * 32;
        const bool vec = ((a.lddx & 3) == 0) && ((reinterpret_cast<uintptr_t>(a.dx) & 15) == 0) && (c * 32 + 32 <= a.n_in);
        if (vec) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                __stcs(reinterpret_cast<float4*>(dst + 4 * i), make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (c * 32 + i < a.n_in) dst[i] = v[i];
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<32>(tmem);
}

}  // namespace hcgs

extern "C" int hcgs_tc_pack_nn(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, void* wpack,
                               hcgs_stream stream) {
    HCGS_REQUIRE(plan && w && wpack, "hcgs_tc_pack_nn: NULL argument");
    HCGS_REQUIRE(n_w == 1 || n_w == 4, "hcgs_tc_pack_nn: n_w must be 1 or 4");
    if (plan->nz_tiles == 0) return 0;
    const int64_t total = (int64_t)plan->nz_tiles * 32 * n_w * 4;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_w == 4)
        tc_pack_nn_kernel<4><<<blocks, 256, 0, st>>>(plan->d_csc_rc, plan->nz_tiles, make_pack<const float>(w, 4), ldw,
                                                     plan->n_out, plan->n_in, (__nv_bfloat16*)wpack);
    else
        tc_pack_nn_kernel<1><<<blocks, 256, 0, st>>>(plan->d_csc_rc, plan->nz_tiles, make_pack<const float>(w, 1), ldw,
                                                     plan->n_out, plan->n_in, (__nv_bfloat16*)wpack);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_tc_gemm_nn(const hcgs_plan* plan, int n_w, const void* wpack, const void* dy_bf16, int64_t lddy,
                               int64_t m, float* dx, int64_t lddx, hcgs_stream stream) {
    HCGS_REQUIRE(plan && wpack && dy_bf16 && dx, "hcgs_tc_gemm_nn: NULL argument");
    HCGS_REQUIRE(n_w == 1 || n_w == 4, "hcgs_tc_gemm_nn: n_w must be 1 or 4");
    if (m <= 0) return 0;
    CUtensorMap tm;
    const int kch = n_w == 4 ? 64 : 32;
    // global width = lddy: pad columns of dY are zero by contract, tiles past n_out*n_w read zeros / OOB zero fill
    if (int rc = make_tmap_bf16(&tm, dy_bf16, m, lddy, lddy, kch, 128,
                                n_w == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B))
        return rc;
    TcNnArgs a;
    a.wpack = (const __nv_bfloat16*)wpack;
    a.col_ptr = plan->d_col_ptr;
    a.row_idx = plan->d_row_idx;
    a.dx = dx;
    a.lddx = lddx;
    a.M = m;
    a.n_in = plan->n_in;
    const int64_t mt = (m + 127) / 128;
    HCGS_REQUIRE(mt <= 65535, "hcgs_tc_gemm_nn: too many rows (%lld)", (long long)m);
    dim3 grid(plan->tc, (unsigned)mt, 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_w == 4) {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_nn_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             NnCfg<4>::SMEM));
        tc_gemm_nn_kernel<4><<<grid, 128, NnCfg<4>::SMEM, st>>>(tm, a);
    } else {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_nn_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             NnCfg<1>::SMEM));
        tc_gemm_nn_kernel<1><<<grid, 128, NnCfg<1>::SMEM, st>>>(tm, a);
    }
    HCGS_LAUNCH_CHECK();
    return 0;
}

namespace hcgs {

// ---- sddmm: dW = mask .* (dY^T X) on the kept tiles ------------------------------------------------
// Work item = (tile row r, group of <= 4 kept tile columns, split of the m rows).  Both operands are
// MN-major (the reduction index m is the slow dimension of the row-major activations):
//   A = X^T  : M = 128 = 4 kept column blocks x 32, each block one TMA box [32 cols x RK rows] (SWIZZLE_64B atom);
//              a work item takes up to kSdMaxTiles kept blocks of its tile row = up to 3 MMAs per k-step that
//              share ONE dY^T operand (the kernel is bound by its operand traffic: dY is then read once per tile row)
//   B = dY^T : N = 32*NG (unit, gate) columns of tile row r; NG = 4: two boxes [64 x RK] (SWIZZLE_128B atoms)
//   D[kc, (u,g)] = dW_g[r*32+u, c*32+kc]  (transposed tile) -> fp32 atomics, re-masked on the way out.
constexpr int SD_RK = 64;       // m rows per stage
constexpr int SD_STAGES = 3;
constexpr int SD_A_ATOM = SD_RK * 64;
constexpr int SD_MM = kSdMaxTiles / 4;  // MMAs (accumulators) per k-step

template <int NG>
struct SdCfg {
    static constexpr int BN = 32 * NG;
    static constexpr int B_BYTES = SD_RK * BN * 2;
    static constexpr int STAGE = kSdMaxTiles * SD_A_ATOM + B_BYTES;
    static constexpr int SMEM = 1024 + SD_STAGES * STAGE + 256;
    static constexpr int TCOLS = SD_MM * BN <= 128 ? 128 : (SD_MM * BN <= 256 ? 256 : 512);
};

struct TcSdArgs {
    const int* items;    // [n_items][3] (tile row, first CSR tile, count)
    const int* col_idx;  // [nz]
    PtrPack<float> dw;   // per gate [n_out, lddw] fp32, accumulated atomically
    int64_t lddw;
    const float* mask;   // [n_out, n_in] or NULL
    int64_t M, rows_per_split;
    int n_out, n_in;
};

template <int NG>
__global__ void __launch_bounds__(128) tc_sddmm_kernel(const __grid_constant__ CUtensorMap tmX,
                                                       const __grid_constant__ CUtensorMap tmDY, const TcSdArgs a) {
    using C = SdCfg<NG>;
    constexpr uint32_t IDESC = umma_idesc_bf16(128, C::BN, 1, 1);
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SD_STAGES * C::STAGE);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * SD_STAGES + 1);
    const uint32_t full0 = g_smem_u32(bars), empty0 = g_smem_u32(bars + SD_STAGES),
                   done = g_smem_u32(bars + 2 * SD_STAGES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);  // provably warp-uniform copy (see out_elect_one)
    const int r = a.items[3 * blockIdx.x], t0 = a.items[3 * blockIdx.x + 1], cnt = a.items[3 * blockIdx.x + 2];
    const int64_t mbeg = (int64_t)blockIdx.y * a.rows_per_split;
    const int64_t mend = (mbeg + a.rows_per_split < a.M) ? mbeg + a.rows_per_split : a.M;
    const int nk = mend > mbeg ? (int)((mend - mbeg + SD_RK - 1) / SD_RK) : 0;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < SD_STAGES; ++s) {
            g_mbar_init(full0 + 8 * s, 1);
            g_mbar_init(empty0 + 8 * s, 1);
        }
        g_mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc<C::TCOLS>(g_smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // producer warp: lane j issues the box of kept block j, lanes 12 / 13 the dY boxes (a TMA issue costs its
        // thread a few hundred cycles, one thread issuing 6..14 boxes per stage was the bottleneck of the kernel)
        for (int i = 0; i < nk; ++i) {
            const int s = i % SD_STAGES;
            if (i >= SD_STAGES) g_mbar_wait(empty0 + 8 * s, (uint32_t)(((i / SD_STAGES) - 1) & 1));
            if (lane == 0) g_mbar_expect_tx(full0 + 8 * s, cnt * SD_A_ATOM + C::B_BYTES);
            __syncwarp();
            unsigned char* st = smem + s * C::STAGE;
            const int row = (int)(mbeg + (int64_t)i * SD_RK);
            // splits are multiples of SD_RK rows (host), so only the global tail is ragged: TMA zero-fills it
            if (lane < cnt)
                tma_load_2d(g_smem_u32(st + lane * SD_A_ATOM), &tmX, a.col_idx[t0 + lane] * 32, row, full0 + 8 * s);
            if (NG == 4) {
                if (lane == 12) tma_load_2d(g_smem_u32(st + kSdMaxTiles * SD_A_ATOM), &tmDY, r * 128, row, full0 + 8 * s);
                if (lane == 13)
                    tma_load_2d(g_smem_u32(st + kSdMaxTiles * SD_A_ATOM + SD_RK * 128), &tmDY, r * 128 + 64, row,
                                full0 + 8 * s);
            } else {
                if (lane == 12) tma_load_2d(g_smem_u32(st + kSdMaxTiles * SD_A_ATOM), &tmDY, r * 32, row, full0 + 8 * s);
            }
        }
    } else if (warp_u == 1) {
        const int nk_u = __shfl_sync(0xffffffffu, nk, 0);
        const int nmm = (__shfl_sync(0xffffffffu, cnt, 0) + 3) / 4;  // groups of four kept blocks; a partly filled group
                                            // multiplies stale shared memory into accumulator rows that are never read
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        for (int i = 0; i < nk_u; ++i) {
            const int s = i % SD_STAGES;
            g_mbar_wait(full0 + 8 * s, (uint32_t)((i / SD_STAGES) & 1));
            tc_fence_after();
            const uint32_t abase = g_smem_u32(smem + s * C::STAGE), bbase = abase + kSdMaxTiles * SD_A_ATOM;
            if (out_elect_one()) {
#pragma unroll
                for (int kk = 0; kk < SD_RK / 16; ++kk) {  // 16 m rows per MMA = two 8-row groups
                    const uint64_t bd = (NG == 4) ? umma_smem_desc(bbase + kk * 2048, SD_RK * 128, 1024, UMMA_SW128)
                                                  : umma_smem_desc(bbase + kk * 1024, SD_A_ATOM, 512, UMMA_SW64);
                    for (int j = 0; j < nmm; ++j) {
                        const uint64_t ad =
                            umma_smem_desc(abase + j * 4 * SD_A_ATOM + kk * 1024, SD_A_ATOM, 512, UMMA_SW64);
                        umma_bf16(tmem_u + (uint32_t)(j * C::BN), ad, bd, IDESC, (i > 0 || kk > 0) ? 1u : 0u);
                    }
                }
                umma_commit(empty0 + 8 * s);
            }
            __syncwarp();
        }
        if (out_elect_one()) umma_commit(done);
        __syncwarp();
    }
    __syncwarp();
    if (nk > 0 && warp < cnt) {
        g_mbar_wait(done, 0);
        tc_fence_after();
    }
    // warp w owns the kept column blocks w, w+4, w+8 of the item (TMEM lanes 32w..32w+31 of accumulator j/4)
#pragma unroll 1
    for (int jt = warp; nk > 0 && jt < cnt; jt += 4) {
        const int col = a.col_idx[t0 + jt] * 32 + lane;
#pragma unroll 1
        for (int cb = 0; cb < C::BN / 32; ++cb) {
            float v[32];
            tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)((jt >> 2) * C::BN) + cb * 32, v);
            if (col < a.n_in) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int n = cb * 32 + i, u = n / NG, g = n % NG;
                    const int row = r * 32 + u;
                    if (row < a.n_out) {
                        float o = v[i];
                        if (a.mask) o *= __ldg(a.mask + (int64_t)row * a.n_in + col);
                        if (o != 0.f) atomicAdd(a.dw.p[g] + (int64_t)row * a.lddw + col, o);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<C::TCOLS>(tmem);
}

}  // namespace hcgs

extern "C" int hcgs_tc_sddmm(const hcgs_plan* plan, int n_w, const void* dy_bf16, int64_t lddy, const void* x_bf16,
                             int64_t ldx, int64_t m, float* const* dw, int64_t lddw, const float* mask,
                             hcgs_stream stream) {
    HCGS_REQUIRE(plan && dy_bf16 && x_bf16 && dw, "hcgs_tc_sddmm: NULL argument");
    HCGS_REQUIRE(n_w == 1 || n_w == 4, "hcgs_tc_sddmm: n_w must be 1 or 4");
    if (m <= 0 || plan->n_sd_items == 0) return 0;
    CUtensorMap tmx, tmdy;
    if (int rc = make_tmap_bf16(&tmx, x_bf16, m, plan->n_in, ldx, 32, SD_RK, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    if (int rc = make_tmap_bf16(&tmdy, dy_bf16, m, lddy, lddy, n_w == 4 ? 64 : 32, SD_RK,
                                n_w == 4 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B))
        return rc;
    // one CTA per SM (shared memory, all 512 TMEM columns) and ONE wave: every extra split of the rows costs a
    // full set of fp32 atomics per kept weight (measured 2.70 / 3.06 / 3.62 / 4.77 ms per step for 1 / 2 / 4 / 8
    // waves, HCGS_SD_WAVES); a split is a multiple of SD_RK rows
    static const int waves = [] { const char* e = getenv("HCGS_SD_WAVES"); return e ? std::max(1, atoi(e)) : 1; }();
    int splits = std::max(1, (148 * waves) / plan->n_sd_items);
    int64_t rows = (m + splits - 1) / splits;
    rows = std::max<int64_t>((rows + SD_RK - 1) / SD_RK * SD_RK, 4 * SD_RK);
    splits = (int)((m + rows - 1) / rows);
    HCGS_REQUIRE(splits <= 65535, "hcgs_tc_sddmm: too many splits");
    TcSdArgs a;
    a.items = plan->d_sd_items;
    a.col_idx = plan->d_col_idx;
    a.dw = make_pack<float>(dw, n_w);
    a.lddw = lddw;
    a.mask = mask;
    a.M = m;
    a.rows_per_split = rows;
    a.n_out = plan->n_out;
    a.n_in = plan->n_in;
    dim3 grid(plan->n_sd_items, splits, 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_w == 4) {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_sddmm_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             SdCfg<4>::SMEM));
        tc_sddmm_kernel<4><<<grid, 128, SdCfg<4>::SMEM, st>>>(tmx, tmdy, a);
    } else {
        HCGS_CHECK_CUDA(cudaFuncSetAttribute(tc_sddmm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             SdCfg<1>::SMEM));
        tc_sddmm_kernel<1><<<grid, 128, SdCfg<1>::SMEM, st>>>(tmx, tmdy, a);
    }
    HCGS_LAUNCH_CHECK();
    return 0;
}

// ---- fused output layer: log-posteriors = log_softmax(x W^T + b) - log_prior ---------------------------
// (reference: MLP.forward neural_networks.py:306-317 with act = LogSoftmax(dim=1) :73-74, followed by the prior
//  normalisation of core.py:241-245.)  The logits never touch HBM.  A cluster of C CTAs owns a 128-row tile; CTA j
// accumulates the logits of its S strips of 64 output columns in TMEM (kept 32-wide k tiles only).  The kernel is
// persistent and warp specialised:
//   warp 0      TMA producer: x boxes + packed weight tiles into a ring of shared-memory stages
//   warp 1      tcgen05.mma issuer; accumulators double buffered in TMEM (2 x S*64 columns) when S <= 4, so the
//               products of row tile i+1 run under the epilogue of tile i
//   warps 2..17 epilogue (two groups of 8 on alternate row tiles when the accumulators are double buffered):
//               pass 1 reads the accumulators once for the row max / sum of exponentials, pushes the
//               partial {max, sum} to every CTA of the cluster with st.async (transaction-count mbarrier, no cluster
//               barrier), pass 2 reads them again, subtracts logsumexp + log prior and stores 32x32 fp32 boxes
//               with TMA (128-byte swizzled staging tiles, two per warp)
// HBM traffic per frame: the bf16 input row once (peers hit L2) and the fp32 posterior row once.
namespace hcgs {

constexpr int OUT_A_BYTES = 128 * 64;  // 128 rows x 32 bf16, SWIZZLE_64B
constexpr int OUT_B_TILE = 64 * 64;    // 64 output columns x 32 bf16, SWIZZLE_64B
constexpr int OUT_EPI_WARPS = 16;   // four per TMEM lane quadrant, each takes a quarter of the CTA's columns
constexpr int OUT_MAX_ITEMS = 128;   // 32-wide k tiles of one CTA (n_in <= 4096)
constexpr int OUT_THREADS = 32 * (2 + OUT_EPI_WARPS);
constexpr float OUT_L2E = 1.4426950408889634f;

template <int TC>
struct OutCfg {
    static constexpr int SMAX = TC / 64;
    static constexpr int NBUF = TC <= 256 ? 2 : 1;           // TMEM accumulator buffers
    static constexpr int TALLOC = NBUF * TC < 32 ? 32 : NBUF * TC;
    static constexpr int STAGE = OUT_A_BYTES + SMAX * OUT_B_TILE;
    static constexpr int STG_OFF = 0;                                // one 4 KB store staging tile per epilogue warp
    static constexpr int PART_OFF = STG_OFF + OUT_EPI_WARPS * 4096;  // [2][8][128] float2 partials of the cluster
    static constexpr int LP_OFF = PART_OFF + 2 * 8 * 128 * 8;        // [4][128] float2 partials of this CTA's warps
    static constexpr int SUB_OFF = LP_OFF + 4 * 128 * 8;             // [TC] bias, [TC] bias - log_prior
    static constexpr int ITEM_OFF = SUB_OFF + 2 * TC * 4;            // [OUT_MAX_ITEMS] int4 (k tile, mask, tile, n)
    static constexpr int BAR_OFF = ITEM_OFF + OUT_MAX_ITEMS * 16;    // barriers, TMEM slot, strip mask
    static constexpr int RING_OFF = (BAR_OFF + 256 + 1023) & ~1023;
    static constexpr int STAGES = (227 * 1024 - 1024 - RING_OFF) / STAGE < 8 ? (227 * 1024 - 1024 - RING_OFF) / STAGE : 8;
    static constexpr int SMEM = 1024 + RING_OFF + STAGES * STAGE;
};

__global__ void out_pack_kernel(const int* __restrict__ tile_sk, int n_tiles, const float* __restrict__ W, int ldw,
                                int n_out, int n_in, unsigned char* __restrict__ out) {
    const int64_t total = (int64_t)n_tiles * 64 * 4;  // 16-byte chunks
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int c16 = (int)(idx & 3);
        const int n = (int)((idx >> 2) & 63);
        const int t = (int)(idx >> 8);
        const int row = tile_sk[2 * t] * 64 + n, col0 = tile_sk[2 * t + 1] * 32 + c16 * 8;
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo = 0.f, hi = 0.f;
            if (row < n_out) {
                if (col0 + 2 * j < n_in) lo = __ldg(W + (int64_t)row * ldw + col0 + 2 * j);
                if (col0 + 2 * j + 1 < n_in) hi = __ldg(W + (int64_t)row * ldw + col0 + 2 * j + 1);
            }
            const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
            w[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
        *reinterpret_cast<uint4*>(out + (size_t)t * OUT_B_TILE + n * 64 + ((c16 ^ ((n >> 1) & 3)) << 4)) =
            make_uint4(w[0], w[1], w[2], w[3]);
    }
}

struct OutArgs {
    const unsigned char* wpack;
    const int* slice_ptr;    // [C+1]
    const int* items;        // [n][3] (k tile, strip mask, first packed tile)
    const float* bias;       // [P] or NULL
    const float* log_prior;  // [P] or NULL
    float* out;              // [M, ldo]
    int64_t ldo, M;
    int P, S, C;
    int n_mtiles;            // row tiles of 128
    int use_tma_store;       // out is 16-byte aligned with ldo % 4 == 0
    long long* trace;        // tuning aid (HCGS_OUT_TRACE=1): per-phase clock totals of CTA (0,0), else NULL
};

__device__ __forceinline__ float ex2_fast(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ uint32_t out_mapa(uint32_t addr, int rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void out_mbar_arrive(uint32_t addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}

// one 32x32 fp32 box of posteriors (thread = row): 128-byte swizzled staging tile -> TMA store; rows / columns past
// the tensor are clipped by the TMA.  Without a 16-byte aligned output the rows are written directly.
__device__ __forceinline__ void out_store_chunk(const float (&v)[32], const OutArgs& a, const CUtensorMap& tmOut,
                                                unsigned char* tile, int lane, int col0, int m0, int q) {
    if (a.use_tma_store) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i)  // 128-byte swizzle: 16-byte chunk i of row `lane` lands at chunk i ^ (lane & 7)
            *reinterpret_cast<float4*>(tile + lane * 128 + ((i ^ (lane & 7)) << 4)) =
                make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(&tmOut),
                         "r"(g_smem_u32(tile)), "r"(col0), "r"(m0 + q * 32)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    } else {
        const int64_t r = (int64_t)m0 + q * 32 + lane;
        if (r < a.M) {
            float* dst = a.out + r * a.ldo + col0;
            const int nv = a.P - col0;
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (i < nv) dst[i] = v[i];
        }
    }
}

template <int TC>
__global__ void __launch_bounds__(OUT_THREADS, 1)
out_lsm_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmOut, const OutArgs a) {
    using C = OutCfg<TC>;
    constexpr int ST = C::STAGES, NBUF = C::NBUF;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF);
    // barrier map: full[ST] | empty[ST] | tfull[2] | tempty[2] | pfull[2]
    const uint32_t full0 = g_smem_u32(bars), empty0 = full0 + 8 * ST, tfull0 = empty0 + 8 * ST, tempty0 = tfull0 + 16,
                   pfull0 = tempty0 + 16;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * ST + 6);
    uint32_t* any_slot = tmem_slot + 1;
    float* s_bias = reinterpret_cast<float*>(smem + C::SUB_OFF);
    float* s_sub = s_bias + TC;
    float2* s_part = reinterpret_cast<float2*>(smem + C::PART_OFF);
    unsigned char* ring = smem + C::RING_OFF;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);  // the same value, provably warp-uniform for the compiler
    const int j = blockIdx.x;  // slice of the output columns = rank in the cluster (grid.x == cluster.x == C)
    const int i0 = a.slice_ptr[j], ni = a.slice_ptr[j + 1] - i0;
    const int NC = a.S * 64, col_base = j * NC;
    const int n_my = (a.n_mtiles - (int)blockIdx.y + (int)gridDim.y - 1) / (int)gridDim.y;  // row tiles of this cluster

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < ST; ++s) {
            g_mbar_init(full0 + 8 * s, 1);
            g_mbar_init(empty0 + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            g_mbar_init(tfull0 + 8 * b, 1);
            g_mbar_init(tempty0 + 8 * b, OUT_EPI_WARPS / NBUF);
            g_mbar_init(pfull0 + 8 * b, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc<C::TALLOC>(g_smem_u32(tmem_slot));
    int4* s_items = reinterpret_cast<int4*>(smem + C::ITEM_OFF);
    float2* s_lp = reinterpret_cast<float2*>(smem + C::LP_OFF);
    if (warp == 2) {
        uint32_t m = 0;
        for (int i = lane; i < ni; i += 32) {
            const int mk = a.items[3 * (i0 + i) + 1];
            m |= (uint32_t)mk;
            s_items[i] = make_int4(a.items[3 * (i0 + i)], mk, a.items[3 * (i0 + i) + 2], __popc((uint32_t)mk));
        }
        m = __reduce_or_sync(0xffffffffu, m);
        if (lane == 0) *any_slot = m;
    }
    if (warp >= 2)
        for (int c = threadIdx.x - 64; c < NC; c += 32 * OUT_EPI_WARPS) {
            const int col = col_base + c;
            const float b = (a.bias && col < a.P) ? __ldg(a.bias + col) : 0.f;
            const float lp = (a.log_prior && col < a.P) ? __ldg(a.log_prior + col) : 0.f;
            s_bias[c] = b;
            s_sub[c] = b - lp;
        }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // every CTA of the cluster has initialised its barriers before any peer pushes partials into it
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    const uint32_t any = *any_slot;
    long long* tr = (a.trace && blockIdx.x == 0 && blockIdx.y == 0) ? a.trace : nullptr;
    const long long t_begin = tr ? clock64() : 0;
    long long t_acc0 = 0, t_acc1 = 0, t_acc2 = 0, t_acc3 = 0;  // per-thread phase totals, written once at the end

    if (warp == 0) {
        if (lane == 0) {
            int g = 0;  // running stage counter over all tiles
            for (int it = 0; it < n_my; ++it) {
                const int m0 = ((int)blockIdx.y + it * (int)gridDim.y) * 128;
                for (int i = 0; i < ni; ++i, ++g) {
                    const int s = g % ST;
                    const long long tq0 = tr ? clock64() : 0;
                    if (g >= ST) g_mbar_wait(empty0 + 8 * s, (uint32_t)(((g / ST) - 1) & 1));
                    if (tr) t_acc0 += clock64() - tq0;
                    const int4 item = s_items[i];
                    const int kt = item.x, nb = item.w, t0 = item.z;
                    g_mbar_expect_tx(full0 + 8 * s, OUT_A_BYTES + nb * OUT_B_TILE);
                    unsigned char* st = ring + s * C::STAGE;
                    tma_load_2d(g_smem_u32(st), &tmX, kt * 32, m0, full0 + 8 * s);
                    bulk_load_1d(g_smem_u32(st + OUT_A_BYTES), a.wpack + (size_t)t0 * OUT_B_TILE, nb * OUT_B_TILE,
                                 full0 + 8 * s);
                }
            }
        }
    } else if (warp_u == 1) {
        // The whole warp runs this loop convergently on warp-uniform values (warp_u and the TMEM base come from a
        // shuffle, everything else from loop counters and constants) and ONE elected lane issues: the descriptors
        // then live in uniform registers.  Issued from a `lane == 0` branch instead, every tcgen05.mma was wrapped in
        // an ELECT / R2UR.BROADCAST loop by the compiler and cost ~180 cycles of the issuing thread.
        int g = 0;
        const uint64_t desc0 = umma_smem_desc(g_smem_u32(ring), 16, 512, UMMA_SW64);
        const uint32_t idesc0 = umma_idesc_bf16(128, 0, 0, 0);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
        for (int it = 0; it < n_my; ++it) {
            const int b = it % NBUF, k = it / NBUF;
            const long long tq0 = tr ? clock64() : 0;
            if (k >= 1) {
                g_mbar_wait(tempty0 + 8 * b, (uint32_t)((k - 1) & 1));
                tc_fence_after();
            }
            if (tr) t_acc0 += clock64() - tq0;
            const uint32_t acc = tmem_u + (uint32_t)(b * TC);
            for (int i = 0; i < ni; ++i, ++g) {
                const int s = g % ST;
                const long long tq1 = tr ? clock64() : 0;
                g_mbar_wait(full0 + 8 * s, (uint32_t)((g / ST) & 1));
                tc_fence_after();
                if (tr) t_acc1 += clock64() - tq1;
                // descriptors differ only in the start-address field (bits 0..13, address >> 4).  Every run of
                // adjacent kept strips is one instruction of N = 64 * run (<= 256); the first item of a tile
                // covers all strips (it initialises the accumulators), so every later one accumulates
                const uint64_t ad = desc0 + (uint64_t)((s * C::STAGE) >> 4);
                uint64_t bd = ad + (OUT_A_BYTES >> 4);
                const uint32_t accf = i > 0 ? 1u : 0u;
                uint32_t m = __shfl_sync(0xffffffffu, (uint32_t)s_items[i].y, 0);
                const bool issuer = out_elect_one();
                while (m) {
                    const int lo = __ffs(m) - 1;
                    int len = __ffs(~(m >> lo)) - 1;
                    len = len > 4 ? 4 : len;
                    const uint32_t idesc = idesc0 | ((uint32_t)(len * 8) << 17);
                    const uint32_t d = acc + (uint32_t)(lo * 64);
                    if (issuer) {
                        umma_bf16(d, ad, bd, idesc, accf);
                        umma_bf16(d, ad + 2, bd + 2, idesc, 1u);
                    }
                    bd += (uint64_t)len * (OUT_B_TILE >> 4);
                    m &= ~(((1u << len) - 1u) << lo);
                }
                if (issuer) umma_commit(empty0 + 8 * s);
                __syncwarp();
            }
            if (out_elect_one()) umma_commit(tfull0 + 8 * b);  // arrives once every product of this tile has completed
            __syncwarp();
        }
    } else {
        // ---- epilogue warps: TMEM lane quadrant q (hardware: warp id % 4), column half hh ----
        // With two accumulator buffers the 16 warps form two groups of 8 that take alternate row tiles: their
        // phases (TMEM read, exponentials, exchange, stores) then overlap instead of adding up.  Inside a group
        // PARTS warps share a TMEM lane quadrant, each takes 1/PARTS of the CTA's columns.
        constexpr int GROUPS = NBUF, PARTS = 4 / GROUPS;
        const int ew = warp - 2;
        const int q = warp & 3, grp = (ew >> 2) / PARTS, hh = (ew >> 2) % PARTS;
        const bool leader = (ew == grp * PARTS * 4) && lane == 0;
        const int row = q * 32 + lane;
        const int ncb = a.S * 2;                               // 32-column chunks of this CTA
        const int cb_lo = (ncb * hh) / PARTS, cb_hi = (ncb * (hh + 1)) / PARTS;
        unsigned char* tile = smem + C::STG_OFF + ew * 4096;
        const uint32_t my_part = g_smem_u32(&s_part[j * 128 + row]);
        float2* lp_grp = s_lp + grp * PARTS * 128;
        for (int it = grp; it < n_my; it += GROUPS) {
            const int b = it % NBUF, k = it / NBUF, par = it & 1;
            const int m0 = ((int)blockIdx.y + it * (int)gridDim.y) * 128;
            const uint32_t trow = tmem + (uint32_t)(b * TC) + ((uint32_t)(q * 32) << 16);
            if (leader) g_mbar_expect_tx(pfull0 + 8 * par, (uint32_t)a.C * 128 * 8);
            long long tq = (tr && threadIdx.x == 64) ? clock64() : 0;
            g_mbar_wait(tfull0 + 8 * b, (uint32_t)(k & 1));
            tc_fence_after();
            if (tr && threadIdx.x == 64) { const long long n = clock64(); t_acc0 += n - tq; tq = n; }
            // ---- pass 1: row max and sum of exponentials over this warp's columns ----
            float mx = -INFINITY, sum = 0.f;
#pragma unroll 1
            for (int cb = cb_lo; cb < cb_hi; ++cb) {
                const int c0 = cb * 32;
                if (col_base + c0 >= a.P) break;
                float v[32];
                if (any >> (cb >> 1) & 1) {
                    tmem_ld32(trow + c0, v);
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = 0.f;
                }
                const int nv = a.P - col_base - c0;  // valid columns of this chunk (>= 32: all)
                float cm = -INFINITY;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 bb = *reinterpret_cast<const float4*>(s_bias + c0 + i);
                    v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
                }
                if (nv >= 32) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) cm = fmaxf(cm, v[i]);
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        if (i >= nv) v[i] = -INFINITY;
                        cm = fmaxf(cm, v[i]);
                    }
                }
                const float mn = fmaxf(mx, cm);
                const float off = -mn * OUT_L2E;
                float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    acc0 += ex2_fast(fmaf(v[i], OUT_L2E, off));
                    acc1 += ex2_fast(fmaf(v[i + 1], OUT_L2E, off));
                }
                sum = sum * ex2_fast((mx - mn) * OUT_L2E) + (acc0 + acc1);
                mx = mn;
            }
            // ---- the four warps of a lane quadrant merge their partial {max, sum}; the merged pair goes to every CTA
            //      of the cluster (the four warps share the peers) ----
            {
                lp_grp[hh * 128 + row] = make_float2(mx, sum);
                asm volatile("bar.sync %0, %1;" ::"r"(1 + grp * 4 + q), "n"(32 * PARTS) : "memory");
                float gm = -INFINITY;
#pragma unroll
                for (int p = 0; p < PARTS; ++p) gm = fmaxf(gm, lp_grp[p * 128 + row].x);
                float gs = 0.f;
#pragma unroll
                for (int p = 0; p < PARTS; ++p) {
                    const float2 e = lp_grp[p * 128 + row];
                    gs += e.y * ex2_fast((e.x - gm) * OUT_L2E);
                }
                const uint32_t slot = my_part + (uint32_t)(par * 8 * 128 * 8);
                const uint64_t pv = ((uint64_t)__float_as_uint(gs) << 32) | (uint64_t)__float_as_uint(gm);
                for (int peer = hh; peer < a.C; peer += PARTS)
                    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(
                                     out_mapa(slot, peer)),
                                 "l"(pv), "r"(out_mapa(pfull0 + 8 * par, peer))
                                 : "memory");
            }
            if (tr && threadIdx.x == 64) { const long long n = clock64(); t_acc1 += n - tq; tq = n; }
            g_mbar_wait(pfull0 + 8 * par, (uint32_t)((it >> 1) & 1));
            if (tr && threadIdx.x == 64) { const long long n = clock64(); t_acc2 += n - tq; tq = n; }
            float lse;
            {
                const float2* pp = s_part + par * 8 * 128 + row;
                float gm = -INFINITY;
                for (int p = 0; p < a.C; ++p) gm = fmaxf(gm, pp[p * 128].x);
                float gs = 0.f;
                for (int p = 0; p < a.C; ++p) {
                    const float2 e = pp[p * 128];
                    gs += e.y * ex2_fast((e.x - gm) * OUT_L2E);
                }
                lse = gm + logf(gs);
            }
            // ---- pass 2: z - logsumexp - log prior -> 32x32 boxes ----
            {
#pragma unroll 1
                for (int cb = cb_lo; cb < cb_hi; ++cb) {
                    const int c0 = cb * 32;
                    if (col_base + c0 >= a.P) break;
                    float v[32];
                    if (any >> (cb >> 1) & 1) {
                        tmem_ld32(trow + c0, v);
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) v[i] = 0.f;
                    }
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const float4 sb = *reinterpret_cast<const float4*>(s_sub + c0 + i);
                        v[i] += sb.x - lse; v[i + 1] += sb.y - lse; v[i + 2] += sb.z - lse; v[i + 3] += sb.w - lse;
                    }
                    out_store_chunk(v, a, tmOut, tile, lane, col_base + c0, m0, q);
                }
                // this warp has drained its part of accumulator buffer b
                tc_fence_before();
                __syncwarp();
                if (lane == 0) out_mbar_arrive(tempty0 + 8 * b);
            }
            if (tr && threadIdx.x == 64) { const long long n = clock64(); t_acc3 += n - tq; tq = n; }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    if (tr && lane == 0) {
        if (warp == 0) tr[0] = t_acc0;
        if (warp == 1) { tr[1] = t_acc0; tr[2] = t_acc1; tr[9] = clock64() - t_begin; }
        if (warp == 2) { tr[3] = t_acc0; tr[4] = t_acc1; tr[5] = t_acc2; tr[6] = t_acc3; tr[7] = clock64() - t_begin; tr[8] = (n_my + NBUF - 1) / NBUF; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<C::TALLOC>(tmem);
    // a CTA must not exit while a peer may still push partials into its shared memory
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int TC>
static int out_launch(const CUtensorMap& tm, const CUtensorMap& tmo, OutArgs& a, cudaStream_t st) {
    using C = OutCfg<TC>;
    static_assert(C::STAGES >= 2, "output kernel: ring too small");
    HCGS_CHECK_CUDA(cudaFuncSetAttribute(out_lsm_kernel<TC>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.blockDim = dim3(OUT_THREADS, 1, 1);
    cfg.dynamicSmemBytes = C::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)a.C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    static int max_clusters[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // per cluster size (same footprint for every TC: 1 CTA / SM)
    if (max_clusters[a.C] == 0) {
        cfg.gridDim = dim3((unsigned)a.C, 1024, 1);
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, out_lsm_kernel<TC>, &cfg) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            n = 148 / a.C / 2;
        }
        max_clusters[a.C] = n;
    }
    cfg.gridDim = dim3((unsigned)a.C, (unsigned)std::min(a.n_mtiles, max_clusters[a.C]), 1);
    HCGS_CHECK_CUDA(cudaLaunchKernelEx(&cfg, out_lsm_kernel<TC>, tm, tmo, a));
    return 0;
}

static int out_tmem_cols(int S) { return S <= 1 ? 64 : (S <= 2 ? 128 : (S <= 4 ? 256 : 512)); }

}  // namespace hcgs

extern "C" int hcgs_out_info(const hcgs_plan* plan, int64_t* info) {
    HCGS_REQUIRE(plan && info, "hcgs_out_info: NULL argument");
    for (int i = 0; i < 8; ++i) info[i] = 0;
    if (plan->out_C == 0) return 0;
    const int tc = out_tmem_cols(plan->out_S);
    info[0] = plan->out_C;
    info[1] = plan->out_S;
    info[2] = plan->out_tiles;
    info[3] = tc;
    info[4] = tc == 64 ? OutCfg<64>::SMEM : tc == 128 ? OutCfg<128>::SMEM : tc == 256 ? OutCfg<256>::SMEM : OutCfg<512>::SMEM;
    info[5] = plan->out_items;
    info[6] = (int64_t)plan->out_tiles * OUT_B_TILE;  // bytes of the packed weights
    info[7] = tc == 64 ? OutCfg<64>::STAGES : tc == 128 ? OutCfg<128>::STAGES : tc == 256 ? OutCfg<256>::STAGES : OutCfg<512>::STAGES;
    return 0;
}

extern "C" int hcgs_out_pack(const hcgs_plan* plan, const float* w, int ldw, void* wpack, hcgs_stream stream) {
    HCGS_REQUIRE(plan && w && wpack, "hcgs_out_pack: NULL argument");
    HCGS_REQUIRE(plan->out_C > 0, "hcgs_out_pack: %d output columns exceed the fused output kernel (max 4096)", plan->n_out);
    if (plan->out_tiles == 0) return 0;
    const int64_t total = (int64_t)plan->out_tiles * 256;
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
    out_pack_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(plan->d_out_tile, plan->out_tiles, w, ldw, plan->n_out,
                                                              plan->n_in, (unsigned char*)wpack);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_out_logsoftmax_prior_fwd(const hcgs_plan* plan, const void* wpack, const void* x_bf16, int64_t ldx,
                                             int64_t m, const float* bias, const float* log_prior, float* out,
                                             int64_t ldo, hcgs_stream stream) {
    HCGS_REQUIRE(plan && wpack && x_bf16 && out, "hcgs_out_logsoftmax_prior_fwd: NULL argument");
    HCGS_REQUIRE(plan->out_C > 0, "hcgs_out_logsoftmax_prior_fwd: %d output columns exceed the fused kernel (max 4096)",
                 plan->n_out);
    HCGS_REQUIRE(ldo >= plan->n_out, "hcgs_out_logsoftmax_prior_fwd: ldo too small");
    HCGS_REQUIRE(plan->tc <= OUT_MAX_ITEMS, "hcgs_out_logsoftmax_prior_fwd: n_in = %d exceeds %d", plan->n_in,
                 OUT_MAX_ITEMS * 32);
    if (m <= 0) return 0;
    HCGS_REQUIRE(m <= (int64_t)16777216 * 128 / 2, "hcgs_out_logsoftmax_prior_fwd: too many rows (%lld)", (long long)m);
    CUtensorMap tm, tmo;
    if (int rc = make_tmap_bf16(&tm, x_bf16, m, plan->n_in, ldx, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    OutArgs a;
    a.use_tma_store = (((uintptr_t)out & 15) == 0 && (ldo & 3) == 0) ? 1 : 0;
    memset(&tmo, 0, sizeof(tmo));
    if (a.use_tma_store) {
        EncodeTiledFn fn = get_encode_fn();
        HCGS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
        cuuint64_t gdim[2] = {(cuuint64_t)plan->n_out, (cuuint64_t)m};
        cuuint64_t gstr[1] = {(cuuint64_t)ldo * 4};
        cuuint32_t box[2] = {32, 32};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = fn(&tmo, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        HCGS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (posteriors) failed (%d)", (int)r);
    }
    a.wpack = (const unsigned char*)wpack;
    a.slice_ptr = plan->d_out_ptr;
    a.items = plan->d_out_items;
    a.bias = bias;
    a.log_prior = log_prior;
    a.out = out;
    a.ldo = ldo;
    a.M = m;
    a.P = plan->n_out;
    a.S = plan->out_S;
    a.C = plan->out_C;
    a.n_mtiles = (int)((m + 127) / 128);
    cudaStream_t st = (cudaStream_t)stream;
    static const bool want_trace = getenv("HCGS_OUT_TRACE") != nullptr;
    static long long* d_trace = nullptr;
    a.trace = nullptr;
    if (want_trace) {
        if (!d_trace) HCGS_CHECK_CUDA(cudaMalloc((void**)&d_trace, 16 * sizeof(long long)));
        HCGS_CHECK_CUDA(cudaMemsetAsync(d_trace, 0, 16 * sizeof(long long), st));
        a.trace = d_trace;
    }
    int rc;
    switch (out_tmem_cols(plan->out_S)) {
        case 64: rc = out_launch<64>(tm, tmo, a, st); break;
        case 128: rc = out_launch<128>(tm, tmo, a, st); break;
        case 256: rc = out_launch<256>(tm, tmo, a, st); break;
        default: rc = out_launch<512>(tm, tmo, a, st); break;
    }
    if (want_trace && rc == 0) {
        long long h[16];
        HCGS_CHECK_CUDA(cudaStreamSynchronize(st));
        HCGS_CHECK_CUDA(cudaMemcpy(h, d_trace, sizeof(h), cudaMemcpyDeviceToHost));
        const double n = h[8] > 0 ? (double)h[8] : 1.0;
        fprintf(stderr, "[out trace] tiles %lld, cycles/tile: total %.0f | producer wait-empty %.0f | mma wait-tempty %.0f "
                        "wait-full %.0f (loop total %.0f) | epilogue wait-tfull %.0f pass1+push %.0f wait-pfull %.0f combine+pass2 %.0f\n",
                h[8], h[7] / n, h[0] / n, h[1] / n, h[2] / n, h[9] / n, h[3] / n, h[4] / n, h[5] / n, h[6] / n);
    }
    return rc;
}
