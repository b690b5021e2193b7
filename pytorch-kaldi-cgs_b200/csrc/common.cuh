// Shared helpers for the HCGS B200 (sm_100a) kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/hcgs_b200.h"

namespace hcgs {

// ---- error plumbing (no exception crosses the C ABI) -------------------------------------
void set_error(const char* fmt, ...);

#define HCGS_CHECK_CUDA(expr)                                                              \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            hcgs::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,           \
                            cudaGetErrorString(_e));                                       \
            return 1;                                                                      \
        }                                                                                  \
    } while (0)

#define HCGS_REQUIRE(cond, ...)                                                            \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            hcgs::set_error(__VA_ARGS__);                                                  \
            return 2;                                                                      \
        }                                                                                  \
    } while (0)

#define HCGS_LAUNCH_CHECK() HCGS_CHECK_CUDA(cudaGetLastError())

// A handful of device pointers passed by value into a kernel (the gates of one layer).
template <typename T>
struct PtrPack {
    T* p[HCGS_MAX_GROUPS];
};

template <typename T>
static inline PtrPack<T> make_pack(T* const* src, int n) {
    PtrPack<T> r;
    for (int i = 0; i < HCGS_MAX_GROUPS; ++i) r.p[i] = (src != nullptr && i < n) ? src[i] : nullptr;
    return r;
}

constexpr int kSdMaxTiles = 12;  // kept tiles of one tile row handled by one sddmm work item (3 MMAs of M = 128)
constexpr int kTile = 32;  // weight sparsity tile (rows x cols) used by the block-sparse GEMMs

// ---- the opaque plan --------------------------------------------------------------------
struct RecurrentPlan {
    int CL = 0;        // CTAs per cluster (each cluster runs 8 sequences over all H hidden units)
    int Hc = 0;        // hidden units (rows of U) owned by one CTA, multiple of 4
    int HcP = 0;       // Hc padded so that HcP % 4 == 1 (gate buffer stride)
    int RG = 0;        // 4-row groups per CTA = Hc/4
    int L = 0;         // BELL width: max non-empty 4x4 blocks in any row group
    int NCG = 0;       // column groups = ceil(H/4)
    int max_cta_blocks = 0;  // largest number of real blocks owned by one CTA
    int n_blocks = 0;        // real blocks in total
    int threads = 0;         // block size of both recurrent kernels (16 threads per row group)
    bool w_in_smem = true;   // false: compressed weights stay in global memory (too large for smem)
    uint16_t* d_bidx = nullptr;     // [CL][L][RG] column group of block j of row group a (0xFFFF = padding)
    int* d_cta_blk_off = nullptr;   // [CL+1] offsets into the flat real-block list
    uint16_t* d_blk_a = nullptr;    // [n_blocks] local row group of real block s
    uint16_t* d_blk_c = nullptr;    // [n_blocks] column group of real block s
    int* d_cg_ptr = nullptr;        // [CL][NCG+1] per-CTA CSR: column group -> block slots
    uint16_t* d_cg_list = nullptr;  // [n_blocks] local block slots grouped by column group
    int64_t fwd_pack_floats = 0;    // CL*L*RG*64
    int64_t bwd_pack_floats = 0;    // n_blocks*64
    size_t smem_fwd = 0, smem_bwd = 0;
};

struct TcPlan;
// gemm_tc.cu: CUtensorMap of a [d2][d1][d0] tensor with box {box0, box1, 1} (no swizzle)
int make_tmap_3d(void* map, int elem_bytes, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                 uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0, uint32_t box1);
// tensor-core recurrent layouts of one cluster size (lstm_tc.cu)
void tc_plan_free(TcPlan* t);  // defined in lstm_tc.cu

}  // namespace hcgs

struct hcgs_plan {
    int n_out = 0, n_in = 0;
    int tr = 0, tc = 0;  // tile rows / cols
    int nz_tiles = 0;
    int64_t nnz = 0;
    int* d_row_ptr = nullptr;  // [tr+1]   CSR over tiles
    int* d_col_idx = nullptr;  // [nz]
    int* d_col_ptr = nullptr;  // [tc+1]   CSC over tiles
    int* d_row_idx = nullptr;  // [nz]
    int* d_tile_rc = nullptr;  // [nz*2]   (tile row, tile col) flat list for the SDDMM
    int* d_csc_rc = nullptr;   // [nz*2]   (tile row, tile col) in CSC order (tensor-core nn GEMM weight tiles)
    int* d_sd_items = nullptr; // [n_sd_items*3] (tile row, first CSR tile, count<=4): tensor-core SDDMM work items
    int n_sd_items = 0;
    int max_row_tiles = 0, max_col_tiles = 0;
    bool recurrent = false;
    hcgs::RecurrentPlan rp;
    // recurrent plans keep the 4x4-block occupancy on the host (ns44 x ns44 bytes) so that the
    // tensor-core layouts can be derived lazily for whichever cluster size a launch asks for
    uint8_t* h_blk44 = nullptr;
    int ns44 = 0;
    hcgs::TcPlan* tcp[3] = {nullptr, nullptr, nullptr};  // cluster size 4, 8, 16
    // fused output layer (gemm_tc.cu, out_lsm_kernel): the n_out columns are cut into strips of 64; a cluster of
    // out_C CTAs owns a 128-row tile, CTA j the strips [j*out_S, (j+1)*out_S).  Per CTA the kept 32-wide k tiles
    // (union over 32x32 tile rows of a strip) are listed as items (k tile, strip bit mask, first packed tile).
    int out_C = 0, out_S = 0;   // 0 = n_out too wide for the fused kernel (> 4096 columns)
    int out_tiles = 0;          // packed [64 n x 32 k] bf16 tiles
    int out_items = 0;
    uint32_t out_any[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // per CTA: strips with at least one kept tile
    int* d_out_ptr = nullptr;    // [out_C+1] item offsets
    int* d_out_items = nullptr;  // [out_items*3]
    int* d_out_tile = nullptr;   // [out_tiles*2] (global strip, k tile), pack order
};

namespace hcgs {

// ---- device math ------------------------------------------------------------------------
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float act_apply(int act, float x) {
    switch (act) {
        case HCGS_ACT_RELU: return x > 0.f ? x : 0.f;
        case HCGS_ACT_TANH: return tanhf(x);
        case HCGS_ACT_HTANH: return fminf(fmaxf(x, -1.f), 1.f);
        case HCGS_ACT_SIGMOID: return sigmoidf_(x);
        case HCGS_ACT_LEAKY_RELU: return x > 0.f ? x : 0.2f * x;
        case HCGS_ACT_ELU: return x > 0.f ? x : expm1f(x);
        default: return x;
    }
}

// derivative expressed through input x and output y = act(x) (either may be used)
__device__ __forceinline__ float act_grad(int act, float x, float y) {
    switch (act) {
        case HCGS_ACT_RELU: return x > 0.f ? 1.f : 0.f;
        case HCGS_ACT_TANH: return 1.f - y * y;
        case HCGS_ACT_HTANH: return (x > -1.f && x < 1.f) ? 1.f : 0.f;
        case HCGS_ACT_SIGMOID: return y * (1.f - y);
        case HCGS_ACT_LEAKY_RELU: return x > 0.f ? 1.f : 0.2f;
        case HCGS_ACT_ELU: return x > 0.f ? 1.f : y + 1.f;
        default: return 1.f;
    }
}

// derivative from the output only (what the LSTM/MLP backward has saved)
__device__ __forceinline__ float act_grad_from_out(int act, float y) {
    switch (act) {
        case HCGS_ACT_RELU: return y > 0.f ? 1.f : 0.f;
        case HCGS_ACT_TANH: return 1.f - y * y;
        // hardtanh: gradient is 1 strictly inside (-1,1); the saturated outputs are exactly +-1
        case HCGS_ACT_HTANH: return (y > -1.f && y < 1.f) ? 1.f : 0.f;
        case HCGS_ACT_SIGMOID: return y * (1.f - y);
        case HCGS_ACT_LEAKY_RELU: return y > 0.f ? 1.f : 0.2f;
        case HCGS_ACT_ELU: return y > 0.f ? 1.f : y + 1.f;
        default: return 1.f;
    }
}

// one application of Quantize_inp to an element, given the tensor's max |x| (quantized_modules.py:109-118):
// inp.div(var).abs().mul(2^(B-1)).ceil().div(2^(B-1)).mul(var) * sign
__device__ __forceinline__ float quant_act_once(float v, float var, float scale, float inv_scale) {
    const float sgn = (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f);
    float t = v / var;
    t = ceilf(fabsf(t) * scale) * inv_scale;
    t = t * var;
    return t * sgn;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace hcgs
