// Block-sparse fp32 GEMM family on the 32x32 tile index of a plan (parity build: fp32 FMA).
//
//   nt   : Y[m, g*N + n]  = sum_k X[m,k] W_g[n,k] (+bias)      forward projections
//   nn   : dX[m,k]        = sum_g sum_n dY[m, g*N+n] W_g[n,k]   data gradient
//   sddmm: dW_g[n,k]      = mask[n,k] * sum_m dY[m,g*N+n] X[m,k] weight gradient, non-empty tiles only
//
// Weights are read straight from the dense (masked) fp32 parameter matrices PyTorch owns; the
// plan only says which 32x32 tiles can be skipped.  The bf16 tcgen05 variants live in
// gemm_tc.cu; these kernels are the fp32 (rel 1e-5) parity path and the fallback for shapes the
// tensor-core path does not cover.
#include "common.cuh"

namespace hcgs {

constexpr int BM = 128;  // rows of X / dY per CTA
constexpr int XS = 33;   // smem stride of the [m][inner] operand tile
constexpr int WS = 36;   // smem stride of the [inner][out] weight tile (16 B aligned rows)

// acc[i][j] += sum_inner Xs[ty*4+i][inner] * Ws[inner][tx*4+j]
__device__ __forceinline__ void tile_mac(const float (*Xs)[XS], const float (*Ws)[WS], int ty, int tx,
                                         float acc[4][4]) {
#pragma unroll 8
    for (int k = 0; k < kTile; ++k) {
        const float4 w4 = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
        float xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xv[i] = Xs[ty * 4 + i][k];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[i][0] = fmaf(xv[i], w4.x, acc[i][0]);
            acc[i][1] = fmaf(xv[i], w4.y, acc[i][1]);
            acc[i][2] = fmaf(xv[i], w4.z, acc[i][2]);
            acc[i][3] = fmaf(xv[i], w4.w, acc[i][3]);
        }
    }
}

// Load a [BM x 32] operand tile (rows m0.., cols c0..) with zero fill.
__device__ __forceinline__ void load_rows(float (*Xs)[XS], const float* __restrict__ X, int ldx, int64_t m0,
                                          int64_t M, int c0, int c_limit, int warp, int lane) {
    const int c = c0 + lane;
    const bool cok = c < c_limit;
#pragma unroll 4
    for (int r = 0; r < BM / 8; ++r) {
        const int row = warp * (BM / 8) + r;
        const int64_t m = m0 + row;
        Xs[row][lane] = (cok && m < M) ? __ldg(X + m * ldx + c) : 0.f;
    }
}

__global__ void __launch_bounds__(256) bsgemm_nt_kernel(const int* __restrict__ row_ptr,
                                                        const int* __restrict__ col_idx,
                                                        PtrPack<const float> W, int ldw,
                                                        const float* __restrict__ X, int ldx, int64_t M,
                                                        int n_out, int n_in, PtrPack<const float> bias,
                                                        float* __restrict__ Y, int ldy) {
    __shared__ float Xs[BM][XS];
    __shared__ __align__(16) float Ws[kTile][WS];
    const int tn = blockIdx.x, g = blockIdx.z;
    const int64_t m0 = (int64_t)blockIdx.y * BM;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & 7, ty = tid >> 3;
    const float* __restrict__ Wg = W.p[g];
    float acc[4][4] = {};
    const int t0 = row_ptr[tn], t1 = row_ptr[tn + 1];
    for (int t = t0; t < t1; ++t) {
        const int k0 = col_idx[t] * kTile;
        load_rows(Xs, X, ldx, m0, M, k0, n_in, warp, lane);
        // W tile [32 n][32 k] stored transposed as Ws[k][n]
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int nl = warp * 4 + r;
            const int n = tn * kTile + nl;
            const int k = k0 + lane;
            Ws[lane][nl] = (n < n_out && k < n_in) ? __ldg(Wg + (int64_t)n * ldw + k) : 0.f;
        }
        __syncthreads();
        tile_mac(Xs, Ws, ty, tx, acc);
        __syncthreads();
    }
    const float* bg = bias.p[g];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = tn * kTile + tx * 4 + j;
            if (n < n_out) Y[m * ldy + (int64_t)g * n_out + n] = acc[i][j] + (bg ? __ldg(bg + n) : 0.f);
        }
    }
}

__global__ void __launch_bounds__(256) bsgemm_nn_kernel(const int* __restrict__ col_ptr,
                                                        const int* __restrict__ row_idx,
                                                        PtrPack<const float> W, int n_w, int ldw,
                                                        const float* __restrict__ dY, int lddy, int64_t M,
                                                        int n_out, int n_in, float* __restrict__ dX,
                                                        int lddx) {
    __shared__ float Xs[BM][XS];
    __shared__ __align__(16) float Ws[kTile][WS];
    const int tk = blockIdx.x;
    const int64_t m0 = (int64_t)blockIdx.y * BM;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & 7, ty = tid >> 3;
    float acc[4][4] = {};
    const int t0 = col_ptr[tk], t1 = col_ptr[tk + 1];
    for (int g = 0; g < n_w; ++g) {
        const float* __restrict__ Wg = W.p[g];
        for (int t = t0; t < t1; ++t) {
            const int n0 = row_idx[t] * kTile;
            // dY tile: rows m, cols g*n_out + n0 + lane (valid while n0+lane < n_out)
            {
                const int n = n0 + lane;
                const bool ok = n < n_out;
#pragma unroll 4
                for (int r = 0; r < BM / 8; ++r) {
                    const int row = warp * (BM / 8) + r;
                    const int64_t m = m0 + row;
                    Xs[row][lane] = (ok && m < M) ? __ldg(dY + m * lddy + (int64_t)g * n_out + n) : 0.f;
                }
            }
            // W tile [32 n][32 k] stored as Ws[n][k]
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int nl = warp * 4 + r;
                const int n = n0 + nl;
                const int k = tk * kTile + lane;
                Ws[nl][lane] = (n < n_out && k < n_in) ? __ldg(Wg + (int64_t)n * ldw + k) : 0.f;
            }
            __syncthreads();
            tile_mac(Xs, Ws, ty, tx, acc);
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = tk * kTile + tx * 4 + j;
            if (k < n_in) dX[m * lddx + k] = acc[i][j];
        }
    }
}

// SDDMM: one CTA per (non-empty tile, M-split).  256 threads = 4 groups of 64; with 4 weight
// matrices each group owns one gate (the X tile is shared by the four gates), with 1 weight matrix
// the groups split the rows of every 32-row chunk and are reduced through shared memory.
constexpr int SM_ROWS = 32;

__global__ void __launch_bounds__(256) bs_sddmm_kernel(const int* __restrict__ tile_rc, int n_w,
                                                       const float* __restrict__ dY, int lddy,
                                                       const float* __restrict__ X, int ldx, int64_t M,
                                                       int64_t rows_per_split, int n_out, int n_in,
                                                       PtrPack<float> dW, int lddw,
                                                       const float* __restrict__ mask, int use_atomic) {
    __shared__ __align__(16) float As[4][SM_ROWS][WS];  // dY tiles [gate][m][n]
    __shared__ __align__(16) float Bs[SM_ROWS][WS];     // X tile [m][k]
    const int tile = blockIdx.x;
    const int tn = tile_rc[2 * tile], tk = tile_rc[2 * tile + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q = tid >> 6;          // thread group
    const int ty = (tid >> 3) & 7;   // n rows ty*4..+3
    const int tx = tid & 7;          // k cols tx*4..+3
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    float acc[4][4] = {};
    const int n_lane = tn * kTile + lane, k_lane = tk * kTile + lane;
    for (int64_t mc = mbeg; mc < mend; mc += SM_ROWS) {
        // X tile: 32 rows x 32 k ; warp w loads rows w*4..w*4+3
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int row = warp * 4 + r;
            const int64_t m = mc + row;
            Bs[row][lane] = (m < mend && k_lane < n_in) ? __ldg(X + m * ldx + k_lane) : 0.f;
        }
        for (int g = 0; g < n_w; ++g) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int row = warp * 4 + r;
                const int64_t m = mc + row;
                As[g][row][lane] =
                    (m < mend && n_lane < n_out) ? __ldg(dY + m * lddy + (int64_t)g * n_out + n_lane) : 0.f;
            }
        }
        __syncthreads();
        if (n_w == 4) {
#pragma unroll 8
            for (int m = 0; m < SM_ROWS; ++m) {
                const float4 a4 = *reinterpret_cast<const float4*>(&As[q][m][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&Bs[m][tx * 4]);
                const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    acc[i][0] = fmaf(av[i], b4.x, acc[i][0]);
                    acc[i][1] = fmaf(av[i], b4.y, acc[i][1]);
                    acc[i][2] = fmaf(av[i], b4.z, acc[i][2]);
                    acc[i][3] = fmaf(av[i], b4.w, acc[i][3]);
                }
            }
        } else {  // n_w == 1: group q takes rows q*8..q*8+7 of the chunk
#pragma unroll
            for (int mm = 0; mm < SM_ROWS / 4; ++mm) {
                const int m = q * (SM_ROWS / 4) + mm;
                const float4 a4 = *reinterpret_cast<const float4*>(&As[0][m][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&Bs[m][tx * 4]);
                const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    acc[i][0] = fmaf(av[i], b4.x, acc[i][0]);
                    acc[i][1] = fmaf(av[i], b4.y, acc[i][1]);
                    acc[i][2] = fmaf(av[i], b4.z, acc[i][2]);
                    acc[i][3] = fmaf(av[i], b4.w, acc[i][3]);
                }
            }
        }
        __syncthreads();
    }
    int g_out = q;
    if (n_w == 1) {
        // reduce the four row-subsets through shared memory (reuse As as [4][32][36] scratch)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) As[q][ty * 4 + i][tx * 4 + j] = acc[i][j];
        __syncthreads();
        if (q != 0) return;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                acc[i][j] = As[0][ty * 4 + i][tx * 4 + j] + As[1][ty * 4 + i][tx * 4 + j] +
                            As[2][ty * 4 + i][tx * 4 + j] + As[3][ty * 4 + i][tx * 4 + j];
        g_out = 0;
    }
    float* __restrict__ out = dW.p[g_out];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = tn * kTile + ty * 4 + i;
        if (n >= n_out) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = tk * kTile + tx * 4 + j;
            if (k >= n_in) continue;
            float v = acc[i][j];
            if (mask) v *= __ldg(mask + (int64_t)n * n_in + k);
            if (use_atomic)
                atomicAdd(out + (int64_t)n * lddw + k, v);
            else
                out[(int64_t)n * lddw + k] = v;
        }
    }
}

// column sums of a [M, G*C] matrix: out_g[c] = sum_m dY[m, g*C + c]
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ dY, int lddy, int64_t M,
                                                     int C, int64_t rows_per_split, PtrPack<float> out,
                                                     int use_atomic) {
    __shared__ float red[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = blockIdx.z;
    const int c = blockIdx.x * 32 + lane;
    const int64_t mbeg = (int64_t)blockIdx.y * rows_per_split;
    const int64_t mend = (mbeg + rows_per_split < M) ? mbeg + rows_per_split : M;
    float s = 0.f;
    if (c < C)
        for (int64_t m = mbeg + warp; m < mend; m += 8) s += __ldg(dY + m * lddy + (int64_t)g * C + c);
    red[warp][lane] = s;
    __syncthreads();
    if (warp == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][lane];
        if (use_atomic)
            atomicAdd(out.p[g] + c, t);
        else
            out.p[g][c] = t;
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_bsgemm_nt(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, const float* x,
                              int ldx, int64_t m, const float* const* bias, float* y, int ldy,
                              hcgs_stream stream) {
    HCGS_REQUIRE(plan && w && x && y, "hcgs_bsgemm_nt: NULL argument");
    HCGS_REQUIRE(n_w >= 1 && n_w <= HCGS_MAX_GROUPS, "hcgs_bsgemm_nt: n_w=%d out of range", n_w);
    if (m <= 0) return 0;
    const int64_t mblocks = (m + BM - 1) / BM;
    HCGS_REQUIRE(mblocks <= 65535, "hcgs_bsgemm_nt: too many rows (%lld)", (long long)m);
    dim3 grid(plan->tr, (unsigned)mblocks, n_w);
    bsgemm_nt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(plan->d_row_ptr, plan->d_col_idx,
                                                            make_pack<const float>(w, n_w), ldw, x, ldx, m,
                                                            plan->n_out, plan->n_in,
                                                            make_pack<const float>(bias, n_w), y, ldy);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_bsgemm_nn(const hcgs_plan* plan, int n_w, const float* const* w, int ldw, const float* dy,
                              int lddy, int64_t m, float* dx, int lddx, hcgs_stream stream) {
    HCGS_REQUIRE(plan && w && dy && dx, "hcgs_bsgemm_nn: NULL argument");
    HCGS_REQUIRE(n_w >= 1 && n_w <= HCGS_MAX_GROUPS, "hcgs_bsgemm_nn: n_w=%d out of range", n_w);
    if (m <= 0) return 0;
    const int64_t mblocks = (m + BM - 1) / BM;
    HCGS_REQUIRE(mblocks <= 65535, "hcgs_bsgemm_nn: too many rows (%lld)", (long long)m);
    dim3 grid(plan->tc, (unsigned)mblocks, 1);
    bsgemm_nn_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(plan->d_col_ptr, plan->d_row_idx,
                                                            make_pack<const float>(w, n_w), n_w, ldw, dy, lddy,
                                                            m, plan->n_out, plan->n_in, dx, lddx);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_bs_sddmm(const hcgs_plan* plan, int n_w, const float* dy, int lddy, const float* x, int ldx,
                             int64_t m, float* const* dw, int lddw, const float* mask, hcgs_stream stream) {
    HCGS_REQUIRE(plan && dy && x && dw, "hcgs_bs_sddmm: NULL argument");
    HCGS_REQUIRE(n_w >= 1 && n_w <= HCGS_MAX_GROUPS, "hcgs_bs_sddmm: n_w=%d out of range", n_w);
    if (m <= 0 || plan->nz_tiles == 0) return 0;
    // enough CTAs to fill the machine: split M when there are few tiles
    int splits = 1;
    const int target = 148 * 4;
    if (plan->nz_tiles < target) splits = (target + plan->nz_tiles - 1) / plan->nz_tiles;
    int64_t rows = (m + splits - 1) / splits;
    if (rows < 256) rows = 256;
    rows = (rows + SM_ROWS - 1) / SM_ROWS * SM_ROWS;
    splits = (int)((m + rows - 1) / rows);
    HCGS_REQUIRE(splits <= 65535, "hcgs_bs_sddmm: too many splits");
    if (n_w == 4 || n_w == 1) {
        dim3 grid(plan->nz_tiles, splits, 1);
        bs_sddmm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(plan->d_tile_rc, n_w, dy, lddy, x, ldx, m, rows,
                                                               plan->n_out, plan->n_in,
                                                               make_pack<float>(dw, n_w), lddw, mask,
                                                               splits > 1 ? 1 : 0);
        HCGS_LAUNCH_CHECK();
    } else {
        for (int g = 0; g < n_w; ++g) {
            dim3 grid(plan->nz_tiles, splits, 1);
            float* one[1] = {dw[g]};
            bs_sddmm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
                plan->d_tile_rc, 1, dy + (int64_t)g * plan->n_out, lddy, x, ldx, m, rows, plan->n_out,
                plan->n_in, make_pack<float>(one, 1), lddw, mask, splits > 1 ? 1 : 0);
            HCGS_LAUNCH_CHECK();
        }
    }
    return 0;
}

extern "C" int hcgs_colsum(const float* dy, int lddy, int64_t m, int n_groups, int group_cols, float* const* out,
                           hcgs_stream stream) {
    HCGS_REQUIRE(dy && out, "hcgs_colsum: NULL argument");
    HCGS_REQUIRE(n_groups >= 1 && n_groups <= HCGS_MAX_GROUPS, "hcgs_colsum: n_groups out of range");
    if (m <= 0) return 0;
    int cb = (group_cols + 31) / 32;
    int splits = 1;
    const int target = 148 * 2;
    if (cb * n_groups < target) splits = (target + cb * n_groups - 1) / (cb * n_groups);
    int64_t rows = (m + splits - 1) / splits;
    if (rows < 64) rows = 64;
    splits = (int)((m + rows - 1) / rows);
    if (splits > 1)
        for (int g = 0; g < n_groups; ++g)
            HCGS_CHECK_CUDA(cudaMemsetAsync(out[g], 0, sizeof(float) * group_cols, (cudaStream_t)stream));
    dim3 grid(cb, splits, n_groups);
    colsum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dy, lddy, m, group_cols, rows,
                                                         make_pack<float>(out, n_groups), splits > 1 ? 1 : 0);
    HCGS_LAUNCH_CHECK();
    return 0;
}
