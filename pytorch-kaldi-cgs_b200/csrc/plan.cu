// Plan = compressed block index built once from an HCGS / pattern / arbitrary 0-1 mask.
//
//  (a) 32x32-tile CSR + CSC + flat list  -> block-sparse GEMM family (gemm.cu)
//  (b) 4x4-block ELL ("BELL-4") per cluster CTA -> persistent recurrent kernels (lstm.cu)
//      HCGS masks in every shipped cfg end in an inner block of 4 (hcgsh_block = 32,4 / 64,4 /
//      128,4), so a 4x4 block is either fully kept or fully dropped; other masks are covered
//      with zero fill (correct, just less efficient).
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace hcgs {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

template <typename T>
static int upload(const std::vector<T>& v, T** dptr) {
    *dptr = nullptr;
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    HCGS_CHECK_CUDA(cudaMalloc((void**)dptr, bytes));
    if (!v.empty()) HCGS_CHECK_CUDA(cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return 0;
}

size_t lstm_fwd_smem_bytes(int H, int CL, int Hc, int HcP, int L, bool w_in_smem);
size_t lstm_bwd_smem_bytes(int H, int CL, int Hc, int max_blocks, int n_colgroups, bool w_in_smem);

static int round_up(int a, int b) { return (a + b - 1) / b * b; }

// Build the recurrent (BELL-4) part for a given cluster size.  Returns host vectors.
struct RecHost {
    int CL, Hc, HcP, RG, L;
    std::vector<uint16_t> bidx;      // [CL][L][RG]  column group or 0xFFFF
    std::vector<int> cta_blk_off;    // [CL+1] offsets into the flat real-block list
    std::vector<uint16_t> blk_a;     // [nblk] local row group of real block s
    std::vector<uint16_t> blk_c;     // [nblk] column group of real block s
    std::vector<int> cg_ptr;         // [CL][NCG+1] per CTA CSR over column groups -> block slots (local)
    std::vector<uint16_t> cg_list;   // [nblk] local block slot ids grouped by column group
    int max_cta_blocks;
    int NCG;
};

static void build_rec(const float* mask, int H, int CL, RecHost& r) {
    r.CL = CL;
    int per = (H + CL - 1) / CL;
    r.Hc = round_up(per, 4);
    r.HcP = r.Hc;
    while (r.HcP % 4 != 1) ++r.HcP;
    r.RG = r.Hc / 4;
    r.NCG = (H + 3) / 4;
    // block lists per global row group
    int nRGtot = CL * r.RG;
    std::vector<std::vector<uint16_t>> lists(nRGtot);
    int L = 0;
    for (int A = 0; A < nRGtot; ++A) {
        int r0 = A * 4;
        if (r0 >= H) continue;
        int r1 = std::min(H, r0 + 4);
        for (int c = 0; c < r.NCG; ++c) {
            int c0 = c * 4, c1 = std::min(H, c0 + 4);
            bool any = false;
            for (int i = r0; i < r1 && !any; ++i)
                for (int j = c0; j < c1; ++j)
                    if (mask[(size_t)i * H + j] != 0.f) { any = true; break; }
            if (any) lists[A].push_back((uint16_t)c);
        }
        L = std::max<int>(L, (int)lists[A].size());
    }
    r.L = std::max(L, 1);
    r.bidx.assign((size_t)CL * r.L * r.RG, 0xFFFF);
    r.cta_blk_off.assign(CL + 1, 0);
    r.blk_a.clear();
    r.blk_c.clear();
    r.cg_ptr.assign((size_t)CL * (r.NCG + 1), 0);
    r.cg_list.clear();
    r.max_cta_blocks = 0;
    for (int cta = 0; cta < CL; ++cta) {
        r.cta_blk_off[cta] = (int)r.blk_a.size();
        std::vector<std::vector<uint16_t>> by_cg(r.NCG);
        int slot = 0;
        for (int a = 0; a < r.RG; ++a) {
            const auto& l = lists[cta * r.RG + a];
            for (int j = 0; j < (int)l.size(); ++j) {
                r.bidx[((size_t)cta * r.L + j) * r.RG + a] = l[j];
                r.blk_a.push_back((uint16_t)a);
                r.blk_c.push_back(l[j]);
                by_cg[l[j]].push_back((uint16_t)slot);
                ++slot;
            }
        }
        r.max_cta_blocks = std::max(r.max_cta_blocks, slot);
        int* ptr = &r.cg_ptr[(size_t)cta * (r.NCG + 1)];
        int run = 0;
        for (int c = 0; c < r.NCG; ++c) {
            ptr[c] = run;
            for (uint16_t s : by_cg[c]) r.cg_list.push_back(s);
            run += (int)by_cg[c].size();
        }
        ptr[r.NCG] = run;
    }
    r.cta_blk_off[CL] = (int)r.blk_a.size();
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_abi_version(void) { return HCGS_ABI_VERSION; }

extern "C" const char* hcgs_last_error(void) { return hcgs::g_err; }

extern "C" int hcgs_device_info(int device, int* info) {
    HCGS_REQUIRE(info != nullptr, "hcgs_device_info: info is NULL");
    cudaDeviceProp p;
    HCGS_CHECK_CUDA(cudaGetDeviceProperties(&p, device));
    info[0] = p.multiProcessorCount;
    info[1] = p.major;
    info[2] = p.minor;
    info[3] = (int)p.sharedMemPerBlockOptin;
    int cl = 0;
    HCGS_CHECK_CUDA(cudaDeviceGetAttribute(&cl, cudaDevAttrClusterLaunch, device));
    info[4] = cl;
    return 0;
}

extern "C" int hcgs_plan_create(const float* mask, int n_out, int n_in, int recurrent, int cluster_size,
                                hcgs_plan** out) {
    HCGS_REQUIRE(out != nullptr && mask != nullptr, "hcgs_plan_create: NULL argument");
    HCGS_REQUIRE(n_out > 0 && n_in > 0, "hcgs_plan_create: bad shape %d x %d", n_out, n_in);
    HCGS_REQUIRE(!recurrent || n_out == n_in, "hcgs_plan_create: recurrent plan needs a square mask");
    HCGS_REQUIRE(!recurrent || n_out <= 65532, "hcgs_plan_create: hidden size too large");
    hcgs_plan* p = new hcgs_plan();
    p->n_out = n_out;
    p->n_in = n_in;
    p->tr = (n_out + kTile - 1) / kTile;
    p->tc = (n_in + kTile - 1) / kTile;
    std::vector<char> nz((size_t)p->tr * p->tc, 0);
    int64_t nnz = 0;
    for (int i = 0; i < n_out; ++i)
        for (int j = 0; j < n_in; ++j)
            if (mask[(size_t)i * n_in + j] != 0.f) {
                nz[(size_t)(i / kTile) * p->tc + j / kTile] = 1;
                ++nnz;
            }
    p->nnz = nnz;
    std::vector<int> row_ptr(p->tr + 1, 0), col_idx, col_ptr(p->tc + 1, 0), row_idx, tile_rc;
    for (int i = 0; i < p->tr; ++i) {
        row_ptr[i] = (int)col_idx.size();
        for (int j = 0; j < p->tc; ++j)
            if (nz[(size_t)i * p->tc + j]) {
                col_idx.push_back(j);
                tile_rc.push_back(i);
                tile_rc.push_back(j);
            }
        p->max_row_tiles = std::max(p->max_row_tiles, (int)col_idx.size() - row_ptr[i]);
    }
    row_ptr[p->tr] = (int)col_idx.size();
    for (int j = 0; j < p->tc; ++j) {
        col_ptr[j] = (int)row_idx.size();
        for (int i = 0; i < p->tr; ++i)
            if (nz[(size_t)i * p->tc + j]) row_idx.push_back(i);
        p->max_col_tiles = std::max(p->max_col_tiles, (int)row_idx.size() - col_ptr[j]);
    }
    col_ptr[p->tc] = (int)row_idx.size();
    p->nz_tiles = (int)col_idx.size();
    std::vector<int> csc_rc, sd_items;
    for (int j = 0; j < p->tc; ++j)
        for (int e = col_ptr[j]; e < col_ptr[j + 1]; ++e) {
            csc_rc.push_back(row_idx[e]);
            csc_rc.push_back(j);
        }
    for (int i = 0; i < p->tr; ++i)
        for (int t0 = row_ptr[i]; t0 < row_ptr[i + 1]; t0 += kSdMaxTiles) {
            sd_items.push_back(i);
            sd_items.push_back(t0);
            sd_items.push_back(std::min(kSdMaxTiles, row_ptr[i + 1] - t0));
        }
    p->n_sd_items = (int)sd_items.size() / 3;
    int rc = 0;
    rc |= upload(csc_rc, &p->d_csc_rc);
    rc |= upload(sd_items, &p->d_sd_items);
    rc |= upload(row_ptr, &p->d_row_ptr);
    rc |= upload(col_idx, &p->d_col_idx);
    rc |= upload(col_ptr, &p->d_col_ptr);
    rc |= upload(row_idx, &p->d_row_idx);
    rc |= upload(tile_rc, &p->d_tile_rc);
    {
        // fused output-layer index (strips of 64 output columns x 32-wide k tiles)
        const int strips = (n_out + 63) / 64;
        int C = 1;
        while (C < 8 && C * 4 < strips) C *= 2;
        const int S = (strips + C - 1) / C;
        // n_in <= 4096 (128 k tiles per CTA, OUT_MAX_ITEMS in gemm_tc.cu) and n_out <= 4096; anything wider keeps
        // out_C = 0 and runs as hcgs_tc_gemm_nt + hcgs_logsoftmax_fwd
        if (S <= 8 && p->tc <= 128) {
            std::vector<int> optr(C + 1, 0), oitems, otile;
            for (int j = 0; j < C; ++j) {
                optr[j] = (int)oitems.size() / 3;
                // k tiles kept by any strip of this CTA.  The FIRST item multiplies all strips 0..hi (masked weights
                // inside are zeros) so that every accumulator strip is initialised by it; the following items carry
                // only their kept strips, and the kernel multiplies each run of adjacent kept strips with one wide MMA
                // (the issue rate of the single MMA thread and the bytes streamed per row tile bound this kernel)
                std::vector<int> kts;
                std::vector<uint32_t> kms;
                int hi = -1;
                for (int kt = 0; kt < p->tc; ++kt) {
                    uint32_t km = 0;
                    for (int s = 0; s < S; ++s)
                        for (int h = 0; h < 2; ++h) {
                            const int trow = (j * S + s) * 2 + h;
                            if (trow < p->tr && nz[(size_t)trow * p->tc + kt]) {
                                km |= 1u << s;
                                hi = std::max(hi, s);
                            }
                        }
                    if (km) {
                        kts.push_back(kt);
                        kms.push_back(km);
                    }
                }
                const uint32_t m = hi < 0 ? 0u : (uint32_t)((1u << (hi + 1)) - 1u);
                for (size_t q = 0; q < kts.size(); ++q) {
                    const uint32_t im = q == 0 ? m : kms[q];
                    oitems.push_back(kts[q]);
                    oitems.push_back((int)im);
                    oitems.push_back((int)otile.size() / 2);
                    for (int s = 0; s <= hi; ++s)
                        if (im >> s & 1) {
                            otile.push_back(j * S + s);
                            otile.push_back(kts[q]);
                        }
                }
                p->out_any[j] = m;
            }
            optr[C] = (int)oitems.size() / 3;
            p->out_C = C;
            p->out_S = S;
            p->out_items = optr[C];
            p->out_tiles = (int)otile.size() / 2;
            rc |= upload(optr, &p->d_out_ptr);
            rc |= upload(oitems, &p->d_out_items);
            rc |= upload(otile, &p->d_out_tile);
        }
    }
    if (rc) {
        hcgs_plan_destroy(p);
        return 1;
    }

    if (recurrent) {
        const int H = n_out;
        int dev = 0, smem_max = 232448;
        if (cudaGetDevice(&dev) == cudaSuccess)
            cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        RecHost rh;
        bool fits = false;
        const int cands[5] = {1, 2, 4, 8, 16};
        if (cluster_size > 0) {
            HCGS_REQUIRE(cluster_size == 1 || cluster_size == 2 || cluster_size == 4 || cluster_size == 8 ||
                             cluster_size == 16,
                         "hcgs_plan_create: cluster_size must be 1,2,4,8 or 16");
            build_rec(mask, H, cluster_size, rh);
            fits = lstm_fwd_smem_bytes(H, rh.CL, rh.Hc, rh.HcP, rh.L, true) <= (size_t)smem_max &&
                   lstm_bwd_smem_bytes(H, rh.CL, rh.Hc, rh.max_cta_blocks, rh.NCG, true) <= (size_t)smem_max &&
                   rh.Hc * 4 <= 512;
        } else {
            for (int ci = 0; ci < 5 && !fits; ++ci) {
                build_rec(mask, H, cands[ci], rh);
                fits = lstm_fwd_smem_bytes(H, rh.CL, rh.Hc, rh.HcP, rh.L, true) <= (size_t)smem_max &&
                       lstm_bwd_smem_bytes(H, rh.CL, rh.Hc, rh.max_cta_blocks, rh.NCG, true) <= (size_t)smem_max &&
                       rh.Hc * 4 <= 512;
            }
            if (!fits) build_rec(mask, H, 8, rh);  // weights stay in global memory (L2-resident)
        }
        HCGS_REQUIRE(rh.Hc * 4 <= 512, "hcgs_plan_create: %d rows per CTA exceed the block size; use a larger cluster",
                     rh.Hc);
        RecurrentPlan& rp = p->rp;
        rp.CL = rh.CL;
        rp.Hc = rh.Hc;
        rp.HcP = rh.HcP;
        rp.RG = rh.RG;
        rp.L = rh.L;
        rp.NCG = rh.NCG;
        rp.max_cta_blocks = rh.max_cta_blocks;
        rp.n_blocks = (int)rh.blk_a.size();
        rp.w_in_smem = fits;
        rp.fwd_pack_floats = (int64_t)rh.CL * rh.L * rh.RG * 64;
        rp.bwd_pack_floats = (int64_t)rp.n_blocks * 64;
        rp.smem_fwd = lstm_fwd_smem_bytes(H, rh.CL, rh.Hc, rh.HcP, rh.L, fits);
        rp.smem_bwd = lstm_bwd_smem_bytes(H, rh.CL, rh.Hc, rh.max_cta_blocks, rh.NCG, fits);
        HCGS_REQUIRE(rp.smem_fwd <= (size_t)smem_max && rp.smem_bwd <= (size_t)smem_max,
                     "hcgs_plan_create: recurrent state does not fit shared memory (H=%d)", H);
        rp.threads = std::max(round_up(rh.RG * 16, 32), 64);
        rc |= upload(rh.bidx, &rp.d_bidx);
        rc |= upload(rh.cta_blk_off, &rp.d_cta_blk_off);
        rc |= upload(rh.blk_a, &rp.d_blk_a);
        rc |= upload(rh.blk_c, &rp.d_blk_c);
        rc |= upload(rh.cg_ptr, &rp.d_cg_ptr);
        rc |= upload(rh.cg_list, &rp.d_cg_list);
        if (rc) {
            hcgs_plan_destroy(p);
            return 1;
        }
        p->ns44 = (H + 3) / 4;
        p->h_blk44 = (uint8_t*)calloc((size_t)p->ns44 * p->ns44, 1);
        HCGS_REQUIRE(p->h_blk44 != nullptr, "hcgs_plan_create: out of host memory");
        for (int i = 0; i < H; ++i)
            for (int j = 0; j < H; ++j)
                if (mask[(size_t)i * H + j] != 0.f) p->h_blk44[(size_t)(i / 4) * p->ns44 + j / 4] = 1;
        p->recurrent = true;
    }
    *out = p;
    return 0;
}

extern "C" int hcgs_plan_destroy(hcgs_plan* p) {
    if (!p) return 0;
    cudaFree(p->d_row_ptr);
    cudaFree(p->d_col_idx);
    cudaFree(p->d_col_ptr);
    cudaFree(p->d_row_idx);
    cudaFree(p->d_tile_rc);
    cudaFree(p->d_csc_rc);
    cudaFree(p->d_sd_items);
    cudaFree(p->d_out_ptr);
    cudaFree(p->d_out_items);
    cudaFree(p->d_out_tile);
    cudaFree(p->rp.d_bidx);
    cudaFree(p->rp.d_cta_blk_off);
    cudaFree(p->rp.d_blk_a);
    cudaFree(p->rp.d_blk_c);
    cudaFree(p->rp.d_cg_ptr);
    cudaFree(p->rp.d_cg_list);
    for (int i = 0; i < 3; ++i)
        if (p->tcp[i]) tc_plan_free(p->tcp[i]);
    free(p->h_blk44);
    delete p;
    return 0;
}

extern "C" int hcgs_plan_info(const hcgs_plan* p, int64_t* info) {
    HCGS_REQUIRE(p && info, "hcgs_plan_info: NULL argument");
    for (int i = 0; i < 16; ++i) info[i] = 0;
    info[0] = p->n_out;
    info[1] = p->n_in;
    info[2] = p->nz_tiles;
    info[3] = (int64_t)p->tr * p->tc;
    info[4] = p->nnz;
    if (p->recurrent) {
        info[5] = p->rp.CL;
        info[6] = p->rp.Hc;
        info[7] = p->rp.L;
        info[8] = p->rp.fwd_pack_floats;
        info[9] = p->rp.bwd_pack_floats;
        info[10] = (int64_t)p->rp.smem_fwd;
        info[11] = (int64_t)p->rp.smem_bwd;
        info[12] = p->rp.threads;
        info[13] = p->rp.n_blocks;
        info[14] = p->rp.w_in_smem ? 1 : 0;
    }
    return 0;
}
