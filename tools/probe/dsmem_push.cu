// Probe: cost of the reduce-scatter push of the recurrent backward kernel over DSMEM, by store form.
// A cluster of 4 CTAs; per iteration every CTA sends (CL-1) * SPC strips of 256 B (16 rows x 4 fp32) to its
// peers (26 KB for SPC = 36, the H = 550 case), completes transaction bytes on the peers' mbarrier, waits for
// its own slots to fill and hands them back (consumed handshake), like lstm_tc_bwd.  33 clusters run at once.
//   mode 0: st.async.v2.b32 (8 B per lane, 16 lanes per instruction, 16-byte row pieces: the round-1 kernel)
//   mode 1: st.async.v4.b32 (16 B per lane, 16 lanes, one 256-byte strip per instruction)
//   mode 2: st.async.v4.b32 (16 B per lane, 32 lanes, two adjacent strips = 512 contiguous bytes)
//   mode 3: staging in local shared memory + one cp.async.bulk shared::cta -> shared::cluster per warp
//   mode 4: staging + one cp.async.bulk per destination (SPC * 256 B each)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dsmem_push dsmem_push.cu
#include <cooperative_groups.h>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

constexpr int CL = 4, SPC = 36, CW = 18, THREADS = (CW + 1) * 32;
constexpr int STRIP_B = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t c) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(c) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t a, uint32_t b) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void mbar_arrive_local(uint32_t a) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t a) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity) {
    uint32_t ok = 0;
    for (int i = 0; i < (1 << 22) && !ok; ++i)
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(a), "r"(parity), "r"(20000u)
            : "memory");
    if (!ok) __trap();
}
__device__ __forceinline__ void st_async_v2(uint32_t ra, float x, float y, uint32_t rb) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(ra),
                 "r"(__float_as_uint(x)), "r"(__float_as_uint(y)), "r"(rb)
                 : "memory");
}
__device__ __forceinline__ void st_async_v4(uint32_t ra, float4 v, uint32_t rb) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(ra),
                 "r"(__float_as_uint(v.x)), "r"(__float_as_uint(v.y)), "r"(__float_as_uint(v.z)),
                 "r"(__float_as_uint(v.w)), "r"(rb)
                 : "memory");
}
__device__ __forceinline__ void bulk_s2c(uint32_t dst, uint32_t src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "r"(src), "r"(bytes), "r"(mbar)
                 : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) push_kernel(int iters, long long* cycles, float* sink) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    extern __shared__ __align__(128) unsigned char smem[];
    float* slots = reinterpret_cast<float*>(smem);                                    // [CL][SPC][64]
    float* stage = reinterpret_cast<float*>(smem + (size_t)CL * SPC * STRIP_B);       // [CL-1][SPC][64]
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + (size_t)(2 * CL - 1) * SPC * STRIP_B);
    const uint32_t mb_full = smem_u32(mbar), mb_free = mb_full + 8;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < (2 * CL - 1) * SPC * 64; i += THREADS) slots[i] = 1.f;
    if (tid == 0) {
        mbar_init(mb_full, CW + 1);
        mbar_init(mb_free, CL - 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    cluster.sync();
    const uint32_t slots_addr = smem_u32(slots), stage_addr = smem_u32(stage);
    const int br = lane >> 2, q = lane & 3;
    float acc = (float)tid;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (warp < CW) {
            if (it > 0) mbar_wait(mb_free, (uint32_t)((it - 1) & 1));
            // foreign strips f = 0 .. (CL-1)*SPC-1; destination d = (rank + 1 + f / SPC) % CL, local strip f % SPC
            constexpr int NF = (CL - 1) * SPC;
            constexpr int PER = (NF + CW - 1) / CW;  // 6 strips per warp
            const int f0 = warp * PER, f1 = min(NF, f0 + PER);
            if (MODE == 0) {
                for (int f = f0; f < f1; ++f) {
                    const int d = (rank + 1 + f / SPC) % CL, sl = f % SPC;
                    if (q < 2) {
                        const uint32_t off = (uint32_t)(((rank * 16 + br) * (SPC * 4) + sl * 4 + q * 2) * 4);  // row-major [16][Hc]
                        const uint32_t ra = mapa_u32(slots_addr + off, d), rb = mapa_u32(mb_full, d);
                        st_async_v2(ra, acc, acc + 1.f, rb);
                        st_async_v2(ra + 8 * SPC * 4 * 4, acc + 2.f, acc + 3.f, rb);
                    }
                }
            } else if (MODE == 1) {
                for (int f = f0; f < f1; ++f) {
                    const int d = (rank + 1 + f / SPC) % CL, sl = f % SPC;
                    // the two shuffles that turn the accumulator layout into 16-byte row pieces
                    const float x = __shfl_xor_sync(0xffffffffu, acc, 1), y = __shfl_xor_sync(0xffffffffu, acc + 1.f, 1);
                    if (q < 2) {
                        const uint32_t off = (uint32_t)((rank * SPC + sl) * STRIP_B + (br + 8 * q) * 16);
                        st_async_v4(mapa_u32(slots_addr + off, d), make_float4(acc, acc + 1.f, x, y), mapa_u32(mb_full, d));
                    }
                }
            } else if (MODE == 2) {
                for (int f = f0; f < f1; f += 2) {
                    const float x = __shfl_xor_sync(0xffffffffu, acc, 1), y = __shfl_xor_sync(0xffffffffu, acc + 1.f, 1);
                    const float x2 = __shfl_xor_sync(0xffffffffu, acc, 2), y2 = __shfl_xor_sync(0xffffffffu, acc + 1.f, 2);
                    const int ff = f + (q >> 1);
                    if (ff < f1) {
                        const int d = (rank + 1 + ff / SPC) % CL, sl = ff % SPC;
                        const uint32_t off = (uint32_t)((rank * SPC + sl) * STRIP_B + (br + 8 * (q & 1)) * 16);
                        st_async_v4(mapa_u32(slots_addr + off, d), make_float4(acc, x2, x, y + y2), mapa_u32(mb_full, d));
                    }
                }
            } else {
                // stage: two 8-byte stores per strip from 16 lanes (the accumulator layout), then bulk copies
                for (int f = f0; f < f1; ++f)
                    if (q < 2) {
                        float* p = stage + (size_t)f * 64 + br * 4 + q * 2;
                        *reinterpret_cast<float2*>(p) = make_float2(acc, acc + 1.f);
                        *reinterpret_cast<float2*>(p + 32) = make_float2(acc + 2.f, acc + 3.f);
                    }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                if (MODE == 3) {
                    __syncwarp();
                    if (lane == 0 && f1 > f0) {
                        // PER divides SPC here, so a warp's strips have one destination
                        const int d = (rank + 1 + f0 / SPC) % CL, sl = f0 % SPC;
                        bulk_s2c(mapa_u32(slots_addr + (uint32_t)((rank * SPC + sl) * STRIP_B), d),
                                 stage_addr + (uint32_t)f0 * STRIP_B, (uint32_t)(f1 - f0) * STRIP_B, mapa_u32(mb_full, d));
                    }
                } else {
                    asm volatile("bar.sync 1, %0;" ::"r"(CW * 32) : "memory");
                    if (warp == 0 && lane < CL - 1) {
                        const int d = (rank + 1 + lane) % CL;
                        bulk_s2c(mapa_u32(slots_addr + (uint32_t)(rank * SPC * STRIP_B), d),
                                 stage_addr + (uint32_t)(lane * SPC) * STRIP_B, (uint32_t)SPC * STRIP_B, mapa_u32(mb_full, d));
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive_local(mb_full);
            mbar_wait(mb_full, (uint32_t)(it & 1));
            // "consume": one read per thread
            acc += slots[(tid * 7 + it) % (CL * SPC * 64)] * 1e-9f;
        } else if (lane == 0) {
            mbar_expect_tx(mb_full, (uint32_t)(CL - 1) * SPC * STRIP_B);
            mbar_wait(mb_full, (uint32_t)(it & 1));
            // everything of this iteration has landed and (approximately) been read: hand the slots back
            for (int d = 1; d < CL; ++d) mbar_arrive_remote(mapa_u32(mb_free, (uint32_t)((rank + d) % CL)));
        }
    }
    long long t1 = clock64();
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
    if (acc == 12345.678f) sink[0] = acc;
    cluster.sync();
}

template <int MODE>
static void run(int iters, const char* name) {
    const size_t smem = (size_t)(2 * CL - 1) * SPC * STRIP_B + 64;
    cudaFuncSetAttribute(push_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024));
    long long* d_cyc;
    float* d_sink;
    const int n_clusters = 33;
    cudaMalloc(&d_cyc, sizeof(long long) * n_clusters * CL);
    cudaMalloc(&d_sink, 4);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(n_clusters * CL, 1, 1);
    cfg.blockDim = dim3(THREADS, 1, 1);
    cfg.dynamicSmemBytes = 200 * 1024;  // one CTA per SM, like the real kernel
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    (void)smem;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        cudaError_t e = cudaLaunchKernelEx(&cfg, push_kernel<MODE>, iters, d_cyc, d_sink);
        cudaEventRecord(e1);
        cudaError_t e2 = cudaDeviceSynchronize();
        if (e != cudaSuccess || e2 != cudaSuccess) {
            printf("%s: %s / %s\n", name, cudaGetErrorString(e), cudaGetErrorString(e2));
            return;
        }
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[33 * CL];
    cudaMemcpy(h, d_cyc, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0, mn = 1LL << 60;
    for (int i = 0; i < n_clusters * CL; ++i) {
        mx = h[i] > mx ? h[i] : mx;
        mn = h[i] < mn ? h[i] : mn;
    }
    const double bytes = (double)(CL - 1) * SPC * STRIP_B;
    printf("%-44s %7.0f cycles/iter (min %7.0f)  %.2f us/iter  %.1f B/cycle/CTA sent\n", name, (double)mx / iters,
           (double)mn / iters, ms * 1e3 / iters, bytes / ((double)mx / iters));
    cudaFree(d_cyc);
    cudaFree(d_sink);
}

int main() {
    const int iters = 2000;
    run<0>(iters, "0 st.async.v2 8B x16 lanes (row pieces)");
    run<1>(iters, "1 st.async.v4 16B x16 lanes (256B strip)");
    run<2>(iters, "2 st.async.v4 16B x32 lanes (512B)");
    run<3>(iters, "3 stage + cp.async.bulk per warp (1.5KB)");
    run<4>(iters, "4 stage + cp.async.bulk per dest (9KB)");
    return 0;
}
