// Probe: issue rate and latency of mma.sync.m16n8k16 (bf16, fp32 accumulate) and ldmatrix.x4 on sm_100a, the two
// instructions the recurrent kernels are built from.  One CTA per SM, W warps; every warp runs `iters` rounds of
// K independent mma (different accumulators) or one dependent chain.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hmma_rate hmma_rate.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_pred(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, uint32_t on) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %10, 0;\n\t"
        "@p mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
        "{%0, %1, %2, %3};\n\t}"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(on));
}
__device__ __forceinline__ void ldsm(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}

template <int K, int MODE>  // MODE 0: K independent mma per round; 1: K ldmatrix per round; 2: ldmatrix -> mma pairs
__global__ void __launch_bounds__(1024, 1) rate_kernel(int iters, long long* out, float* sink) {
    __shared__ __align__(128) unsigned char tile[16 * 1168];
    for (int i = threadIdx.x; i < 16 * 1168 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(tile)[i] = 0x3c003c00u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(tile) + (uint32_t)((lane & 15) * 1168 + (lane >> 4) * 16);
    float acc[K][4];
    uint32_t a[4] = {0x3c003c00u, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u};
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int k = 0; k < K; ++k) mma(acc[k], a, 0x3c003c00u, 0x3c003c00u);
        } else if (MODE == 3) {  // K predicated mma, the predicate is false at run time
#pragma unroll
            for (int k = 0; k < K; ++k) mma_pred(acc[k], a, 0x3c003c00u, 0x3c003c00u, (uint32_t)(iters < 0));
        } else if (MODE == 4) {  // K predicated mma, predicate true
#pragma unroll
            for (int k = 0; k < K; ++k) mma_pred(acc[k], a, 0x3c003c00u, 0x3c003c00u, (uint32_t)(iters > 0));
        } else if (MODE == 1) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                uint32_t r[4];
                ldsm(r, base + (uint32_t)(((it + k) & 31) * 32));
                a[0] ^= r[0] ^ r[1] ^ r[2] ^ r[3];
            }
        } else {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                uint32_t r[4];
                ldsm(r, base + (uint32_t)(((it + k) & 31) * 32));
                mma(acc[k], r, 0x3c003c00u, 0x3c003c00u);
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) s += acc[k][0] + acc[k][3];
    if (s == 123.456f || a[0] == 77u) sink[0] = s;
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int K, int MODE>
static void run(int warps, const char* what) {
    long long* d;
    float* sink;
    cudaMalloc(&d, 148 * sizeof(long long));
    cudaMalloc(&sink, 4);
    const int iters = 2000;
    rate_kernel<K, MODE><<<148, warps * 32>>>(iters, d, sink);
    rate_kernel<K, MODE><<<148, warps * 32>>>(iters, d, sink);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
    const double per_round = (double)mx / iters;
    const double per_smsp = per_round / ((double)K * warps / 4.0);
    printf("%-28s K=%d warps=%2d: %7.1f cycles/round, %6.2f cycles per instruction per SM sub-partition %s\n", what, K,
           warps, per_round, per_smsp, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(d);
    cudaFree(sink);
}

int main() {
    run<1, 0>(4, "mma dependent chain");
    run<8, 0>(4, "mma independent");
    run<8, 0>(8, "mma independent");
    run<8, 0>(16, "mma independent");
    run<4, 0>(20, "mma independent");
    run<8, 3>(16, "mma predicated off");
    run<8, 4>(16, "mma predicated on");
    run<1, 1>(4, "ldmatrix.x4 (dependent xor)");
    run<4, 1>(4, "ldmatrix.x4");
    run<4, 1>(16, "ldmatrix.x4");
    run<4, 2>(4, "ldmatrix.x4 -> mma");
    run<4, 2>(16, "ldmatrix.x4 -> mma");
    return 0;
}
