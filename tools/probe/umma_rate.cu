// Probe: cycles per tcgen05.mma (M=128, K=16, bf16, SS operands) as a function of N and of the shared-memory
// layout of the operands (SWIZZLE_64B rows of 32 k vs SWIZZLE_128B rows of 64 k), issued back to back by one
// elected lane of a convergent warp.  One CTA per SM; operands are zeros (layout / timing only).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t addr, uint32_t sbo, uint32_t layout) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((16 >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(layout & 7) << 61;
    return d;
}
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

template <int N, int SW128, int ALT>
__global__ void __launch_bounds__(128) probe(long long* out, int reps) {
    extern __shared__ unsigned char raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < (64 * 1024) / 4; i += 128) ((uint32_t*)smem)[i] = 0;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = __shfl_sync(0xffffffffu, slot, 0);
    if (warp == 1) {
        const uint32_t sbo = SW128 ? 1024 : 512, lay = SW128 ? 2 : 4;
        const uint64_t ad = desc(smem_u32(smem), sbo, lay), bd = desc(smem_u32(smem) + 32768, sbo, lay);
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const bool issuer = elect_one();
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            if (issuer) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {  // 4 instructions per trip; ALT: alternate two accumulators
                    const uint32_t d = tmem + (ALT ? (uint32_t)((k & 1) * 256) : 0u);
                    const uint64_t a2 = ad + 2 * (SW128 ? k : (k & 1)), b2 = bd + 2 * (SW128 ? k : (k & 1));
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a2), "l"(b2),
                                 "r"(idesc), "r"(1u)
                                 : "memory");
                }
            }
            __syncwarp();
        }
        if (issuer)
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar))
                         : "memory");
        __syncwarp();
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar)) : "memory");
        const long long t1 = clock64();
        if (issuer && blockIdx.x == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int N, int SW128, int ALT>
static void run(const char* name) {
    long long* d;
    cudaMalloc(&d, 8);
    const int reps = 2000, smem = 66 * 1024;
    cudaFuncSetAttribute(probe<N, SW128, ALT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int w = 0; w < 2; ++w) probe<N, SW128, ALT><<<148, 128, smem>>>(d, reps);
    long long h = 0;
    cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %7.1f cycles per tcgen05.mma (floor 128*N/256 = %d) %s\n", name, (double)h / (reps * 4.0), N / 2,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    run<64, 0, 0>("N=64  SWIZZLE_64B  (32-wide k tiles)");
    run<64, 1, 0>("N=64  SWIZZLE_128B (64-wide k tiles)");
    run<64, 0, 1>("N=64  SWIZZLE_64B  two accumulators");
    run<128, 0, 0>("N=128 SWIZZLE_64B");
    run<128, 1, 0>("N=128 SWIZZLE_128B");
    run<256, 0, 0>("N=256 SWIZZLE_64B");
    run<256, 1, 0>("N=256 SWIZZLE_128B");
    run<32, 0, 0>("N=32  SWIZZLE_64B");
    run<32, 1, 0>("N=32  SWIZZLE_128B");
    return 0;
}
