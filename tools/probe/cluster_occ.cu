// Probe: how many clusters of a 128-thread kernel are co-resident as a function of cluster size and dynamic smem.
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(128) k(int* p) { extern __shared__ int s[]; if (p) p[0] = s[0]; }
int main() {
    const int sizes[] = {16, 48, 56, 72, 100, 108, 112, 160, 200};
    for (int cl : {1, 2, 4, 8}) {
        for (int kb : sizes) {
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024);
            cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof(cfg));
            cfg.gridDim = dim3(cl * 1024, 1, 1); cfg.blockDim = dim3(128, 1, 1); cfg.dynamicSmemBytes = kb * 1024;
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = -1; cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
            int nb = -1; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, 128, kb * 1024);
            printf("cluster %d smem %3d KB: max active clusters %d (CTAs %d), blocks/SM %d %s\n", cl, kb, n, n * cl, nb,
                   e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
    }
    return 0;
}
