// probe: how many clusters of size 2/4/8/16 of a 1-CTA-per-SM kernel (200 KB smem, 320 threads) can be co-resident
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* p) { extern __shared__ unsigned char s[]; if (p) p[0] = s[0]; }
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cs * 64); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute a[1];
        a[0].id = cudaLaunchAttributeClusterDimension; a[0].val.clusterDim.x = cs; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
        cfg.attrs = a; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
        printf("cluster %2d: max active clusters %d (%d CTAs) %s\n", cs, n, n * cs, cudaGetErrorString(e));
    }
    return 0;
}
