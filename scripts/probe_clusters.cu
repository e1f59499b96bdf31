// How many 8-CTA clusters with ~220 KB of shared memory per CTA can be resident at once on this device?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    for (int cs : {2, 4, 8}) for (int kb : {200, 216, 224}) {
        cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024);
        cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(128); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = kb * 1024;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int nc = -1; cudaError_t e = cudaOccupancyMaxActiveClusters(&nc, k, &cfg);
        printf("cluster %d smem %d KB: max active clusters %d (%s) -> %d CTAs\n", cs, kb, nc, cudaGetErrorString(e), nc * cs);
    }
    return 0;
}
