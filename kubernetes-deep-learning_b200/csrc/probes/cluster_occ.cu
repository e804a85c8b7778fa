// Probe: how many clusters of 2 / 3 / 4 / 8 CTAs (one CTA per SM: 220 KB dynamic smem, 512 threads) can be co-resident?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 probes/cluster_occ.cu -o probes/cluster_occ
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int *p) { extern __shared__ char s[]; if (p && threadIdx.x == 9999) p[0] = s[0]; }
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {1, 2, 3, 4, 6, 8}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cs * 64); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = 220 * 1024;
        cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension; at.val.clusterDim.x = cs; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
        cfg.attrs = &at; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
        printf("cluster size %d: max active clusters %d (%d SMs)  %s\n", cs, n, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
    return 0;
}
