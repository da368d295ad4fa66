// How many thread-block clusters of a given size can be resident on this GPU with the SYRK kernel's footprint
// (256 threads, ~195 KB dynamic smem)?  Answers whether a 4-CTA multicast variant could still cover all 148 SMs.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256, 1) probe_kernel(int* out) {
    extern __shared__ unsigned char smem[];
    if (out != nullptr && threadIdx.x == 0) out[blockIdx.x] = smem[0];
}
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int smem = 6 * 32768 + 1280;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, p.multiProcessorCount);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t lc{};
        lc.gridDim = dim3(cs * 64);
        lc.blockDim = dim3(256);
        lc.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = cs;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        lc.attrs = attr;
        lc.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, probe_kernel, &lc);
        printf(", \"clusters_of_%d\": %d, \"ctas_%d\": %d", cs, e == cudaSuccess ? n : -1, cs, e == cudaSuccess ? n * cs : -1);
        cudaGetLastError();
    }
    printf("}\n");
    return 0;
}
