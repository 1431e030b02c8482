#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* p) { extern __shared__ float s[]; if (p) p[0] = s[0]; }
int main() {
  const int smem = 226 * 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int C : {2, 4, 8}) {
    for (int pol = 0; pol < 3; ++pol) {
      cudaLaunchConfig_t cfg = {};
      cfg.blockDim = dim3(576); cfg.gridDim = dim3(C); cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute a[2];
      a[0].id = cudaLaunchAttributeClusterDimension; a[0].val.clusterDim.x = C; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
      a[1].id = cudaLaunchAttributeClusterSchedulingPolicyPreference;
      a[1].val.clusterSchedulingPolicyPreference = (cudaClusterSchedulingPolicy)pol;
      cfg.attrs = a; cfg.numAttrs = pol == 0 ? 1 : 2;
      int n = -1;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
      printf("C=%d policy=%d -> max active clusters %d (%s) = %d SMs\n", C, pol, n, cudaGetErrorString(e), n * C);
    }
  }
  return 0;
}
