// Where does the block scheduler put the CTAs of a 148 x 3 grid (128 threads, 44 KB dynamic smem)?
// Prints the SM id of CTA b for b = 0..gridDim-1 and the number of distinct SMs among CTAs 0..147.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void where(int* smid, long long spin) {
  extern __shared__ char s[];
  if (threadIdx.x == 0) {
    unsigned id;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(id));
    smid[blockIdx.x] = (int)id;
  }
  long long t0 = clock64();
  while (clock64() - t0 < spin) {}
  if (spin < 0) s[threadIdx.x] = 1;
}
int main() {
  int nsm; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  const int grid = nsm * 3; const size_t smem = 44 * 1024;
  cudaFuncSetAttribute(where, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int* d; cudaMalloc(&d, grid * sizeof(int));
  where<<<grid, 128, smem>>>(d, 2000000);
  cudaDeviceSynchronize();
  int h[1024]; cudaMemcpy(h, d, grid * sizeof(int), cudaMemcpyDeviceToHost);
  int seen[1024] = {0}, distinct = 0, maxid = 0;
  for (int b = 0; b < nsm; b++) { if (!seen[h[b]]++) distinct++; }
  for (int b = 0; b < grid; b++) if (h[b] > maxid) maxid = h[b];
  printf("SMs %d, grid %d, distinct SMs among CTAs 0..%d: %d, max smid %d\n", nsm, grid, nsm - 1, distinct, maxid);
  int same = 0; for (int b = 0; b + nsm < grid; b++) same += h[b] == h[b + nsm];
  printf("CTA b and b+S on the same SM: %d of %d\n", same, grid - nsm);
  for (int b = 0; b < 32; b++) printf("%d ", h[b]); printf("...\n");
  for (int b = nsm; b < nsm + 32; b++) printf("%d ", h[b]); printf("...\n");
  return 0;
}
