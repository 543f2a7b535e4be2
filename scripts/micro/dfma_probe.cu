// FP64 issue-rate / latency probe: cycles per DFMA for one warp with C independent chains, with 1, 2
// or 4 warps per SM active (threads per block 32 / 64 / 128, one block per SM).
#include <cstdio>
#include <cuda_runtime.h>
template <int C>
__global__ void probe(double* out, long long* cyc, int iters, double a, double b) {
  double v[C];
#pragma unroll
  for (int i = 0; i < C; i++) v[i] = threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
      for (int i = 0; i < C; i++) v[i] = fma(v[i], a, b);
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < C; i++) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int C>
void run(int threads, int blocks) {
  double* out; long long* cyc; long long h;
  cudaMalloc(&out, sizeof(double) * threads * blocks); cudaMalloc(&cyc, 8);
  const int iters = 2000;
  probe<C><<<blocks, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
  probe<C><<<blocks, threads>>>(out, cyc, iters, 1.0000001, 1e-9);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("chains %d, warps/block %d, blocks %d: %.2f cycles per DFMA per warp\n", C, threads / 32, blocks,
         (double)h / (iters * 8.0 * C));
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int t : {32, 64, 128, 256, 512}) { run<1>(t, 148); run<2>(t, 148); run<4>(t, 148); run<8>(t, 148); }
  return 0;
}
