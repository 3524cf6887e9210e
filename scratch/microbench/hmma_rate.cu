// HMMA.16816.F32 (mma.sync m16n8k16 f16) issue rate / dependent latency on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hmma_rate hmma_rate.cu && ./hmma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <int C>
__global__ void k(float* out, long long* cyc, int iters) {
  uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 9u};
  float acc[C][4];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int c = 0; c < C; ++c) mma(acc[c], a, 0x3c003c00u + r, 0x3c003c00u + c);
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < C; ++c) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int C>
void run(int threads) {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 200;
  k<C><<<148, threads>>>(out, cyc, iters); cudaDeviceSynchronize();
  k<C><<<148, threads>>>(out, cyc, iters); cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double per_warp = double(h) / (iters * 8 * C);
  const int wps = threads / 128;
  printf("warps/SMSP %d chains/warp %d : %.1f cycles per HMMA per warp, %.1f cycles per HMMA per SMSP\n", wps, C, per_warp,
         per_warp / wps);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int th : {128, 256, 512, 1024}) { run<1>(th); run<2>(th); run<4>(th); run<8>(th); }
  return 0;
}
