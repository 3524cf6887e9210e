#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
constexpr int KB = 8, HS = 544;   // halves
template <int MODE, int CH>
__global__ void __launch_bounds__(512, 1) k(float* out, long long* cyc, int iters, const uint32_t* wsrc) {
  __shared__ __align__(16) unsigned short h_s[16 * HS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, q = lane & 3, fr = lane >> 2, kq = warp & 1;
  for (int i = tid; i < 16 * HS; i += blockDim.x) h_s[i] = 0x3c00;
  uint32_t w[KB * 2][4];
#pragma unroll
  for (int i = 0; i < KB * 2; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) w[i][j] = wsrc[(tid * 64 + i * 4 + j) & 1023];
  __syncthreads();
  float total = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float acc[CH][4];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
    const unsigned short* hrow = h_s + ((it & 1) * 8 + fr) * HS + kq * 256 + 8 * q;
    if (MODE == 2) {
      uint4 hv[KB];
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) hv[kb] = *reinterpret_cast<const uint4*>(hrow + kb * 32);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        mma(acc[(2 * kb) % CH], w[kb * 2], hv[kb].x, hv[kb].y);
        mma(acc[(2 * kb + 1) % CH], w[kb * 2 + 1], hv[kb].z, hv[kb].w);
      }
    } else {
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        uint4 hv;
        if (MODE == 1) hv = *reinterpret_cast<const uint4*>(hrow + kb * 32);
        else hv = make_uint4(0x3c003c00u + kb, 0x3c003c00u, 0x3c003c01u, 0x3c003c00u + it);
        mma(acc[(2 * kb) % CH], w[kb * 2], hv.x, hv.y);
        mma(acc[(2 * kb + 1) % CH], w[kb * 2 + 1], hv.z, hv.w);
      }
    }
#pragma unroll
    for (int c = 0; c < CH; ++c) total += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
    asm volatile("bar.sync 0;" ::: "memory");
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + tid] = total;
  if (tid == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE, int CH>
void run(const uint32_t* wsrc, const char* name) {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 8);
  const int iters = 400;
  k<MODE, CH><<<148, 512>>>(out, cyc, iters, wsrc); cudaDeviceSynchronize();
  k<MODE, CH><<<148, 512>>>(out, cyc, iters, wsrc); cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-28s chains %d : %.0f cycles per group-MMA (16 HMMA/warp, 64/SMSP) = %.1f cycles per HMMA per SMSP  [%s]\n", name, CH,
         double(h) / iters, double(h) / iters / 64, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}
int main() {
  uint32_t* wsrc; cudaMalloc(&wsrc, 4096); cudaMemset(wsrc, 0x3c, 4096);
  run<0, 2>(wsrc, "B constant"); run<0, 4>(wsrc, "B constant");
  run<1, 2>(wsrc, "LDS.128 in loop"); run<1, 4>(wsrc, "LDS.128 in loop");
  run<2, 2>(wsrc, "LDS.128 hoisted"); run<2, 4>(wsrc, "LDS.128 hoisted");
  return 0;
}
