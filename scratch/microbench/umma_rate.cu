// tcgen05.mma kind::f16 per-instruction cost at SMALL N (the recurrence's [G*U x H] x [H x NB] product), A from shared
// memory (SS) or tensor memory (TS), on sm_100a.  One thread issues `cnt` UMMAs (K = 16 each) + one commit and waits
// on the mbarrier; cycles = clock64 around issue..completion, slope over cnt = cost per instruction.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_rate umma_rate.cu && ./umma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t byte_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((byte_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
__device__ __forceinline__ uint32_t idesc_f16(uint32_t m, uint32_t n) { return (1u << 4) | ((n >> 3) << 17) | ((m >> 4) << 24); }

template <bool TS>
__device__ __forceinline__ void umma(uint32_t d, uint32_t a_t, uint64_t a_d, uint64_t b_d, uint32_t idesc, uint32_t acc) {
  if (TS)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
                 ::"r"(d), "r"(a_t), "l"(b_d), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
                 ::"r"(d), "l"(a_d), "l"(b_d), "r"(idesc), "r"(acc) : "memory");
}

// nacc: number of independent accumulators the K range is spread over (1 = one dependent chain)
template <bool TS, int M, int N, int CNT, int NACC, int NW>
__global__ void __launch_bounds__(128) k(long long* cyc, int reps) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  // A: 8 K-chunks of 64 halves x 128 rows (16 KB each) ; B: 8 K-chunks x N rows (N/8 KB each) after A
  for (int i = threadIdx.x; i < (128 * 1024 + 64 * 1024) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003800u;      // (0.5, 1.0) halves
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(NW) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = slot;
  const uint32_t a_addr = smem_u32(smem), b_addr = a_addr + 128 * 1024;
  const uint32_t idesc = idesc_f16(M, N);
  constexpr int bmask = N >= 256 ? 1 : (N >= 128 ? 3 : 7);       // B region is 64 KB
  const uint64_t ad0 = desc_sw128(a_addr), bd0 = desc_sw128(b_addr);
  // TMEM: columns [0,256) = A (fp16 pairs, 512 K), columns [256, 512) = accumulators
  if (TS) {                      // fill A region of TMEM so the values are defined
    uint32_t r[8] = {0x3c003800u, 0x3c003800u, 0x3c003800u, 0x3c003800u, 0x3c003800u, 0x3c003800u, 0x3c003800u, 0x3c003800u};
    const uint32_t lane_base = tbase + ((threadIdx.x / 32) * 32 << 16);
    for (int c = 0; c < 256; c += 8)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                   ::"r"(lane_base + c), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  {
    const int w = threadIdx.x / 32;
    uint32_t phase = 0;
    long long best = 1ll << 60;
    for (int rep = 0; rep < reps; ++rep) {
      __syncthreads();
      const long long t0 = clock64();
      if (w < NW && (threadIdx.x & 31) == 0) {
#pragma unroll
        for (int i = 0; i < CNT / NW; ++i) {
          const int kk = (w * (CNT / NW) + i) & 31;             // K position inside the 512-wide operand
          const uint64_t ad = ad0 + (((kk >> 2) * 16384 + (kk & 3) * 32) >> 4);
          const uint64_t bd = bd0 + ((((kk >> 2) & bmask) * (N / 8) * 1024 + (kk & 3) * 32) >> 4);
          umma<TS>(tbase + 256 + (w * NACC + i % NACC) * N, tbase + kk * 8, ad, bd, idesc, i >= NACC ? 1u : 0u);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      }
      __syncwarp();
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
      const long long t1 = clock64();
      phase ^= 1u;
      if (t1 - t0 < best) best = t1 - t0;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *cyc = best;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(512) : "memory");
}

template <bool TS, int M, int N, int NACC, int NW>
void run(int grid) {
  long long* cyc; cudaMalloc(&cyc, 8);
  const int smem = 128 * 1024 + 64 * 1024 + 1024;
  long long c[3];
  auto go = [&](auto kern, int j) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    kern<<<grid, 128, smem>>>(cyc, 20);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    cudaMemcpy(&c[j], cyc, 8, cudaMemcpyDeviceToHost);
  };
  go(k<TS, M, N, 8, NACC, NW>, 0); go(k<TS, M, N, 32, NACC, NW>, 1); go(k<TS, M, N, 64, NACC, NW>, 2);
  printf("%s M=%3d N=%3d nacc=%d issuers=%d grid=%3d : cnt 8 -> %5lld, 32 -> %5lld, 64 -> %6lld cycles ; slope %.1f cyc/UMMA\n", TS ? "TS" : "SS", M, N,
         NACC, NW, grid, c[0], c[1], c[2], double(c[2] - c[1]) / 32.0);
  cudaFree(cyc);
}

int main() {
  const int grid = 148;
  run<false, 128, 16, 1, 1>(grid); run<false, 128, 16, 1, 2>(grid); run<false, 128, 16, 1, 4>(grid);
  run<true, 128, 16, 1, 1>(grid); run<true, 128, 16, 1, 2>(grid); run<true, 128, 16, 1, 4>(grid);
  run<false, 128, 32, 1, 1>(grid); run<false, 128, 32, 1, 2>(grid); run<false, 128, 32, 1, 4>(grid);
  run<true, 128, 32, 1, 4>(grid);
  run<false, 128, 64, 1, 1>(grid); run<false, 128, 64, 1, 2>(grid); run<false, 128, 64, 1, 4>(grid);
  run<true, 128, 64, 1, 1>(grid); run<true, 128, 64, 1, 2>(grid); run<true, 128, 64, 1, 4>(grid);
  run<false, 128, 128, 1, 1>(grid); run<false, 128, 256, 1, 1>(grid);
  return 0;
}
