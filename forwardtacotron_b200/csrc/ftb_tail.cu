// K2f: the row-local tail of a CBHG as ONE persistent tcgen05 kernel:
//
//     pre_highway Linear (common_layers.py:118) -> NH x HighwayNetwork (:27-40, :119-120) -> W_ih half of the BiGRU (:123)
//
// Every layer of this chain maps row r to row r, so a 128-row tile can run all of it without leaving the SM: the
// activation tile lives in shared memory as TWO fp16 tiles, hi = fp16(x) and lo = fp16(x - hi), both in the K-major
// 128B-swizzled layout tcgen05.mma reads.  `hi` is the A operand of every layer (the same 11-bit operand the unfused
// kernels use); hi + lo is the residual term (1-g)*x of the highway at ~2^-22 relative precision -- what the unfused
// path keeps as an fp32 twin in HBM.  Weights stream through a 3-stage TMA ring (they are L2 resident: 1.9 MB per CBHG),
// accumulators live in TMEM, the epilogue of a layer rewrites hi / lo IN PLACE once all MMAs of that layer have retired,
// and only the x-projection [rows, 2*3*H] leaves the chip.
//
//   warp 0      : TMA producer  (A0 tile of the next row tile; weight blocks of all layers in issue order)
//   warp 1      : MMA issuer    (one elected lane; tcgen05.mma.cta_group::1.kind::f16, M = 128, N = 256)
//   warps 2..17 : epilogue      (four warps per TMEM lane quarter: warp slice h owns units / columns [64h, 64h+64);
//                                the epilogue is a latency chain TMEM -> math -> smem, so it is spread over 16 warps and
//                                every TMEM load is issued one 16-column chunk ahead of the math that consumes it)
//
// Replaces 6 launches per CBHG (pre_highway, 4 highways, gru_ih: 0.17 ms token side + 0.19 ms frame side at config 2,
// each one short tile per CTA paced by re-fetching activations, residual twins and weights from L2) by one.
#include <cstdio>
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int T_BM = 128;                    // rows per tile
constexpr int T_CH = 256;                    // highway width (CBHG channels): fixed
constexpr int T_KB = T_CH / 64;              // 64-element (128-byte) k-blocks of an activation tile
constexpr int T_ATILE = T_BM * 128;          // bytes of one k-block of the activation tile (128 rows x 128 B)
constexpr int T_ABYTES = T_KB * T_ATILE;     // 64 KB: hi (and lo) tile
constexpr int T_WSTAGE = 256 * 128;          // 32 KB: 256 weight rows x one k-block
constexpr int T_STAGES = 3;
constexpr int T_EPI_WARPS = 16;              // four per TMEM lane quarter: warp slice h owns units / columns [64h, 64h + 64)
constexpr int T_THREADS = (2 + T_EPI_WARPS) * 32;
constexpr int T_SLICE = T_CH / (T_EPI_WARPS / 4);     // 64

struct TailSmem {
  uint64_t w_full[T_STAGES];
  uint64_t w_empty[T_STAGES];
  uint64_t a_full;            // A0 of this tile has landed in `hi`
  uint64_t a_free;            // every MMA that reads `hi` for this tile has retired
  uint64_t acc_full;          // the MMAs of one (pre / highway) layer have retired
  uint64_t epi_done;          // the epilogue has rewritten hi / lo (8 warp arrivals)
  uint64_t ih_full[2];        // x-projection chunk in TMEM buffer b
  uint64_t ih_empty[2];       // ... drained (8 warp arrivals)
  uint32_t tmem_base;
  uint32_t pad;
};
static_assert(2 * T_ABYTES + T_STAGES * T_WSTAGE + sizeof(TailSmem) + 1024 <= 232448, "over the 227 KB shared-memory limit");
constexpr int T_SMEM_BYTES = 2 * T_ABYTES + T_STAGES * T_WSTAGE + sizeof(TailSmem) + 1024;

// byte offset of the 16-byte chunk holding elements [k8*8, k8*8+8) of row r inside a [T_KB][128 rows][64] fp16 tile in
// the 128B-swizzle layout TMA writes and tcgen05.mma reads (8-row groups of 1024 B; chunk index XOR row%8)
__device__ __forceinline__ uint32_t tile_chunk_off(int r, int k8) {
  const int kb = k8 >> 3, c = k8 & 7;
  return static_cast<uint32_t>(kb * T_ATILE + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

__device__ __forceinline__ void split8(const float (&v)[8], uint4* hi, uint4* lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half2 hh = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    const float2 f = __half22float2(hh);
    const __half2 ll = __floats2half2_rn(v[2 * i] - f.x, v[2 * i + 1] - f.y);
    h[i] = *reinterpret_cast<const uint32_t*>(&hh);
    l[i] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  *hi = make_uint4(h[0], h[1], h[2], h[3]);
  *lo = make_uint4(l[0], l[1], l[2], l[3]);
}

__device__ __forceinline__ void join8(const uint4& hi, const uint4& lo, float (&x)[8]) {
  const uint32_t h[4] = {hi.x, hi.y, hi.z, hi.w}, l[4] = {lo.x, lo.y, lo.z, lo.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&h[i]));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&l[i]));
    x[2 * i] = a.x + b.x;
    x[2 * i + 1] = a.y + b.y;
  }
}

// g * relu(x1) + (1 - g) * x, g = sigmoid(x2)   (common_layers.py:36-39; same form as the unfused epilogue)
__device__ __forceinline__ float tail_highway(float x1, float x2, float x) {
  float e, g;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x2));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(g) : "f"(1.0f + e));
  return fmaf(g, fmaxf(x1, 0.0f) - x, x);
}

__global__ void __launch_bounds__(T_THREADS, 1)
cbhg_tail_kernel(const TailProblemDev* __restrict__ prob) {
  extern __shared__ uint8_t tail_smem_raw[];
  uint8_t* smem = tail_smem_raw + ((1024u - (smem_u32(tail_smem_raw) & 1023u)) & 1023u);
  uint8_t* hi_s = smem;
  uint8_t* lo_s = smem + T_ABYTES;
  uint8_t* w_s = smem + 2 * T_ABYTES;
  TailSmem* ts = reinterpret_cast<TailSmem*>(smem + 2 * T_ABYTES + T_STAGES * T_WSTAGE);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const TailProblemDev& p = *prob;
  // plan fields -> registers once (the asm memory clobbers would otherwise re-load them from global memory in every loop)
  const int m_tiles = p.m_tiles, nh = p.n_highways, kb_in = p.kb_in, ih_chunks = p.ih_chunks;
  const long long P_rows = p.rows;
  const int P_seq_len = p.seq_len, P_seq_stride = p.seq_stride, P_out_ld = p.out_ld, P_n_ih = p.n_ih;
  const float* const P_b_hw = p.b_hw;
  const float* const P_b_ih = p.b_ih;
  void* const P_out = p.out;

#ifdef FTB_TAIL_PROFILE
  long long tp[6] = {0, 0, 0, 0, 0, 0}, tp_t0 = 0;
  const long long tp_start = clock64();
#define TP_T0 tp_t0 = clock64();
#define TP_ACC(i) tp[i] += clock64() - tp_t0;
#else
#define TP_T0
#define TP_ACC(i)
#endif
  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < T_STAGES; ++s) { mbar_init(&ts->w_full[s], 1); mbar_init(&ts->w_empty[s], 1); }
    mbar_init(&ts->a_full, 1);
    mbar_init(&ts->a_free, 1);
    mbar_init(&ts->acc_full, 1);
    mbar_init(&ts->epi_done, T_EPI_WARPS);
    for (int b = 0; b < 2; ++b) { mbar_init(&ts->ih_full[b], 1); mbar_init(&ts->ih_empty[b], T_EPI_WARPS); }
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(&ts->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ts->tmem_base;
  pdl_wait();

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      tma_desc_fence_acquire(&p.tm_a);
      tma_desc_fence_acquire(&p.tm_wpre);
      tma_desc_fence_acquire(&p.tm_whw);
      tma_desc_fence_acquire(&p.tm_wih);
      int stage = 0;
      uint32_t phase = 0, it = 0;
      auto w_block = [&](const CUtensorMap* tm, int k0, int row0) {
        TP_T0 mbar_wait(&ts->w_empty[stage], phase ^ 1u); TP_ACC(0)
        mbar_expect_tx(&ts->w_full[stage], T_WSTAGE);
        tma_load_2d(w_s + stage * T_WSTAGE, tm, &ts->w_full[stage], k0, row0);
        if (++stage == T_STAGES) { stage = 0; phase ^= 1u; }
      };
      for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x, ++it) {
        // the activation tile of the previous row tile is dead once its last x-projection MMA has retired
        mbar_wait(&ts->a_free, (it & 1u) ^ 1u);
        mbar_expect_tx(&ts->a_full, static_cast<uint32_t>(kb_in) * T_ATILE);
        for (int kb = 0; kb < kb_in; ++kb) tma_load_2d(hi_s + kb * T_ATILE, &p.tm_a, &ts->a_full, kb * 64, tile * T_BM);
        for (int kb = 0; kb < kb_in; ++kb) w_block(&p.tm_wpre, kb * 64, 0);
        for (int l = 0; l < nh; ++l)
          for (int half = 0; half < 2; ++half)
            for (int kb = 0; kb < T_KB; ++kb) w_block(&p.tm_whw, kb * 64, l * 2 * T_CH + half * 256);
        for (int c = 0; c < ih_chunks; ++c)
          for (int kb = 0; kb < T_KB; ++kb) w_block(&p.tm_wih, kb * 64, c * 256);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_f16(T_BM, 256);
      int stage = 0;
      uint32_t phase = 0, it = 0, epi_phase = 0, ih_it[2] = {0, 0};
      auto mma_group = [&](uint32_t d_tmem, int kblocks) {       // one N = 256 chunk over `kblocks` k-blocks of `hi`
        for (int kb = 0; kb < kblocks; ++kb) {
          TP_T0 mbar_wait(&ts->w_full[stage], phase); TP_ACC(1)
          tc_fence_after();
          const uint64_t da = umma_desc_sw128(smem_u32(hi_s + kb * T_ATILE));
          const uint64_t db = umma_desc_sw128(smem_u32(w_s + stage * T_WSTAGE));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16_ss(d_tmem, da + 2u * k, db + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&ts->w_empty[stage]);
          if (++stage == T_STAGES) { stage = 0; phase ^= 1u; }
        }
      };
      for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x, ++it) {
        mbar_wait(&ts->a_full, it & 1u);
        // both TMEM buffers are drained (the previous row tile's last x-projection chunks)
        for (int b = 0; b < 2; ++b) mbar_wait(&ts->ih_empty[b], (ih_it[b] & 1u) ^ 1u);
        tc_fence_after();
        // pre_highway: x0 = A0 W_pre^T  -> TMEM columns [0, 256)
        mma_group(tmem_base, kb_in);
        umma_commit(&ts->acc_full);
        for (int l = 0; l < nh; ++l) {
          TP_T0 mbar_wait(&ts->epi_done, epi_phase); TP_ACC(2)       // hi / lo hold this layer's input
          epi_phase ^= 1u;
          tc_fence_after();
          mma_group(tmem_base, T_KB);                 // [x1 | x2] of units 0..127
          mma_group(tmem_base + 256, T_KB);           // [x1 | x2] of units 128..255
          umma_commit(&ts->acc_full);
        }
        TP_T0 mbar_wait(&ts->epi_done, epi_phase); TP_ACC(2)         // hi holds the last highway's output
        epi_phase ^= 1u;
        tc_fence_after();
        for (int c = 0; c < ih_chunks; ++c) {
          const int b = c & 1;
          TP_T0 mbar_wait(&ts->ih_empty[b], (ih_it[b] & 1u) ^ 1u); TP_ACC(3)
          tc_fence_after();
          mma_group(tmem_base + b * 256, T_KB);
          umma_commit(&ts->ih_full[b]);
          ++ih_it[b];
        }
        umma_commit(&ts->a_free);                     // all reads of `hi` for this tile retired -> next A0 may land
      }
    }
  } else {
    // ================================ epilogue ====================================
    // lane = row of the tile (TMEM lane quarter q = warp % 4); slice h = (warp - 2) / 4 owns units [64h, 64h + 64)
    const int q = warp & 3, h = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const uint32_t tlane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t acc_phase = 0, ih_it[2] = {0, 0};
    const bool o16 = p.out_f16 != 0;
    constexpr int NSUB = T_SLICE / 16;                       // 16-column chunks per warp and layer
    for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x) {
      const long long r = static_cast<long long>(tile) * T_BM + row;
      bool valid = r < P_rows;
      if (valid) valid = static_cast<int>(r % P_seq_stride) < P_seq_len;
      // ---- pre_highway: TMEM [0,256) -> hi / lo  (no bias, no activation: common_layers.py:118)
      TP_T0 mbar_wait(&ts->acc_full, acc_phase); TP_ACC(0)
      acc_phase ^= 1u;
      tc_fence_after();
      TP_T0
      {
        float v[2][16];
        tmem_ld_32x16(tlane + h * T_SLICE, v[0]);
#pragma unroll
        for (int c = 0; c < NSUB; ++c) {
          const int j0 = h * T_SLICE + 16 * c;
          tmem_ld_wait();
          if (c + 1 < NSUB) tmem_ld_32x16(tlane + j0 + 16, v[(c + 1) & 1]);
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            float y[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) y[i] = v[c & 1][8 * s + i];
            uint4 ph, pl;
            split8(y, &ph, &pl);
            const uint32_t off = tile_chunk_off(row, (j0 >> 3) + s);
            *reinterpret_cast<uint4*>(hi_s + off) = ph;
            *reinterpret_cast<uint4*>(lo_s + off) = pl;
          }
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ts->epi_done);
      TP_ACC(1)
      // ---- highways: units [64h, 64h + 64) live in TMEM half hh = h / 2 at columns hh*256 + (64 (h % 2) + j) (x1) and
      //      + 128 (x2): n-half hh of the packed weight holds [W1 rows | W2 rows] of units [128 hh, 128 hh + 128)
      const int hh = h >> 1, jbase = (h & 1) * 64;
      for (int l = 0; l < nh; ++l) {
        const float* bl = P_b_hw + l * 2 * T_CH;        // b1 | b2 of this layer: warp-uniform 16-byte loads (L1 broadcast)
        TP_T0 mbar_wait(&ts->acc_full, acc_phase); TP_ACC(0)
        acc_phase ^= 1u;
        tc_fence_after();
        TP_T0
        float x1[2][16], x2[2][16];
        tmem_ld_32x16(tlane + hh * 256 + jbase, x1[0]);
        tmem_ld_32x16(tlane + hh * 256 + 128 + jbase, x2[0]);
#pragma unroll
        for (int c = 0; c < NSUB; ++c) {
          const int u0 = h * T_SLICE + 16 * c;                 // first unit of this chunk
          uint4 rh[2], rl[2];
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            const uint32_t off = tile_chunk_off(row, (u0 >> 3) + s);
            rh[s] = *reinterpret_cast<const uint4*>(hi_s + off);
            rl[s] = *reinterpret_cast<const uint4*>(lo_s + off);
          }
          tmem_ld_wait();
          if (c + 1 < NSUB) {
            tmem_ld_32x16(tlane + hh * 256 + jbase + 16 * (c + 1), x1[(c + 1) & 1]);
            tmem_ld_32x16(tlane + hh * 256 + 128 + jbase + 16 * (c + 1), x2[(c + 1) & 1]);
          }
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            float x[8], y[8];
            join8(rh[s], rl[s], x);
#pragma unroll
            for (int i = 0; i < 8; i += 4) {
              const float4 t1 = __ldg(reinterpret_cast<const float4*>(bl + u0 + 8 * s + i));
              const float4 t2 = __ldg(reinterpret_cast<const float4*>(bl + T_CH + u0 + 8 * s + i));
              y[i] = tail_highway(x1[c & 1][8 * s + i] + t1.x, x2[c & 1][8 * s + i] + t2.x, x[i]);
              y[i + 1] = tail_highway(x1[c & 1][8 * s + i + 1] + t1.y, x2[c & 1][8 * s + i + 1] + t2.y, x[i + 1]);
              y[i + 2] = tail_highway(x1[c & 1][8 * s + i + 2] + t1.z, x2[c & 1][8 * s + i + 2] + t2.z, x[i + 2]);
              y[i + 3] = tail_highway(x1[c & 1][8 * s + i + 3] + t1.w, x2[c & 1][8 * s + i + 3] + t2.w, x[i + 3]);
            }
            uint4 ph, pl;
            split8(y, &ph, &pl);
            const uint32_t off = tile_chunk_off(row, (u0 >> 3) + s);
            *reinterpret_cast<uint4*>(hi_s + off) = ph;
            *reinterpret_cast<uint4*>(lo_s + off) = pl;
          }
        }
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&ts->epi_done);
        TP_ACC(2)
      }
      // ---- x-projection chunks: TMEM buffer c % 2 -> + b_ih -> staging -> global.
      // A lane owns a ROW of the accumulator; storing from that layout makes every store instruction touch 32 different
      // 128-byte lines (measured: 45 000 cycles per tile, 9 B/clk).  The `lo` tile is dead after the last highway, so it
      // serves as a per-warp 4 KB staging block [32 rows][128 B] (16-byte chunks XOR-swizzled by row % 8, conflict-free
      // both ways): written row-per-lane, read back as 4 rows x 128 contiguous bytes per store instruction.
      uint8_t* stg = lo_s + (warp - 2) * 4096;
      const unsigned vmask = __ballot_sync(0xffffffffu, valid);
      const long long r_warp = static_cast<long long>(tile) * T_BM + q * 32;
      const int srow = lane >> 3, schunk = lane & 7;
      for (int c = 0; c < ih_chunks; ++c) {
        const int b = c & 1;
        TP_T0 mbar_wait(&ts->ih_full[b], ih_it[b] & 1u); TP_ACC(3)
        ++ih_it[b];
        tc_fence_after();
        TP_T0
        float v[2][16];
        tmem_ld_32x16(tlane + b * 256 + h * T_SLICE, v[0]);
        // one staging round = 128 bytes per row: 64 columns of fp16 (the whole slice) or 32 columns of fp32 (half of it)
        const int sub_per_round = o16 ? 4 : 2;
#pragma unroll
        for (int sc = 0; sc < NSUB; ++sc) {
          const int j0 = h * T_SLICE + 16 * sc;
          const int n0 = c * 256 + j0;
          tmem_ld_wait();
          if (sc + 1 < NSUB) tmem_ld_32x16(tlane + b * 256 + j0 + 16, v[(sc + 1) & 1]);
          float y[16];
#pragma unroll
          for (int i = 0; i < 16; i += 4) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (n0 < P_n_ih) t = __ldg(reinterpret_cast<const float4*>(P_b_ih + n0 + i));
            y[i] = v[sc & 1][i] + t.x; y[i + 1] = v[sc & 1][i + 1] + t.y;
            y[i + 2] = v[sc & 1][i + 2] + t.z; y[i + 3] = v[sc & 1][i + 3] + t.w;
          }
          const int within = sc % sub_per_round;            // position of this sub-chunk inside the staging round
          if (o16) {
            uint32_t w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const __half2 t = __floats2half2_rn(y[2 * i], y[2 * i + 1]);
              w[i] = *reinterpret_cast<const uint32_t*>(&t);
            }
            *reinterpret_cast<uint4*>(stg + lane * 128 + (((2 * within) ^ (lane & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<uint4*>(stg + lane * 128 + (((2 * within + 1) ^ (lane & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<float4*>(stg + lane * 128 + (((4 * within + i) ^ (lane & 7)) << 4)) =
                  make_float4(y[4 * i], y[4 * i + 1], y[4 * i + 2], y[4 * i + 3]);
          }
          if (within == sub_per_round - 1) {
            __syncwarp();
            // first column of this staging round, and the column of this lane's 16-byte chunk
            const int nr0 = c * 256 + h * T_SLICE + (sc + 1 - sub_per_round) * 16;
            const int ncol = nr0 + schunk * (o16 ? 8 : 4);
            uint4 d[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int R = i * 4 + srow;
              d[i] = *reinterpret_cast<const uint4*>(stg + R * 128 + ((schunk ^ (R & 7)) << 4));
            }
            if (ncol < P_n_ih) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int R = i * 4 + srow;
                if ((vmask >> R) & 1u) {
                  uint8_t* o = reinterpret_cast<uint8_t*>(P_out) +
                               ((r_warp + R) * P_out_ld + ncol) * static_cast<long long>(o16 ? 2 : 4);
                  *reinterpret_cast<uint4*>(o) = d[i];
                }
              }
            }
            __syncwarp();
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&ts->ih_empty[b]);
        TP_ACC(4)
      }
    }
  }
#ifdef FTB_TAIL_PROFILE
  if (blockIdx.x == 0 && lane == 0 && (warp <= 2 || warp == 10)) {
    const long long tot = clock64() - tp_start;
    if (warp == 0) printf("tail prof producer: total %lld wait_w_empty %lld\n", tot, tp[0]);
    else if (warp == 1) printf("tail prof mma: total %lld wait_w_full %lld wait_epi_done %lld wait_ih_empty %lld\n", tot, tp[1], tp[2], tp[3]);
    else printf("tail prof epilogue warp %d: total %lld wait_acc_full %lld pre_work %lld hw_work %lld wait_ih_full %lld ih_work %lld\n",
                warp, tot, tp[0], tp[1], tp[2], tp[3], tp[4]);
  }
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------ host side
int tail_plan_build(const ftb_cbhg_tail_problem* q, int num_sms, TailPlan* plan) {
  if (!q || !q->a || !q->w_pre || !q->w_hw || !q->b_hw || !q->w_ih || !q->b_ih || !q->out) return FTB_E_ARG;
  if (q->channels != T_CH || q->n_highways < 1 || q->n_highways > 8) return FTB_E_SHAPE;
  if (q->cin < 8 || q->cin > T_CH || (q->cin % 8) || (q->lda % 8) || q->cin > q->lda) return FTB_E_ALIGN;
  if (q->n_ih < 16 || (q->n_ih % 16) || (q->out_ld % 8) || q->n_ih > q->out_ld) return FTB_E_SHAPE;
  if (q->batch < 1 || q->seq_len < 1 || q->seq_stride < q->seq_len) return FTB_E_ARG;
  const uintptr_t al = reinterpret_cast<uintptr_t>(q->a) | reinterpret_cast<uintptr_t>(q->w_pre) |
                       reinterpret_cast<uintptr_t>(q->w_hw) | reinterpret_cast<uintptr_t>(q->w_ih) |
                       reinterpret_cast<uintptr_t>(q->out);
  if (al & 15u) return FTB_E_ALIGN;
  TailProblemDev d{};
  d.rows = static_cast<long long>(q->batch) * q->seq_stride;
  d.cap_rows = d.rows;
  d.batch = q->batch; d.seq_len = q->seq_len; d.seq_stride = q->seq_stride;
  d.m_tiles = static_cast<int>((d.rows + T_BM - 1) / T_BM);
  d.n_highways = q->n_highways;
  d.kb_in = (q->cin + 63) / 64;
  d.n_ih = q->n_ih;
  d.ih_chunks = (q->n_ih + 255) / 256;
  d.b_hw = q->b_hw; d.b_ih = q->b_ih; d.out = q->out; d.out_ld = q->out_ld; d.out_f16 = q->out_f16;
  int rc = tensor_map_2d(&d.tm_a, q->a, d.rows, q->cin, q->lda, T_BM, true);
  if (rc) return rc;
  rc = tensor_map_2d(&d.tm_wpre, q->w_pre, T_CH, q->cin, q->cin, 256, true);
  if (rc) return rc;
  rc = tensor_map_2d(&d.tm_whw, q->w_hw, static_cast<long long>(q->n_highways) * 2 * T_CH, T_CH, T_CH, 256, true);
  if (rc) return rc;
  rc = tensor_map_2d(&d.tm_wih, q->w_ih, static_cast<long long>(d.ih_chunks) * 256, T_CH, T_CH, 256, true);
  if (rc) return rc;
  cudaError_t ce = cudaMalloc(&plan->dev, sizeof(TailProblemDev));
  if (ce != cudaSuccess) return static_cast<int>(ce);
  ce = cudaMemcpy(plan->dev, &d, sizeof(d), cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  plan->host = d;
  plan->cap_tiles = d.m_tiles;
  (void)num_sms;
  return 0;
}

void tail_plan_free(TailPlan& plan) {
  cudaFree(plan.dev);
  plan = TailPlan();
}

int tail_plan_set_geometry(TailPlan& plan, int seq_len, int seq_stride, cudaStream_t stream) {
  if (seq_len < 1 || seq_stride < seq_len) return FTB_E_ARG;
  TailProblemDev& d = plan.host;
  const long long rows = static_cast<long long>(d.batch) * seq_stride;
  if (rows > d.cap_rows) return FTB_E_LIMIT;
  d.rows = rows; d.seq_len = seq_len; d.seq_stride = seq_stride;
  d.m_tiles = static_cast<int>((rows + T_BM - 1) / T_BM);
  return static_cast<int>(cudaMemcpyAsync(plan.dev, &d, sizeof(d), cudaMemcpyHostToDevice, stream));
}

int tail_plan_launch(const TailPlan& plan, int num_sms, cudaStream_t stream) {
  static bool attr_set[64] = {};
  int dev_id = 0;
  cudaError_t ce = cudaGetDevice(&dev_id);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  if (dev_id < 0 || dev_id >= 64 || !attr_set[dev_id]) {
    ce = cudaFuncSetAttribute(cbhg_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, T_SMEM_BYTES);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    if (dev_id >= 0 && dev_id < 64) attr_set[dev_id] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(plan.cap_tiles < num_sms ? plan.cap_tiles : num_sms);
  cfg.blockDim = dim3(T_THREADS);
  cfg.dynamicSmemBytes = T_SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  const TailProblemDev* dev = plan.dev;
  return static_cast<int>(cudaLaunchKernelEx(&cfg, cbhg_tail_kernel, dev));
}

}  // namespace ftb
