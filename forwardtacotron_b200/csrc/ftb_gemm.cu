// K1/K2: persistent, warp-specialised implicit-GEMM convolution / linear kernel for sm_100a.
//
//   warp 0      : TMA producer   (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx)
//   warp 1      : MMA issuer     (one elected lane, tcgen05.mma kind::tf32, fp32 accum in TMEM)
//   warps 2..9  : epilogue       (tcgen05.ld -> fused ReLU/BN-affine/maxpool/highway -> global), two sets of four
//                 warps: set 0 drains TMEM accumulator 0 (even tiles of this CTA), set 1 accumulator 1 (odd tiles)
//
// A convolution with k taps is k shifted GEMMs over the same channel-last activation
// matrix: tap j loads the A tile at row (m0 + tap0 + j).  Sentence edges need no masks
// because the padded-flat layout keeps >= 8 zero rows between sentences and TMA zero-fills
// rows outside the tensor (include/ftb200.h).  Two 256-column TMEM accumulators let the
// epilogue of tile i overlap the MMAs of tile i+1 AND the epilogue of tile i-1: an epilogue warp is a
// latency chain (one warp per SM sub-partition ran at ~0.2 IPC, profiles/r1m_gemm_epilogue.txt), so two
// warps per sub-partition on different tiles nearly double the drain rate of the small-K layers.
//
// Precision: operands are fp32 in memory, pre-rounded to tf32 (cvt.rna) by their
// producers, so the tensor core's mantissa truncation is exact; accumulation is fp32.
// FTB_F_A_F16 problems keep both operands in fp16 instead (same 11-bit significand; kind::f16 runs at twice the
// tf32 rate on half the bytes): a 128-byte swizzle row then holds 64 channels and one MMA consumes K = 16.
// FTB_F_OUT_F16 problems store their result as fp16 for such a consumer (next to any fp32 twin / transposed output).
//
// Two forms of the same body (template <bool PAIR>):
//   conv_gemm_tc_kernel       one CTA per tile, tcgen05.mma.cta_group::1 (M = 128), 4-stage ring of 48 KB
//   conv_gemm_tc_pair_kernel  the two CTAs of a 2-CTA cluster take two consecutive m-tiles of one n-tile: each stages its
//                             own A tile and HALF of the W tile, the even CTA issues tcgen05.mma.cta_group::2 (M = 256)
//                             for both, tcgen05.commit multicasts the "slot free" / "accumulator ready" arrives to both
//                             CTAs; 6-stage ring of 32 KB.  Picked per launch (gemm_plan_build) for tiles of >= 8
//                             k-blocks; the large-K layers are bound by how fast stages land, and a stage that carries
//                             the same MMA work in 2/3 of the bytes makes the ring 1.5x deeper (profiles/r1t_gemm_pair.txt).
// Both start under programmatic dependent launch: barrier set-up and the TMEM allocation run before griddepcontrol.wait.
#include <cstdio>
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"


namespace ftb {

constexpr int BM = 128;                 // tile rows  (UMMA M)
constexpr int BK = 32;                  // fp32 elements per k-block = 128 B = one swizzle row
constexpr int UMMA_K = 8;               // tf32: 32 bytes per MMA
constexpr int STAGES = 4;
constexpr int A_STAGE_BYTES = BM * BK * 4;        // 16 KB
constexpr int B_STAGE_BYTES = 256 * BK * 4;       // 32 KB (bn <= 256)
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
// CTA-pair mode: a CTA stages only half of the W tile, so the same 192 KB ring holds six 32 KB stages
constexpr int PAIR_STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES / 2;
constexpr int PAIR_STAGES = (STAGES * STAGE_BYTES) / PAIR_STAGE_BYTES;
constexpr double GEMM_PAIR_MIN_KITERS = 8.0;      // mean k-blocks per tile from which a launch runs as CTA pairs
constexpr int TMEM_COLS = 512;          // 2 accumulators x 256 fp32 columns
constexpr int NUM_THREADS = 320;
constexpr int EPI_THREADS = 128;         // per epilogue set
constexpr int CHUNK = 32;               // accumulator columns per epilogue chunk (one tcgen05.ld .x32)
constexpr int STG_LD = 36;              // staging row stride (floats): conflict-free 128-bit row writes AND reads
constexpr int STG_ROWS = 16;            // the 32 rows of a chunk pass through the staging tile in two halves

struct SmemTail {
  uint64_t full[PAIR_STAGES];
  uint64_t empty[PAIR_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad;
  alignas(16) float exch[2][2][4][CHUNK];           // per set: maxpool last row of each 32-row group, double buffered
  alignas(16) float params[2][3][256];              // per set: bias0 / scale / shift of the current tile's bn accumulator columns
  alignas(16) float stg[8][STG_ROWS][STG_LD];       // per epilogue warp: 16 rows x 32 columns staging tile (row<->lane transpose)
};
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + sizeof(SmemTail) + 1024;   // +1024: manual alignment

struct TileCoord {
  const GemmProblemDev* p;
  int m0, n0;
};

// Static schedule: CTA c runs tile i*G + c in even waves i and tile i*G + (G-1-c) in odd waves ("snake"), so with the
// problems sorted heaviest-first (host side) no CTA collects the heaviest tile of every wave.  A CTA's tile of wave i is
// out of range only in the last, partial wave, so every role of the CTA stops at the same i.
__device__ __forceinline__ int sched_tile(int i, int c, int G) { return i * G + ((i & 1) ? (G - 1 - c) : c); }

// pair_rank < 0: one tile per CTA.  Otherwise (CTA-pair plans) `tile` counts PAIRS of consecutive m-tiles of one n-tile
// and the CTA of rank r takes the r-th of the two (the second may lie behind the last row: it then stores nothing).
__device__ __forceinline__ TileCoord decode_tile(const GemmProblemDev* probs, int nprob, int tile, int pair_rank) {
  int pi = 0;
  while (pi + 1 < nprob && tile >= probs[pi + 1].tile_base) ++pi;
  const GemmProblemDev* p = probs + pi;
  const int local = tile - p->tile_base;
  const int nt = local % p->n_tiles;
  const int mt = pair_rank < 0 ? local / p->n_tiles : 2 * (local / p->n_tiles) + pair_rank;
  return {p, mt * p->m_step + p->m_shift, nt * p->bn};
}

// ------------------------------------------------------------------------------------------
// Epilogue math shared by the tensor-core kernel and the SIMT cross-check kernel.
__device__ __forceinline__ float epi_affine(const GemmProblemDev* p, float acc, int n) {
  float v = acc + (p->bias0 ? __ldg(p->bias0 + n) : 0.0f);
  if (p->flags & FTB_F_RELU) v = fmaxf(v, 0.0f);
  const float s = p->scale ? __ldg(p->scale + n) : 1.0f;
  const float b = p->shift ? __ldg(p->shift + n) : 0.0f;
  return v * s + b;     // contracted to one FMA: y = relu(acc)*s + b, s/b precomputed on the host
}
// y = g*relu(x1) + (1-g)*x with g = sigmoid(x2) (common_layers.py:36-39), as x + g*(relu(x1) - x): 7 instructions per
// output (FMUL, MUFU.EX2, FADD, MUFU.RCP, FMNMX, FADD, FFMA).  The highway epilogue is issue-bound (~430 SASS
// instructions per 16-unit chunk before this form, profiles/r1r_highway_ncu.txt); __expf / __fdividef carried range
// fix-ups that the bare ex2 / rcp (~2 ulp, far below the tf32 operand rounding) do not need: |x2| < 88 here.
__device__ __forceinline__ float epi_highway(float x1, float x2, float x) {
  float e, g;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x2));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(g) : "f"(1.0f + e));
  return fmaf(g, fmaxf(x1, 0.0f) - x, x);
}
__device__ __forceinline__ float epi_finish(const GemmProblemDev* p, float v) {
  return (p->flags & FTB_F_ROUND_TF32) ? round_tf32(v) : v;
}

// ------------------------------------------------------------------------------------------
// Affine / max-pool epilogue of one 128 x bn accumulator tile, as seen by ONE epilogue warp (32 rows).
// Everything that varies per problem is a template flag, so each instantiation carries only the
// instructions its layer needs: the epilogue runs with one warp per SMSP and is issue-bound (the
// all-runtime-branches version executed ~1400 SASS instructions per 32-column chunk, see
// profiles/r1e_gemm_roles.txt).
struct EpiCtx {
  const float (*prm)[256];       // [3][256] bias0 / scale / shift of this tile's columns
  float (*stg)[STG_LD];          // this warp's 16 x 32 staging tile
  float (*exch)[4][CHUNK];       // [2][4][CHUNK] max-pool row hand-over between the four warps of this set
  uint32_t pool_bar;             // named barrier of this set for the hand-over
  uint32_t taddr;                // TMEM address of this warp's lanes, column 0 of the tile
  int bn, n0, N, g, lane, t, b, seq_len, res_ld, out_ld;
  unsigned smask;                // rows of this warp that are stored
  bool store;                    // this lane's row is stored
  long long r, r_warp;           // this lane's row / the warp's first row
  float relu_lo;                 // 0 (ReLU) or -inf
  bool plain;                    // bias only: no ReLU, no scale / shift (linear layers, the W_ih projections)
  const float* residual;
  float* out; float* out_full; float* out_t;    // out / out_full already offset by out_col0
  __half* out16;                                // FTB_F_OUT_F16: the fp16 output (offset by out_col0), `out` is null
  bool split;                                   // FTB_F_OUT_SPLIT: out16 receives [hi | lo | hi] column blocks of width N
};

template <bool POOL, bool RES, bool FULL, bool OUTT, bool RND, bool OUT16 = false, bool SPLIT = false>
__device__ __forceinline__ void epi_affine_tile(const EpiCtx& c) {
  const int lane = c.lane, sub = lane >> 3, part = lane & 7;
  // phase-B base pointers: row (r_warp + sub), column n0 + 4*part; +32 columns per chunk, +4 rows per i8
  const long long orow = (c.r_warp + sub) * c.out_ld + c.n0 + 4 * part;
  // fp16 output next to a residual / unrounded twin / transposed output: phase-B layout, 8-byte stores of 4 halves
  constexpr bool O16_FAST = OUT16 && !RES && !FULL && !OUTT;     // (SPLIT included: hi / lo / hi as 16-byte row segments)
  float* optr = c.out ? c.out + orow : nullptr;
  __half* hptr = (OUT16 && !O16_FAST) ? c.out16 + orow : nullptr;
  float* fptr = FULL ? c.out_full + orow : nullptr;
  const float* rptr = RES ? c.residual + (c.r_warp + sub) * c.res_ld + c.n0 + 4 * part : nullptr;
  const long long ostep = 4ll * c.out_ld, rstep = 4ll * c.res_ld;
  // fp16 output: lane = (row r8 of 8, 8 consecutive columns c8): 16-byte stores, 64-byte row segments
  const int r8 = lane >> 2, c8 = lane & 3;
  __half* o16 = O16_FAST ? c.out16 + (c.r_warp + r8) * c.out_ld + c.n0 + 8 * c8 : nullptr;
  float4 xr_cur[RES ? 8 : 1], xr_nxt[RES ? 8 : 1];
  auto load_res = [&](float4 (&dst)[RES ? 8 : 1], int c0) {
    if (RES) {
#pragma unroll
      for (int i8 = 0; i8 < 8; ++i8) {
        dst[i8] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c0 + 4 * part < c.bn && c.n0 + c0 + 4 * part < c.N && ((c.smask >> (i8 * 4 + sub)) & 1u))
          dst[i8] = *reinterpret_cast<const float4*>(rptr + i8 * rstep + c0);
      }
    }
  };
  load_res(xr_cur, 0);
  int chunk_i = 0;
  for (int c0 = 0; c0 < c.bn; c0 += CHUNK, ++chunk_i) {
    load_res(xr_nxt, c0 + CHUNK);
    float v[CHUNK];
    if (c0 + CHUNK <= c.bn) {
      tmem_ld_32x32(c.taddr + c0, v);
    } else {                                              // bn is a multiple of 16: last half chunk
      float lo[16];
      tmem_ld_32x16(c.taddr + c0, lo);
#pragma unroll
      for (int j = 0; j < 16; ++j) { v[j] = lo[j]; v[16 + j] = 0.0f; }
    }
    tmem_ld_wait();
    const int nb = c.n0 + c0;
    if (c.plain) {                                        // warp-uniform: a third of the parameter loads and math
#pragma unroll
      for (int j = 0; j < CHUNK; j += 4) {
        const float4 b0 = *reinterpret_cast<const float4*>(&c.prm[0][(c0 + j) & 255]);
        v[j] += b0.x; v[j + 1] += b0.y; v[j + 2] += b0.z; v[j + 3] += b0.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < CHUNK; j += 4) {
        const float4 b0 = *reinterpret_cast<const float4*>(&c.prm[0][(c0 + j) & 255]);
        const float4 sc = *reinterpret_cast<const float4*>(&c.prm[1][(c0 + j) & 255]);
        const float4 sh = *reinterpret_cast<const float4*>(&c.prm[2][(c0 + j) & 255]);
        v[j] = fmaxf(v[j] + b0.x, c.relu_lo) * sc.x + sh.x;
        v[j + 1] = fmaxf(v[j + 1] + b0.y, c.relu_lo) * sc.y + sh.y;
        v[j + 2] = fmaxf(v[j + 2] + b0.z, c.relu_lo) * sc.z + sh.z;
        v[j + 3] = fmaxf(v[j + 3] + b0.w, c.relu_lo) * sc.w + sh.w;
      }
    }
    if (POOL) {
      float (*ex)[CHUNK] = c.exch[chunk_i & 1];
      if (lane == 31) {
#pragma unroll
        for (int j = 0; j < CHUNK; j += 4)
          *reinterpret_cast<float4*>(&ex[c.g][j]) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
      named_bar_sync(c.pool_bar, EPI_THREADS);
      const bool first_t = (c.t == 0);
#pragma unroll
      for (int j = 0; j < CHUNK; ++j) {
        float prev = __shfl_up_sync(0xffffffffu, v[j], 1);
        if (lane == 0) prev = (c.g > 0) ? ex[c.g - 1][j] : v[j];
        v[j] = first_t ? v[j] : fmaxf(v[j], prev);
      }
    }
    if (OUTT) {
      if (c.store) {
        if (c.residual) {                                 // rare: residual together with a transposed output
          const float* res = c.residual + c.r * c.res_ld + nb;
#pragma unroll
          for (int j = 0; j < CHUNK; ++j)
            if (c0 + j < c.bn && nb + j < c.N) v[j] += __ldg(res + j);
        }
        float* ot = c.out_t + (static_cast<long long>(c.b) * c.N + nb) * c.seq_len + c.t;
#pragma unroll
        for (int j = 0; j < CHUNK; ++j)
          if (c0 + j < c.bn && nb + j < c.N) ot[static_cast<long long>(j) * c.seq_len] = v[j];
      }
    }
    if (O16_FAST) {
      const bool cols_ok = c0 + 8 * c8 < c.bn && nb + 8 * c8 < c.N;          // N, nb and bn are multiples of 8
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if ((lane >> 4) == h) {
#pragma unroll
          for (int j = 0; j < CHUNK; j += 4)
            *reinterpret_cast<float4*>(&c.stg[lane & 15][j]) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        __syncwarp();
        if (cols_ok) {
          // all staging reads first, then the predicated stores: one shared-memory latency per half, not one per row
          float4 ya[2], yb[2];
#pragma unroll
          for (int i2 = 0; i2 < 2; ++i2) {
            ya[i2] = *reinterpret_cast<const float4*>(&c.stg[i2 * 8 + r8][8 * c8]);
            yb[i2] = *reinterpret_cast<const float4*>(&c.stg[i2 * 8 + r8][8 * c8 + 4]);
          }
#pragma unroll
          for (int i2 = 0; i2 < 2; ++i2) {
            const int row = h * 16 + i2 * 8 + r8;
            if ((c.smask >> row) & 1u) {
              const float4 y0 = ya[i2], y1 = yb[i2];
              uint4 pk;
              __half2 t;
              t = __floats2half2_rn(y0.x, y0.y); pk.x = *reinterpret_cast<const uint32_t*>(&t);
              t = __floats2half2_rn(y0.z, y0.w); pk.y = *reinterpret_cast<const uint32_t*>(&t);
              t = __floats2half2_rn(y1.x, y1.y); pk.z = *reinterpret_cast<const uint32_t*>(&t);
              t = __floats2half2_rn(y1.z, y1.w); pk.w = *reinterpret_cast<const uint32_t*>(&t);
              __half* dst = o16 + static_cast<long long>(h * 16 + i2 * 8) * c.out_ld + c0;
              *reinterpret_cast<uint4*>(dst) = pk;
              if (SPLIT) {
                // FTB_F_OUT_SPLIT: [hi | lo | hi] column blocks of width out_ld / 3 -- the K-concatenated operand of a
                // split-precision consumer (weights packed as [W_hi | W_hi | W_lo]): hi*W_hi + lo*W_hi + hi*W_lo
                const float yv[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
                const uint32_t hw[4] = {pk.x, pk.y, pk.z, pk.w};
                uint32_t lw[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&hw[u]));
                  const __half2 r = __floats2half2_rn(yv[2 * u] - f.x, yv[2 * u + 1] - f.y);
                  lw[u] = *reinterpret_cast<const uint32_t*>(&r);
                }
                const int seg = c.out_ld / 3;
                *reinterpret_cast<uint4*>(dst + seg) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
                *reinterpret_cast<uint4*>(dst + 2 * seg) = pk;
              }
            }
          }
        }
        __syncwarp();
      }
    } else if (optr || hptr) {
      // a half chunk (bn = 16 mod 32) must not spill into the columns of the next n-tile
      const bool cols_ok = c0 + 4 * part < c.bn && nb + 4 * part < c.N;      // N, nb and bn are multiples of 4
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if ((lane >> 4) == h) {
#pragma unroll
          for (int j = 0; j < CHUNK; j += 4)
            *reinterpret_cast<float4*>(&c.stg[lane & 15][j]) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        __syncwarp();
        if (cols_ok) {
          float4 ys[4];
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) ys[i4] = *reinterpret_cast<const float4*>(&c.stg[i4 * 4 + sub][4 * part]);
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const int i8 = h * 4 + i4;
            const int row = i8 * 4 + sub;
            if ((c.smask >> row) & 1u) {
              float4 y = ys[i4];
              if (RES) { y.x += xr_cur[i8].x; y.y += xr_cur[i8].y; y.z += xr_cur[i8].z; y.w += xr_cur[i8].w; }
              if (FULL) *reinterpret_cast<float4*>(fptr + i8 * ostep + c0) = y;
              if (OUT16) {
                const __half2 lo = __floats2half2_rn(y.x, y.y), hi = __floats2half2_rn(y.z, y.w);
                uint2 pk;
                pk.x = *reinterpret_cast<const uint32_t*>(&lo); pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(hptr + i8 * ostep + c0) = pk;
                if (SPLIT) {
                  // FTB_F_OUT_SPLIT: [hi | lo | hi] column blocks of width out_ld / 3 -- the K-concatenated operand of a
                  // split-precision consumer (weights packed as [W_hi | W_hi | W_lo]): hi*W_hi + lo*W_hi + hi*W_lo
                  const float2 f0 = __half22float2(lo), f1 = __half22float2(hi);
                  const __half2 r0 = __floats2half2_rn(y.x - f0.x, y.y - f0.y), r1 = __floats2half2_rn(y.z - f1.x, y.w - f1.y);
                  uint2 pr;
                  pr.x = *reinterpret_cast<const uint32_t*>(&r0); pr.y = *reinterpret_cast<const uint32_t*>(&r1);
                  *reinterpret_cast<uint2*>(hptr + i8 * ostep + c0 + c.out_ld / 3) = pr;
                  *reinterpret_cast<uint2*>(hptr + i8 * ostep + c0 + 2 * (c.out_ld / 3)) = pk;
                }
              } else {
                if (RND) { y.x = round_tf32(y.x); y.y = round_tf32(y.y); y.z = round_tf32(y.z); y.w = round_tf32(y.w); }
                *reinterpret_cast<float4*>(optr + i8 * ostep + c0) = y;
              }
            }
          }
        }
        __syncwarp();
      }
    }
    if (RES) {
#pragma unroll
      for (int i8 = 0; i8 < 8; ++i8) xr_cur[i8] = xr_nxt[i8];
    }
  }
}

template <bool POOL, bool RES, bool FULL, bool OUTT>
__device__ __forceinline__ void epi_disp_rnd(const EpiCtx& c, bool rnd) {
  if (rnd) epi_affine_tile<POOL, RES, FULL, OUTT, true>(c);
  else epi_affine_tile<POOL, RES, FULL, OUTT, false>(c);
}
template <bool POOL, bool RES, bool FULL>
__device__ __forceinline__ void epi_disp_outt(const EpiCtx& c, bool outt, bool rnd) {
  if (outt) epi_disp_rnd<POOL, false, FULL, true>(c, rnd);      // with out_t the residual (if any) is added in phase A
  else epi_disp_rnd<POOL, RES, FULL, false>(c, rnd);
}
template <bool POOL, bool RES>
__device__ __forceinline__ void epi_disp_full(const EpiCtx& c, bool full, bool outt, bool rnd) {
  if (full) epi_disp_outt<POOL, RES, true>(c, outt, rnd);
  else epi_disp_outt<POOL, RES, false>(c, outt, rnd);
}
__device__ __forceinline__ void epi_dispatch(const EpiCtx& c, bool pool, bool res, bool full, bool outt, bool rnd) {
  if (c.out16) {                                                  // fp16 operand copy as `out`
    if (c.split) {
      if (pool) epi_affine_tile<true, false, false, false, false, true, true>(c);
      else epi_affine_tile<false, false, false, false, false, true, true>(c);
      return;
    }
    if (pool) epi_affine_tile<true, false, false, false, false, true>(c);        // CBHG bank: never residual / twin
    else if (outt) {
      if (full) epi_affine_tile<false, false, true, true, false, true>(c);
      else epi_affine_tile<false, false, false, true, false, true>(c);
    } else if (res) {
      if (full) epi_affine_tile<false, true, true, false, false, true>(c);
      else epi_affine_tile<false, true, false, false, false, true>(c);
    } else {
      if (full) epi_affine_tile<false, false, true, false, false, true>(c);
      else epi_affine_tile<false, false, false, false, false, true>(c);
    }
    return;
  }
  if (pool) epi_disp_full<true, false>(c, full, outt, rnd);     // pooled layers (CBHG bank) never carry a residual
  else if (res) epi_disp_full<false, true>(c, full, outt, rnd);
  else epi_disp_full<false, false>(c, full, outt, rnd);
}

// ------------------------------------------------------------------------------------------
template <bool PAIR>
__device__ __forceinline__ void conv_gemm_tc_body(const GemmProblemDev* __restrict__ probs, int nprob) {
  extern __shared__ uint8_t smem_raw[];
  // schedule slots at the plan's CURRENT geometry: plan metadata the host uploads (ftb_gemm_plan_set_geometry), not an
  // output of the preceding kernel, and not a kernel argument -- a captured CUDA graph replays with whatever sequence
  // length the records hold
  const int total_tiles = probs[0].plan_tiles;
  // 1024-byte alignment by pointer arithmetic on the __shared__ array (an integer round trip would demote every
  // later access through `tail` to generic LD/ST, which the epilogue pays for in latency)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  SmemTail* tail = reinterpret_cast<SmemTail*>(smem + STAGES * STAGE_BYTES);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
#ifdef FTB_GEMM_PROFILE
  long long gp[5] = {0, 0, 0, 0, 0}, gp_t0 = 0;
  const long long gp_start = clock64();
#define GP_T0 gp_t0 = clock64();
#define GP_ACC(i) gp[i] += clock64() - gp_t0;
#else
#define GP_T0
#define GP_ACC(i)
#endif

  // CTA-pair mode: the two CTAs of a cluster share one schedule slot and run M = 256 MMAs (cta_group::2) issued by the
  // even CTA; each loads its own A tile and HALF of the W tile, which is what lifts the 64 B/clk per-SM operand ingest
  // bound of the 128 x 256 single-CTA tile (profiles/r1t_gemm_ingest.txt).
  constexpr int NSTAGE = PAIR ? PAIR_STAGES : STAGES;
  constexpr int SBYTES = PAIR ? PAIR_STAGE_BYTES : STAGE_BYTES;
  const int rank = PAIR ? static_cast<int>(cta_rank_in_cluster()) : 0;
  const bool leader = rank == 0;
  const int pair_rank = PAIR ? rank : -1;
  const int sched_c = PAIR ? (blockIdx.x >> 1) : blockIdx.x;
  const int sched_g = PAIR ? (gridDim.x >> 1) : gridDim.x;
  pdl_launch_dependents();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < NSTAGE; ++s) {
      mbar_init(&tail->full[s], 1);
      mbar_init(&tail->empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tail->tmem_full[b], 1);
      mbar_init(&tail->tmem_empty[b], PAIR ? 8 : 4);   // one arrive per epilogue warp (of both CTAs) draining buffer b
    }
    fence_mbar_init();
  }
  if (warp == 1) {                              // whole warp: .sync.aligned
    if (PAIR) { tmem_alloc_pair(&tail->tmem_base, TMEM_COLS); tmem_relinquish_pair(); }
    else { tmem_alloc(&tail->tmem_base, TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  if (PAIR) cluster_barrier();      // the peer's barriers exist before any remote arrive / multicast commit
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;
  // everything above (barriers, TMEM) overlapped the tail of the previous kernel; its outputs are read from here on
  pdl_wait();

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const GemmProblemDev* last_p = nullptr;
      for (int it = 0;; ++it) {
        const int tile = sched_tile(it, sched_c, sched_g);
        if (tile >= total_tiles) break;
        const TileCoord tc = decode_tile(probs, nprob, tile, pair_rank);
        const GemmProblemDev* p = tc.p;
        // pair: the leader's barrier collects both CTAs' A tiles and W halves
        const uint32_t tx = PAIR ? 2u * (A_STAGE_BYTES + static_cast<uint32_t>(p->bn / 2) * BK * 4)
                                 : A_STAGE_BYTES + static_cast<uint32_t>(p->bn) * BK * 4;
        const int w_row0 = tc.n0 + (PAIR ? rank * (p->bn / 2) : 0);
        if (p != last_p) {
          tma_desc_fence_acquire(&p->tm_a);
          tma_desc_fence_acquire(&p->tm_w);
          last_p = p;
        }
        // problem fields -> registers: the "memory" clobbers of the barrier/TMA asm would otherwise force a
        // re-load of every p->field from global memory inside the loops
        const int taps = p->taps, kblocks = p->kblocks, tap0 = p->tap0, w_tap_rows = p->w_tap_rows;
        const int kelems = p->kelems;
        const CUtensorMap* tm_a = &p->tm_a;
        const CUtensorMap* tm_w = &p->tm_w;
        // channel block outer, taps inner: the taps of one channel block re-read the same activation rows shifted by
        // one, so they hit in L2 back to back (taps-outer re-read the whole [128 x C_in] slab from DRAM per tap:
        // postnet.proj1 moved 533 MB for a 214 MB input, profiles/r1j_launch_summary.txt)
        for (int kb = 0; kb < kblocks; ++kb) {
          for (int tap = 0; tap < taps; ++tap) {
            GP_T0 mbar_wait(&tail->empty[stage], phase ^ 1u); GP_ACC(0)
            uint8_t* sa = smem + stage * SBYTES;
            uint8_t* sb = sa + A_STAGE_BYTES;
            if (PAIR) {
              if (leader) mbar_expect_tx(&tail->full[stage], tx);
              tma_load_2d_pair(sa, tm_a, &tail->full[stage], kb * kelems, tc.m0 + tap0 + tap);
              tma_load_2d_pair(sb, tm_w, &tail->full[stage], kb * kelems, tap * w_tap_rows + w_row0);
            } else {
              mbar_expect_tx(&tail->full[stage], tx);
              tma_load_2d(sa, tm_a, &tail->full[stage], kb * kelems, tc.m0 + tap0 + tap);
              tma_load_2d(sb, tm_w, &tail->full[stage], kb * kelems, tap * w_tap_rows + tc.n0);
            }
            if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0;; ++it) {
        const int tile = sched_tile(it, sched_c, sched_g);
        if (tile >= total_tiles) break;
        const TileCoord tc = decode_tile(probs, nprob, tile, pair_rank);
        const GemmProblemDev* p = tc.p;
        const int buf = it & 1;
        GP_T0 mbar_wait(&tail->tmem_empty[buf], ((it >> 1) & 1u) ^ 1u); GP_ACC(1)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        const bool f16 = (p->flags & FTB_F_A_F16) != 0;
        const uint32_t mma_m = PAIR ? 2 * BM : BM;
        const uint32_t idesc = f16 ? umma_idesc_f16(mma_m, p->bn) : umma_idesc_tf32(mma_m, p->bn);
        const int kiters = p->taps * p->kblocks;
        for (int ki = 0; ki < kiters; ++ki) {
          GP_T0 mbar_wait(&tail->full[stage], phase); GP_ACC(2)
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * SBYTES);
          const uint64_t da = umma_desc_sw128(sa);
          const uint64_t db = umma_desc_sw128(sa + A_STAGE_BYTES);
          // +32 bytes per MMA (8 tf32 / 16 fp16 channels) inside the 128-byte swizzle row: +2 in (addr >> 4) units
          if (PAIR) {
            if (f16) {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k)
                umma_f16_ss_pair(d_tmem, da + 2u * k, db + 2u * k, idesc, (ki | k) != 0 ? 1u : 0u);
            } else {
#pragma unroll
              for (int k = 0; k < BK / UMMA_K; ++k)
                umma_tf32_pair(d_tmem, da + 2u * k, db + 2u * k, idesc, (ki | k) != 0 ? 1u : 0u);
            }
            umma_commit_pair(&tail->empty[stage]);      // frees the slot in BOTH CTAs
          } else {
          if (f16) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) umma_f16_ss(d_tmem, da + 2u * k, db + 2u * k, idesc, (ki | k) != 0 ? 1u : 0u);
          } else {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d_tmem, da + 2u * k, db + 2u * k, idesc, (ki | k) != 0 ? 1u : 0u);
          }
          umma_commit(&tail->empty[stage]);     // frees the smem slot when these MMAs retire
          }
          if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
        }
        if (PAIR) umma_commit_pair(&tail->tmem_full[buf]);      // both CTAs' epilogues
        else umma_commit(&tail->tmem_full[buf]);     // accumulator complete -> epilogue
      }
    }
  } else {
    // ================================ epilogue ====================================
    // Phase A (lane = accumulator row): tcgen05.ld 32 columns -> bias/ReLU/affine (+ max-pool) -> staging tile.
    // Phase B (lane = 4 consecutive columns of a row): staging tile -> (+ residual) -> 128-byte coalesced
    // row-segment stores of `out` (tf32-rounded) / `out_full` (unrounded).  `out_t` ([B, n, L]) is written
    // in phase A, where consecutive lanes are consecutive t.
    const int g = warp & 3;                       // TMEM lane quarter this warp may read
    const int set = (warp - 2) >> 2;              // 0: warps 2..5 (even tiles), 1: warps 6..9 (odd tiles)
    const int etid = g * 32 + lane;
    const uint32_t bar_prm = 1 + 2 * set, bar_pool = 2 + 2 * set;      // named barriers of this set
    float (*stg)[STG_LD] = tail->stg[set * 4 + g];
    for (int it = set;; it += 2) {
      const int tile = sched_tile(it, sched_c, sched_g);
      if (tile >= total_tiles) break;
      const TileCoord tc = decode_tile(probs, nprob, tile, pair_rank);
      const GemmProblemDev* p = tc.p;
      const int buf = set;                        // == it & 1
      // problem fields -> registers ONCE per tile: this warp stores to global memory, so without the copies the
      // compiler re-loads p->out / p->residual / ... from global memory (a long-scoreboard stall) in every
      // iteration of the store loops -- measured as the dominant epilogue stall (profiles/r1e_gemm_roles.txt)
      const int N = p->n, P_bn = p->bn, P_res_ld = p->res_ld, P_out_ld = p->out_ld, P_out_col0 = p->out_col0;
      const int P_seq_len = p->seq_len, P_seq_stride = p->seq_stride, P_flags = p->flags, P_epi = p->epilogue;
      const long long P_rows = p->rows;
      const float* const P_bias0 = p->bias0;
      const float* const P_scale = p->scale;
      const float* const P_shift = p->shift;
      const float* const P_residual = p->residual;
      float* const P_out = p->out;
      float* const P_out_full = p->out_full;
      float* const P_out_t = p->out_t;
      const bool highway = (P_epi == FTB_EPI_HIGHWAY);
      const bool pool = (P_epi == FTB_EPI_MAXPOOL);

      // ---- per-tile column parameters -> smem (overlaps the tail of this tile's MMAs)
      float (*prm)[256] = tail->params[set];
      named_bar_sync(bar_prm, EPI_THREADS);       // every warp of the set is done with the previous tile's parameters
      for (int c = etid; c < P_bn; c += EPI_THREADS) {
        float b0 = 0.0f, sc = 1.0f, sh = 0.0f;
        if (highway) {
          const int U = P_bn >> 1;
          const int u = (tc.n0 >> 1) + (c < U ? c : c - U);
          b0 = __ldg(P_bias0 + (c < U ? u : (N >> 1) + u));
        } else if (tc.n0 + c < N) {
          if (P_bias0) b0 = __ldg(P_bias0 + tc.n0 + c);
          if (P_scale) sc = __ldg(P_scale + tc.n0 + c);
          if (P_shift) sh = __ldg(P_shift + tc.n0 + c);
        }
        prm[0][c] = b0; prm[1][c] = sc; prm[2][c] = sh;
      }
      named_bar_sync(bar_prm, EPI_THREADS);

      GP_T0 mbar_wait(&tail->tmem_full[buf], (it >> 1) & 1u); GP_ACC(3)
      tc_fence_after();
      GP_T0

      const int lrow = g * 32 + lane;
      const long long r = static_cast<long long>(tc.m0) + lrow;
      int t = 0, b = 0;
      bool valid = (r >= 0) && (r < P_rows);
      if (valid) {
        b = static_cast<int>(r / P_seq_stride);
        t = static_cast<int>(r - static_cast<long long>(b) * P_seq_stride);
        valid = t < P_seq_len;
      }
      const bool store = valid && !(pool && lrow == 0);        // pooled tiles: row 0 is the halo row
      const unsigned smask = __ballot_sync(0xffffffffu, store);
      const long long r_warp = static_cast<long long>(tc.m0) + g * 32;     // first row of this warp
      const uint32_t taddr = tmem_base + buf * 256 + (static_cast<uint32_t>(g * 32) << 16);
      const bool relu = (P_flags & FTB_F_RELU) != 0;
      const bool rnd = (P_flags & FTB_F_ROUND_TF32) != 0;

      if (highway) {
        // Phase A (lane = row): x1/x2 chunks of 16 units from TMEM, this row's residual (prefetched one chunk
        // ahead), the gate math on all 32 lanes, result staged.  Phase B: 64-byte row segments, coalesced.
        const int U = P_bn >> 1;
        const int u0 = tc.n0 >> 1;
        const float* rrow = P_residual + r * P_res_ld + u0;                  // this lane's residual row
        const int brow = lane >> 2, c4 = lane & 3;                           // phase-B coordinates: 8 rows x 4 float4
        const bool h16 = (P_flags & FTB_F_OUT_F16) != 0;                     // `out` is the fp16 operand copy
        const long long ooff = (r_warp + brow) * P_out_ld + P_out_col0 + u0 + 4 * c4;
        float* optr = P_out + ooff;
        __half* hptr = reinterpret_cast<__half*>(P_out) + ooff;
        float* fptr = P_out_full ? P_out_full + (r_warp + brow) * P_out_ld + P_out_col0 + u0 + 4 * c4 : nullptr;
        const long long ostep = 8ll * P_out_ld;
        float* hstg = &stg[0][0];       // [32 rows][16 floats], float4 slots XOR-swizzled by (row >> 1) & 3: conflict-free
        float4 xr_cur[4], xr_nxt[4];
        auto load_res = [&](float4 (&dst)[4], int j0) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            dst[q] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (store && j0 < U) dst[q] = *reinterpret_cast<const float4*>(rrow + j0 + 4 * q);
          }
        };
        load_res(xr_cur, 0);
        for (int j0 = 0; j0 < U; j0 += 16) {
          load_res(xr_nxt, j0 + 16);
          float x1[16], x2[16];
          tmem_ld_32x16(taddr + j0, x1);
          tmem_ld_32x16(taddr + U + j0, x2);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 b1 = *reinterpret_cast<const float4*>(&prm[0][j0 + 4 * q]);
            const float4 b2 = *reinterpret_cast<const float4*>(&prm[0][U + j0 + 4 * q]);
            float4 y;
            y.x = epi_highway(x1[4 * q] + b1.x, x2[4 * q] + b2.x, xr_cur[q].x);
            y.y = epi_highway(x1[4 * q + 1] + b1.y, x2[4 * q + 1] + b2.y, xr_cur[q].y);
            y.z = epi_highway(x1[4 * q + 2] + b1.z, x2[4 * q + 2] + b2.z, xr_cur[q].z);
            y.w = epi_highway(x1[4 * q + 3] + b1.w, x2[4 * q + 3] + b2.w, xr_cur[q].w);
            *reinterpret_cast<float4*>(hstg + lane * 16 + 4 * (q ^ ((lane >> 1) & 3))) = y;
          }
          __syncwarp();
          float4 ys[4];
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const int row = i4 * 8 + brow;
            ys[i4] = *reinterpret_cast<const float4*>(hstg + row * 16 + 4 * (c4 ^ ((row >> 1) & 3)));
          }
#pragma unroll
          for (int i4 = 0; i4 < 4; ++i4) {
            const int row = i4 * 8 + brow;
            if ((smask >> row) & 1u) {
              float4 y = ys[i4];
              if (fptr) *reinterpret_cast<float4*>(fptr + i4 * ostep + j0) = y;
              if (h16) {
                const __half2 lo = __floats2half2_rn(y.x, y.y), hi = __floats2half2_rn(y.z, y.w);
                uint2 pk;
                pk.x = *reinterpret_cast<const uint32_t*>(&lo); pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(hptr + i4 * ostep + j0) = pk;
              } else {
                if (rnd) { y.x = round_tf32(y.x); y.y = round_tf32(y.y); y.z = round_tf32(y.z); y.w = round_tf32(y.w); }
                *reinterpret_cast<float4*>(optr + i4 * ostep + j0) = y;
              }
            }
          }
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 4; ++q) xr_cur[q] = xr_nxt[q];
        }
      } else {
        EpiCtx c;
        c.prm = prm; c.stg = stg; c.exch = tail->exch[set]; c.pool_bar = bar_pool; c.taddr = taddr;
        c.bn = P_bn; c.n0 = tc.n0; c.N = N; c.g = g; c.lane = lane; c.t = t; c.b = b; c.seq_len = P_seq_len;
        c.res_ld = P_res_ld; c.out_ld = P_out_ld; c.smask = smask; c.store = store; c.r = r; c.r_warp = r_warp;
        c.relu_lo = relu ? 0.0f : -INFINITY;
        c.plain = !relu && P_scale == nullptr && P_shift == nullptr;
        c.residual = P_residual;
        const bool out16 = (P_flags & FTB_F_OUT_F16) != 0;
        c.out = (P_out && !out16) ? P_out + P_out_col0 : nullptr;
        c.out16 = out16 ? reinterpret_cast<__half*>(P_out) + P_out_col0 : nullptr;
        c.out_full = P_out_full ? P_out_full + P_out_col0 : nullptr;
        c.out_t = P_out_t;
        c.split = (P_flags & FTB_F_OUT_SPLIT) != 0;
        epi_dispatch(c, pool, P_residual != nullptr && P_out_t == nullptr, P_out_full != nullptr, P_out_t != nullptr, rnd);
      }
      tc_fence_before();
      __syncwarp();
      GP_ACC(4)
      if (lane == 0) {
        if (PAIR) mbar_arrive_leader(&tail->tmem_empty[buf]);
        else mbar_arrive(&tail->tmem_empty[buf]);
      }
    }
  }
#ifdef FTB_GEMM_PROFILE
  if (blockIdx.x == 0 && lane == 0 && (warp <= 2 || warp == 6)) {
    const long long tot = clock64() - gp_start;
    printf("gemm prof warp %d tiles %d total %lld | wait_empty %lld wait_tmem_empty %lld wait_full %lld wait_tmem_full %lld epi_work %lld\n",
           warp, (total_tiles + gridDim.x - 1) / gridDim.x, tot, gp[0], gp[1], gp[2], gp[3], gp[4]);
  }
#endif

  tc_fence_before();
  if (PAIR) cluster_barrier();      // neither CTA frees tensor memory (or leaves) while the pair's MMAs may still use it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_gemm_tc_kernel(const GemmProblemDev* __restrict__ probs, int nprob) {
  conv_gemm_tc_body<false>(probs, nprob);
}
// CTA-pair form: launched with a cluster dimension of 2; `total_tiles` counts tile PAIRS.
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_gemm_tc_pair_kernel(const GemmProblemDev* __restrict__ probs, int nprob) {
  conv_gemm_tc_body<true>(probs, nprob);
}

// ------------------------------------------------------------------------------------------
// Plain fp32 SIMT kernel over the same problem records: one thread per output element.
// TEST-ONLY cross-check for the descriptor/swizzle/pipeline logic above (impl = 1).
__device__ __forceinline__ float simt_acc(const GemmProblemDev* p, long long r, int wrow) {
  float acc = 0.0f;
  for (int tap = 0; tap < p->taps; ++tap) {
    const long long ra = r + p->tap0 + tap;
    // (folded problems, FTB_F_TAPS_IN_K: `a` starts FTB_FOLD_LEAD_ROWS zero rows early, tap0 is shifted by as much, and
    //  the single tap walks taps*cin contiguous elements = the original rows, so the same loop below covers them)
    if (ra < 0 || ra >= p->rows + ((p->flags & FTB_F_TAPS_IN_K) ? FTB_FOLD_LEAD_ROWS : 0)) continue;
    if (p->flags & FTB_F_A_F16) {
      const __half* a = reinterpret_cast<const __half*>(p->a) + ra * p->lda;
      const __half* w = reinterpret_cast<const __half*>(p->w) + (static_cast<long long>(tap) * p->w_tap_rows + wrow) * p->cin;
      for (int c = 0; c < p->cin; ++c) acc = fmaf(__half2float(a[c]), __half2float(w[c]), acc);
      continue;
    }
    const float* a = p->a + ra * p->lda;
    const float* w = p->w + (static_cast<long long>(tap) * p->w_tap_rows + wrow) * p->cin;
    for (int c = 0; c < p->cin; ++c) acc = fmaf(a[c], w[c], acc);
  }
  return acc;
}

__global__ void conv_gemm_simt_kernel(const GemmProblemDev* __restrict__ probs, int nprob) {
  const GemmProblemDev* p = probs + blockIdx.y;
  const int N = (p->epilogue == FTB_EPI_HIGHWAY) ? p->n / 2 : p->n;
  const long long total = p->rows * N;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / N;
    const int n = static_cast<int>(i - r * N);
    const int b = static_cast<int>(r / p->seq_stride);
    const int t = static_cast<int>(r - static_cast<long long>(b) * p->seq_stride);
    if (t >= p->seq_len) continue;
    float v;
    if (p->epilogue == FTB_EPI_HIGHWAY) {
      const int U = p->bn / 2;
      const int tile = n / U, j = n % U;
      const float x1 = simt_acc(p, r, tile * p->bn + j) + p->bias0[n];
      const float x2 = simt_acc(p, r, tile * p->bn + U + j) + p->bias0[N + n];
      v = epi_highway(x1, x2, p->residual[r * p->res_ld + n]);
    } else {
      v = epi_affine(p, simt_acc(p, r, n), n);
      if (p->epilogue == FTB_EPI_MAXPOOL && t > 0) v = fmaxf(v, epi_affine(p, simt_acc(p, r - 1, n), n));
      if (p->residual) v += p->residual[r * p->res_ld + n];
      if (p->out_t) p->out_t[(static_cast<long long>(b) * N + n) * p->seq_len + t] = v;
    }
    if (p->out_full) p->out_full[r * p->out_ld + p->out_col0 + n] = v;
    if (p->out && (p->flags & FTB_F_OUT_SPLIT)) {
      __half* o = reinterpret_cast<__half*>(p->out) + r * p->out_ld + p->out_col0 + n;
      const __half hi = __float2half_rn(v);
      const int seg = p->out_ld / 3;
      o[0] = hi; o[seg] = __float2half_rn(v - __half2float(hi)); o[2 * seg] = hi;
    } else if (p->out && (p->flags & FTB_F_OUT_F16))
      reinterpret_cast<__half*>(p->out)[r * p->out_ld + p->out_col0 + n] = __float2half_rn(v);
    else if (p->out) p->out[r * p->out_ld + p->out_col0 + n] = epi_finish(p, v);
  }
}

// ------------------------------------------------------------------------------------------
// Host side: tensor maps and plans.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// 2-D row-major [rows, cols] fp32 or fp16 (leading dimension ld elements), box {128 bytes, box_rows}, 128B swizzle.
int tensor_map_2d(CUtensorMap* m, const void* base, long long rows, int cols, int ld, int box_rows, bool f16) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return FTB_E_DRIVER;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * (f16 ? 2 : 4)};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(f16 ? 2 * BK : BK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult res = enc(m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                     const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return res == CUDA_SUCCESS ? 0 : FTB_E_SHAPE;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int gemm_plan_build(const ftb_gemm_problem* problems, int nproblems, GemmPlan* plan) {
  if (!problems || nproblems < 1 || nproblems > FTB_MAX_PROBLEMS) return FTB_E_ARG;
  std::vector<GemmProblemDev> host(nproblems);
  // CTA pairs (cta_group::2 MMAs; each CTA stages half of the W tile, so the ring holds six stages instead of four).
  // Measured (profiles/r1t_gemm_pair.txt): the deeper ring pays on launches whose tiles run many k-blocks (banks,
  // proj1: -12..15 %) and the cluster launch + cross-CTA barriers cost 2-3 us on launches of one short tile per CTA.
  // Default: by the work-weighted mean k-block count per tile; FTB_GEMM_PAIR=0/1 forces it.
  const char* pair_env = getenv("FTB_GEMM_PAIR");
  bool pair;
  if (pair_env && (pair_env[0] == '0' || pair_env[0] == '1')) {
    pair = pair_env[0] == '1';
  } else {
    double tiles = 0.0, kiters = 0.0;
    for (int i = 0; i < nproblems; ++i) {
      const ftb_gemm_problem& q = problems[i];
      if (q.bn < 16 || q.rows < 1 || q.cin < 1) continue;         // rejected below
      const int kel = (q.flags & FTB_F_A_F16) ? 2 * BK : BK;
      const double t = static_cast<double>((q.rows + BM - 1) / BM) * ((q.n + q.bn - 1) / q.bn);
      tiles += t;
      kiters += t * ((q.flags & FTB_F_TAPS_IN_K) ? (q.taps * q.cin + kel - 1) / kel : q.taps * ((q.cin + kel - 1) / kel));
    }
    pair = tiles > 0.0 && kiters / tiles >= GEMM_PAIR_MIN_KITERS;
  }
  int tile_base = 0;
  long long simt_elems = 0;
  for (int i = 0; i < nproblems; ++i) {
    const ftb_gemm_problem& q = problems[i];
    GemmProblemDev& d = host[i];
    if (!q.a || !q.w || (!q.out && !q.out_t)) return FTB_E_ARG;
    if (q.rows < 1 || q.cin < 1 || q.taps < 1 || q.n < 1) return FTB_E_ARG;
    if (q.bn < 16 || q.bn > 256 || (q.bn % 16) != 0) return FTB_E_SHAPE;
    if (q.w_tap_rows < q.n || (q.w_tap_rows % q.bn) != 0) return FTB_E_SHAPE;
    const bool a16 = (q.flags & FTB_F_A_F16) != 0, o16 = (q.flags & FTB_F_OUT_F16) != 0;
    const int amul = a16 ? 8 : 4, omul = o16 ? 8 : 4;        // elements per 16 bytes
    if ((q.lda % amul) != 0 || (q.cin % amul) != 0 || q.cin > q.lda) return FTB_E_ALIGN;
    if (!aligned16(q.a) || !aligned16(q.w)) return FTB_E_ALIGN;
    if (q.out && (!aligned16(q.out) || (q.out_ld % omul) != 0 || (q.out_col0 % omul) != 0)) return FTB_E_ALIGN;
    if (q.out && (q.n % omul) != 0) return FTB_E_SHAPE;       // row segments are stored as 16-byte vectors
    if (o16 && (!q.out || (q.epilogue == FTB_EPI_MAXPOOL && (q.residual || q.out_t || q.out_full)))) return FTB_E_ARG;
    if (q.flags & FTB_F_OUT_SPLIT) {
      if (!o16 || q.epilogue == FTB_EPI_HIGHWAY || q.residual || q.out_t || q.out_full) return FTB_E_ARG;
      if ((q.out_ld % 24) != 0 || q.out_col0 + q.n > q.out_ld / 3 || (q.n % 8) != 0) return FTB_E_SHAPE;
    }
    if (q.residual && ((q.res_ld % 4) != 0 || !aligned16(q.residual))) return FTB_E_ALIGN;
    if (q.batch < 1 || q.seq_len < 1 || q.seq_stride < q.seq_len) return FTB_E_ARG;
    if (q.rows != static_cast<long long>(q.batch) * q.seq_stride) return FTB_E_ARG;
    if (q.epilogue == FTB_EPI_HIGHWAY) {
      if (!q.residual || !q.bias0 || !q.out || (q.n % 2) != 0 || (q.bn % 32) != 0 || (q.n % q.bn) != 0)
        return FTB_E_SHAPE;
      if ((q.res_ld % 4) != 0 || !aligned16(q.residual)) return FTB_E_ALIGN;
    } else if (q.epilogue != FTB_EPI_AFFINE && q.epilogue != FTB_EPI_MAXPOOL) {
      return FTB_E_ARG;
    }
    const bool fold = (q.flags & FTB_F_TAPS_IN_K) != 0;
    if (fold && (!a16 || q.lda != q.cin || q.taps > FTB_FOLD_LEAD_ROWS || q.tap0 < -FTB_FOLD_LEAD_ROWS || q.tap0 > 0)) return FTB_E_ARG;
    d.a = static_cast<const float*>(q.a); d.w = static_cast<const float*>(q.w); d.bias0 = q.bias0; d.scale = q.scale; d.shift = q.shift;
    d.residual = q.residual; d.out = static_cast<float*>(q.out); d.out_t = q.out_t; d.out_full = q.out_full;
    if (q.out_full && (!q.out || !aligned16(q.out_full))) return FTB_E_ALIGN;
    d.rows = q.rows; d.lda = q.lda; d.cin = q.cin; d.taps = q.taps; d.tap0 = q.tap0;
    d.w_tap_rows = q.w_tap_rows; d.n = q.n; d.bn = q.bn; d.epilogue = q.epilogue; d.flags = q.flags;
    d.res_ld = q.res_ld; d.out_ld = q.out_ld; d.out_col0 = q.out_col0;
    d.batch = q.batch; d.seq_len = q.seq_len; d.seq_stride = q.seq_stride;
    d.kelems = a16 ? 2 * BK : BK;
    d.kblocks = (q.cin + d.kelems - 1) / d.kelems;
    d.n_tiles = (q.n + q.bn - 1) / q.bn;
    if (q.epilogue == FTB_EPI_MAXPOOL) {
      d.m_step = BM - 1; d.m_shift = -1;
    } else {
      d.m_step = BM; d.m_shift = 0;
    }
    d.m_tiles = static_cast<int>((q.rows + d.m_step - 1) / d.m_step);
    d.cap_rows = q.rows;
    d.tile_base = tile_base;
    tile_base += (pair ? (d.m_tiles + 1) / 2 : d.m_tiles) * d.n_tiles;
    int rc;
    if (fold) {
      // Taps folded into K (FTB_F_TAPS_IN_K): ONE tap over an overlapping-row view of A.  Virtual row v starts at real row
      // v - LEAD of `a` (the LEAD zero rows in front of `a` make v >= 0 for every output row) and extends over taps*cin
      // contiguous elements, i.e. the real rows v - LEAD .. v - LEAD + taps - 1; columns past taps*cin and virtual rows
      // past the end are zero-filled by TMA, and the rows behind the last sentence are gap rows (zero) anyway.
      const int kext = q.taps * q.cin;
      const __half* base = static_cast<const __half*>(q.a) - static_cast<long long>(FTB_FOLD_LEAD_ROWS) * q.lda;
      d.a = reinterpret_cast<const float*>(base);
      d.tap0 = q.tap0 + FTB_FOLD_LEAD_ROWS;
      d.taps = 1;
      d.cin = kext;
      d.kblocks = (kext + d.kelems - 1) / d.kelems;
      d.lda = q.lda;                                  // real row pitch (the SIMT cross-check walks the same memory)
      rc = tensor_map_2d(&d.tm_a, base, q.rows + FTB_FOLD_LEAD_ROWS - q.taps + 1, kext, q.lda, BM, true);
      if (rc) return rc;
      rc = tensor_map_2d(&d.tm_w, q.w, q.w_tap_rows, kext, kext, pair ? q.bn / 2 : q.bn, true);
      if (rc) return rc;
    } else {
      rc = tensor_map_2d(&d.tm_a, q.a, q.rows, q.cin, q.lda, BM, a16);
      if (rc) return rc;
      rc = tensor_map_2d(&d.tm_w, q.w, static_cast<long long>(q.taps) * q.w_tap_rows, q.cin, q.cin, pair ? q.bn / 2 : q.bn, a16);
      if (rc) return rc;
    }
    const long long e = q.rows * q.n;
    if (e > simt_elems) simt_elems = e;
  }
  for (GemmProblemDev& d : host) d.plan_tiles = tile_base;
  GemmProblemDev* dev = nullptr;
  cudaError_t ce = cudaMalloc(&dev, sizeof(GemmProblemDev) * nproblems);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  ce = cudaMemcpy(dev, host.data(), sizeof(GemmProblemDev) * nproblems, cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) { cudaFree(dev); return static_cast<int>(ce); }
  plan->dev = dev; plan->nprob = nproblems; plan->total_tiles = tile_base; plan->simt_elems = simt_elems;
  plan->pair = pair;
  plan->cap_tiles = tile_base;
  plan->host = new std::vector<GemmProblemDev>(std::move(host));
  return 0;
}

void gemm_plan_free(GemmPlan& plan) {
  cudaFree(plan.dev);
  delete plan.host;
  plan = GemmPlan();
}

// Launch-time geometry: the same buffers (allocated for cap_rows) hold `batch` sentences of `seq_len` rows at stride
// `seq_stride` -- rows, tile grid and schedule are recomputed and the records re-uploaded in stream order, so a program
// built once per capacity bucket serves every sequence length inside it (generate() sees a new length on almost every
// call with real text).  The caller keeps the rows t in [seq_len, seq_stride) of every sentence zero.
int gemm_plan_set_geometry(GemmPlan& plan, int seq_len, int seq_stride, cudaStream_t stream) {
  if (!plan.host || seq_len < 1 || seq_stride < seq_len) return FTB_E_ARG;
  std::vector<GemmProblemDev>& host = *plan.host;
  int tile_base = 0;
  for (GemmProblemDev& d : host) {
    const long long rows = static_cast<long long>(d.batch) * seq_stride;
    if (rows > d.cap_rows) return FTB_E_LIMIT;
    d.rows = rows; d.seq_len = seq_len; d.seq_stride = seq_stride;
    d.m_tiles = static_cast<int>((rows + d.m_step - 1) / d.m_step);
    d.tile_base = tile_base;
    tile_base += (plan.pair ? (d.m_tiles + 1) / 2 : d.m_tiles) * d.n_tiles;
  }
  if (tile_base > plan.cap_tiles) return FTB_E_LIMIT;
  for (GemmProblemDev& d : host) d.plan_tiles = tile_base;
  plan.total_tiles = tile_base;
  // pageable source: the runtime stages the bytes before the call returns, so `host` may change again afterwards
  return static_cast<int>(cudaMemcpyAsync(plan.dev, host.data(), sizeof(GemmProblemDev) * host.size(),
                                          cudaMemcpyHostToDevice, stream));
}

int gemm_plan_launch(const GemmPlan& plan, int impl, int num_sms, cudaStream_t stream) {
  if (impl == 1) {
    long long blocks = (plan.simt_elems + 255) / 256;
    if (blocks > 65535) blocks = 65535;
    if (blocks < 1) blocks = 1;
    dim3 grid(static_cast<unsigned>(blocks), static_cast<unsigned>(plan.nprob));
    conv_gemm_simt_kernel<<<grid, 256, 0, stream>>>(plan.dev, plan.nprob);
    return static_cast<int>(cudaGetLastError());
  }
  // function attributes are per device: set them once for every device a plan is launched on
  static bool attr_set[64] = {};
  int dev_id = 0;
  cudaError_t ce = cudaGetDevice(&dev_id);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  if (dev_id < 0 || dev_id >= 64 || !attr_set[dev_id]) {
    ce = cudaFuncSetAttribute(conv_gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    ce = cudaFuncSetAttribute(conv_gemm_tc_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    if (dev_id >= 0 && dev_id < 64) attr_set[dev_id] = true;
  }
  const int slots = plan.pair ? num_sms / 2 : num_sms;          // schedule slots: CTAs or CTA pairs
  // sized by the capacity geometry: the grid is baked into captured graphs, the tile count is not (a CTA without a
  // tile at the current geometry sets up, finds its first schedule slot past plan_tiles and leaves)
  const int used = plan.cap_tiles < slots ? plan.cap_tiles : slots;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(plan.pair ? 2 * used : used);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (plan.pair) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  const GemmProblemDev* dev = plan.dev;
  if (plan.pair)
    return static_cast<int>(cudaLaunchKernelEx(&cfg, conv_gemm_tc_pair_kernel, dev, plan.nprob));
  return static_cast<int>(cudaLaunchKernelEx(&cfg, conv_gemm_tc_kernel, dev, plan.nprob));
}

}  // namespace ftb
