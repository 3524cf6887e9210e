// K8u: the split-precision attention of ftb_attn.cu on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), head width 128
// (FastPitch pre / postnet: d_model 256, 2 heads; common_layers.py:173-175).
//
// Same arithmetic contract as mha_tc_kernel: S = Qh Kh^T + Qh Kl^T + Ql Kh^T and O = Ph Vh + Ph Vl + Pl Vh on fp16 hi / lo
// operand pairs with fp32 accumulation, fp32 softmax.  What changes is where it runs.  mha_tc_kernel sits at 67 % of the
// legacy HMMA pipe (profiles/r2d_ncu_summaries.txt): 0.73 ms per FastPitch-postnet layer, 44 % of generate().  Here
//
//   warp 0      : TMA producer   Q (hi, lo) once; per key tile of 64 keys K (hi, lo) [+ V (hi, lo)] into a 2-stage ring,
//                                straight from the in_proj GEMM's split output [rows, 9e] (128B-swizzled boxes)
//   warp 1      : MMA issuer     S[128 q x 64 keys] into one of two TMEM buffers (24 x tcgen05.mma M128 N64 K16, all K-major);
//                                O[128 q x 128 d] += P V (12 x M128 N128 K16; A = the P tile the softmax warps wrote to shared
//                                memory, B = the V tile as loaded: keys x d with d contiguous = an MN-major operand)
//   warps 2..9  : softmax        thread = (query row = TMEM lane, 32 of the tile's 64 keys): row shift / row sum need no shuffles
//
// TWO PASSES over the keys instead of an online softmax: pass 1 keeps a row-wise shift m, pass 2 computes S, P =
// exp2(S - m), O += P V.  The accumulator O then never has to be rescaled -- in TMEM that would be a load / multiply / store
// round trip of 128 x 128 fp32 per key tile.  softmax is invariant under the shift, so m need not be the exact maximum:
// pass 1 runs the Qh Kh^T product alone (8 MMAs and the K hi boxes per tile; its maximum is within ~1e-2 of the true one,
// which only has to keep exp2 in range), pass 2 the three products: 8 + 24 + 12 MMAs per key tile against the 384 HMMAs per
// warp of the old kernel.  Two softmax warpgroups split the 64 keys of a tile (both read the same TMEM lanes); their
// partial row sums and shifts meet in shared memory.
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int A5_QT = 128;                   // queries per CTA (UMMA M)
constexpr int A5_KT = 64;                    // keys per tile
constexpr int A5_D = 128;                    // head width
constexpr int A5_QBOX = A5_QT * 128;         // one [128 rows x 64 fp16] box: 16 KB
constexpr int A5_KBOX = A5_KT * 128;         // one [64 rows x 64 fp16] box: 8 KB
constexpr int A5_Q_BYTES = 4 * A5_QBOX;      // Q hi (d 0-63, d 64-127), Q lo (same)
constexpr int A5_STAGE = 4 * A5_KBOX;        // one ring slot: hi (2 boxes) + lo (2 boxes) of a K tile, or of a V tile: 32 KB
constexpr int A5_P_BYTES = 2 * A5_QBOX;      // P hi, P lo: [128 q x 64 keys] each
constexpr int A5_THREADS = 352;            // K producer, MMA issuer, 2 softmax warpgroups of 4 warps, V producer

struct A5Smem {
  uint64_t q_full;
  uint64_t k_full[2], k_empty[2];            // K ring: a slot is free once the S MMAs that read it have retired
  uint64_t v_full[2], v_empty[2];            // V ring: ... once the P V MMAs have (a tile later), so K runs ahead of V
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full, p_free;
  uint64_t o_full;
  uint32_t tmem_base;
  uint32_t pad;
  float xch[2][A5_QT];                       // row shift (pass 1) / row sum (pass 2) of each softmax warpgroup
};
constexpr int A5_SMEM_BYTES = A5_Q_BYTES + 4 * A5_STAGE + A5_P_BYTES + static_cast<int>(sizeof(A5Smem)) + 1024;
static_assert(A5_SMEM_BYTES <= 232448, "over the 227 KB shared-memory limit");

// MN-major shared-memory descriptor, 128B swizzle (cute::UMMA canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte
// units): 64 elements of the N extent are one 128-byte row, the next K index is the next row, 8 rows make a 1024-byte
// group (SBO), the next 64 elements of N start LBO bytes on.
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t byte_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((byte_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__device__ __forceinline__ void a5_split2(float a, float b, uint32_t* hi, uint32_t* lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
  *hi = *reinterpret_cast<const uint32_t*>(&h);
  *lo = *reinterpret_cast<const uint32_t*>(&l);
}

__global__ void __launch_bounds__(A5_THREADS, 1)
mha_tc5_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_kv, int e,
               const int64_t* __restrict__ key_tok, int seq_len, int seq_stride, float* __restrict__ out, int round) {
  extern __shared__ uint8_t a5_smem_raw[];
  uint8_t* smem = a5_smem_raw + ((1024u - (smem_u32(a5_smem_raw) & 1023u)) & 1023u);
  uint8_t* q_s = smem;                                   // [hi0 | hi1 | lo0 | lo1]
  uint8_t* k_s = smem + A5_Q_BYTES;                      // 2 slots x [Kh0 Kh1 Kl0 Kl1]
  uint8_t* v_s = k_s + 2 * A5_STAGE;                     // 2 slots x [Vh0 Vh1 Vl0 Vl1]
  uint8_t* p_s = v_s + 2 * A5_STAGE;                     // [P hi | P lo]
  A5Smem* sm = reinterpret_cast<A5Smem*>(p_s + A5_P_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * A5_QT, head = blockIdx.y, b = blockIdx.z;
  const int row0 = b * seq_stride;                       // first row of this sentence in the padded-flat buffer
  const int n_tiles = (seq_len + A5_KT - 1) / A5_KT;
  const int col_q = head * A5_D, col_k = e + head * A5_D, col_v = 2 * e + head * A5_D, lo_off = 3 * e;

  if (warp == 0 && lane == 0) {
    mbar_init(&sm->q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm->k_full[i], 1); mbar_init(&sm->k_empty[i], 1);
      mbar_init(&sm->v_full[i], 1); mbar_init(&sm->v_empty[i], 1);
      mbar_init(&sm->s_full[i], 1); mbar_init(&sm->s_empty[i], 8);
    }
    mbar_init(&sm->p_full, 8);
    mbar_init(&sm->p_free, 1);
    mbar_init(&sm->o_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(&sm->tmem_base, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      mbar_expect_tx(&sm->q_full, A5_Q_BYTES);
      for (int half = 0; half < 2; ++half)
        for (int box = 0; box < 2; ++box)
          tma_load_2d(q_s + (half * 2 + box) * A5_QBOX, &tm_q, &sm->q_full, col_q + half * lo_off + box * 64, row0 + q0);
      uint32_t it = 0;
      for (int pass = 0; pass < 2; ++pass) {
        for (int j = 0; j < n_tiles; ++j, ++it) {
          const int slot = it & 1;
          mbar_wait(&sm->k_empty[slot], ((it >> 1) & 1u) ^ 1u);
          uint8_t* st = k_s + slot * A5_STAGE;
          mbar_expect_tx(&sm->k_full[slot], (pass ? 4 : 2) * A5_KBOX);
          for (int half = 0; half < (pass ? 2 : 1); ++half)                  // pass 1: the hi boxes of K alone
            for (int box = 0; box < 2; ++box)
              tma_load_2d(st + (half * 2 + box) * A5_KBOX, &tm_kv, &sm->k_full[slot], col_k + half * lo_off + box * 64,
                          row0 + j * A5_KT);
        }
      }
    }
  } else if (warp == 10) {
    // ================================ TMA producer of V (pass 2) ===================
    if (lane == 0) {
      for (int j = 0; j < n_tiles; ++j) {
        const int slot = j & 1;
        mbar_wait(&sm->v_empty[slot], ((j >> 1) & 1u) ^ 1u);
        uint8_t* st = v_s + slot * A5_STAGE;
        mbar_expect_tx(&sm->v_full[slot], 4 * A5_KBOX);
        for (int half = 0; half < 2; ++half)
          for (int box = 0; box < 2; ++box)
            tma_load_2d(st + (half * 2 + box) * A5_KBOX, &tm_kv, &sm->v_full[slot], col_v + half * lo_off + box * 64,
                        row0 + j * A5_KT);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_f16(A5_QT, A5_KT);
      const uint32_t idesc_o = umma_idesc_f16(A5_QT, A5_D) | (1u << 16);        // B (the V tile) is MN-major
      const uint32_t q_base = smem_u32(q_s), p_base = smem_u32(p_s);
      // S[sb] = Qh Kh^T + Qh Kl^T + Ql Kh^T over d = 2 boxes x 4 k-steps of 16
      auto issue_s = [&](int sb, int stage, int nprod) {
        const uint32_t k_base = smem_u32(k_s + stage * A5_STAGE);
        const uint32_t d_tmem = tmem_base + sb * A5_KT;
        uint32_t acc = 0;
        for (int prod = 0; prod < nprod; ++prod) {
          const uint32_t qa = q_base + (prod == 2 ? 2 * A5_QBOX : 0);           // Q lo for the third product
          const uint32_t kb = k_base + (prod == 1 ? 2 * A5_KBOX : 0);           // K lo for the second
#pragma unroll
          for (int box = 0; box < 2; ++box) {
            const uint64_t da = umma_desc_sw128(qa + box * A5_QBOX);
            const uint64_t db = umma_desc_sw128(kb + box * A5_KBOX);
#pragma unroll
            for (int k = 0; k < 4; ++k) { umma_f16_ss(d_tmem, da + 2u * k, db + 2u * k, idesc_s, acc); acc = 1; }
          }
        }
      };
      // O += Ph Vh + Ph Vl + Pl Vh over the 64 keys of the tile = 4 k-steps of 16
      auto issue_pv = [&](int stage, bool first) {
        const uint32_t v_base = smem_u32(v_s + stage * A5_STAGE);
        const uint32_t d_tmem = tmem_base + 2 * A5_KT;
        uint32_t acc = first ? 0u : 1u;
#pragma unroll
        for (int prod = 0; prod < 3; ++prod) {
          const uint32_t pa = p_base + (prod == 2 ? A5_QBOX : 0);               // P lo for the third product
          const uint32_t vb = v_base + (prod == 1 ? 2 * A5_KBOX : 0);           // V lo for the second
          const uint64_t da = umma_desc_sw128(pa);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // 16 keys = two 8-row groups of 1024 bytes; the second 64 columns of d are the next box (LBO = 8 KB)
            const uint64_t db = umma_desc_sw128_mn(vb + k * 2048, A5_KBOX, 1024);
            umma_f16_ss(d_tmem, da + 2u * k, db, idesc_o, acc);
            acc = 1;
          }
        }
      };
      mbar_wait(&sm->q_full, 0);
      uint32_t it = 0, sit = 0;
      // ---- pass 1: Qh Kh^T only (the softmax warps keep a row shift)
      for (int j = 0; j < n_tiles; ++j, ++it, ++sit) {
        const int slot = it & 1, sb = sit & 1;
        mbar_wait(&sm->k_full[slot], (it >> 1) & 1u);
        mbar_wait(&sm->s_empty[sb], ((sit >> 1) & 1u) ^ 1u);
        tc_fence_after();
        issue_s(sb, slot, 1);
        umma_commit(&sm->k_empty[slot]);
        umma_commit(&sm->s_full[sb]);
      }
      // ---- pass 2: S(j) first, then P V of the previous tile (the softmax of tile j runs meanwhile)
      for (int j = 0; j <= n_tiles; ++j) {
        if (j < n_tiles) {
          const int slot = it & 1, sb = sit & 1;
          mbar_wait(&sm->k_full[slot], (it >> 1) & 1u);
          mbar_wait(&sm->s_empty[sb], ((sit >> 1) & 1u) ^ 1u);
          tc_fence_after();
          issue_s(sb, slot, 3);
          umma_commit(&sm->k_empty[slot]);
          umma_commit(&sm->s_full[sb]);
          ++it; ++sit;
        }
        if (j >= 1) {
          const uint32_t pj = static_cast<uint32_t>(j - 1);
          const int vslot = pj & 1;
          mbar_wait(&sm->v_full[vslot], (pj >> 1) & 1u);
          mbar_wait(&sm->p_full, pj & 1u);
          tc_fence_after();
          issue_pv(vslot, pj == 0);
          umma_commit(&sm->p_free);
          umma_commit(&sm->v_empty[vslot]);
        }
      }
      umma_commit(&sm->o_full);
    }
  } else {
    // ================================ softmax =====================================
    // warps 2..5 = warpgroup 0 (keys 0-31 of every tile), warps 6..9 = warpgroup 1 (keys 32-63); warp w of either group reads
    // TMEM lane quarter w % 4.  A thread owns one query row and 32 keys per tile.
    const int w = warp & 3;                                 // TMEM lane quarter of this warp
    const int wg = (warp - 2) >> 2;                         // softmax warpgroup = key half
    const int row = w * 32 + lane;                          // query row of the tile = TMEM lane
    const uint32_t tlane = tmem_base + (static_cast<uint32_t>(w * 32) << 16) + wg * 32;
    const float scale_log2 = rsqrtf(static_cast<float>(A5_D)) * 1.4426950408889634f;
    const int64_t* tok = key_tok ? key_tok + static_cast<long long>(b) * seq_len : nullptr;
    auto key_mask = [&](int j) {                            // bit i: key j*64 + 32*wg + i takes part
      const int k = j * A5_KT + wg * 32 + lane;
      bool ok = k < seq_len;
      if (tok && ok) ok = tok[k] != 0;
      return __ballot_sync(0xffffffffu, ok);
    };
    uint32_t sit = 0;
    float mx = -INFINITY;
    // ---- pass 1: row shift (the maximum of the hi x hi product)
    for (int j = 0; j < n_tiles; ++j, ++sit) {
      const int sb = sit & 1;
      const unsigned mk = key_mask(j);
      mbar_wait(&sm->s_full[sb], (sit >> 1) & 1u);
      tc_fence_after();
      float v[32];
      tmem_ld_32x32(tlane + sb * A5_KT, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->s_empty[sb]);
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if ((mk >> i) & 1u) mx = fmaxf(mx, v[i]);
    }
    sm->xch[wg][row] = mx;
    named_bar_sync(1, 256);                                 // the eight softmax warps
    mx = fmaxf(mx, sm->xch[wg ^ 1][row]);
    const float mu = (mx == -INFINITY) ? 0.0f : mx * scale_log2;
    named_bar_sync(1, 256);                                 // everyone has read the shifts: xch is reused for the sums
    // ---- pass 2: P = exp2(S * scale - shift) as hi / lo fp16 into the swizzled K-major P tile
    float l_sum = 0.0f;
    uint8_t* p_hi = p_s;
    uint8_t* p_lo = p_s + A5_QBOX;
    const uint32_t prow = static_cast<uint32_t>((row >> 3) * 1024 + (row & 7) * 128);
    for (int j = 0; j < n_tiles; ++j, ++sit) {
      const int sb = sit & 1;
      const unsigned mk = key_mask(j);
      mbar_wait(&sm->s_full[sb], (sit >> 1) & 1u);
      tc_fence_after();
      float v[32];
      tmem_ld_32x32(tlane + sb * A5_KT, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->s_empty[sb]);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float p = ((mk >> i) & 1u) ? exp2f(fmaf(v[i], scale_log2, -mu)) : 0.0f;
        v[i] = p;
        l_sum += p;
      }
      if (j >= 1) mbar_wait(&sm->p_free, static_cast<uint32_t>(j - 1) & 1u);      // P V of the previous tile has retired
#pragma unroll
      for (int c = 0; c < 4; ++c) {                          // 16-byte chunk 4*wg + c = keys 32*wg + 8c .. + 7 of this row
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a5_split2(v[8 * c + 2 * i], v[8 * c + 2 * i + 1], &h[i], &l[i]);
        const uint32_t off = prow + ((static_cast<uint32_t>(4 * wg + c) ^ static_cast<uint32_t>(row & 7)) << 4);
        *reinterpret_cast<uint4*>(p_hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(p_lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->p_full);
    }
    sm->xch[wg][row] = l_sum;
    named_bar_sync(1, 256);
    l_sum += sm->xch[wg ^ 1][row];
    // ---- O / l -> out: warpgroup wg stores d columns [64 wg, 64 wg + 64)
    mbar_wait(&sm->o_full, 0);
    tc_fence_after();
    const int t = q0 + row;
    const float inv = 1.0f / l_sum;
    const long long orow = static_cast<long long>(row0) + t;
    const uint32_t olane = tmem_base + (static_cast<uint32_t>(w * 32) << 16) + 2 * A5_KT;
#pragma unroll
    for (int c = 2 * wg; c < 2 * wg + 2; ++c) {
      float o[32];
      tmem_ld_32x32(olane + c * 32, o);
      tmem_ld_wait();
      if (t < seq_len) {
        const int col = head * A5_D + c * 32;
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] *= inv;
        if (round == 3) {
          __half* dst = reinterpret_cast<__half*>(out) + orow * 3 * e + col;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint32_t h[4], l[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) a5_split2(o[i + 2 * u], o[i + 2 * u + 1], &h[u], &l[u]);
            const uint4 ph = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(dst + i) = ph;
            *reinterpret_cast<uint4*>(dst + e + i) = make_uint4(l[0], l[1], l[2], l[3]);
            *reinterpret_cast<uint4*>(dst + 2 * e + i) = ph;
          }
        } else if (round == 2) {
          __half* dst = reinterpret_cast<__half*>(out) + orow * e + col;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint32_t h[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const __half2 hh = __floats2half2_rn(o[i + 2 * u], o[i + 2 * u + 1]);
              h[u] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            *reinterpret_cast<uint4*>(dst + i) = make_uint4(h[0], h[1], h[2], h[3]);
          }
        } else {
          float* dst = out + orow * e + col;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            float4 y = make_float4(o[i], o[i + 1], o[i + 2], o[i + 3]);
            if (round) { y.x = round_tf32(y.x); y.y = round_tf32(y.y); y.z = round_tf32(y.z); y.w = round_tf32(y.w); }
            *reinterpret_cast<float4*>(dst + i) = y;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// Launch: one CTA per (128 queries, head, sentence).  The two tensor maps describe the split in_proj output [rows, 9e] with
// 128-row (Q) and 64-row (K, V) boxes of 64 columns; they travel as kernel parameters.
int mha_tc5_launch(const void* qkv16, int e, int heads, const int64_t* key_tok, int batch, int seq_len, int seq_stride,
                   float* out, int round, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev_id = 0;
  cudaError_t ce = cudaGetDevice(&dev_id);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  if (dev_id < 0 || dev_id >= 64 || !attr_set[dev_id]) {
    ce = cudaFuncSetAttribute(mha_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, A5_SMEM_BYTES);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    if (dev_id >= 0 && dev_id < 64) attr_set[dev_id] = true;
  }
  const long long rows = static_cast<long long>(batch) * seq_stride;
  CUtensorMap tm_q, tm_kv;
  int rc = tensor_map_2d(&tm_q, qkv16, rows, 9 * e, 9 * e, A5_QT, true);
  if (rc) return rc;
  rc = tensor_map_2d(&tm_kv, qkv16, rows, 9 * e, 9 * e, A5_KT, true);
  if (rc) return rc;
  dim3 grid((seq_len + A5_QT - 1) / A5_QT, heads, batch);
  mha_tc5_kernel<<<grid, A5_THREADS, A5_SMEM_BYTES, st>>>(tm_q, tm_kv, e, key_tok, seq_len, seq_stride, out, round);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace ftb
