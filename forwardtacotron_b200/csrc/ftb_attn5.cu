// K8u: the split-precision attention of ftb_attn.cu on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), head width 128
// (FastPitch pre / postnet: d_model 256, 2 heads; common_layers.py:173-175).
//
// Same arithmetic contract as mha_tc_kernel: S = Qh Kh^T + Qh Kl^T + Ql Kh^T and O = Ph Vh + Ph Vl + Pl Vh on fp16 hi / lo
// operand pairs with fp32 accumulation, fp32 softmax.  What changes is where it runs.  mha_tc_kernel sits at 67 % of the
// legacy HMMA pipe (profiles/r2d_ncu_summaries.txt): 0.73 ms per FastPitch-postnet layer, 44 % of generate().  Here
//
//   warps 0, 10 : TMA producers  per key tile of 64 keys: K (hi, lo) and V (hi, lo) into two 3-slot rings of 32 KB, straight
//                                from the in_proj GEMM's split output [rows, 9e] (128B-swizzled boxes).  A K slot is free once
//                                the S MMAs that read it have retired, a V slot a tile later: K runs ahead of V.
//   warp 1      : MMA issuer     S[128 q x 64 keys] into one of two TMEM buffers; O[128 q x 128 d] += P V.  Both products take
//                                their A operand FROM TENSOR MEMORY (tcgen05.mma with A in TMEM): Q hi / lo are stored there
//                                once per CTA, P hi / lo by the softmax warps per tile -- so an MMA reads only its B operand
//                                (a K box, K-major; a V box as loaded: keys x d with d contiguous = MN-major) from shared
//                                memory.  With both operands in shared memory an M128 N64 K16 MMA is bound by the 6 KB it
//                                reads (48 cycles at 128 B/clk): 0.39 ms per layer; with A in TMEM 16.
//   warps 2..9  : softmax        thread = (query row = TMEM lane, 32 of the tile's 64 keys): row shift / row sum need no
//                                shuffles; P leaves as packed fp16 pairs through tcgen05.st (no shared-memory tile, no proxy fence)
//
// TWO PASSES over the keys instead of an online softmax: pass 1 keeps a row-wise shift m, pass 2 computes S, P =
// exp2(S - m), O += P V.  The accumulator O then never has to be rescaled -- in TMEM that would be a load / multiply / store
// round trip of 128 x 128 fp32 per key tile.  softmax is invariant under the shift, so m need not be the exact maximum:
// pass 1 runs the Qh Kh^T product alone (8 MMAs and the K hi boxes per tile; its maximum is within ~1e-2 of the true one,
// which only has to keep exp2 in range), pass 2 the three products: 8 + 24 + 12 MMAs per key tile against the 384 HMMAs per
// warp of the old kernel.  Two softmax warpgroups split the 64 keys of a tile (both read the same TMEM lanes); their
// partial row sums and shifts meet in shared memory.
//
// TMEM columns: S0 [0,64) S1 [64,128) O [128,256) Q hi [256,320) Q lo [320,384) P0 [384,448) P1 [448,512) (hi | lo, 32 each).
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int A5_QT = 128;                   // queries per CTA (UMMA M)
constexpr int A5_KT = 128;                   // keys per S tile (UMMA N of the S product: an M128 x K16 MMA costs ~55 cycles to
                                             // issue whatever its N, so N = 128 covers twice the keys of N = 64 per instruction)
constexpr int A5_VT = 64;                    // keys per V slot: the P V product of a tile runs as two halves
constexpr int A5_D = 128;                    // head width
constexpr int A5_KBOX = A5_KT * 128;         // one [128 keys x 64 fp16] box of K: 16 KB
constexpr int A5_VBOX = A5_VT * 128;         // one [64 keys x 64 fp16] box of V: 8 KB
constexpr int A5_KSLOT = 4 * A5_KBOX;        // K hi (2 boxes) + K lo (2 boxes): 64 KB
constexpr int A5_VSLOT = 4 * A5_VBOX;        // V hi (2 boxes) + V lo (2 boxes): 32 KB
constexpr int A5_NK = 2, A5_NV = 3;          // ring slots
constexpr int A5_THREADS = 384;              // K producer, S issuer, 2 softmax warpgroups of 4 warps, V producer, P V issuer
constexpr uint32_t A5_COL_O = 128, A5_COL_Q = 256, A5_COL_P = 384;

struct A5Smem {
  uint64_t q_full;
  uint64_t k_full[A5_NK], k_empty[A5_NK];
  uint64_t v_full[A5_NV], v_empty[A5_NV];
  uint64_t s_full, s_empty;
  uint64_t p_full, p_empty;
  uint64_t o_full;
  uint32_t tmem_base;
  uint32_t pad;
  float xch[2][A5_QT];                       // row shift (pass 1) / row sum (pass 2) of each softmax warpgroup
};
constexpr int A5_SMEM_BYTES = A5_NK * A5_KSLOT + A5_NV * A5_VSLOT + static_cast<int>(sizeof(A5Smem)) + 1024;
static_assert(A5_SMEM_BYTES <= 232448, "over the 227 KB shared-memory limit");

// registers -> TMEM, 16 consecutive 32-bit columns of this warp's lanes
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

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

// tcgen05.mma with A in TMEM, the B descriptor passed as two 32-bit halves: the issuing thread is the bottleneck of this
// kernel (measured: ~50 cycles per MMA with descriptors rebuilt per instruction), so the per-instruction work is one
// 32-bit add on the address half.
__device__ __forceinline__ void a5_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 bd;\n\t"
      "mov.b64 bd, {%2, %3};\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float a5_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void a5_split2(float a, float b, uint32_t* hi, uint32_t* lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
  *hi = *reinterpret_cast<const uint32_t*>(&h);
  *lo = *reinterpret_cast<const uint32_t*>(&l);
}

#ifdef FTB_A5_PROFILE
#define A5_T0 a5_t0 = clock64();
#define A5_ACC(i) a5_p[i] += clock64() - a5_t0;
#else
#define A5_T0
#define A5_ACC(i)
#endif

__global__ void __launch_bounds__(A5_THREADS, 1)
mha_tc5_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
               const __half* __restrict__ qkv16, long long total_rows, int e, const int64_t* __restrict__ key_tok,
               int seq_len, int seq_stride, float* __restrict__ out, int round) {
  extern __shared__ uint8_t a5_smem_raw[];
  uint8_t* smem = a5_smem_raw + ((1024u - (smem_u32(a5_smem_raw) & 1023u)) & 1023u);
  uint8_t* k_s = smem;                                   // A5_NK slots x [Kh0 Kh1 Kl0 Kl1], 128 keys each
  uint8_t* v_s = k_s + A5_NK * A5_KSLOT;                 // A5_NV slots x [Vh0 Vh1 Vl0 Vl1], 64 keys each
  A5Smem* sm = reinterpret_cast<A5Smem*>(v_s + A5_NV * A5_VSLOT);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * A5_QT, head = blockIdx.y, b = blockIdx.z;
  const int row0 = b * seq_stride;                       // first row of this sentence in the padded-flat buffer
  const int n_tiles = (seq_len + A5_KT - 1) / A5_KT;     // S tiles of 128 keys
  const int n_vt = (seq_len + A5_VT - 1) / A5_VT;        // V half tiles of 64 keys (the last S tile may have one only)
  const int col_q = head * A5_D, col_k = e + head * A5_D, col_v = 2 * e + head * A5_D, lo_off = 3 * e;
#ifdef FTB_A5_PROFILE
  long long a5_p[8] = {0, 0, 0, 0, 0, 0, 0, 0}, a5_t0 = 0;
  const long long a5_start = clock64();
#endif

  if (warp == 0 && lane == 0) {
    mbar_init(&sm->q_full, 8);
    for (int i = 0; i < A5_NK; ++i) { mbar_init(&sm->k_full[i], 1); mbar_init(&sm->k_empty[i], 1); }
    for (int i = 0; i < A5_NV; ++i) { mbar_init(&sm->v_full[i], 1); mbar_init(&sm->v_empty[i], 1); }
    mbar_init(&sm->s_full, 1); mbar_init(&sm->s_empty, 8);
    mbar_init(&sm->p_full, 8); mbar_init(&sm->p_empty, 1);
    mbar_init(&sm->o_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) { tmem_alloc(&sm->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm->tmem_base;

  if (warp == 0) {
    // ================================ TMA producer of K ===========================
    if (lane == 0) {
      uint32_t it = 0;
      for (int pass = 0; pass < 2; ++pass) {
        for (int j = 0; j < n_tiles; ++j, ++it) {
          const uint32_t slot = it % A5_NK;
          mbar_wait(&sm->k_empty[slot], ((it / A5_NK) & 1u) ^ 1u);
          uint8_t* st = k_s + slot * A5_KSLOT;
          mbar_expect_tx(&sm->k_full[slot], (pass ? 4 : 2) * A5_KBOX);
          for (int half = 0; half < (pass ? 2 : 1); ++half)                  // pass 1: the hi boxes of K alone
            for (int box = 0; box < 2; ++box)
              tma_load_2d(st + (half * 2 + box) * A5_KBOX, &tm_k, &sm->k_full[slot], col_k + half * lo_off + box * 64,
                          row0 + j * A5_KT);
        }
      }
    }
  } else if (warp == 10) {
    // ================================ TMA producer of V (pass 2) ===================
    if (lane == 0) {
      for (uint32_t j = 0; j < static_cast<uint32_t>(n_vt); ++j) {
        const uint32_t slot = j % A5_NV;
        mbar_wait(&sm->v_empty[slot], ((j / A5_NV) & 1u) ^ 1u);
        uint8_t* st = v_s + slot * A5_VSLOT;
        mbar_expect_tx(&sm->v_full[slot], 4 * A5_VBOX);
        for (int half = 0; half < 2; ++half)
          for (int box = 0; box < 2; ++box)
            tma_load_2d(st + (half * 2 + box) * A5_VBOX, &tm_v, &sm->v_full[slot], col_v + half * lo_off + box * 64,
                        row0 + static_cast<int>(j) * A5_VT);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer: S ===============================
    // Two issuing threads (this one for S, warp 11 for P V): their MMAs go to different accumulators, and one thread issuing
    // all MMAs of a tile was the bound of the kernel.
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_f16(A5_QT, A5_KT);
      const uint64_t kd0 = umma_desc_sw128(smem_u32(k_s));
      const uint32_t kd_lo = static_cast<uint32_t>(kd0), kd_hi = static_cast<uint32_t>(kd0 >> 32);
      mbar_wait(&sm->q_full, 0);
      tc_fence_after();
      uint32_t it = 0;
      for (int pass = 0; pass < 2; ++pass) {
        // pass 1: Qh Kh^T only (the softmax warps keep a row shift); pass 2: Qh Kh^T + Qh Kl^T + Ql Kh^T
        for (int j = 0; j < n_tiles; ++j, ++it) {
          const uint32_t slot = it % A5_NK;
          A5_T0 mbar_wait(&sm->k_full[slot], (it / A5_NK) & 1u); A5_ACC(0)
          A5_T0 mbar_wait(&sm->s_empty, (it & 1u) ^ 1u); A5_ACC(1)
          tc_fence_after();
          A5_T0
          const uint32_t k_lo = kd_lo + ((slot * A5_KSLOT) >> 4);
          uint32_t acc = 0;
          for (int prod = 0; prod < (pass ? 3 : 1); ++prod) {
            const uint32_t qa = tmem_base + A5_COL_Q + (prod == 2 ? 64u : 0u);  // Q lo for the third product
            const uint32_t kb = k_lo + (prod == 1 ? (2 * A5_KBOX) >> 4 : 0);    // K lo for the second
#pragma unroll
            for (int i = 0; i < 8; ++i) {                                       // d box i / 4, k-step i % 4 (32 bytes each)
              a5_mma_ts(tmem_base, qa + i * 8u, kb + (i >> 2) * (A5_KBOX >> 4) + (i & 3) * 2, kd_hi, idesc_s, acc);
              acc = 1;
            }
          }
          A5_ACC(2)
          umma_commit(&sm->k_empty[slot]);
          umma_commit(&sm->s_full);
        }
      }
#ifdef FTB_A5_PROFILE
      if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
        printf("a5 S issuer: total %lld | wait k_full %lld s_empty %lld issue %lld (tiles %d)\n", clock64() - a5_start,
               a5_p[0], a5_p[1], a5_p[2], n_tiles);
#endif
    }
  } else if (warp == 11) {
    // ================================ MMA issuer: P V =============================
    if (lane == 0) {
#ifdef FTB_A5_EXP_KMAJOR_V       // timing experiment only (wrong results): V read as if it were K-major
      const uint32_t idesc_o = umma_idesc_f16(A5_QT, A5_D);
      const uint64_t vd0 = umma_desc_sw128(smem_u32(v_s));
#else
      const uint32_t idesc_o = umma_idesc_f16(A5_QT, A5_D) | (1u << 16);        // B (the V tile) is MN-major
      // 16 keys = two 8-row groups of 1024 bytes; the second 64 columns of d are the next box (LBO = 8 KB)
      const uint64_t vd0 = umma_desc_sw128_mn(smem_u32(v_s), A5_VBOX, 1024);
#endif
      const uint32_t vd_lo = static_cast<uint32_t>(vd0), vd_hi = static_cast<uint32_t>(vd0 >> 32);
      const uint32_t d_tmem = tmem_base + A5_COL_O;
      uint32_t vj = 0;
      for (uint32_t j = 0; j < static_cast<uint32_t>(n_tiles); ++j) {
        A5_T0 mbar_wait(&sm->p_full, j & 1u); A5_ACC(1)
        for (uint32_t h = 0; h < 2 && vj < static_cast<uint32_t>(n_vt); ++h, ++vj) {
          const uint32_t vslot = vj % A5_NV;
          A5_T0 mbar_wait(&sm->v_full[vslot], (vj / A5_NV) & 1u); A5_ACC(0)
          tc_fence_after();
          A5_T0
          // O += Ph Vh + Ph Vl + Pl Vh over the 64 keys of this half = 4 k-steps of 16; A = P from TMEM
          const uint32_t v_lo = vd_lo + ((vslot * A5_VSLOT) >> 4);
          uint32_t acc = vj ? 1u : 0u;
#pragma unroll
          for (int prod = 0; prod < 3; ++prod) {
            const uint32_t pa = tmem_base + A5_COL_P + h * 32u + (prod == 2 ? 64u : 0u);    // P lo for the third product
            const uint32_t vb = v_lo + (prod == 1 ? (2 * A5_VBOX) >> 4 : 0);                // V lo for the second
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              a5_mma_ts(d_tmem, pa + k * 8u, vb + k * (2048 >> 4), vd_hi, idesc_o, acc);
              acc = 1;
            }
          }
          A5_ACC(2)
          umma_commit(&sm->v_empty[vslot]);
        }
        umma_commit(&sm->p_empty);
      }
      umma_commit(&sm->o_full);
#ifdef FTB_A5_PROFILE
      if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
        printf("a5 PV issuer: total %lld | wait v_full %lld p_full %lld issue %lld\n", clock64() - a5_start, a5_p[0], a5_p[1],
               a5_p[2]);
#endif
    }
  } else {
    // ================================ softmax =====================================
    // warps 2..5 = warpgroup 0 (keys 0-63 of every tile), warps 6..9 = warpgroup 1 (keys 64-127); warp w of either group reads
    // TMEM lane quarter w % 4.  A thread owns one query row and 64 keys per tile.
    const int w = warp & 3;                                 // TMEM lane quarter of this warp
    const int wg = (warp - 2) >> 2;                         // softmax warpgroup = key half = P V half
    const int row = w * 32 + lane;                          // query row of the tile = TMEM lane
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(w * 32) << 16);
    const uint32_t tlane = lane_base + wg * 64;             // this thread's 64 columns of S
    const float scale_log2 = rsqrtf(static_cast<float>(A5_D)) * 1.4426950408889634f;
    const int64_t* tok = key_tok ? key_tok + static_cast<long long>(b) * seq_len : nullptr;
    auto key_mask = [&](int j, int c) {                     // bit i: key j*128 + 64*wg + 32*c + i takes part
      const int k = j * A5_KT + wg * 64 + c * 32 + lane;
      bool ok = k < seq_len;
      if (tok && ok) ok = tok[k] != 0;
      return __ballot_sync(0xffffffffu, ok);
    };
    // ---- Q -> TMEM (once): warpgroup 0 the hi halves, warpgroup 1 the lo halves of this thread's query row; column c of the
    //      operand holds d = (2c, 2c+1).  Rows behind the end of the buffer read as zero (they are never stored either).
    {
      const long long qrow = static_cast<long long>(row0) + q0 + row;
      const uint4* src = reinterpret_cast<const uint4*>(qkv16 + qrow * (9ll * e) + col_q + wg * lo_off);
      const uint32_t qcol = lane_base + A5_COL_Q + wg * 64;
#pragma unroll
      for (int c = 0; c < 4; ++c) {                          // 16 columns = 32 halves = four 16-byte loads
        uint32_t r[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 x = make_uint4(0u, 0u, 0u, 0u);
          if (qrow < total_rows) x = __ldg(src + c * 4 + i);
          r[4 * i] = x.x; r[4 * i + 1] = x.y; r[4 * i + 2] = x.z; r[4 * i + 3] = x.w;
        }
        tmem_st_32x16(qcol + c * 16, r);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->q_full);
    }
    uint32_t sit = 0;
    float mx = -INFINITY;
    // ---- pass 1: row shift (the maximum of the hi x hi product)
    for (int j = 0; j < n_tiles; ++j, ++sit) {
      const unsigned m0 = key_mask(j, 0), m1 = key_mask(j, 1);
      A5_T0 mbar_wait(&sm->s_full, sit & 1u); A5_ACC(0)
      tc_fence_after();
      A5_T0
      float v0[32], v1[32];
      tmem_ld_32x32(tlane, v0);
      tmem_ld_32x32(tlane + 32, v1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->s_empty);
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if ((m0 >> i) & 1u) mx = fmaxf(mx, v0[i]);
        if ((m1 >> i) & 1u) mx = fmaxf(mx, v1[i]);
      }
      A5_ACC(1)
    }
    sm->xch[wg][row] = mx;
    named_bar_sync(1, 256);                                 // the eight softmax warps
    mx = fmaxf(mx, sm->xch[wg ^ 1][row]);
    const float mu = (mx == -INFINITY) ? 0.0f : mx * scale_log2;
    named_bar_sync(1, 256);                                 // everyone has read the shifts: xch is reused for the sums
    // ---- pass 2: P = exp2(S * scale - shift) as packed hi / lo fp16 pairs into the TMEM operand buffer of P V
    float l_sum = 0.0f;
    for (int j = 0; j < n_tiles; ++j, ++sit) {
      const unsigned m0 = key_mask(j, 0), m1 = key_mask(j, 1);
      A5_T0 mbar_wait(&sm->s_full, sit & 1u); A5_ACC(2)
      tc_fence_after();
      A5_T0
      float v[2][32];
      tmem_ld_32x32(tlane, v[0]);
      tmem_ld_32x32(tlane + 32, v[1]);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->s_empty);               // S is free again: the next tile's MMAs overlap the exponentials
      uint32_t ph[2][16], pl[2][16];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const unsigned mk = c ? m1 : m0;
        if (mk == 0xffffffffu) {                             // every key of this chunk takes part (all but the last tile)
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float p0 = a5_ex2(fmaf(v[c][2 * i], scale_log2, -mu));
            const float p1 = a5_ex2(fmaf(v[c][2 * i + 1], scale_log2, -mu));
            l_sum += p0 + p1;
            a5_split2(p0, p1, &ph[c][i], &pl[c][i]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float p0 = ((mk >> (2 * i)) & 1u) ? a5_ex2(fmaf(v[c][2 * i], scale_log2, -mu)) : 0.0f;
            const float p1 = ((mk >> (2 * i + 1)) & 1u) ? a5_ex2(fmaf(v[c][2 * i + 1], scale_log2, -mu)) : 0.0f;
            l_sum += p0 + p1;
            a5_split2(p0, p1, &ph[c][i], &pl[c][i]);
          }
        }
      }
      A5_ACC(3)
      A5_T0 mbar_wait(&sm->p_empty, (static_cast<uint32_t>(j) & 1u) ^ 1u); A5_ACC(4)      // P V of the previous tile has retired
      tc_fence_after();
      A5_T0
      const uint32_t pcol = lane_base + A5_COL_P + wg * 32;  // keys 64 wg .. + 63 = columns 32 wg .. + 31 of the hi operand
      tmem_st_32x16(pcol, ph[0]);
      tmem_st_32x16(pcol + 16, ph[1]);
      tmem_st_32x16(pcol + 64, pl[0]);
      tmem_st_32x16(pcol + 80, pl[1]);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm->p_full);
      A5_ACC(5)
    }
#ifdef FTB_A5_PROFILE
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x == 64 || threadIdx.x == 192))   // warps 2, 6
      printf("a5 softmax warp %d: total %lld | pass1 wait s_full %lld work %lld | pass2 wait s_full %lld exp+split %lld wait p_empty %lld store %lld\n",
             warp, clock64() - a5_start, a5_p[0], a5_p[1], a5_p[2], a5_p[3], a5_p[4], a5_p[5]);
#endif
    sm->xch[wg][row] = l_sum;
    named_bar_sync(1, 256);
    l_sum += sm->xch[wg ^ 1][row];
    // ---- O / l -> out: warpgroup wg stores d columns [64 wg, 64 wg + 64)
    mbar_wait(&sm->o_full, 0);
    tc_fence_after();
    const int t = q0 + row;
    const float inv = 1.0f / l_sum;
    const long long orow = static_cast<long long>(row0) + t;
    const uint32_t olane = lane_base + A5_COL_O;
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
            const uint4 ph4 = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(dst + i) = ph4;
            *reinterpret_cast<uint4*>(dst + e + i) = make_uint4(l[0], l[1], l[2], l[3]);
            *reinterpret_cast<uint4*>(dst + 2 * e + i) = ph4;
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
    tmem_dealloc(tmem_base, 512);
  }
}

// Launch: one CTA per (128 queries, head, sentence).  The tensor maps describe the split in_proj output [rows, 9e] with
// 128-row (K) and 64-row (V) boxes of 64 columns; they travel as kernel parameters.
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
  CUtensorMap tm_k, tm_v;
  int rc = tensor_map_2d(&tm_k, qkv16, rows, 9 * e, 9 * e, A5_KT, true);
  if (rc) return rc;
  rc = tensor_map_2d(&tm_v, qkv16, rows, 9 * e, 9 * e, A5_VT, true);
  if (rc) return rc;
  dim3 grid((seq_len + A5_QT - 1) / A5_QT, heads, batch);
  mha_tc5_kernel<<<grid, A5_THREADS, A5_SMEM_BYTES, st>>>(tm_k, tm_v, static_cast<const __half*>(qkv16), rows, e, key_tok,
                                                          seq_len, seq_stride, out, round);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace ftb
