// K8t: multi-head self attention on the tensor cores with SPLIT-PRECISION operands (FastPitch / FFT blocks,
// common_layers.py:173-175: nn.MultiheadAttention, softmax over keys, key_padding_mask -> -inf).
//
// The parity budget of FastPitch (mel rel-L2 6.7e-4 of 1e-3) leaves no room for single-pass tf32 / fp16 scores, so both
// contractions run as three fp16 MMAs on hi / lo operand pairs (x = hi + lo, hi = fp16(x), lo = fp16(x - hi)):
//     S = Qh Kh^T + Qh Kl^T + Ql Kh^T          O += Ph Vh + Ph Vl + Pl Vh          (fp32 accumulation, ~2^-20 operands)
// Q, K, V arrive ALREADY split: the in_proj GEMM writes its output with FTB_F_OUT_SPLIT ([hi | lo | hi] column blocks of
// width 3E), so tiles are staged with plain 16-byte cp.async copies.  Flash-attention-2 tiling: a CTA of 4 warps owns 64
// queries of one (sentence, head); each warp keeps 16 query rows, streams key tiles of 64 through shared memory, does
// the online softmax in fp32 on the accumulator fragments and feeds P straight back as the A operand of P V.
// mma.sync.m16n8k16 (legacy tensor-core path): the operands of both products are produced by the same thread in
// registers (P) or change every tile (K, V) and the M tile is 16 rows per warp, which is what this instruction is for;
// a tcgen05 version would need P staged through shared memory per key tile.
#include <cstdlib>
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// 16-byte async copy; `ok` false -> the 16 bytes are zero-filled (src-size 0)
__device__ __forceinline__ void cp16_zfill(uint32_t dst, const void* src, bool ok) {
  const int n = ok ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void split2(float a, float b, uint32_t* hi, uint32_t* lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
  *hi = *reinterpret_cast<const uint32_t*>(&h);
  *lo = *reinterpret_cast<const uint32_t*>(&l);
}

// qkv16: fp16 [rows, 9e] = [hi(3e) | lo(3e) | hi(3e)] of the packed in_proj output (q | k | v, each e wide).
template <int D>
__global__ void __launch_bounds__(128)
mha_tc_kernel(const __half* __restrict__ qkv16, int e, const int64_t* __restrict__ key_tok, int seq_len, int seq_stride,
              float* __restrict__ out, int round) {
  constexpr int QT = 64, KT = 64, LDS = D + 8;            // row stride in halves: 16 bytes of padding, ldmatrix conflict-free
  constexpr int KS = D / 16, DT = D / 8;
  extern __shared__ __align__(16) __half attn_smem[];
  __half* q_hi = attn_smem;
  __half* q_lo = q_hi + QT * LDS;
  __half* k_hi = q_lo + QT * LDS;
  __half* k_lo = k_hi + KT * LDS;
  __half* v_hi = k_lo + KT * LDS;
  __half* v_lo = v_hi + KT * LDS;
  __shared__ int kmask[KT];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * QT, head = blockIdx.y, b = blockIdx.z;
  const long long rowbase = static_cast<long long>(b) * seq_stride;
  const int ld = 9 * e, lo_off = 3 * e;
  const float scale_log2 = rsqrtf(static_cast<float>(D)) * 1.4426950408889634f;     // softmax in base 2

  // ---- stage Q (hi, lo): 64 rows x D halves each; rows past seq_len are zero
  constexpr int CPR = D / 8;                               // 16-byte chunks per row
  for (int i = tid; i < QT * CPR; i += 128) {
    const int r = i / CPR, c = i % CPR;
    const bool ok = q0 + r < seq_len;
    const __half* src = qkv16 + (rowbase + (ok ? q0 + r : 0)) * ld + head * D + c * 8;
    cp16_zfill(smem_u32(q_hi + r * LDS + c * 8), src, ok);
    cp16_zfill(smem_u32(q_lo + r * LDS + c * 8), src + lo_off, ok);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");

  float o_acc[DT][4];
#pragma unroll
  for (int i = 0; i < DT; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o_acc[i][j] = 0.0f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.0f, 0.0f};

  // ldmatrix lane addressing: matrices 0..3 = (rows 0-7, k 0-7), (rows 8-15, k 0-7), (rows 0-7, k 8-15), (rows 8-15, k 8-15)
  const int lm_row = (lane & 7) + ((lane >> 3) & 1) * 8, lm_col = (lane >> 4) * 8;
  const uint32_t qa_hi = smem_u32(q_hi + (warp * 16 + lm_row) * LDS + lm_col);
  const uint32_t qa_lo = smem_u32(q_lo + (warp * 16 + lm_row) * LDS + lm_col);
  // K as the B operand of S = Q K^T (n = key, k = d): x4 = (keys 0-7, d 0-7), (keys 0-7, d 8-15), (keys 8-15, d 0-7), (keys 8-15, d 8-15)
  const int kb_row = (lane & 7) + (lane >> 4) * 8, kb_col = ((lane >> 3) & 1) * 8;
  // V as the B operand of O = P V (k = key, n = d), transposed load: (keys 0-7, d 0-7), (keys 8-15, d 0-7), (keys 0-7, d 8-15), (keys 8-15, d 8-15)
  const int vb_row = (lane & 7) + ((lane >> 3) & 1) * 8, vb_col = (lane >> 4) * 8;

  for (int k0 = 0; k0 < seq_len; k0 += KT) {
    __syncthreads();                                       // everyone is done with the previous key tile
    for (int i = tid; i < KT * CPR; i += 128) {
      const int r = i / CPR, c = i % CPR;
      const bool ok = k0 + r < seq_len;
      const __half* src = qkv16 + (rowbase + (ok ? k0 + r : 0)) * ld + head * D + c * 8;
      cp16_zfill(smem_u32(k_hi + r * LDS + c * 8), src + e, ok);
      cp16_zfill(smem_u32(k_lo + r * LDS + c * 8), src + e + lo_off, ok);
      cp16_zfill(smem_u32(v_hi + r * LDS + c * 8), src + 2 * e, ok);
      cp16_zfill(smem_u32(v_lo + r * LDS + c * 8), src + 2 * e + lo_off, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (tid < KT) {
      const int t = k0 + tid;
      int ok = t < seq_len;
      if (ok && key_tok) ok = key_tok[static_cast<long long>(b) * seq_len + t] != 0;
      kmask[tid] = ok;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- S = Q K^T  (16 x 64 per warp), three MMAs per product
    float s[KT / 8][4];
#pragma unroll
    for (int i = 0; i < KT / 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.0f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t ah[4], al[4];
      ldsm_x4(ah, qa_hi + ks * 32);
      ldsm_x4(al, qa_lo + ks * 32);
#pragma unroll
      for (int np = 0; np < KT / 16; ++np) {
        uint32_t bh[4], bl[4];
        const uint32_t off = ((np * 16 + kb_row) * LDS + ks * 16 + kb_col) * 2;
        ldsm_x4(bh, smem_u32(k_hi) + off);
        ldsm_x4(bl, smem_u32(k_lo) + off);
        mma16816(s[2 * np], ah, bh[0], bh[1]);
        mma16816(s[2 * np + 1], ah, bh[2], bh[3]);
        mma16816(s[2 * np], ah, bl[0], bl[1]);
        mma16816(s[2 * np + 1], ah, bl[2], bl[3]);
        mma16816(s[2 * np], al, bh[0], bh[1]);
        mma16816(s[2 * np + 1], al, bh[2], bh[3]);
      }
    }
    // ---- online softmax (fp32): this thread holds rows g = lane/4 (regs 0,1) and g+8 (regs 2,3), keys nt*8 + 2*(lane%4) + {0,1}
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < KT / 8; ++nt) {
      const int kc = nt * 8 + 2 * (lane & 3);
      const bool ok0 = kmask[kc] != 0, ok1 = kmask[kc + 1] != 0;
      s[nt][0] = ok0 ? s[nt][0] * scale_log2 : -INFINITY;
      s[nt][1] = ok1 ? s[nt][1] * scale_log2 : -INFINITY;
      s[nt][2] = ok0 ? s[nt][2] * scale_log2 : -INFINITY;
      s[nt][3] = ok1 ? s[nt][3] * scale_log2 : -INFINITY;
      mx[0] = fmaxf(mx[0], fmaxf(s[nt][0], s[nt][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[nt][2], s[nt][3]));
    }
    float corr[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      corr[r] = (m_new == -INFINITY) ? 1.0f : exp2f(m_run[r] - m_new);
      m_run[r] = m_new;
    }
    float rs[2] = {0.0f, 0.0f};
    const float mu0 = (m_run[0] == -INFINITY) ? 0.0f : m_run[0], mu1 = (m_run[1] == -INFINITY) ? 0.0f : m_run[1];
#pragma unroll
    for (int nt = 0; nt < KT / 8; ++nt) {
      s[nt][0] = exp2f(s[nt][0] - mu0); s[nt][1] = exp2f(s[nt][1] - mu0);
      s[nt][2] = exp2f(s[nt][2] - mu1); s[nt][3] = exp2f(s[nt][3] - mu1);
      rs[0] += s[nt][0] + s[nt][1];
      rs[1] += s[nt][2] + s[nt][3];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      rs[r] += __shfl_xor_sync(0xffffffffu, rs[r], 1);
      rs[r] += __shfl_xor_sync(0xffffffffu, rs[r], 2);
      l_run[r] = l_run[r] * corr[r] + rs[r];
    }
#pragma unroll
    for (int dt = 0; dt < DT; ++dt) {
      o_acc[dt][0] *= corr[0]; o_acc[dt][1] *= corr[0];
      o_acc[dt][2] *= corr[1]; o_acc[dt][3] *= corr[1];
    }
    // ---- O += P V: the S accumulator fragments ARE the A fragments of the next product (m16n8 C layout -> m16k16 A)
#pragma unroll
    for (int kk = 0; kk < KT / 16; ++kk) {
      uint32_t ph[4], pl[4];
      split2(s[2 * kk][0], s[2 * kk][1], &ph[0], &pl[0]);
      split2(s[2 * kk][2], s[2 * kk][3], &ph[1], &pl[1]);
      split2(s[2 * kk + 1][0], s[2 * kk + 1][1], &ph[2], &pl[2]);
      split2(s[2 * kk + 1][2], s[2 * kk + 1][3], &ph[3], &pl[3]);
#pragma unroll
      for (int dp = 0; dp < DT / 2; ++dp) {
        uint32_t vh[4], vl[4];
        const uint32_t off = ((kk * 16 + vb_row) * LDS + dp * 16 + vb_col) * 2;
        ldsm_x4_t(vh, smem_u32(v_hi) + off);
        ldsm_x4_t(vl, smem_u32(v_lo) + off);
        mma16816(o_acc[2 * dp], ph, vh[0], vh[1]);
        mma16816(o_acc[2 * dp + 1], ph, vh[2], vh[3]);
        mma16816(o_acc[2 * dp], ph, vl[0], vl[1]);
        mma16816(o_acc[2 * dp + 1], ph, vl[2], vl[3]);
        mma16816(o_acc[2 * dp], pl, vh[0], vh[1]);
        mma16816(o_acc[2 * dp + 1], pl, vh[2], vh[3]);
      }
    }
  }
  // ---- normalise and store: rows g and g + 8 of this warp's strip, columns dt*8 + 2*(lane%4) + {0,1}
  const int g = lane >> 2, cq = 2 * (lane & 3);
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int t = q0 + warp * 16 + g + 8 * r;
    if (t >= seq_len) continue;
    const float inv = 1.0f / l_run[r];
    const long long row = rowbase + t;
#pragma unroll
    for (int dt = 0; dt < DT; ++dt) {
      float a = o_acc[dt][2 * r] * inv, c2 = o_acc[dt][2 * r + 1] * inv;
      const int col = head * D + dt * 8 + cq;
      if (round == 3) {
        __half* o = reinterpret_cast<__half*>(out) + row * 3 * e + col;
        uint32_t hi, lo;
        split2(a, c2, &hi, &lo);
        *reinterpret_cast<uint32_t*>(o) = hi;
        *reinterpret_cast<uint32_t*>(o + e) = lo;
        *reinterpret_cast<uint32_t*>(o + 2 * e) = hi;
      } else if (round == 2) {
        const __half2 hh = __floats2half2_rn(a, c2);
        *reinterpret_cast<uint32_t*>(reinterpret_cast<__half*>(out) + row * e + col) = *reinterpret_cast<const uint32_t*>(&hh);
      } else {
        if (round) { a = round_tf32(a); c2 = round_tf32(c2); }
        *reinterpret_cast<float2*>(out + row * e + col) = make_float2(a, c2);
      }
    }
  }
}

template <int D>
static constexpr int mha_tc_smem() { return 6 * 64 * (D + 8) * 2; }

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_mha_split(ftb_handle h, const void* qkv16, int e, int heads, const int64_t* key_tok, int batch,
                             int seq_len, int seq_stride, float* out, int round_tf32, void* stream) {
  if (!h || !qkv16 || !out || heads < 1 || batch < 1 || seq_len < 1 || e % heads || (e % 8)) return FTB_E_ARG;
  if ((reinterpret_cast<uintptr_t>(qkv16) | reinterpret_cast<uintptr_t>(out)) & 15u) return FTB_E_ALIGN;
  const int d = e / heads;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((seq_len + 63) / 64, heads, batch);
  cudaError_t ce = cudaSuccess;
  static bool attr64 = false, attr128 = false;
  static const bool tc5 = [] { const char* v = getenv("FTB_ATTN_TC5"); return !(v && v[0] == '0'); }();
  if (d == 64) {
    if (!attr64) {
      ce = cudaFuncSetAttribute(mha_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, mha_tc_smem<64>());
      if (ce != cudaSuccess) return static_cast<int>(ce);
      attr64 = true;
    }
    mha_tc_kernel<64><<<grid, 128, mha_tc_smem<64>(), st>>>(static_cast<const __half*>(qkv16), e, key_tok, seq_len,
                                                            seq_stride, out, round_tf32);
  } else if (d == 128 && tc5) {
    // head width 128 (FastPitch pre / postnet): the tcgen05 kernel (ftb_attn5.cu); FTB_ATTN_TC5=0 keeps mma.sync
    const int rc = mha_tc5_launch(qkv16, e, heads, key_tok, batch, seq_len, seq_stride, out, round_tf32, st);
    if (rc) return rc;
  } else if (d == 128) {
    if (!attr128) {
      ce = cudaFuncSetAttribute(mha_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, mha_tc_smem<128>());
      if (ce != cudaSuccess) return static_cast<int>(ce);
      attr128 = true;
    }
    mha_tc_kernel<128><<<grid, 128, mha_tc_smem<128>(), st>>>(static_cast<const __half*>(qkv16), e, key_tok, seq_len,
                                                              seq_stride, out, round_tf32);
  } else {
    return FTB_E_SHAPE;          // other head widths: ftb_mha (SIMT fp32)
  }
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}
