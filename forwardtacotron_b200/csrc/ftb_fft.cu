// K7/K8: FFT-block glue for FastPitch (common_layers.py:127-223): positional encoding,
// residual + LayerNorm (optionally chained with the transformer's final norm), and a
// flash-style fp32 self-attention for 2..8 heads.  The QKV / out_proj / conv GEMMs run on
// the tcgen05 kernel (ftb_gemm.cu); these kernels are the HBM-bound / SIMT remainder.
#include <cstdlib>
#include <cstring>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

#include <cuda_fp16.h>

namespace ftb {

// round mode 3 of K7/K8: `out` is an fp16 array of THREE column blocks of width e (row stride 3e halves), [hi | lo | hi]
// with hi = fp16(v), lo = fp16(v - hi): the split-precision GEMM operand (include/ftb200.h, FTB_F_OUT_SPLIT).
__device__ __forceinline__ void store_split1(float* out, long long row_off3, int e, int c, float v) {
  __half* o = reinterpret_cast<__half*>(out) + row_off3 + c;
  const __half hi = __float2half_rn(v);
  o[0] = hi; o[e] = __float2half_rn(v - __half2float(hi)); o[2 * e] = hi;
}
__device__ __forceinline__ void store_split4(float* out, long long row_off3, int e, int c, float4 v) {
  __half* o = reinterpret_cast<__half*>(out) + row_off3 + c;
  const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
  const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
  const __half2 l0 = __floats2half2_rn(v.x - f0.x, v.y - f0.y), l1 = __floats2half2_rn(v.z - f1.x, v.w - f1.y);
  const uint2 ph = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
  const uint2 pl = make_uint2(*reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
  *reinterpret_cast<uint2*>(o) = ph;
  *reinterpret_cast<uint2*>(o + e) = pl;
  *reinterpret_cast<uint2*>(o + 2 * e) = ph;
}

// ---------------------------------------------------------------- residual + LayerNorm (vectorised)
// One warp per row, P4 float4 chunks per lane (e <= 128 * P4, e % 4 == 0, 16-byte aligned rows): every load / store is a
// 16-byte access and the whole row is requested before the first reduction.  The scalar kernel below (32 predicated
// 4-byte accesses per lane and tensor, 117 registers -> two blocks per SM) ran at 1.0 TB/s on the FastPitch postnet
// (72 000 rows x 256: 0.22 ms per LayerNorm, eight of them per generate()); round == 2 writes `out` as the plain fp16
// operand copy of the following GEMM (FTB_F_A_F16), 3 the split [hi | lo | hi] form.
template <int P4>
__global__ void __launch_bounds__(256)
add_layernorm_vec_kernel(const float* __restrict__ x, const float* __restrict__ res, int ld, int e,
                         const float* __restrict__ g1, const float* __restrict__ b1, const float* __restrict__ g2,
                         const float* __restrict__ b2, int batch, int seq_len, int seq_stride, float* __restrict__ out,
                         float* __restrict__ out_full, int round) {
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * 8ll + (threadIdx.x >> 5);
  if (row >= static_cast<long long>(batch) * seq_len) return;
  const int b = static_cast<int>(row / seq_len), t = static_cast<int>(row % seq_len);
  const long long prow = static_cast<long long>(b) * seq_stride + t;
  const long long off = prow * ld;
  float4 v[P4];
#pragma unroll
  for (int i = 0; i < P4; ++i) {
    const int c = 4 * lane + 128 * i;
    v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < e) v[i] = *reinterpret_cast<const float4*>(x + off + c);
  }
  if (res) {
#pragma unroll
    for (int i = 0; i < P4; ++i) {
      const int c = 4 * lane + 128 * i;
      if (c < e) {
        const float4 r = *reinterpret_cast<const float4*>(res + off + c);
        v[i].x += r.x; v[i].y += r.y; v[i].z += r.z; v[i].w += r.w;
      }
    }
  }
  const float inv_e = 1.0f / e;
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    const float* gw = pass == 0 ? g1 : g2;
    const float* bw = pass == 0 ? b1 : b2;
    if (!gw) break;
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < P4; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * inv_e;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < P4; ++i) {
      if (4 * lane + 128 * i < e) {
        const float d0 = v[i].x - mean, d1 = v[i].y - mean, d2 = v[i].z - mean, d3 = v[i].w - mean;
        q += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * inv_e + 1e-5f);
#pragma unroll
    for (int i = 0; i < P4; ++i) {
      const int c = 4 * lane + 128 * i;
      if (c < e) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(gw + c));
        const float4 bb = __ldg(reinterpret_cast<const float4*>(bw + c));
        v[i].x = (v[i].x - mean) * rstd * g.x + bb.x;
        v[i].y = (v[i].y - mean) * rstd * g.y + bb.y;
        v[i].z = (v[i].z - mean) * rstd * g.z + bb.z;
        v[i].w = (v[i].w - mean) * rstd * g.w + bb.w;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < P4; ++i) {
    const int c = 4 * lane + 128 * i;
    if (c < e) {
      if (out_full) *reinterpret_cast<float4*>(out_full + off + c) = v[i];
      if (round == 3) {
        store_split4(out, prow * 3 * e, e, c, v[i]);
      } else if (round == 2) {
        const __half2 h0 = __floats2half2_rn(v[i].x, v[i].y), h1 = __floats2half2_rn(v[i].z, v[i].w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(out) + off + c) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
      } else {
        float4 y = v[i];
        if (round) { y.x = round_tf32(y.x); y.y = round_tf32(y.y); y.z = round_tf32(y.z); y.w = round_tf32(y.w); }
        *reinterpret_cast<float4*>(out + off + c) = y;
      }
    }
  }
}

// ---------------------------------------------------------------- residual + LayerNorm
// One warp per row; e <= 1024 -> up to 32 values per lane held in registers.
__global__ void __launch_bounds__(256)
add_layernorm_kernel(const float* __restrict__ x, const float* __restrict__ res, int ld, int e,
                     const float* __restrict__ g1, const float* __restrict__ b1, const float* __restrict__ g2,
                     const float* __restrict__ b2, int batch, int seq_len, int seq_stride, float* __restrict__ out,
                     float* __restrict__ out_full, int round) {
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * 8ll + (threadIdx.x >> 5);
  if (row >= static_cast<long long>(batch) * seq_len) return;
  const int b = static_cast<int>(row / seq_len), t = static_cast<int>(row % seq_len);
  const long long off = (static_cast<long long>(b) * seq_stride + t) * ld;
  float v[32];
  const int per = (e + 31) / 32;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const int c = lane + i * 32;
    v[i] = 0.0f;
    if (i < per && c < e) v[i] = x[off + c] + (res ? res[off + c] : 0.0f);
  }
  for (int pass = 0; pass < 2; ++pass) {
    const float* gw = pass == 0 ? g1 : g2;
    const float* bw = pass == 0 ? b1 : b2;
    if (!gw) break;
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / e;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int c = lane + i * 32;
      if (i < per && c < e) { const float d = v[i] - mean; q += d * d; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / e + 1e-5f);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const int c = lane + i * 32;
      if (i < per && c < e) v[i] = (v[i] - mean) * rstd * __ldg(gw + c) + __ldg(bw + c);
    }
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const int c = lane + i * 32;
    if (i < per && c < e) {
      if (round == 3) store_split1(out, (static_cast<long long>(b) * seq_stride + t) * 3 * e, e, c, v[i]);
      else if (round == 2) reinterpret_cast<__half*>(out)[off + c] = __float2half_rn(v[i]);
      else out[off + c] = round ? round_tf32(v[i]) : v[i];
      if (out_full) out_full[off + c] = v[i];
    }
  }
}

// ---------------------------------------------------------------- embedding/x + scale * pe
__global__ void __launch_bounds__(256)
posenc_kernel(const int64_t* __restrict__ tok, const float* __restrict__ table, const float* __restrict__ x_in,
              const float* __restrict__ pe, const float* __restrict__ scale_ptr, int e, int batch, int seq_len,
              int seq_stride, float* __restrict__ out, float* __restrict__ out_full, int round, int table_rows,
              int* __restrict__ err) {
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * 8ll + (threadIdx.x >> 5);
  if (row >= static_cast<long long>(batch) * seq_len) return;
  const int b = static_cast<int>(row / seq_len), t = static_cast<int>(row % seq_len);
  const long long off = (static_cast<long long>(b) * seq_stride + t) * e;
  const float sc = __ldg(scale_ptr);
  long long id = tok ? tok[row] : 0;
  if (tok && (id < 0 || id >= table_rows)) {       // nn.Embedding would raise IndexError: clamp + flag (see ftb_misc.cu)
    if (err) atomicOr(err, 1);
    id = 0;
  }
  const float* src = tok ? (table + id * e) : (x_in + off);
  const float* per = pe + static_cast<long long>(t) * e;
  for (int c = lane; c < e; c += 32) {
    const float v = src[c] + sc * per[c];
    if (round == 3) store_split1(out, off * 3, e, c, v);
    else if (round == 2) reinterpret_cast<__half*>(out)[off + c] = __float2half_rn(v);
    else out[off + c] = round ? round_tf32(v) : v;
    if (out_full) out_full[off + c] = v;
  }
}

// ---------------------------------------------------------------- self attention
// grid = (query tiles of QT = 32 (D=64) / 16 (D=128), heads, batch); block = 8 warps, each warp
// owns QT/8 query rows; K/V tiles of 32 keys staged in smem; online softmax in fp32.
// D = head dim (64 or 128): each lane holds D/32 output columns per query row.
template <int D>
__global__ void __launch_bounds__(256)
mha_kernel(const float* __restrict__ qkv, int e, const int64_t* __restrict__ key_tok, int seq_len, int seq_stride,
           float* __restrict__ out, int round) {
  constexpr int KT = 32, QT = (D > 64) ? 16 : 32, QPW = QT / 8, DPL = D / 32;
  __shared__ float sk[KT][D + 1];      // read as sk[lane][c]: padded against bank conflicts
  __shared__ float sv[KT][D];          // read row-wise
  __shared__ float sq[QT][D];          // read as a broadcast
  __shared__ int smask[KT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q0 = blockIdx.x * QT, head = blockIdx.y, b = blockIdx.z;
  const long long rowbase = static_cast<long long>(b) * seq_stride;
  const int ld = 3 * e;
  const float scale = rsqrtf(static_cast<float>(D));
  for (int i = threadIdx.x; i < QT * D; i += 256) {
    const int r = i / D, c = i % D;
    const int t = q0 + r;
    sq[r][c] = (t < seq_len) ? qkv[(rowbase + t) * ld + head * D + c] * scale : 0.0f;
  }
  float m[QPW], l[QPW], acc[QPW][DPL];
#pragma unroll
  for (int j = 0; j < QPW; ++j) {
    m[j] = -INFINITY; l[j] = 0.0f;
#pragma unroll
    for (int d = 0; d < DPL; ++d) acc[j][d] = 0.0f;
  }
  for (int k0 = 0; k0 < seq_len; k0 += KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < KT * D; i += 256) {
      const int r = i / D, c = i % D;
      const int t = k0 + r;
      const bool ok = t < seq_len;
      sk[r][c] = ok ? qkv[(rowbase + t) * ld + e + head * D + c] : 0.0f;
      sv[r][c] = ok ? qkv[(rowbase + t) * ld + 2 * e + head * D + c] : 0.0f;
    }
    if (threadIdx.x < KT) {
      const int t = k0 + threadIdx.x;
      int ok = t < seq_len;
      if (ok && key_tok) ok = key_tok[static_cast<long long>(b) * seq_len + t] != 0;
      smask[threadIdx.x] = ok;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < QPW; ++j) {
      const int qr = warp * QPW + j;
      // lane = key index within the tile
      float s = 0.0f;
#pragma unroll 8
      for (int c = 0; c < D; ++c) s = fmaf(sq[qr][c], sk[lane][c], s);
      if (!smask[lane]) s = -INFINITY;
      float tmax = s;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
      const float mnew = fmaxf(m[j], tmax);
      if (mnew == -INFINITY) continue;                 // every key so far masked
      const float corr = (m[j] == -INFINITY) ? 0.0f : expf(m[j] - mnew);
      const float p = (s == -INFINITY) ? 0.0f : expf(s - mnew);
      float psum = p;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, o);
      l[j] = l[j] * corr + psum;
      m[j] = mnew;
#pragma unroll
      for (int d = 0; d < DPL; ++d) acc[j][d] *= corr;
      for (int kk = 0; kk < KT; ++kk) {
        const float pk = __shfl_sync(0xffffffffu, p, kk);
#pragma unroll
        for (int d = 0; d < DPL; ++d) acc[j][d] = fmaf(pk, sv[kk][lane + 32 * d], acc[j][d]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < QPW; ++j) {
    const int t = q0 + warp * QPW + j;
    if (t < seq_len) {
      const float inv = 1.0f / l[j];
#pragma unroll
      for (int d = 0; d < DPL; ++d) {
        const float v = acc[j][d] * inv;
        if (round == 2) reinterpret_cast<__half*>(out)[(rowbase + t) * e + head * D + lane + 32 * d] = __float2half_rn(v);
        else out[(rowbase + t) * e + head * D + lane + 32 * d] = round ? round_tf32(v) : v;
      }
    }
  }
}


// ---------------------------------------------------------------- self attention, register-tiled
// Same exact-fp32 arithmetic (the FastPitch parity margin leaves no room for tf32 scores), but both products are
// register-tiled SIMT GEMMs instead of one dot product per lane:
//   grid = (query tiles of 64, heads, batch), 256 threads; thread (ty, tx) = (tid / 16, tid % 16) owns query rows
//   4ty..4ty+3.  Per key tile of KT keys: S = Q K^T with a 4 x KT/16 micro-tile per thread (float4 operand reads, the
//   q rows are a quarter-warp broadcast), scores -> smem, online softmax by one warp per 8 rows (P overwrites S, the
//   rescale factor goes to smem), then O = O*corr + P V with a 4 x D/16 micro-tile per thread.
template <int D>
__global__ void __launch_bounds__(256)
mha_tiled_kernel(const float* __restrict__ qkv, int e, const int64_t* __restrict__ key_tok, int seq_len,
                 int seq_stride, float* __restrict__ out, int round) {
  // D % 4 == 0 (float4 operand reads).  Row stride LDK = odd multiple of 4 floats (mod 32) keeps the 8 key rows a
  // quarter-warp reads at once on distinct bank groups.  A thread's output columns are the float4 groups tx + 16h.
  constexpr int QT = 64, KT = (D > 64) ? 32 : 64, NJ = KT / 16, D4 = D / 4, NH = (D4 + 15) / 16;
  constexpr int LDK = D + ((D % 8 == 0) ? 4 : 8), LDP = KT + 4;
  static_assert(D % 4 == 0, "head dim must be a multiple of 4");
  extern __shared__ __align__(16) float mha_sm[];
  float* sq = mha_sm;                  // [QT][LDK]  q * 1/sqrt(D)
  float* sk = sq + QT * LDK;           // [KT][LDK]
  float* sv = sk + KT * LDK;           // [KT][D]
  float* ss = sv + KT * D;             // [QT][LDP]  scores, then probabilities
  float* scorr = ss + QT * LDP;        // [QT] rescale factor of this key tile, then 1 / row sum
  int* smask = reinterpret_cast<int*>(scorr + QT);     // [KT]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, ty = tid >> 4, tx = tid & 15;
  const int q0 = blockIdx.x * QT, head = blockIdx.y, b = blockIdx.z;
  const long long rowbase = static_cast<long long>(b) * seq_stride;
  const int ld = 3 * e;
  const float scale = rsqrtf(static_cast<float>(D));
  for (int i = tid; i < QT * (D / 4); i += 256) {
    const int r = i / (D / 4), c4 = i % (D / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q0 + r < seq_len) {
      v = *reinterpret_cast<const float4*>(qkv + (rowbase + q0 + r) * ld + head * D + 4 * c4);
      v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    }
    *reinterpret_cast<float4*>(sq + r * LDK + 4 * c4) = v;
  }
  float m_run[8], l_run[8];            // softmax state of rows 8*warp + j (uniform over the warp)
#pragma unroll
  for (int j = 0; j < 8; ++j) { m_run[j] = -INFINITY; l_run[j] = 0.0f; }
  float oacc[4][4 * NH];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int c = 0; c < 4 * NH; ++c) oacc[i][c] = 0.0f;

  for (int k0 = 0; k0 < seq_len; k0 += KT) {
    __syncthreads();                   // previous tile fully consumed (and sq written, first iteration)
    for (int i = tid; i < KT * (D / 4); i += 256) {
      const int r = i / (D / 4), c4 = i % (D / 4);
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < seq_len) {
        const float* src = qkv + (rowbase + k0 + r) * ld + head * D + 4 * c4;
        kv = *reinterpret_cast<const float4*>(src + e);
        vv = *reinterpret_cast<const float4*>(src + 2 * e);
      }
      *reinterpret_cast<float4*>(sk + r * LDK + 4 * c4) = kv;
      *reinterpret_cast<float4*>(sv + r * D + 4 * c4) = vv;
    }
    if (tid < KT) {
      const int t = k0 + tid;
      int ok = t < seq_len;
      if (ok && key_tok) ok = key_tok[static_cast<long long>(b) * seq_len + t] != 0;
      smask[tid] = ok;
    }
    __syncthreads();
    // ---- S = Q K^T
    {
      float sacc[4][NJ];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) sacc[i][j] = 0.0f;
#pragma unroll 4
      for (int d = 0; d < D; d += 4) {
        float4 qv[4], kv[NJ];
#pragma unroll
        for (int i = 0; i < 4; ++i) qv[i] = *reinterpret_cast<const float4*>(sq + (4 * ty + i) * LDK + d);
#pragma unroll
        for (int j = 0; j < NJ; ++j) kv[j] = *reinterpret_cast<const float4*>(sk + (tx + 16 * j) * LDK + d);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < NJ; ++j) {
            sacc[i][j] = fmaf(qv[i].x, kv[j].x, sacc[i][j]);
            sacc[i][j] = fmaf(qv[i].y, kv[j].y, sacc[i][j]);
            sacc[i][j] = fmaf(qv[i].z, kv[j].z, sacc[i][j]);
            sacc[i][j] = fmaf(qv[i].w, kv[j].w, sacc[i][j]);
          }
      }
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const bool ok = smask[tx + 16 * j] != 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) ss[(4 * ty + i) * LDP + tx + 16 * j] = ok ? sacc[i][j] : -INFINITY;
      }
    }
    __syncthreads();
    // ---- online softmax: warp w owns rows 8w .. 8w+7
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float* row = ss + (8 * warp + j) * LDP;
      float sv_[KT / 32];
      float tmax = -INFINITY;
#pragma unroll
      for (int c = 0; c < KT / 32; ++c) { sv_[c] = row[lane + 32 * c]; tmax = fmaxf(tmax, sv_[c]); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
      const float mnew = fmaxf(m_run[j], tmax);
      const float corr = (m_run[j] == -INFINITY) ? 0.0f : expf(m_run[j] - mnew);   // mnew = -inf only if m_run is
      float psum = 0.0f;
#pragma unroll
      for (int c = 0; c < KT / 32; ++c) {
        const float pexp = (sv_[c] == -INFINITY) ? 0.0f : expf(sv_[c] - mnew);
        row[lane + 32 * c] = pexp;
        psum += pexp;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, o);
      l_run[j] = l_run[j] * corr + psum;
      m_run[j] = mnew;
      if (lane == 0) scorr[8 * warp + j] = corr;
    }
    __syncthreads();
    // ---- O = O * corr + P V
    {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float cr = scorr[4 * ty + i];
#pragma unroll
        for (int c = 0; c < 4 * NH; ++c) oacc[i][c] *= cr;
      }
#pragma unroll 2
      for (int k = 0; k < KT; k += 4) {
        float4 pv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) pv[i] = *reinterpret_cast<const float4*>(ss + (4 * ty + i) * LDP + k);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
          for (int h = 0; h < NH; ++h) {
            if ((D4 % 16) != 0 && tx + 16 * h >= D4) continue;
            const float4 vv = *reinterpret_cast<const float4*>(sv + (k + kk) * D + 4 * tx + 64 * h);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float pk = kk == 0 ? pv[i].x : kk == 1 ? pv[i].y : kk == 2 ? pv[i].z : pv[i].w;
              oacc[i][4 * h + 0] = fmaf(pk, vv.x, oacc[i][4 * h + 0]);
              oacc[i][4 * h + 1] = fmaf(pk, vv.y, oacc[i][4 * h + 1]);
              oacc[i][4 * h + 2] = fmaf(pk, vv.z, oacc[i][4 * h + 2]);
              oacc[i][4 * h + 3] = fmaf(pk, vv.w, oacc[i][4 * h + 3]);
            }
          }
        }
      }
    }
  }
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) scorr[8 * warp + j] = 1.0f / l_run[j];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = q0 + 4 * ty + i;
    if (t < seq_len) {
      const float inv = scorr[4 * ty + i];
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        if ((D4 % 16) != 0 && tx + 16 * h >= D4) continue;
        float4 v = make_float4(oacc[i][4 * h] * inv, oacc[i][4 * h + 1] * inv, oacc[i][4 * h + 2] * inv,
                               oacc[i][4 * h + 3] * inv);
        if (round == 3) { store_split4(out, (rowbase + t) * 3 * e, e, head * D + 4 * tx + 64 * h, v); continue; }
        if (round == 2) {
          const __half2 h0 = __floats2half2_rn(v.x, v.y), h1 = __floats2half2_rn(v.z, v.w);
          *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(out) + (rowbase + t) * e + head * D + 4 * tx + 64 * h) =
              make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
          continue;
        }
        if (round) { v.x = round_tf32(v.x); v.y = round_tf32(v.y); v.z = round_tf32(v.z); v.w = round_tf32(v.w); }
        *reinterpret_cast<float4*>(out + (rowbase + t) * e + head * D + 4 * tx + 64 * h) = v;
      }
    }
  }
}

template <int D>
static constexpr int mha_tiled_smem() {
  constexpr int QT = 64, KT = (D > 64) ? 32 : 64, LDK = D + ((D % 8 == 0) ? 4 : 8), LDP = KT + 4;
  return (QT * LDK + KT * LDK + KT * D + QT * LDP + QT + KT) * 4;
}

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_add_layernorm(ftb_handle h, const float* x, const float* res, int ld, int e, const float* gamma,
                                 const float* beta, const float* gamma2, const float* beta2, int batch, int seq_len,
                                 int seq_stride, float* out, float* out_full, int round_tf32, void* stream) {
  if (!h || !x || !gamma || !beta || !out || e < 1 || e > 1024 || e > ld || batch < 1 || seq_len < 1)
    return FTB_E_ARG;
  if (round_tf32 == 3 && (e % 8)) return FTB_E_SHAPE;
  if (round_tf32 == 2 && ((e % 8) || (ld % 8))) return FTB_E_SHAPE;
  const long long rows = static_cast<long long>(batch) * seq_len;
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  const bool vec = (e % 4) == 0 && (ld % 4) == 0 && al16(x) && al16(res) && al16(out) && al16(out_full) && al16(gamma) &&
                   al16(beta) && al16(gamma2) && al16(beta2);
#define FTB_LN_CASE(PP)                                                                                              \
  add_layernorm_vec_kernel<PP><<<grid, 256, 0, st>>>(x, res, ld, e, gamma, beta, gamma2, beta2, batch, seq_len,      \
                                                     seq_stride, out, out_full, round_tf32)
  if (!vec) add_layernorm_kernel<<<grid, 256, 0, st>>>(x, res, ld, e, gamma, beta, gamma2, beta2, batch, seq_len, seq_stride,
                                                       out, out_full, round_tf32);
  else if (e <= 128) FTB_LN_CASE(1);
  else if (e <= 256) FTB_LN_CASE(2);
  else if (e <= 384) FTB_LN_CASE(3);
  else if (e <= 512) FTB_LN_CASE(4);
  else FTB_LN_CASE(8);
#undef FTB_LN_CASE
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_posenc(ftb_handle h, const int64_t* tok, const float* table, const float* x_in, const float* pe,
                          const float* scale_ptr, int e, int batch, int seq_len, int seq_stride, float* out,
                          float* out_full, int round_tf32, int table_rows, void* stream) {
  if (!h || !pe || !scale_ptr || !out || e < 1 || batch < 1 || seq_len < 1) return FTB_E_ARG;
  if (tok ? (!table || table_rows < 1) : !x_in) return FTB_E_ARG;
  if ((round_tf32 == 3 || round_tf32 == 2) && (e % 8)) return FTB_E_SHAPE;
  const long long rows = static_cast<long long>(batch) * seq_len;
  posenc_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      tok, table, x_in, pe, scale_ptr, e, batch, seq_len, seq_stride, out, out_full, round_tf32, table_rows, h->err_word);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_mha(ftb_handle h, const float* qkv, int e, int heads, const int64_t* key_tok, int batch,
                       int seq_len, int seq_stride, float* out, int round_tf32, void* stream) {
  if (!h || !qkv || !out || heads < 1 || batch < 1 || seq_len < 1 || e % heads) return FTB_E_ARG;
  const int d = e / heads;
  if ((round_tf32 == 3 || round_tf32 == 2) && (e % 8)) return FTB_E_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static const bool legacy = [] { const char* v = getenv("FTB_MHA_IMPL"); return v && strcmp(v, "legacy") == 0; }();
  if ((e % 4) != 0 || (reinterpret_cast<uintptr_t>(qkv) & 15u) || (reinterpret_cast<uintptr_t>(out) & 15u))
    return FTB_E_ALIGN;
  if (legacy && round_tf32 != 3) {                  // first-build kernel, kept for A/B measurements
    const int qt = d > 64 ? 16 : 32;
    dim3 grid((seq_len + qt - 1) / qt, heads, batch);
    if (d == 64) mha_kernel<64><<<grid, 256, 0, st>>>(qkv, e, key_tok, seq_len, seq_stride, out, round_tf32);
    else if (d == 128) mha_kernel<128><<<grid, 256, 0, st>>>(qkv, e, key_tok, seq_len, seq_stride, out, round_tf32);
    else return FTB_E_SHAPE;
  } else {
    dim3 grid((seq_len + 63) / 64, heads, batch);
    cudaError_t ce = cudaSuccess;
#define FTB_MHA_CASE(DD)                                                                                          \
    case DD: {                                                                                                    \
      static bool attr_##DD = false;                                                                              \
      if (!attr_##DD) {                                                                                           \
        ce = cudaFuncSetAttribute(mha_tiled_kernel<DD>, cudaFuncAttributeMaxDynamicSharedMemorySize,              \
                                  mha_tiled_smem<DD>());                                                          \
        if (ce != cudaSuccess) return static_cast<int>(ce);                                                       \
        attr_##DD = true;                                                                                         \
      }                                                                                                           \
      mha_tiled_kernel<DD><<<grid, 256, mha_tiled_smem<DD>(), st>>>(qkv, e, key_tok, seq_len, seq_stride, out,    \
                                                                    round_tf32);                                  \
      break;                                                                                                      \
    }
    switch (d) {          // 64 / 128: FastPitch; 192 / 196 / 256: MultiFastPitch (d_model + speaker [+ cond] dims)
      FTB_MHA_CASE(64) FTB_MHA_CASE(128) FTB_MHA_CASE(192) FTB_MHA_CASE(196) FTB_MHA_CASE(256)
      default: return FTB_E_SHAPE;
    }
#undef FTB_MHA_CASE
  }
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}
