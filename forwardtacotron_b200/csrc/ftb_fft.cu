// K7/K8: FFT-block glue for FastPitch (common_layers.py:127-223): positional encoding,
// residual + LayerNorm (optionally chained with the transformer's final norm), and a
// flash-style fp32 self-attention for 2..8 heads.  The QKV / out_proj / conv GEMMs run on
// the tcgen05 kernel (ftb_gemm.cu); these kernels are the HBM-bound / SIMT remainder.
#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

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
      out[off + c] = round ? round_tf32(v[i]) : v[i];
      if (out_full) out_full[off + c] = v[i];
    }
  }
}

// ---------------------------------------------------------------- embedding/x + scale * pe
__global__ void __launch_bounds__(256)
posenc_kernel(const int64_t* __restrict__ tok, const float* __restrict__ table, const float* __restrict__ x_in,
              const float* __restrict__ pe, const float* __restrict__ scale_ptr, int e, int batch, int seq_len,
              int seq_stride, float* __restrict__ out, float* __restrict__ out_full, int round) {
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * 8ll + (threadIdx.x >> 5);
  if (row >= static_cast<long long>(batch) * seq_len) return;
  const int b = static_cast<int>(row / seq_len), t = static_cast<int>(row % seq_len);
  const long long off = (static_cast<long long>(b) * seq_stride + t) * e;
  const float sc = __ldg(scale_ptr);
  const float* src = tok ? (table + tok[row] * e) : (x_in + off);
  const float* per = pe + static_cast<long long>(t) * e;
  for (int c = lane; c < e; c += 32) {
    const float v = src[c] + sc * per[c];
    out[off + c] = round ? round_tf32(v) : v;
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
        out[(rowbase + t) * e + head * D + lane + 32 * d] = round ? round_tf32(v) : v;
      }
    }
  }
}

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_add_layernorm(ftb_handle h, const float* x, const float* res, int ld, int e, const float* gamma,
                                 const float* beta, const float* gamma2, const float* beta2, int batch, int seq_len,
                                 int seq_stride, float* out, float* out_full, int round_tf32, void* stream) {
  if (!h || !x || !gamma || !beta || !out || e < 1 || e > 1024 || e > ld || batch < 1 || seq_len < 1)
    return FTB_E_ARG;
  const long long rows = static_cast<long long>(batch) * seq_len;
  add_layernorm_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, res, ld, e, gamma, beta, gamma2, beta2, batch, seq_len, seq_stride, out, out_full, round_tf32);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_posenc(ftb_handle h, const int64_t* tok, const float* table, const float* x_in, const float* pe,
                          const float* scale_ptr, int e, int batch, int seq_len, int seq_stride, float* out,
                          float* out_full, int round_tf32, void* stream) {
  if (!h || !pe || !scale_ptr || !out || e < 1 || batch < 1 || seq_len < 1) return FTB_E_ARG;
  if (tok ? !table : !x_in) return FTB_E_ARG;
  const long long rows = static_cast<long long>(batch) * seq_len;
  posenc_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      tok, table, x_in, pe, scale_ptr, e, batch, seq_len, seq_stride, out, out_full, round_tf32);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_mha(ftb_handle h, const float* qkv, int e, int heads, const int64_t* key_tok, int batch,
                       int seq_len, int seq_stride, float* out, int round_tf32, void* stream) {
  if (!h || !qkv || !out || heads < 1 || batch < 1 || seq_len < 1 || e % heads) return FTB_E_ARG;
  const int d = e / heads;
  const int qt = d > 64 ? 16 : 32;
  dim3 grid((seq_len + qt - 1) / qt, heads, batch);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (d == 64) mha_kernel<64><<<grid, 256, 0, st>>>(qkv, e, key_tok, seq_len, seq_stride, out, round_tf32);
  else if (d == 128) mha_kernel<128><<<grid, 256, 0, st>>>(qkv, e, key_tok, seq_len, seq_stride, out, round_tf32);
  else return FTB_E_SHAPE;
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}
