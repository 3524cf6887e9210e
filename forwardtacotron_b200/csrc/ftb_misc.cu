// Small glue kernels: embedding gather into the padded-flat layout, predictor head
// (row dot products), arg-max over a tiny class axis.  All are bandwidth/latency trivial
// next to the GEMMs and recurrences; one warp per row, coalesced accesses.
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

// mode: 0 = fp32 as stored, 1 = fp32 rounded to tf32, 2 = fp16 (out is a __half array, out_ld in halves): the operand
// copy of a buffer a FTB_F_A_F16 GEMM problem reads; 3 = split fp16 (out_ld = 3*seg halves): hi = fp16(v) at column c,
// lo = fp16(v - hi) at seg + c, hi again at 2*seg + c -- the operand of a split-precision (FTB_F_OUT_SPLIT) consumer.
__device__ __forceinline__ void embed_store(float* out, long long i, float v, int mode, int seg) {
  if (mode == 3) {
    __half* o = reinterpret_cast<__half*>(out);
    const __half hi = __float2half_rn(v);
    o[i] = hi;
    o[i + seg] = __float2half_rn(v - __half2float(hi));
    o[i + 2 * seg] = hi;
  } else if (mode == 2) {
    reinterpret_cast<__half*>(out)[i] = __float2half_rn(v);
  } else {
    out[i] = mode ? round_tf32(v) : v;
  }
}

// nn.Embedding raises IndexError for an id outside [0, rows) (forward_tacotron.py:31,210); here the id is clamped to row 0
// (no out-of-bounds read) and bit 0 of the handle's error word is set -- generate() reads it with the scan result.
__device__ __forceinline__ long long checked_row(long long id, int rows, int* err) {
  if (id < 0 || id >= rows) {
    if (err) atomicOr(err, 1);
    return 0;
  }
  return id;
}

__device__ __forceinline__ void embed_rows(const int64_t* __restrict__ tok, const float* __restrict__ table, int e1,
                                           const int64_t* __restrict__ tok2, const float* __restrict__ table2, int e2,
                                           const float* __restrict__ extra, int e3, int batch, int T, int seq_stride,
                                           float* __restrict__ out, int out_ld, int mode, int rows1, int rows2,
                                           int* __restrict__ err) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);          // batch * T < 2^31 (checked by the callers)
  if (row >= batch * T) return;
  const int b = row / T, t = row - b * T;
  const long long o = (static_cast<long long>(b) * seq_stride + t) * out_ld;
  const float* s1 = table + checked_row(tok[row], rows1, err) * e1;
  const int seg = out_ld / 3;
  if (mode != 3 && e2 == 0 && e3 == 0 && (e1 & 3) == 0 && (out_ld & 3) == 0 &&
      ((reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0) {
    // the common case (one table, width a multiple of 4): 16-byte reads, 16- / 8-byte stores
    for (int c = lane * 4; c < e1; c += 128) {
      const float4 v = *reinterpret_cast<const float4*>(s1 + c);
      if (mode == 2) {
        const __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(out) + o + c) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
      } else {
        float4 y = v;
        if (mode) { y.x = round_tf32(y.x); y.y = round_tf32(y.y); y.z = round_tf32(y.z); y.w = round_tf32(y.w); }
        *reinterpret_cast<float4*>(out + o + c) = y;
      }
    }
    return;
  }
  for (int c = lane; c < e1; c += 32) embed_store(out, o + c, s1[c], mode, seg);
  if (e2 > 0) {
    const float* s2 = table2 + checked_row(tok2[row], rows2, err) * e2;
    for (int c = lane; c < e2; c += 32) embed_store(out, o + e1 + c, s2[c], mode, seg);
  }
  if (e3 > 0) {
    const float* s3 = extra + static_cast<long long>(b) * e3;
    for (int c = lane; c < e3; c += 32) embed_store(out, o + e1 + e2 + c, s3[c], mode, seg);
  }
}

__global__ void __launch_bounds__(256)
embed_kernel(const int64_t* __restrict__ tok, const float* __restrict__ table, int e1,
             const int64_t* __restrict__ tok2, const float* __restrict__ table2, int e2,
             const float* __restrict__ extra, int e3, int batch, int T, int seq_stride,
             float* __restrict__ out, int out_ld, int mode, int rows1, int rows2, int* __restrict__ err) {
  embed_rows(tok, table, e1, tok2, table2, e2, extra, e3, batch, T, seq_stride, out, out_ld, mode, rows1, rows2, err);
}

struct EmbedBatch { ftb_embed_item item[FTB_MAX_GLUE_ITEMS]; };
__global__ void __launch_bounds__(256)
embed_batch_kernel(const __grid_constant__ EmbedBatch args, int batch, int T, int seq_stride, int* __restrict__ err) {
  const ftb_embed_item& it = args.item[blockIdx.y];
  embed_rows(it.tok, it.table, it.e1, it.tok2, it.table2, it.e2, it.extra, it.e3, batch, T, seq_stride, it.out, it.out_ld,
             it.round_tf32, it.rows1, it.rows2, err);
}

__device__ __forceinline__ void rowdot_rows(const float* __restrict__ x, int x_ld, int k, const float* __restrict__ w,
                                            const float* __restrict__ bias, int n_out, float alpha, int batch, int T,
                                            int seq_stride, float* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);          // batch * T < 2^31 (checked by the callers)
  if (row >= batch * T) return;
  const int b = row / T, t = row - b * T;
  const float* xr = x + (static_cast<long long>(b) * seq_stride + t) * x_ld;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  int c_done = 0;
  if ((k & 3) == 0 && (x_ld & 3) == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) & 15u) == 0) {
    for (int c = lane * 4; c < k; c += 128) {          // 16-byte reads of the row and of every weight vector
      const float4 xv = *reinterpret_cast<const float4*>(xr + c);
#pragma unroll
      for (int o = 0; o < 4; ++o) {
        if (o < n_out) {
          const float4 wv = __ldg(reinterpret_cast<const float4*>(w + o * k + c));
          acc[o] = fmaf(xv.w, wv.w, fmaf(xv.z, wv.z, fmaf(xv.y, wv.y, fmaf(xv.x, wv.x, acc[o]))));
        }
      }
    }
    c_done = k;
  }
  for (int c = c_done + lane; c < k; c += 32) {
    const float xv = xr[c];
#pragma unroll
    for (int o = 0; o < 4; ++o)
      if (o < n_out) acc[o] = fmaf(xv, __ldg(w + o * k + c), acc[o]);
  }
#pragma unroll
  for (int o = 0; o < 4; ++o) {
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], s);
  }
  if (lane == 0) {
    for (int o = 0; o < n_out; ++o)
      y[static_cast<long long>(row) * n_out + o] = __fdiv_rn(acc[o] + bias[o], alpha);   // `/ alpha`: a true division
  }
}

__global__ void __launch_bounds__(256)
rowdot_kernel(const float* __restrict__ x, int x_ld, int k, const float* __restrict__ w,
              const float* __restrict__ bias, int n_out, float alpha, int batch, int T, int seq_stride,
              float* __restrict__ y) {
  rowdot_rows(x, x_ld, k, w, bias, n_out, alpha, batch, T, seq_stride, y);
}

struct RowdotBatch { ftb_rowdot_item item[FTB_MAX_GLUE_ITEMS]; };
__global__ void __launch_bounds__(256)
rowdot_batch_kernel(const __grid_constant__ RowdotBatch args, int batch, int T, int seq_stride) {
  const ftb_rowdot_item& it = args.item[blockIdx.y];
  rowdot_rows(it.x, it.x_ld, it.k, it.w, it.bias, it.n_out, it.alpha, batch, T, seq_stride, it.y);
}

__global__ void argmax_kernel(const float* __restrict__ x, long long n, int c, int64_t* __restrict__ out) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const float* r = x + i * c;
  int best = 0;
  float bv = r[0];
  for (int j = 1; j < c; ++j) {
    if (r[j] > bv) { bv = r[j]; best = j; }     // first maximum wins, like torch.argmax
  }
  out[i] = best;
}

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_embed(ftb_handle h, const int64_t* tok, const float* table, int e1, const int64_t* tok2,
                         const float* table2, int e2, const float* extra, int e3, int batch, int T, int seq_stride,
                         float* out, int out_ld, int round_tf32, int rows1, int rows2, void* stream) {
  if (!h || !tok || !table || !out || e1 < 1 || batch < 1 || T < 1 || rows1 < 1 || (e2 > 0 && rows2 < 1)) return FTB_E_ARG;
  if (e2 > 0 && (!tok2 || !table2)) return FTB_E_ARG;
  if (e3 > 0 && !extra) return FTB_E_ARG;
  if (seq_stride < T || round_tf32 < 0 || round_tf32 > 3) return FTB_E_ARG;
  if (e1 + e2 + e3 > (round_tf32 == 3 ? out_ld / 3 : out_ld) || (round_tf32 == 3 && out_ld % 3)) return FTB_E_ARG;
  const long long rows = static_cast<long long>(batch) * T;
  if (rows > 0x7fffffffll) return FTB_E_ARG;
  embed_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      tok, table, e1, tok2, table2, e2, extra, e3, batch, T, seq_stride, out, out_ld, round_tf32, rows1, rows2, h->err_word);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_embed_batch(ftb_handle h, const ftb_embed_item* items, int nitems, int batch, int T, int seq_stride,
                               void* stream) {
  if (!h || !items || nitems < 1 || nitems > FTB_MAX_GLUE_ITEMS || batch < 1 || T < 1 || seq_stride < T) return FTB_E_ARG;
  EmbedBatch args{};
  for (int i = 0; i < nitems; ++i) {
    const ftb_embed_item& q = items[i];
    if (!q.tok || !q.table || !q.out || q.e1 < 1 || q.rows1 < 1 || (q.e2 > 0 && q.rows2 < 1)) return FTB_E_ARG;
    if (q.e2 > 0 && (!q.tok2 || !q.table2)) return FTB_E_ARG;
    if (q.e3 > 0 && !q.extra) return FTB_E_ARG;
    if (q.round_tf32 < 0 || q.round_tf32 > 3) return FTB_E_ARG;
    if (q.e1 + q.e2 + q.e3 > (q.round_tf32 == 3 ? q.out_ld / 3 : q.out_ld) || (q.round_tf32 == 3 && q.out_ld % 3)) return FTB_E_ARG;
    args.item[i] = q;
  }
  const long long rows = static_cast<long long>(batch) * T;
  if (rows > 0x7fffffffll) return FTB_E_ARG;
  embed_batch_kernel<<<dim3(static_cast<unsigned>((rows + 7) / 8), nitems), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      args, batch, T, seq_stride, h->err_word);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_rowdot_batch(ftb_handle h, const ftb_rowdot_item* items, int nitems, int batch, int T, int seq_stride,
                                void* stream) {
  if (!h || !items || nitems < 1 || nitems > FTB_MAX_GLUE_ITEMS || batch < 1 || T < 1 || seq_stride < T) return FTB_E_ARG;
  RowdotBatch args{};
  for (int i = 0; i < nitems; ++i) {
    const ftb_rowdot_item& q = items[i];
    if (!q.x || !q.w || !q.bias || !q.y || q.n_out < 1 || q.n_out > 4 || q.k < 1 || q.alpha == 0.0f) return FTB_E_ARG;
    args.item[i] = q;
  }
  const long long rows = static_cast<long long>(batch) * T;
  if (rows > 0x7fffffffll) return FTB_E_ARG;
  rowdot_batch_kernel<<<dim3(static_cast<unsigned>((rows + 7) / 8), nitems), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      args, batch, T, seq_stride);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_rowdot(ftb_handle h, const float* x, int x_ld, int k, const float* w, const float* bias,
                          int n_out, float alpha, int batch, int T, int seq_stride, float* y, void* stream) {
  if (!h || !x || !w || !bias || !y || n_out < 1 || n_out > 4 || k < 1 || batch < 1 || T < 1 || alpha == 0.0f) return FTB_E_ARG;
  const long long rows = static_cast<long long>(batch) * T;
  if (rows > 0x7fffffffll) return FTB_E_ARG;
  rowdot_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, x_ld, k, w, bias, n_out, alpha, batch, T, seq_stride, y);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_argmax_lastdim(ftb_handle h, const float* x, int64_t n, int c, int64_t* out, void* stream) {
  if (!h || !x || !out || n < 1 || c < 1 || c > 8) return FTB_E_ARG;
  argmax_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, c, out);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}
