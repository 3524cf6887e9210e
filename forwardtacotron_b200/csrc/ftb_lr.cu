// K5/K6: duration rounding + warp-scan cumsum, and the LengthRegulator gather with the
// pitch/energy projection add fused into the source read.  HBM-bound integer/copy work:
// 128-bit accesses, one warp per output row, no tensor cores.
#include <cuda_fp16.h>

#include <mutex>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int SCAN_THREADS = 1024;

// One block per sentence.  Every block recomputes the batch-global fallback predicate
// (`torch.sum(dur_hat.long()) <= 0`, forward_tacotron.py:176; B*T is a few thousand floats,
// L2-resident) so no grid-wide dependency is needed.
__global__ void __launch_bounds__(SCAN_THREADS)
lr_scan_kernel(float* __restrict__ dur, int batch, int T, int apply_fallback, int32_t* __restrict__ cum,
               int32_t* __restrict__ lens, int32_t* __restrict__ meta, int* __restrict__ err) {
  __shared__ long long red[32];
  __shared__ int warp_tot[32];
  __shared__ int carry_s;
  __shared__ int flag_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;

  // ---- fallback predicate: sum over the WHOLE batch of trunc-toward-zero(dur)
  int flag = 0;
  if (apply_fallback) {
    long long s = 0;
    const long long total = static_cast<long long>(batch) * T;
    for (long long i = tid; i < total; i += SCAN_THREADS) s += static_cast<long long>(dur[i]);   // (long) truncates
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (warp == 0) {
      long long v = red[lane];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) flag_s = (v <= 0) ? 1 : 0;
    }
    __syncthreads();
    flag = flag_s;
  }
  // dur itself is NOT rewritten here: another block may still be summing it.  The in-place
  // fill/clamp the reference performs is done by lr_finalize_kernel in the next launch.
  if (tid == 0) carry_s = 0;
  __syncthreads();

  float* d = dur + static_cast<long long>(b) * T;
  int32_t* c = cum + static_cast<long long>(b) * T;
  for (int base = 0; base < T; base += SCAN_THREADS) {
    const int t = base + tid;
    int rep = 0;
    if (t < T) {
      float v = flag ? 2.0f : d[t];
      if (v < 0.0f) v = 0.0f;                       // dur[dur < 0] = 0      (common_layers.py:18)
      rep = static_cast<int>(__fadd_rn(v, 0.5f));    // (dur + 0.5).long()    (common_layers.py:21)
    }
    int x = rep;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      warp_tot[lane] = w;                            // inclusive over warps
    }
    __syncthreads();
    const int before = carry_s + (warp > 0 ? warp_tot[warp - 1] : 0);
    if (t < T) c[t] = before + x;
    __syncthreads();
    if (tid == 0) carry_s += warp_tot[31];
    __syncthreads();
  }
  if (tid == 0) {
    const int len = carry_s;
    lens[b] = len;
    atomicMax(meta + 0, len);
    atomicAdd(meta + 2, len);
    if (b == 0) {
      meta[1] = flag;
      // error word of the embedding kernels that ran before this scan on the stream (bit 0: id outside its table):
      // handed over in meta[3] and cleared for the next call
      meta[3] = err ? atomicExch(err, 0) : 0;
    }
  }
}

// In-place part of the reference semantics, run after every block of lr_scan_kernel is done:
// the fallback fill (forward_tacotron.py:177) and the clamp (common_layers.py:18) -- the tensor
// returned under 'dur' is this one.
__global__ void lr_finalize_kernel(float* __restrict__ dur, long long n, const int32_t* __restrict__ meta) {
  const int flag = meta[1];
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v = flag ? 2.0f : dur[i];
    if (v < 0.0f) v = 0.0f;
    dur[i] = v;
  }
}

struct GatherArgs {
  ftb_lr_gather_args a;
};

// Source-centric expansion: one warp per TOKEN (b, t).  The token's row (and the speaker tail) is read once, the
// pitch / energy projection add is evaluated once (reference operation order, no contraction across the adds,
// forward_tacotron.py:214-220), and the result is stored to the token's frames m in [cum[t-1], cum[t]) -- no binary
// search, no per-frame re-evaluation.  The per-channel projection coefficients are staged once per (persistent) block
// in shared memory as [8][C] = wp0, wp1, wp2, bp, we0, we1, we2, be.  Blocks past the token range zero-fill the tail
// rows m >= lens[b] of one sentence each (pad_sequence(padding_value=0.), common_layers.py:23).
constexpr int LR_MAX_GROUPS = 8;          // float4 channel groups per lane: C <= 1024

__global__ void __launch_bounds__(256)
lr_gather_kernel(const GatherArgs g, int token_blocks) {
  extern __shared__ __align__(16) float coef[];
  const ftb_lr_gather_args& a = g.a;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int C = a.c1 + a.c2;
  if (static_cast<int>(blockIdx.x) >= token_blocks) {
    // ---- zero tail of sentence b
    const int b = blockIdx.x - token_blocks;
    const int len = min(a.lens[b], a.M);
    for (int m = len + warp; m < a.M; m += 8) {
      const long long o0 = (static_cast<long long>(b) * a.out_stride + m) * a.out_ld;
      if (a.round_tf32 == 2) {
        __half* o = reinterpret_cast<__half*>(a.out) + o0;
        for (int c = lane * 4; c < C; c += 128) *reinterpret_cast<uint2*>(o + c) = make_uint2(0u, 0u);
      } else {
        float* o = a.out + o0;
        for (int c = lane * 4; c < C; c += 128) *reinterpret_cast<float4*>(o + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (a.idx && lane == 0) a.idx[static_cast<long long>(b) * a.M + m] = -1;
    }
    return;
  }
  const bool proj = (a.pitch != nullptr);
  if (proj) {
    for (int i = threadIdx.x; i < C; i += 256) {
      coef[0 * C + i] = a.wp[i * 3 + 0]; coef[1 * C + i] = a.wp[i * 3 + 1]; coef[2 * C + i] = a.wp[i * 3 + 2];
      coef[3 * C + i] = a.bp[i];
      coef[4 * C + i] = a.we[i * 3 + 0]; coef[5 * C + i] = a.we[i * 3 + 1]; coef[6 * C + i] = a.we[i * 3 + 2];
      coef[7 * C + i] = a.be[i];
    }
    __syncthreads();
  }
  const long long ntok = static_cast<long long>(a.batch) * a.T;
  for (long long tokid = blockIdx.x * 8ll + warp; tokid < ntok; tokid += static_cast<long long>(token_blocks) * 8) {
    const int b = static_cast<int>(tokid / a.T);
    const int t = static_cast<int>(tokid - static_cast<long long>(b) * a.T);
    const int32_t* cum = a.cum + static_cast<long long>(b) * a.T;
    const int m0 = t > 0 ? cum[t - 1] : 0;
    const int m1 = min(cum[t], a.M);
    if (m1 <= m0) continue;                            // duration 0 (or beyond M): the token has no frame
    const float* xs = a.x + (static_cast<long long>(b) * a.x_stride + t) * a.x_ld;
    float pm = 0.f, p0 = 0.f, pp = 0.f, em = 0.f, e0 = 0.f, ep = 0.f;
    if (proj) {
      const float* pr = a.pitch + static_cast<long long>(b) * a.T;
      const float* er = a.energy + static_cast<long long>(b) * a.T;
      p0 = pr[t]; e0 = er[t];
      if (t > 0) { pm = pr[t - 1]; em = er[t - 1]; }
      if (t + 1 < a.T) { pp = pr[t + 1]; ep = er[t + 1]; }
    }
    float4 v[LR_MAX_GROUPS];
#pragma unroll
    for (int q = 0; q < LR_MAX_GROUPS; ++q) {
      const int c = lane * 4 + q * 128;
      if (c < C) {
        if (c < a.c1) v[q] = *reinterpret_cast<const float4*>(xs + c);
        else v[q] = *reinterpret_cast<const float4*>(a.extra + static_cast<long long>(b) * a.c2 + (c - a.c1));
        if (proj) {
          const float4 w0 = *reinterpret_cast<const float4*>(coef + 0 * C + c);
          const float4 w1 = *reinterpret_cast<const float4*>(coef + 1 * C + c);
          const float4 w2 = *reinterpret_cast<const float4*>(coef + 2 * C + c);
          const float4 wb = *reinterpret_cast<const float4*>(coef + 3 * C + c);
          const float4 u0 = *reinterpret_cast<const float4*>(coef + 4 * C + c);
          const float4 u1 = *reinterpret_cast<const float4*>(coef + 5 * C + c);
          const float4 u2 = *reinterpret_cast<const float4*>(coef + 6 * C + c);
          const float4 ub = *reinterpret_cast<const float4*>(coef + 7 * C + c);
          // Conv1d(1->C, k=3, pad=1) + bias, scaled add (forward_tacotron.py:214-220)
#define FTB_LR_PROJ(X)                                                                                               \
          {                                                                                                          \
            const float pc = __fadd_rn(__fmaf_rn(w2.X, pp, __fmaf_rn(w1.X, p0, __fmul_rn(w0.X, pm))), wb.X);         \
            const float ec = __fadd_rn(__fmaf_rn(u2.X, ep, __fmaf_rn(u1.X, e0, __fmul_rn(u0.X, em))), ub.X);         \
            v[q].X = __fadd_rn(__fadd_rn(v[q].X, __fmul_rn(pc, a.ps)), __fmul_rn(ec, a.es));                         \
          }
          FTB_LR_PROJ(x) FTB_LR_PROJ(y) FTB_LR_PROJ(z) FTB_LR_PROJ(w)
#undef FTB_LR_PROJ
        }
        if (a.round_tf32 == 1) {
          v[q].x = round_tf32(v[q].x); v[q].y = round_tf32(v[q].y); v[q].z = round_tf32(v[q].z); v[q].w = round_tf32(v[q].w);
        }
      }
    }
    for (int m = m0; m < m1; ++m) {
      const long long o0 = (static_cast<long long>(b) * a.out_stride + m) * a.out_ld;
      if (a.round_tf32 == 2) {                       // fp16 operand copy for a FTB_F_A_F16 consumer (out_ld in halves)
        __half* o = reinterpret_cast<__half*>(a.out) + o0;
#pragma unroll
        for (int q = 0; q < LR_MAX_GROUPS; ++q) {
          const int c = lane * 4 + q * 128;
          if (c < C) {
            const __half2 lo = __floats2half2_rn(v[q].x, v[q].y), hi = __floats2half2_rn(v[q].z, v[q].w);
            *reinterpret_cast<uint2*>(o + c) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo),
                                                          *reinterpret_cast<const uint32_t*>(&hi));
          }
        }
      } else {
        float* o = a.out + o0;
#pragma unroll
        for (int q = 0; q < LR_MAX_GROUPS; ++q) {
          const int c = lane * 4 + q * 128;
          if (c < C) *reinterpret_cast<float4*>(o + c) = v[q];
        }
      }
      if (a.idx && lane == 0) a.idx[static_cast<long long>(b) * a.M + m] = t;
    }
  }
}

// Frame -> phoneme map without the copy: one thread per frame, binary search in the sentence's inclusive cumsum.
__global__ void __launch_bounds__(256)
lr_index_kernel(const int32_t* __restrict__ cum, const int32_t* __restrict__ lens, int batch, int T, int M,
                int row_stride, int pad_row, int32_t* __restrict__ idx) {
  const long long i = blockIdx.x * 256ll + threadIdx.x;
  if (i >= static_cast<long long>(batch) * M) return;
  const int b = static_cast<int>(i / M), m = static_cast<int>(i - static_cast<long long>(b) * M);
  int t = pad_row;
  if (m < min(lens[b], M)) {
    const int32_t* c = cum + static_cast<long long>(b) * T;
    int lo = 0, hi = T - 1;                          // first t with cum[t] > m (exists: m < lens[b] = cum[T-1])
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (c[mid] > m) hi = mid; else lo = mid + 1;
    }
    t = b * row_stride + lo;
  }
  idx[i] = t;
}

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_lr_index(ftb_handle h, const int32_t* cum, const int32_t* lens, int batch, int T, int M,
                            int row_stride, int pad_row, int32_t* idx, void* stream) {
  if (!h || !cum || !lens || !idx || batch < 1 || T < 1 || M < 0 || row_stride < 0) return FTB_E_ARG;
  if (M == 0) return 0;
  const long long n = static_cast<long long>(batch) * M;
  lr_index_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      cum, lens, batch, T, M, row_stride, pad_row, idx);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_lr_scan(ftb_handle h, float* dur, int batch, int T, int apply_fallback, int32_t* cum,
                           int32_t* lens, int32_t* meta, void* stream) {
  if (!h || !dur || !cum || !lens || !meta || batch < 1 || T < 1) return FTB_E_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t ce = cudaMemsetAsync(meta, 0, 4 * sizeof(int32_t), st);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  lr_scan_kernel<<<batch, SCAN_THREADS, 0, st>>>(dur, batch, T, apply_fallback, cum, lens, meta, h->err_word);
  const long long n = static_cast<long long>(batch) * T;
  int blocks = static_cast<int>((n + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  lr_finalize_kernel<<<blocks, 256, 0, st>>>(dur, n, meta);
  h->launches += 2;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_lr_gather(ftb_handle h, const ftb_lr_gather_args* args, void* stream) {
  if (!h || !args || !args->x || !args->cum || !args->lens || !args->out) return FTB_E_ARG;
  const ftb_lr_gather_args& a = *args;
  if (a.batch < 1 || a.T < 1 || a.M < 0) return FTB_E_ARG;
  if ((a.c1 % 4) || (a.c2 % 4) || (a.x_ld % 4) || (a.out_ld % 4)) return FTB_E_ALIGN;
  if ((reinterpret_cast<uintptr_t>(a.x) | reinterpret_cast<uintptr_t>(a.out)) & 15u) return FTB_E_ALIGN;
  if (a.c2 > 0 && (!a.extra || (reinterpret_cast<uintptr_t>(a.extra) & 15u))) return FTB_E_ARG;
  if (a.pitch && (!a.energy || !a.wp || !a.bp || !a.we || !a.be)) return FTB_E_ARG;
  if (a.M == 0) return 0;
  if (a.c1 + a.c2 > 128 * LR_MAX_GROUPS) return FTB_E_SHAPE;
  GatherArgs g{a};
  const long long ntok = static_cast<long long>(a.batch) * a.T;
  long long tb = (ntok + 7) / 8;
  const long long cap = 4ll * h->num_sms;            // persistent: the coefficient table is staged once per block
  if (tb > cap) tb = cap;
  const int token_blocks = static_cast<int>(tb);
  const size_t smem = a.pitch ? static_cast<size_t>(8) * (a.c1 + a.c2) * sizeof(float) : 0;
  lr_gather_kernel<<<token_blocks + a.batch, 256, smem, static_cast<cudaStream_t>(stream)>>>(g, token_blocks);
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------
// ftb_lr_regulate_static: scan + gather with NO allocation anywhere on the call path.
//
// The scan's intermediates (cum [batch*T], lens [batch], meta [4]) normally live in caller-allocated
// scratch tensors (ftb_lr_scan / ftb_lr_gather).  Here they live in ONE file-scope device array that
// every GPU loading the library holds until exit (FTB_LR_STATIC_BYTES, see include/ftb200.h).  All
// calls on a device share it: it keeps the previous call's contents (meta is re-zeroed per call, cum and
// lens are fully overwritten for the [batch, T] in use), and two calls that overlap on the GPU race.
//
// `g_lr_static` in HOST code is only a host-side stand-in for the symbol, never a device address: the
// device address is taken once per device with cudaGetSymbolAddress, with that device current.
// ------------------------------------------------------------------------------------------------
namespace ftb {

constexpr int LR_STATIC_WORDS = FTB_LR_STATIC_MAX_TOKENS + FTB_LR_STATIC_MAX_BATCH + 64;
static_assert(sizeof(int32_t) * LR_STATIC_WORDS == FTB_LR_STATIC_BYTES, "ftb200.h states the footprint");
static __device__ __align__(256) int32_t g_lr_static[LR_STATIC_WORDS];

constexpr int LR_STATIC_MAX_DEVICES = 64;
static int32_t* g_lr_static_addr[LR_STATIC_MAX_DEVICES];      // device address of g_lr_static per device (host table)
static std::mutex g_lr_static_mu;

// Device address of the scratch on `device`, which must be the current device.
static int lr_static_scratch(int device, int32_t** out) {
  if (device < 0 || device >= LR_STATIC_MAX_DEVICES) return FTB_E_ARG;
  std::lock_guard<std::mutex> lock(g_lr_static_mu);
  if (!g_lr_static_addr[device]) {
    int32_t* p = nullptr;
    const cudaError_t ce = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_lr_static);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    if (!p) return FTB_E_ARG;
    g_lr_static_addr[device] = p;
  }
  *out = g_lr_static_addr[device];
  return 0;
}

struct CurrentDevice {                  // make the handle's device current for the call, restore afterwards
  int prev = -1;
  bool switched = false;
  cudaError_t err = cudaSuccess;
  explicit CurrentDevice(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) { err = cudaSetDevice(device); switched = (err == cudaSuccess); }
  }
  ~CurrentDevice() { if (switched) cudaSetDevice(prev); }
};

}  // namespace ftb

extern "C" int ftb_lr_regulate_static(ftb_handle h, const float* x, int x_stride, int x_ld, int c, float* dur,
                                      int batch, int T, int apply_fallback, int M, float* out, int out_stride,
                                      int out_ld, int32_t* idx, int32_t* lens_out, int32_t* meta_out, void* stream) {
  if (!h || !x || !dur || !out || batch < 1 || T < 1 || M < 0 || c < 1) return FTB_E_ARG;
  if (batch > FTB_LR_STATIC_MAX_BATCH || static_cast<long long>(batch) * T > FTB_LR_STATIC_MAX_TOKENS)
    return FTB_E_LIMIT;
  if ((c % 4) || (x_ld % 4) || (out_ld % 4)) return FTB_E_ALIGN;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) return FTB_E_ALIGN;
  if (c > 128 * LR_MAX_GROUPS || c > x_ld || c > out_ld || M > out_stride || T > x_stride) return FTB_E_SHAPE;
  CurrentDevice scope(h->device);
  if (scope.err != cudaSuccess) return static_cast<int>(scope.err);
  int32_t* scratch = nullptr;
  const int rc = lr_static_scratch(h->device, &scratch);
  if (rc) return rc;
  int32_t* cum = scratch;                                                  // [batch*T]
  int32_t* lens = scratch + FTB_LR_STATIC_MAX_TOKENS;                       // [batch]
  int32_t* meta = scratch + FTB_LR_STATIC_MAX_TOKENS + FTB_LR_STATIC_MAX_BATCH;   // [4], must start at zero
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t ce = cudaMemsetAsync(meta, 0, 4 * sizeof(int32_t), st);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  lr_scan_kernel<<<batch, SCAN_THREADS, 0, st>>>(dur, batch, T, apply_fallback, cum, lens, meta, h->err_word);
  const long long n = static_cast<long long>(batch) * T;
  int blocks = static_cast<int>((n + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  lr_finalize_kernel<<<blocks, 256, 0, st>>>(dur, n, meta);
  h->launches += 2;
  if (M > 0) {
    GatherArgs g{};
    g.a.x = x; g.a.x_stride = x_stride; g.a.x_ld = x_ld; g.a.c1 = c;
    g.a.cum = cum; g.a.lens = lens; g.a.batch = batch; g.a.T = T; g.a.M = M;
    g.a.out = out; g.a.out_stride = out_stride; g.a.out_ld = out_ld; g.a.idx = idx; g.a.round_tf32 = 0;
    long long tb = (n + 7) / 8;
    const long long cap = 4ll * h->num_sms;
    if (tb > cap) tb = cap;
    const int token_blocks = static_cast<int>(tb);
    lr_gather_kernel<<<token_blocks + batch, 256, 0, st>>>(g, token_blocks);
    h->launches += 1;
  }
  ce = cudaGetLastError();
  if (ce != cudaSuccess) return static_cast<int>(ce);
  if (lens_out) {
    ce = cudaMemcpyAsync(lens_out, lens, sizeof(int32_t) * batch, cudaMemcpyDeviceToDevice, st);
    if (ce != cudaSuccess) return static_cast<int>(ce);
  }
  if (meta_out) {
    ce = cudaMemcpyAsync(meta_out, meta, sizeof(int32_t) * 4, cudaMemcpyDeviceToDevice, st);
    if (ce != cudaSuccess) return static_cast<int>(ce);
  }
  return 0;
}
