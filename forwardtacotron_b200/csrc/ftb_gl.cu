// K9: mel -> waveform by fast Griffin-Lim on the GPU (reference: DSP.griffinlim, utils/dsp.py:80-94, which
// calls librosa.core.griffinlim -- a third-party dependency that is absent from /root/reference and from this image; the
// algorithm restated here is the published one: Perraudin et al. 2013, "A fast Griffin-Lim algorithm", as librosa
// implements it: momentum 0.99, Hann window, centred frames).
//
//   stft   : one block per (sentence, frame): window * signal segment -> 1024-point FFT in shared memory -> bins 0..n_fft/2
//   istft  : one block per (sentence, frame): Hermitian spectrum -> inverse FFT -> * window -> frame buffer; a second
//            kernel overlap-adds the (<= win/hop) frames that cover each sample and divides by the window sum-square
//   update : angles = rebuilt - mu/(1+mu) * previous; angles /= |angles| + tiny; next = magnitude * angles
// The FFT is a radix-2 Stockham autosort in shared memory (n_fft = 1024: 10 passes, 512 butterflies each, twiddles from
// a table computed in double precision by the host): these transforms are HBM-bound (one read and one write of the
// frame), the arithmetic is noise.
#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int GL_MAX_N = 2048;

// In-place (ping-pong) Stockham FFT of n = 2^p complex points held in shared memory; `tw` = exp(-2 pi i k / n), k < n/2
// (forward) -- the inverse conjugates.  blockDim.x == n / 2.  Returns the buffer that holds the result.
__device__ __forceinline__ float2* fft_stockham(float2* a, float2* b, int n, int logn, const float2* __restrict__ tw,
                                                bool inverse) {
  const int t = threadIdx.x, half = n >> 1;
  float2* src = a;
  float2* dst = b;
  for (int s = 0; s < logn; ++s) {
    const int m = 1 << s;                       // butterflies per group so far
    const int k = t & (m - 1);                  // index inside the group
    const int j = t >> s;                       // group
    float2 w = tw[k * (half >> s)];
    if (inverse) w.y = -w.y;
    const float2 u = src[t];
    const float2 v0 = src[t + half];
    const float2 v = make_float2(v0.x * w.x - v0.y * w.y, v0.x * w.y + v0.y * w.x);
    dst[(j << (s + 1)) + k] = make_float2(u.x + v.x, u.y + v.y);
    dst[(j << (s + 1)) + k + m] = make_float2(u.x - v.x, u.y - v.y);
    __syncthreads();
    float2* tmp = src; src = dst; dst = tmp;
  }
  return src;
}

// y [batch, y_stride] real signal (valid samples [0, y_len)), centred frames with zero padding: frame f covers samples
// [f*hop - n/2, f*hop + n/2).  spec [batch, frames, n/2+1] complex.
__global__ void gl_stft_kernel(const float* __restrict__ y, int y_len, long long y_stride, const float* __restrict__ window,
                               const float2* __restrict__ tw, int n, int logn, int hop, int frames,
                               float2* __restrict__ spec) {
  __shared__ float2 bufa[GL_MAX_N], bufb[GL_MAX_N];
  const int f = blockIdx.x, b = blockIdx.y, t = threadIdx.x, half = n >> 1;
  const float* yb = y + static_cast<long long>(b) * y_stride;
  for (int i = t; i < n; i += half) {
    const int pos = f * hop - half + i;
    const float v = (pos >= 0 && pos < y_len) ? yb[pos] * window[i] : 0.0f;
    bufa[i] = make_float2(v, 0.0f);
  }
  __syncthreads();
  const float2* r = fft_stockham(bufa, bufb, n, logn, tw, false);
  float2* o = spec + (static_cast<long long>(b) * frames + f) * (half + 1);
  for (int i = t; i <= half; i += half) o[i] = r[i];
}

// spec [batch, frames, n/2+1] -> frame buffer [batch, frames, n] = window * irfft(spec)
__global__ void gl_ifft_kernel(const float2* __restrict__ spec, const float* __restrict__ window,
                               const float2* __restrict__ tw, int n, int logn, int frames, float* __restrict__ fbuf) {
  __shared__ float2 bufa[GL_MAX_N], bufb[GL_MAX_N];
  const int f = blockIdx.x, b = blockIdx.y, t = threadIdx.x, half = n >> 1;
  const float2* s = spec + (static_cast<long long>(b) * frames + f) * (half + 1);
  for (int i = t; i < n; i += half) {
    float2 v;
    if (i <= half) v = s[i];
    else { v = s[n - i]; v.y = -v.y; }          // Hermitian extension of a real signal's spectrum
    if (i == 0 || i == half) v.y = 0.0f;        // irfft ignores the imaginary part of the DC / Nyquist bins
    bufa[i] = v;
  }
  __syncthreads();
  const float2* r = fft_stockham(bufa, bufb, n, logn, tw, true);
  float* o = fbuf + (static_cast<long long>(b) * frames + f) * n;
  const float inv = 1.0f / n;
  for (int i = t; i < n; i += half) o[i] = r[i].x * inv * window[i];
}

// y[t] = sum_f fbuf[f][t + n/2 - f*hop] / wss[t + n/2], trimmed to the centred length hop * (frames - 1)
__global__ void gl_ola_kernel(const float* __restrict__ fbuf, const float* __restrict__ window, int n, int hop, int frames,
                              int out_len, long long y_stride, float* __restrict__ y) {
  const int b = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= out_len) return;
  const int p = t + (n >> 1);                                   // position in the padded signal
  const int f_hi = min(p / hop, frames - 1);
  const int f_lo = max(0, (p - n + hop) / hop);                 // smallest f with p - f*hop < n
  float acc = 0.0f, wss = 0.0f;
  for (int f = f_lo; f <= f_hi; ++f) {
    const int i = p - f * hop;
    if (i >= 0 && i < n) {
      acc += fbuf[(static_cast<long long>(b) * frames + f) * n + i];
      wss += window[i] * window[i];
    }
  }
  y[static_cast<long long>(b) * y_stride + t] = (wss > 1.1754944e-38f) ? acc / wss : acc;
}

// Fast Griffin-Lim phase update, elementwise over [count] complex bins:
//   a = rebuilt - coef * tprev;  a /= (|a| + tiny);  next = mag * a;  tprev <- rebuilt
__global__ void gl_update_kernel(const float* __restrict__ mag, const float2* __restrict__ rebuilt, float2* __restrict__ tprev,
                                 float2* __restrict__ next, long long count, float coef) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < count;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float2 r = rebuilt[i], p = tprev[i];
    float2 a = make_float2(r.x - coef * p.x, r.y - coef * p.y);
    const float inv = 1.0f / (sqrtf(a.x * a.x + a.y * a.y) + 1.1754944e-38f);
    const float m = mag[i];
    next[i] = make_float2(m * a.x * inv, m * a.y * inv);
    tprev[i] = r;
  }
}

static int gl_check_n(int n) { return (n >= 64 && n <= GL_MAX_N && (n & (n - 1)) == 0) ? 0 : FTB_E_SHAPE; }
static int ilog2(int n) { int l = 0; while ((1 << l) < n) ++l; return l; }

}  // namespace ftb

using namespace ftb;

extern "C" int ftb_stft(ftb_handle h, const float* y, int y_len, int64_t y_stride, const float* window,
                        const float* twiddles, int n_fft, int hop, int batch, int frames, float* spec, void* stream) {
  if (!h || !y || !window || !twiddles || !spec || batch < 1 || frames < 1 || hop < 1 || y_len < 1) return FTB_E_ARG;
  if (gl_check_n(n_fft)) return FTB_E_SHAPE;
  gl_stft_kernel<<<dim3(frames, batch), n_fft / 2, 0, static_cast<cudaStream_t>(stream)>>>(
      y, y_len, y_stride, window, reinterpret_cast<const float2*>(twiddles), n_fft, ilog2(n_fft), hop, frames,
      reinterpret_cast<float2*>(spec));
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_istft(ftb_handle h, const float* spec, const float* window, const float* twiddles, int n_fft, int hop,
                         int batch, int frames, float* frame_buf, float* y, int64_t y_stride, void* stream) {
  if (!h || !spec || !window || !twiddles || !frame_buf || !y || batch < 1 || frames < 1 || hop < 1) return FTB_E_ARG;
  if (gl_check_n(n_fft)) return FTB_E_SHAPE;
  const int out_len = hop * (frames - 1);
  if (out_len > y_stride) return FTB_E_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  gl_ifft_kernel<<<dim3(frames, batch), n_fft / 2, 0, st>>>(reinterpret_cast<const float2*>(spec), window,
                                                            reinterpret_cast<const float2*>(twiddles), n_fft,
                                                            ilog2(n_fft), frames, frame_buf);
  if (out_len > 0)
    gl_ola_kernel<<<dim3((out_len + 255) / 256, batch), 256, 0, st>>>(frame_buf, window, n_fft, hop, frames, out_len,
                                                                      y_stride, y);
  h->launches += 2;
  return static_cast<int>(cudaGetLastError());
}

extern "C" int ftb_gl_update(ftb_handle h, const float* mag, const float* rebuilt, float* tprev, float* next,
                             int64_t count, float momentum, void* stream) {
  if (!h || !mag || !rebuilt || !tprev || !next || count < 1 || momentum < 0.0f) return FTB_E_ARG;
  long long blocks = (count + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gl_update_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      mag, reinterpret_cast<const float2*>(rebuilt), reinterpret_cast<float2*>(tprev), reinterpret_cast<float2*>(next),
      count, momentum / (1.0f + momentum));
  h->launches += 1;
  return static_cast<int>(cudaGetLastError());
}
