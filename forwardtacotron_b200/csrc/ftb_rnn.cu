// K3/K4: persistent bidirectional GRU / LSTM recurrence for sm_100a.
//
// One cooperative launch runs every (problem, direction, batch-group, unit-slice) CTA of a
// plan.  A CTA owns U hidden units (all gates: G*U rows of W_hh) for NB batch columns and
// keeps its W_hh rows RESIDENT IN REGISTERS as tf32 mma.sync A-fragments for all time steps
// (LSTM H=512: 64 rows x 512 = 128 KB per CTA, half the register file).  Per step:
//     h_{t-1} (published through the layer's own output tensor, L2-resident)
//       -> B-fragments by 128-bit ld.global.cg       [K is permuted so 8 consecutive k feed 4 MMAs]
//       -> mma.sync.m16n8k8 tf32, fp32 accumulate, K split over KS warps
//       -> partial sums through padded smem -> gate math in fp32 (c, h stay in registers)
//       -> h_t stored (tf32-rounded: it is the next step's MMA operand and the next GEMM's A)
//       -> ONE barrier per step among the `slices` CTAs that share (direction, batch group):
//          release-add on a global counter, acquire-spin by one thread.
// The tcgen05 path is deliberately not used here: the per-step GEMM is [G*U x H] x [H x NB]
// with NB <= 32, latency- not throughput-bound, and the operands that change every step (h)
// would have to be re-staged into swizzled smem for a UMMA descriptor; register-resident
// weights + mma.sync keep the critical path to L2-load -> HMMA -> smem reduce.
#include <cstdio>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int RNN_THREADS = 512;
constexpr int RNN_NB = 32;               // batch columns per CTA
constexpr int RNN_U = 16;                // hidden units per CTA

__device__ __forceinline__ void mma_tf32_16x8x8(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

template <int G, int H, int KS>
__device__ void rnn_cta(const RnnProblemDev& p, int local_cta, unsigned* counters, float* smem_p) {
  constexpr int U = RNN_U, NB = RNN_NB;
  constexpr int RC = G * U;              // W_hh rows owned by this CTA
  constexpr int RS = RC / 16;            // 16-row MMA strips
  constexpr int WARPS = RS * KS;         // MMA warps (<= 16)
  constexpr int KW = H / KS;             // k extent per warp (multiple of 32)
  constexpr int KGROUPS = KW / 32;       // 4 k-steps (32 k) per group: two float4 of h per n-tile
  constexpr int NT = NB / 8;
  constexpr int S = RC + 4;              // smem row stride of the partial-sum tile [KS][NB][S]
  constexpr int PAIRS = U * NB;
  constexpr int PPT = (PAIRS + RNN_THREADS - 1) / RNN_THREADS;
  static_assert(RC % 16 == 0 && KW % 32 == 0 && WARPS * 32 <= RNN_THREADS, "bad recurrence tiling");

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slice = local_cta % p.slices;
  const int bg = (local_cta / p.slices) % p.bgroups;
  const int dir = local_cta / (p.slices * p.bgroups);
  const int u0 = slice * U;
  const int b0 = p.b_begin + bg * NB;
  const int nbv = min(NB, p.b_end - b0);            // valid batch columns in this CTA
  const int ntv = (nbv + 7) >> 3;                   // n-tiles that hold any valid column
  unsigned* counter = counters + p.group_base + dir * p.bgroups + bg;
  const int L = p.seq_len;
  const int GH = G * H;
  const float* whh = p.w_hh + static_cast<size_t>(dir) * GH * H;
  const float* bhh = p.b_hh + dir * GH;
  float* out_dir = p.out + p.out_col0 + dir * H;    // column base of this direction's h
  const float* xp_dir = p.xproj + dir * GH;

  // ---- W_hh fragments -> registers (once) -------------------------------------------------
  const int rs = warp / KS, kq = warp % KS;
  const int q = lane & 3, fr = lane >> 2;
  uint32_t wfrag[KGROUPS * 4][4];
  if (warp < WARPS) {
    const int lr_lo = rs * 16 + fr, lr_hi = lr_lo + 8;
    const float* w_lo = whh + static_cast<size_t>((lr_lo / U) * H + u0 + (lr_lo % U)) * H;
    const float* w_hi = whh + static_cast<size_t>((lr_hi / U) * H + u0 + (lr_hi % U)) * H;
#pragma unroll
    for (int kg = 0; kg < KGROUPS; ++kg) {
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int k = kq * KW + kg * 32 + 8 * q + 2 * s;   // slot q -> k, slot q+4 -> k+1
        wfrag[kg * 4 + s][0] = __float_as_uint(w_lo[k]);
        wfrag[kg * 4 + s][1] = __float_as_uint(w_hi[k]);
        wfrag[kg * 4 + s][2] = __float_as_uint(w_lo[k + 1]);
        wfrag[kg * 4 + s][3] = __float_as_uint(w_hi[k + 1]);
      }
    }
  }

  // ---- per-thread (unit, batch) pairs: state + biases ------------------------------------
  float c_state[PPT], h_state[PPT], bias[PPT][G];
  int pu[PPT], pb[PPT];
  bool pv[PPT];
#pragma unroll
  for (int j = 0; j < PPT; ++j) {
    const int pi = tid + j * RNN_THREADS;
    pu[j] = pi % U;
    pb[j] = pi / U;
    pv[j] = (pi < PAIRS) && (pb[j] < nbv);
    c_state[j] = 0.0f;
    h_state[j] = 0.0f;
#pragma unroll
    for (int g = 0; g < G; ++g) bias[j][g] = pv[j] ? bhh[g * H + u0 + pu[j]] : 0.0f;
  }

  for (int step = 0; step < L; ++step) {
    const int t = dir ? (L - 1 - step) : step;
    const int tprev = dir ? (t + 1) : (t - 1);

    // x-projection for this step: independent of h, issued before the barrier wait
    float xg[PPT][G];
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
      if (pv[j]) {
        const float* xr = xp_dir + (static_cast<size_t>(b0 + pb[j]) * p.seq_stride + t) * (2 * GH) + u0 + pu[j];
#pragma unroll
        for (int g = 0; g < G; ++g) xg[j][g] = __ldg(xr + g * H);
      }
    }

    if (step > 0) {
      // ---- wait until every slice of this (dir, batch group) has published h_{t-1}
      if (tid == 0) {
        const unsigned target = static_cast<unsigned>(step) * p.slices;
        unsigned spins = 0;
        while (ld_acquire(counter) < target) {
          if (++spins > (1u << 24)) {
            printf("ftb: recurrence barrier timed out (block %d step %d)\n", blockIdx.x, step);
            __trap();
          }
        }
      }
      __syncthreads();

      // ---- P = W_hh[slice rows] * h_{t-1}^T   (tensor cores, K split over KS warps)
      if (warp < WARPS) {
        float acc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.0f;
#pragma unroll
        for (int kg = 0; kg < KGROUPS; ++kg) {
          float4 hv[NT][2];
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const int bl = nt * 8 + fr;
            if (nt < ntv && bl < nbv) {
              const float* hp = out_dir + (static_cast<size_t>(b0 + bl) * p.seq_stride + tprev) * p.out_ld +
                                kq * KW + kg * 32 + 8 * q;
              hv[nt][0] = ld_cg_f4(hp);
              hv[nt][1] = ld_cg_f4(hp + 4);
            } else {
              hv[nt][0] = hv[nt][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            if (nt < ntv) {
              const float* h8 = reinterpret_cast<const float*>(&hv[nt][0]);
#pragma unroll
              for (int s = 0; s < 4; ++s)
                mma_tf32_16x8x8(acc[nt], wfrag[kg * 4 + s], __float_as_uint(h8[2 * s]), __float_as_uint(h8[2 * s + 1]));
            }
          }
        }
        float* pw = smem_p + static_cast<size_t>(kq) * NB * S;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          if (nt < ntv) {
            const int col = nt * 8 + 2 * q, row = rs * 16 + fr;
            pw[col * S + row] = acc[nt][0];
            pw[(col + 1) * S + row] = acc[nt][1];
            pw[col * S + row + 8] = acc[nt][2];
            pw[(col + 1) * S + row + 8] = acc[nt][3];
          }
        }
      }
      __syncthreads();
    }

    // ---- gate math (fp32), state update, publish h_t
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
      if (pv[j]) {
        float hh[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
          float s = bias[j][g];
          if (step > 0) {
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) s += smem_p[(static_cast<size_t>(ks) * NB + pb[j]) * S + g * U + pu[j]];
          }
          hh[g] = s;
        }
        float hnew;
        if (G == 4) {
          const float ig = sigm(xg[j][0] + hh[0]);
          const float fg = sigm(xg[j][1] + hh[1]);
          const float gg = tanhf(xg[j][2] + hh[2]);
          const float og = sigm(xg[j][3] + hh[3]);
          c_state[j] = fg * c_state[j] + ig * gg;
          hnew = og * tanhf(c_state[j]);
        } else {
          const float r = sigm(xg[j][0] + hh[0]);
          const float z = sigm(xg[j][1] + hh[1]);
          const float n = tanhf(xg[j][2] + r * hh[2]);
          hnew = (1.0f - z) * n + z * h_state[j];
        }
        h_state[j] = hnew;
        out_dir[(static_cast<size_t>(b0 + pb[j]) * p.seq_stride + t) * p.out_ld + u0 + pu[j]] = round_tf32(hnew);
      }
    }
    __syncthreads();                      // all h_t stores issued; smem_p free for the next step
    if (tid == 0) {
      __threadfence();
      atomicAdd(counter, 1u);
    }
  }
}

__global__ void __launch_bounds__(RNN_THREADS, 1)
rnn_persistent_kernel(const RnnProblemDev* __restrict__ probs, int nprob, unsigned* counters) {
  extern __shared__ float rnn_smem[];
  int pi = 0;
  while (pi + 1 < nprob && static_cast<int>(blockIdx.x) >= probs[pi + 1].cta_base) ++pi;
  const RnnProblemDev p = probs[pi];
  const int local = blockIdx.x - p.cta_base;
  if (local >= p.ctas) return;
  if (p.cell == FTB_CELL_LSTM) {
    if (p.hidden == 512) rnn_cta<4, 512, 4>(p, local, counters, rnn_smem);
    else if (p.hidden == 256) rnn_cta<4, 256, 4>(p, local, counters, rnn_smem);
    else if (p.hidden == 128) rnn_cta<4, 128, 4>(p, local, counters, rnn_smem);
    else rnn_cta<4, 64, 2>(p, local, counters, rnn_smem);
  } else {
    if (p.hidden == 512) rnn_cta<3, 512, 4>(p, local, counters, rnn_smem);
    else if (p.hidden == 256) rnn_cta<3, 256, 4>(p, local, counters, rnn_smem);
    else if (p.hidden == 128) rnn_cta<3, 128, 4>(p, local, counters, rnn_smem);
    else rnn_cta<3, 64, 2>(p, local, counters, rnn_smem);
  }
}

constexpr int RNN_SMEM_BYTES = 4 * RNN_NB * (4 * RNN_U + 4) * 4;   // KS=4, G=4 worst case

// ------------------------------------------------------------------------------------------
int rnn_plan_build(const ftb_rnn_problem* problems, int nproblems, int num_sms, RnnPlan* plan) {
  if (!problems || nproblems < 1 || nproblems > FTB_MAX_RNN_PROBLEMS) return FTB_E_ARG;
  std::vector<RnnProblemDev> parts;
  std::vector<RnnLaunch> launches;
  RnnLaunch cur;
  int ngroups = 0;
  for (int i = 0; i < nproblems; ++i) {
    const ftb_rnn_problem& q = problems[i];
    if (!q.xproj || !q.w_hh || !q.b_hh || !q.out) return FTB_E_ARG;
    if (q.cell != FTB_CELL_GRU && q.cell != FTB_CELL_LSTM) return FTB_E_ARG;
    if (q.hidden != 64 && q.hidden != 128 && q.hidden != 256 && q.hidden != 512) return FTB_E_SHAPE;
    if (q.batch < 1 || q.seq_len < 1 || q.seq_stride < q.seq_len) return FTB_E_ARG;
    if ((q.out_ld % 4) != 0 || (q.out_col0 % 4) != 0 || (reinterpret_cast<uintptr_t>(q.out) & 15u)) return FTB_E_ALIGN;
    const int slices = q.hidden / RNN_U;
    const int per_bgroup = 2 * slices;                 // CTAs per batch group (both directions)
    if (per_bgroup > num_sms) return FTB_E_RESIDENT;
    int b = 0;
    while (b < q.batch) {
      int room = (num_sms - cur.ctas) / per_bgroup;    // batch groups that still fit in this launch
      if (room == 0) {
        launches.push_back(cur);
        cur = RnnLaunch();
        cur.first = static_cast<int>(parts.size());
        room = num_sms / per_bgroup;
      }
      const int want = (q.batch - b + RNN_NB - 1) / RNN_NB;
      const int take = want < room ? want : room;
      RnnProblemDev d{};
      d.xproj = q.xproj; d.w_hh = q.w_hh; d.b_hh = q.b_hh; d.out = q.out;
      d.cell = q.cell; d.hidden = q.hidden; d.batch = q.batch; d.seq_len = q.seq_len;
      d.seq_stride = q.seq_stride; d.out_ld = q.out_ld; d.out_col0 = q.out_col0;
      d.gates = q.cell == FTB_CELL_LSTM ? 4 : 3;
      d.units_per_cta = RNN_U; d.slices = slices; d.batch_per_cta = RNN_NB;
      d.bgroups = take;
      d.b_begin = b;
      d.b_end = (b + take * RNN_NB < q.batch) ? b + take * RNN_NB : q.batch;
      d.cta_base = cur.ctas;
      d.ctas = 2 * take * slices;
      d.group_base = ngroups;
      ngroups += 2 * take;
      cur.ctas += d.ctas;
      cur.count += 1;
      parts.push_back(d);
      b = d.b_end;
    }
  }
  if (cur.count) launches.push_back(cur);
  cudaError_t ce = cudaMalloc(&plan->dev, sizeof(RnnProblemDev) * parts.size());
  if (ce != cudaSuccess) return static_cast<int>(ce);
  ce = cudaMemcpy(plan->dev, parts.data(), sizeof(RnnProblemDev) * parts.size(), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMalloc(&plan->counters, sizeof(unsigned) * ngroups);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  plan->ngroups = ngroups;
  plan->launches = launches;
  return 0;
}

int rnn_plan_launch(RnnPlan& plan, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(rnn_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          RNN_SMEM_BYTES);
    if (ce != cudaSuccess) return static_cast<int>(ce);
    attr_set = true;
  }
  cudaError_t ce = cudaMemsetAsync(plan.counters, 0, sizeof(unsigned) * plan.ngroups, stream);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  for (const RnnLaunch& l : plan.launches) {
    const RnnProblemDev* base = plan.dev + l.first;
    int count = l.count;
    unsigned* counters = plan.counters;
    void* args[] = {&base, &count, &counters};
    // cooperative launch: the driver guarantees that all CTAs of the grid are co-resident,
    // which the per-step spin barrier relies on.
    ce = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(rnn_persistent_kernel), dim3(l.ctas), dim3(RNN_THREADS),
                                     args, RNN_SMEM_BYTES, stream);
    if (ce != cudaSuccess) return static_cast<int>(ce);
  }
  return 0;
}

}  // namespace ftb
