// K3/K4: persistent bidirectional GRU / LSTM recurrence for sm_100a.
//
// One launch runs every (problem, direction, batch-group, unit-slice) CTA of a plan.  A CTA owns U = 32 hidden
// units (all gates: G*U rows of W_hh) for NB = 16 batch columns and keeps its W_hh rows RESIDENT IN REGISTERS as
// fp16 mma.sync A-fragments for all time steps (LSTM H=512: 128 rows x 512 = 128 KB per CTA, half the register
// file).  The H/32 CTAs of one (direction, batch group) -- a "unit" -- exchange h_t every step; units are
// independent of each other.  Two hand-off protocols, chosen per plan (rnn_plan_build):
//
// (1) CLUSTER hand-off (rnn_cluster_kernel; the GRUs, and the LSTM when <= 7 units are needed).  The CTAs of a unit
//     sit in one thread-block cluster (size = H/32 = 2..16).  After the gate math every thread owns h_t of one
//     (unit, column) pair; three xor-shuffle rounds gather 8 units into a 16-byte chunk and each lane pushes its
//     chunk straight into the staged h tile of a peer CTA's SHARED MEMORY with
//         st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32
//     The consumer waits on a local mbarrier whose phase completes when all H*32 bytes of the step have landed
//     (one local arrive.expect_tx arms the phase).  Tiles and barriers are double-buffered by step parity; a peer
//     can never be two steps ahead because its step t+1 needs this CTA's h_t.  No global-memory exchange tensor,
//     no sentinel pre-fill, no cooperative launch (units never straddle clusters).  The hand-off costs ~450 cycles
//     (L2 path: ~1800), so the CTA runs ONE group of 16 columns (two n-tiles per A fragment, one CTA barrier per
//     step) instead of interleaving two groups.  Problems are grouped by cluster size into concurrent launch
//     "lanes" (side streams), because a launch has one cluster size and large clusters pack badly into the GPCs:
//     this B200 keeps 15 clusters of 8 and 7 clusters of 16 CTAs resident (7 GPCs x 20 SMs + 1 x 8).  The batch-64
//     LSTM needs 8 units of 16 CTAs and therefore stays on protocol (2).
//
// (2) L2 hand-off (rnn_persistent_kernel, cooperative launch; v4: two interleaved batch groups per CTA, A = columns
//     0-7, B = columns 8-15).
//
// Operand precision.  The recurrent product uses fp16 operands with fp32 accumulation: fp16 carries
// the same 11 significant bits as tf32 (|h| <= 1 and |W_hh| ~ 1/sqrt(H) sit well inside its range; below
// 6e-5 the absolute error is < 3e-8), but a legacy-path HMMA.16816 moves twice the MACs of the tf32
// HMMA.1688 per issue slot -- measured here: tf32 mma.sync sustains ~290 MAC/clk/SM on B200 and was the
// bound of the first version of this kernel (profiles/r1b_rnn_phases.txt).  h_t is therefore published
// twice: as fp32 (exactly the fp16 value, hence tf32-exact) into the layer output the next GEMM reads,
// and as fp16 to the other CTAs of the unit (peer shared memory, or the exchange tensor hx[dir][t][b][H]).
//
// Per step of protocol (2):
//     h_{t-1} (fp16, L2-resident) -> loaded ONCE per CTA and group with ld.relaxed.gpu (16 B per thread, issued one
//         phase ahead, staged into smem when it is needed).  hx is pre-filled with 0xFFFF halves (an fp16 NaN no
//         arithmetic here produces), so "h arrived" == "no sentinel left": each thread checks the 16 bytes it
//         loaded and re-fetches only those that still hold the sentinel.  No counter, no __threadfence, no atomic
//         on the critical path.  The K range is split over KS = 2 warp groups that synchronise with named
//         barriers, so each half starts its MMAs as soon as its own bytes have landed.
//       -> B-fragments from smem (one LDS.128 feeds two k-steps: K is permuted inside 32-wide blocks),
//          mma.sync.m16n8k16 f16 -> fp32 accumulators; ALL MMA warps work on the group whose h has landed
//          while the other group's h is still in flight (the L2 hand-off of A hides behind the MMAs of B).
//       -> partial sums through padded smem -> __syncthreads -> gate math in fp32 by the group's 8 warps
//          (c, h stay in registers; x-projection rows arrive through a 4-stage cp.async ring issued three
//          steps ahead -- always the newest cp.async group, so nothing on the chain ever waits for it)
//       -> h_t stores (fp32 to `out`, fp16 pairs to hx).
// Packed-sequence semantics (optional `lens`): a column whose sentence has ended keeps h = c = 0 and writes
// `pad_value` to `out` -- what pack_padded_sequence / pad_packed_sequence give the reference's eval-mode forward.
//
// tcgen05 is deliberately not used here: the per-step GEMM is [G*U x H] x [H x 16]; a UMMA needs W_hh in shared
// memory (256 KB fp32-class for this slice) or TMEM and h re-staged into a swizzled tile every step.  A TS-mode
// (A from TMEM) variant was built and measured: ~90 cycles per M128 N16 K16 instruction, 3000 cycles of MMA per
// step against ~950 here (profiles/r1g_rnn_tcgen05_experiment.txt; source kept in scratch/experiments/).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_fp16.h>

#include "ftb_internal.h"
#include "ftb_ptx.cuh"

namespace ftb {

constexpr int RNN_THREADS = 512;
constexpr int RNN_U = 32;                // hidden units per CTA
constexpr int RNN_NB = RNN_THREADS / RNN_U;   // 16 batch columns per CTA: one (unit, batch) pair per thread
constexpr int RNN_KS = 2;                // K split (warp groups)
constexpr int RNN_XS = 4;                // cp.async stages of the x-projection ring

// D(16x8, f32) += A(16x16, f16, row) * B(16x8, f16, col)
__device__ __forceinline__ void mma_f16_16x8x16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// Coherent (L2) 128-bit load: the polled value IS the payload, so relaxed ordering is enough.
__device__ __forceinline__ uint4 ld_relaxed_u4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(void* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool has_sentinel(uint32_t w) {      // either fp16 half == 0xFFFF
  return ((w & 0xFFFFu) == 0xFFFFu) || ((w >> 16) == 0xFFFFu);
}
__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
  const __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// exp/rcp at ~2 ulp: absolute gate error ~3e-7, far below the operand rounding of the h @ W_hh products
__device__ __forceinline__ float ex2_ftz(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_ftz(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// ex2 / rcp forms (~2 ulp): the gates of the PRECISE recurrences (duration predictors) and of -DFTB_RNN_EXACT_GATES builds
__device__ __forceinline__ float sigm_x(float x) { return rcp_ftz(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_x(float x) { return fmaf(2.0f, rcp_ftz(1.0f + ex2_ftz(-2.8853900817779268f * x)), -1.0f); }
#ifndef FTB_RNN_EXACT_GATES
// One MUFU.TANH per gate (tanh.approx.f32, max relative error 2^-11 -- the precision h_t is rounded to anyway when it
// is handed to the next step as an fp16 MMA operand) instead of ex2 + rcp: a shorter dependent chain in the gate phase.
// Measured against the reference goldens (scratch/err_report.py): mel rel-L2 5.5-7.0e-4 with it, 5.4-6.9e-4 without
// (-DFTB_RNN_EXACT_GATES); LSTM 0.817 -> 0.796 ms, postnet GRU 0.448 -> 0.434 ms.
__device__ __forceinline__ float tanh_fast(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigm(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }
#else
__device__ __forceinline__ float sigm(float x) { return rcp_ftz(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_fast(float x) { return fmaf(2.0f, rcp_ftz(1.0f + ex2_ftz(-2.8853900817779268f * x)), -1.0f); }
#endif

// ---- thread-block-cluster hand-off (distributed shared memory) ---------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// 16-byte asynchronous store into a peer CTA's shared memory; completes 16 tx-bytes on the peer's mbarrier
__device__ __forceinline__ void st_async_v4(uint32_t raddr, const uint4& v, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(rbar) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// PR (precise; cluster hand-off only): split-fp16 recurrent product.  h_t = hi + lo/2048 with hi = fp16(h_t),
// lo = fp16(2048 (h_t - hi)), W_hh likewise; the step computes W_hi hi + (W_lo hi + W_hi lo)/2048 in two fp32 accumulator
// sets (three MMAs per k-step instead of one): ~2^-20 relative operand precision instead of 2^-11.  h_t is handed over as
// two tiles (hi, lo), stored to `out` as the unrounded fp32 value, and the gates use the ex2/rcp forms.  For the DURATION
// predictors, whose output is rounded to integers (forward_tacotron.py:174-177, common_layers.py:18-21).
constexpr float RNN_LO_SCALE = 2048.0f;
template <int G, int H, bool CL = false, bool PR = false>
struct RnnTiling {
  static constexpr int U = RNN_U, NB = RNN_NB, KS = RNN_KS;
  // L2 hand-off: two independent batch groups of 8 columns (one MMA n-tile each), interleaved to hide the L2 round trip.
  // Cluster hand-off: ONE group of 16 columns (two n-tiles that share every A fragment): the hand-off is short enough
  // that halving the number of phases per step wins over hiding it.
  static constexpr int NG = CL ? 1 : 2, NBG = NB / NG;
  static constexpr int RC = G * U;              // W_hh rows owned by this CTA (128 / 96)
  static constexpr int RS = RC / 16;            // 16-row MMA strips (8 / 6)
  static constexpr int WARPS = RS * KS;         // MMA warps (16 / 12)
  static constexpr int KW = H / KS;             // k extent per warp group (multiple of 32)
  static constexpr int KB = KW / 32;            // 32-wide k blocks: one LDS.128 of h feeds 2 k-steps
  static constexpr int S = RC + 4;              // fp32 row stride of a partial-sum tile [NG][KS][NBG][S] (S = 4 mod 16)
  static constexpr int HS = H + 32;             // fp16 row stride of the staged h tile (HS/2 = 16 mod 32 words)
  static constexpr int XSTR = G * U;            // per-batch stride of an x-projection stage (floats)
  static constexpr int GS = RS * 32;            // threads of one K group
  static constexpr int ITEMS = NBG * KW / 8;    // 16-byte items (8 halves) of one K-half of one batch group of h
  static constexpr int IPT = (ITEMS + GS - 1) / GS;
  static constexpr int SMEM_BYTES = NB * HS * 2 + (NG * KS * NBG * S + RNN_XS * NB * XSTR) * 4;
  // cluster hand-off: h tiles [2 step parities][SL slices][NB columns][32 units] fp16 + 2 mbarriers
  static constexpr int SL = H / U;              // unit slices = CTAs of one (direction, batch group) unit
  static constexpr int HCL_TILE_BYTES = 2 * NG * SL * NBG * U * 2;     // one set of tiles (hi); PR adds the lo set behind it
  static constexpr int HCL_BYTES = (PR ? 2 : 1) * HCL_TILE_BYTES;
  static constexpr int TXB = (PR ? 2 : 1) * SL * NBG * U * 2;  // bytes one (group, step) hand-off delivers to every CTA of the unit
  static constexpr int ND = (SL + 7) / 8;       // peers one lane stores to
  static constexpr int SMEM_BYTES_CL = HCL_BYTES + (NG * KS * NBG * S + RNN_XS * NB * XSTR) * 4 + 2 * NG * 8;
  static_assert(RC % 16 == 0 && KW % 32 == 0 && WARPS * 32 <= RNN_THREADS, "bad recurrence tiling");
  static_assert(!PR || CL, "the precise recurrence exists on the cluster hand-off only");
  static_assert(U * NB == RNN_THREADS && (NB * G * U / 4) <= RNN_THREADS, "one (unit, batch) pair / x chunk per thread");
};

// CL = false: h handed between CTAs through L2 (in-band sentinel polling; cooperative launch).
// CL = true : the SL CTAs of one (direction, batch group) unit sit in one thread-block cluster and push h_t
//             straight into each other's shared memory (st.async + mbarrier complete_tx); see the header.
// XI: the x-projection rows are read through ftb_rnn_problem.x_idx (the LengthRegulator folded into this read).  A
// compile-time switch: the same code as a run-time branch cost every recurrence of the step 7 % (a few more live
// registers in a kernel that sits at the 128-register cap).  Instantiated for the problems that use it: LSTM-512.
// ONE (cluster hand-off): the unit holds at most 8 sentences, so only the first MMA n-tile (columns 0-7) is computed.
// Also compile-time, for the same reason; instantiated for the frame-side recurrences (LSTM-512, GRU-256).
template <int G, int H, bool CL, bool PR = false, bool XI = false, bool ONE = false>
__device__ void rnn_cta(const RnnProblemDev& p, int local_cta, unsigned char* smem) {
  using T = RnnTiling<G, H, CL, PR>;
  constexpr int U = T::U, NB = T::NB, KS = T::KS, WARPS = T::WARPS, KW = T::KW, KB = T::KB, NBG = T::NBG;
  constexpr int S = T::S, HS = T::HS, XSTR = T::XSTR, GS = T::GS, ITEMS = T::ITEMS, IPT = T::IPT;
  constexpr int NG = T::NG, SL = T::SL, ND = T::ND;
  constexpr int TILE = NBG * U;                                        // halves of one slice of one group's h tile
  __half* h_s = reinterpret_cast<__half*>(smem);                       // L2: [NB][HS] fp16 (rows 0-7 group A, 8-15 group B)
                                                                       // CL: [2][NG][SL][NBG][U] fp16
  float* p_s = reinterpret_cast<float*>(smem + (CL ? T::HCL_BYTES : NB * HS * 2));   // [NG][KS][NBG][S]
  float* x_s = p_s + T::NG * KS * NBG * S;                             // [XS][NB][XSTR]
  uint64_t* bars = reinterpret_cast<uint64_t*>(x_s + RNN_XS * NB * XSTR);            // CL: [2][NG] "h tile full"

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slice = local_cta % p.slices;
  const int bg = (local_cta / p.slices) % p.bgroups;
  const int dir = local_cta / (p.slices * p.bgroups);
  const int u0 = slice * U;
  const int b0 = p.b_begin + bg * NB;
  const int nbv = min(NB, p.b_end - b0);            // valid batch columns in this CTA
  const int L = p.seq_len;
  const int GH = G * H;
  const float* whh = p.w_hh + static_cast<size_t>(dir) * GH * H;
  const float* bhh = p.b_hh + dir * GH;
  // x-projection element type: fp32, or (x_f16) the fp16 copy a FTB_F_OUT_F16 GEMM wrote
  const bool x16 = p.x_f16 != 0;
  const int xesz = x16 ? 2 : 4;
  // exchange tensor hx[dir][t][b][H] (fp16)
  __half* hx_dir = p.hx + static_cast<size_t>(dir) * L * p.batch * H;

  // ---- W_hh fragments -> registers (once), rounded to fp16 here --------------------------------
  // k order inside a 32-wide block: lane group q owns k = 8q..8q+7; k-step s uses 8q+4s..8q+4s+3.
  const int rs = warp / KS, kq = warp % KS;
  const int q = lane & 3, fr = lane >> 2;
  const bool mma_warp = warp < WARPS;
  uint32_t wfrag[KB * 2][4];
  uint32_t wfrag_lo[PR ? KB * 2 : 1][4];              // PR: fp16(2048 (w - fp16(w)))
  auto pack_lo2 = [](float a, float b) {
    return pack_half2((a - __half2float(__float2half_rn(a))) * RNN_LO_SCALE, (b - __half2float(__float2half_rn(b))) * RNN_LO_SCALE);
  };
  if (mma_warp) {
    const int lr_lo = rs * 16 + fr, lr_hi = lr_lo + 8;
    const float* w_lo = whh + static_cast<size_t>((lr_lo / U) * H + u0 + (lr_lo % U)) * H;
    const float* w_hi = whh + static_cast<size_t>((lr_hi / U) * H + u0 + (lr_hi % U)) * H;
#pragma unroll
    for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const int k = kq * KW + kb * 32 + 8 * q + 4 * s;
        wfrag[kb * 2 + s][0] = pack_half2(w_lo[k], w_lo[k + 1]);
        wfrag[kb * 2 + s][1] = pack_half2(w_hi[k], w_hi[k + 1]);
        wfrag[kb * 2 + s][2] = pack_half2(w_lo[k + 2], w_lo[k + 3]);
        wfrag[kb * 2 + s][3] = pack_half2(w_hi[k + 2], w_hi[k + 3]);
        if (PR) {
          wfrag_lo[kb * 2 + s][0] = pack_lo2(w_lo[k], w_lo[k + 1]);
          wfrag_lo[kb * 2 + s][1] = pack_lo2(w_hi[k], w_hi[k + 1]);
          wfrag_lo[kb * 2 + s][2] = pack_lo2(w_lo[k + 2], w_lo[k + 3]);
          wfrag_lo[kb * 2 + s][3] = pack_lo2(w_hi[k + 2], w_hi[k + 3]);
        }
      }
    }
  }

  // ---- this thread's (unit, batch) pair: state + biases --------------------------------------
  const int pu = tid % U, pb = tid / U;             // a warp = the 32 units of one batch column
  const int pg = pb / NBG, pbl = pb % NBG;          // batch group (warps 0-7: A, 8-15: B) and column inside it
  const bool pv = pb < nbv;
  float c_state = 0.0f, h_state = 0.0f, bias[G];
#pragma unroll
  for (int g = 0; g < G; ++g) bias[g] = pv ? bhh[g * H + u0 + pu] : 0.0f;
  // element offset into `out`; the layer output is stored as fp32 or (out_f16) as the fp16 value itself
  const size_t out_pair = p.out_col0 + dir * H + static_cast<size_t>(b0 + pb) * p.seq_stride * p.out_ld + u0 + pu;
  __half* hx_pair = hx_dir + static_cast<size_t>(b0 + pb) * H + u0 + pu;       // + t * batch * H

  // ---- x-projection ring: 16-byte chunk `tid` of a step = (batch xb, gate xg_, 4 units) --------
  const int xce = 16 / xesz;                        // elements per 16-byte chunk
  const int XCH = U / xce;                          // chunks per (batch, gate)
  const int xb = tid / (G * XCH), xg_ = (tid / XCH) % G, xpart = tid % XCH;
  const bool xv = (tid < NB * G * XCH) && (xb < nbv);
  const char* x_src = reinterpret_cast<const char*>(p.xproj) +
                      (static_cast<size_t>(b0 + xb) * p.seq_stride * (2 * GH) + dir * GH + xg_ * H + u0 + xpart * xce) * xesz;
  char* x_dst = reinterpret_cast<char*>(x_s) + (xb * XSTR + xg_ * U + xpart * xce) * xesz;
  const int x_stage = NB * XSTR * xesz;             // bytes per ring stage
  // running pointers (strength-reduced: the per-step 64-bit index math sat on the latency chain of the gate phase)
  // XI: step t of sentence b reads row x_idx[b, t] (an absolute row) of a per-PHONEME x-projection; the running pointer
  // walks the index row, and the index of the NEXT prefetch is loaded one prefetch ahead (a whole time step of slack), so
  // the dependent address never waits on the chain.  Row offsets are 32-bit products (x-projections below 4 GB).
  const long long x_step = XI ? (dir ? -1ll : 1ll) * static_cast<long long>(sizeof(int))
                              : (dir ? -static_cast<long long>(2 * GH) : static_cast<long long>(2 * GH)) * xesz;
  const char* x_pf = XI ? reinterpret_cast<const char*>(p.x_idx + static_cast<size_t>(b0 + (xv ? xb : 0)) * L + (dir ? L - 1 : 0))
                        : x_src + static_cast<size_t>(dir ? L - 1 : 0) * (2 * GH) * xesz;
  const char* xi_base = reinterpret_cast<const char*>(p.xproj) + static_cast<size_t>(dir * GH + xg_ * H + u0 + xpart * xce) * xesz;
  const unsigned xi_rowbytes = static_cast<unsigned>(2 * GH * xesz);
  unsigned xi_next = (XI && xv) ? static_cast<unsigned>(__ldg(reinterpret_cast<const int*>(x_pf))) : 0u;
  int pf_step = 0;
  auto prefetch_x = [&]() {                         // x rows of step `pf_step` -> ring slot pf_step % XS
    if (xv && pf_step < L)
      cp_async16(x_dst + (pf_step & (RNN_XS - 1)) * x_stage,
                 XI ? xi_base + static_cast<size_t>(xi_next) * xi_rowbytes : x_pf);
    cp_async_commit();
    x_pf += x_step;
    ++pf_step;
    if (XI && xv && pf_step < L) xi_next = static_cast<unsigned>(__ldg(reinterpret_cast<const int*>(x_pf)));
  };
  // W_hh and the biases above are weights; the x-projection is the previous kernel's output
  pdl_wait();
#pragma unroll
  for (int s = 0; s < RNN_XS - 1; ++s) prefetch_x();

  // ---- h staging: this thread's 16-byte items of (group g, its K half), L2 -> smem by cp.async.cg ----
  const int gtid = rs * 32 + lane;
  const long long h_step = dir ? -static_cast<long long>(p.batch) * H : static_cast<long long>(p.batch) * H;
  const __half* hp[IPT];                            // group-A item j at time row t (advanced every step); B = + NBG rows
  __half* hd[IPT];
  unsigned ok_a = 0, ok_b = 0;
#pragma unroll
  for (int j = 0; j < IPT; ++j) {
    const int i = gtid + j * GS;
    const int b = i / (KW / 8), c8 = i % (KW / 8);
    hp[j] = hx_dir + (static_cast<size_t>(dir ? L - 1 : 0) * p.batch + b0 + b) * H + kq * KW + c8 * 8;
    hd[j] = h_s + b * HS + kq * KW + c8 * 8;
    if (mma_warp && i < ITEMS && b < nbv) ok_a |= 1u << j;
    if (mma_warp && i < ITEMS && b + NBG < nbv) ok_b |= 1u << j;
  }
  // issue / settle take the offset (in halves) of the wanted time row relative to hp[]: 0 = row t, -h_step = row t-1.
  // The items travel through REGISTERS (ld.relaxed.gpu at issue time, st.shared at settle time): an LDGSTS per thread
  // cost ~600 cycles of issue time per group on the step-to-step chain (profiles/r1o_rnn_phases.txt); an LDG issues in
  // a few cycles and its latency is covered by the other group's phases.  Only one group is in flight at a time.
  uint4 hv[IPT];
  auto issue_h = [&](int g, long long toff) {
    const unsigned ok = g ? ok_b : ok_a;
#pragma unroll
    for (int j = 0; j < IPT; ++j)
      if (ok & (1u << j)) hv[j] = ld_relaxed_u4(hp[j] + toff + g * (NBG * H));
  };
  // Opportunistic second poll one phase after the issue: an item that came back with the sentinel (its producer was
  // slower than this CTA) is re-requested now, so the blocking re-fetch in settle_h -- which stalls the whole K group at
  // its named barrier for a full L2 round trip -- becomes the exception.
  auto refresh_h = [&](int g, long long toff) {
    const unsigned ok = g ? ok_b : ok_a;
#pragma unroll
    for (int j = 0; j < IPT; ++j)
      if ((ok & (1u << j)) &&
          (has_sentinel(hv[j].x) || has_sentinel(hv[j].y) || has_sentinel(hv[j].z) || has_sentinel(hv[j].w)))
        hv[j] = ld_relaxed_u4(hp[j] + toff + g * (NBG * H));
  };
  // "h arrived" == "no 0xFFFF half left"; re-fetch the items that still hold the sentinel (their producer had not
  // published yet when the load read L2), then place them in the staged tile
  auto settle_h = [&](int g, long long toff, unsigned& retries) {
    const unsigned ok = g ? ok_b : ok_a;
#pragma unroll
    for (int j = 0; j < IPT; ++j) {
      if (ok & (1u << j)) {
        unsigned spins = 0;
        while (has_sentinel(hv[j].x) || has_sentinel(hv[j].y) || has_sentinel(hv[j].z) || has_sentinel(hv[j].w)) {
          hv[j] = ld_relaxed_u4(hp[j] + toff + g * (NBG * H));
          ++retries;
          if (++spins > (1u << 22)) {
            printf("ftb: recurrence wait for h timed out (block %d)\n", blockIdx.x);
            __trap();
          }
        }
        *reinterpret_cast<uint4*>(hd[j] + g * (NBG * HS)) = hv[j];
      }
    }
  };
  // P[g] = W_hh[slice rows, K-half] * h_{t-1}[group g]^T   (tensor cores, fp16 x fp16 -> fp32)
  auto mma_group = [&](int g, int par) {
    if (CL) {
      // both n-tiles (columns 0-7 / 8-15) per A fragment; slice kb of the K range is a [NB][U] tile, so the two
      // LDS.128 of a warp cover 2 x 512 contiguous bytes
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
      float acs[PR ? 4 : 1][4];              // PR: the two cross terms, scaled by 2048
      if (PR) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acs[PR ? i : 0][j] = 0.0f;
      }
      const __half* hrow = h_s + (par * SL + kq * KB) * TILE + fr * U + 8 * q;
      if (ONE) {
        // at most 8 sentences in this unit (a 4-sentence shard of a long article, a single request): the second n-tile
        // (columns 8-15) holds no sentence; its partial sums stay at the zeros they were initialised with and no gate
        // thread reads them, so half of the step's HMMAs and shared-memory operand reads are skipped
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
          const uint4 h0 = *reinterpret_cast<const uint4*>(hrow + kb * TILE);
          mma_f16_16x8x16(acc[0], wfrag[kb * 2 + 0], h0.x, h0.y);
          mma_f16_16x8x16(acc[1], wfrag[kb * 2 + 1], h0.z, h0.w);
        }
        float* pk1 = p_s + static_cast<size_t>(kq) * NB * S;
        const int col1 = 2 * q, row1 = rs * 16 + fr;
        pk1[col1 * S + row1] = acc[0][0] + acc[1][0];
        pk1[(col1 + 1) * S + row1] = acc[0][1] + acc[1][1];
        pk1[col1 * S + row1 + 8] = acc[0][2] + acc[1][2];
        pk1[(col1 + 1) * S + row1 + 8] = acc[0][3] + acc[1][3];
        return;
      }
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        const uint4 h0 = *reinterpret_cast<const uint4*>(hrow + kb * TILE);
        const uint4 h1 = *reinterpret_cast<const uint4*>(hrow + kb * TILE + 8 * U);
        mma_f16_16x8x16(acc[0], wfrag[kb * 2 + 0], h0.x, h0.y);
        mma_f16_16x8x16(acc[2], wfrag[kb * 2 + 0], h1.x, h1.y);
        mma_f16_16x8x16(acc[1], wfrag[kb * 2 + 1], h0.z, h0.w);
        mma_f16_16x8x16(acc[3], wfrag[kb * 2 + 1], h1.z, h1.w);
        if (PR) {
          const uint4 l0 = *reinterpret_cast<const uint4*>(hrow + kb * TILE + T::HCL_TILE_BYTES / 2);
          const uint4 l1 = *reinterpret_cast<const uint4*>(hrow + kb * TILE + 8 * U + T::HCL_TILE_BYTES / 2);
          mma_f16_16x8x16(acs[0], wfrag_lo[PR ? kb * 2 + 0 : 0], h0.x, h0.y);      // W_lo * h_hi
          mma_f16_16x8x16(acs[PR ? 2 : 0], wfrag_lo[PR ? kb * 2 + 0 : 0], h1.x, h1.y);
          mma_f16_16x8x16(acs[PR ? 1 : 0], wfrag_lo[PR ? kb * 2 + 1 : 0], h0.z, h0.w);
          mma_f16_16x8x16(acs[PR ? 3 : 0], wfrag_lo[PR ? kb * 2 + 1 : 0], h1.z, h1.w);
          mma_f16_16x8x16(acs[0], wfrag[kb * 2 + 0], l0.x, l0.y);                  // W_hi * h_lo
          mma_f16_16x8x16(acs[PR ? 2 : 0], wfrag[kb * 2 + 0], l1.x, l1.y);
          mma_f16_16x8x16(acs[PR ? 1 : 0], wfrag[kb * 2 + 1], l0.z, l0.w);
          mma_f16_16x8x16(acs[PR ? 3 : 0], wfrag[kb * 2 + 1], l1.z, l1.w);
        }
      }
      if (PR) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(acs[PR ? i : 0][j], 1.0f / RNN_LO_SCALE, acc[i][j]);
      }
      float* pk = p_s + static_cast<size_t>(kq) * NB * S;
      const int col = 2 * q, row = rs * 16 + fr;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        pk[(8 * nt + col) * S + row] = acc[2 * nt][0] + acc[2 * nt + 1][0];
        pk[(8 * nt + col + 1) * S + row] = acc[2 * nt][1] + acc[2 * nt + 1][1];
        pk[(8 * nt + col) * S + row + 8] = acc[2 * nt][2] + acc[2 * nt + 1][2];
        pk[(8 * nt + col + 1) * S + row + 8] = acc[2 * nt][3] + acc[2 * nt + 1][3];
      }
      return;
    }
    // NCH independent accumulator chains
    constexpr int NCH = 2;                 // 4 chains spill at the 128-register cap: LSTM 0.86 -> 0.97 ms (measured)
    float acc[NCH][4];
#pragma unroll
    for (int i = 0; i < NCH; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    const __half* hrow = h_s + (g * NBG + fr) * HS + kq * KW + 8 * q;
#pragma unroll
    for (int kb = 0; kb < KB; ++kb) {
      const uint4 hq = *reinterpret_cast<const uint4*>(hrow + kb * 32);
      mma_f16_16x8x16(acc[(2 * kb) % NCH], wfrag[kb * 2 + 0], hq.x, hq.y);
      mma_f16_16x8x16(acc[(2 * kb + 1) % NCH], wfrag[kb * 2 + 1], hq.z, hq.w);
    }
    if (NCH == 4) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { acc[0][j] += acc[2][j]; acc[1][j] += acc[3][j]; }
    }
    float* pk = p_s + static_cast<size_t>(g * KS + kq) * NBG * S;
    const int col = 2 * q, row = rs * 16 + fr;
    pk[col * S + row] = acc[0][0] + acc[1][0];
    pk[(col + 1) * S + row] = acc[0][1] + acc[1][1];
    pk[col * S + row + 8] = acc[0][2] + acc[1][2];
    pk[(col + 1) * S + row + 8] = acc[0][3] + acc[1][3];
  };
  // cluster hand-off: this lane's peers (CTAs of the same unit = cluster ranks [rank - slice, rank - slice + SL))
  uint32_t rdata[ND], rbar[ND], rvalid = 0;
  if (CL) {
    const uint32_t ubase = cluster_ctarank() - slice;
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      const int peer = (lane & 7) + 8 * d;
      rdata[d] = rbar[d] = 0;
      if (peer < SL) {
        rvalid |= 1u << d;
        // tile [par][pg][slice][pbl][unit octet lane/8]
        rdata[d] = mapa_shared(smem_u32(h_s), ubase + peer) + ((pg * SL + slice) * TILE + pbl * U + (lane >> 3) * 8) * 2;
        rbar[d] = mapa_shared(smem_u32(bars), ubase + peer) + pg * 8;
      }
    }
  }
  // gate math (fp32), state update and publication of h_t for this thread's pair (threads of group g only).
  // The partial-sum tiles start zeroed, so step 0 (h_{-1} = 0, no MMA) needs no special case.
  const bool o16 = p.out_f16 != 0;
  const int osz = o16 ? 2 : 4;
  const long long o_step = (dir ? -static_cast<long long>(p.out_ld) : static_cast<long long>(p.out_ld)) * osz;
  // packed-sequence semantics (optional): this pair's sentence is live for t < len_b only
  const int len_b = (p.lens && pv) ? min(p.lens[b0 + pb], L) : L;
  const float pad_value = p.pad_value;
  char* out_p = reinterpret_cast<char*>(p.out) + (out_pair + static_cast<size_t>(dir ? L - 1 : 0) * p.out_ld) * osz;
  auto store_out = [&](bool dead, __half h16, float hfull) {
    if (o16) *reinterpret_cast<__half*>(out_p) = dead ? __float2half_rn(pad_value) : h16;
    else *reinterpret_cast<float*>(out_p) = dead ? pad_value : (PR ? hfull : __half2float(h16));
  };
  __half* hx_p = hx_pair + static_cast<size_t>(dir ? L - 1 : 0) * p.batch * H;
  const float* pw = p_s + static_cast<size_t>(pg * KS) * NBG * S + pbl * S + pu;
  const char* xs0 = reinterpret_cast<const char*>(x_s) + (pb * XSTR + pu) * xesz;
#ifdef FTB_RNN_PROFILE
  // per-phase cycle accounting of warp 0 / lane 0 (diagnostic build only: -DFTB_RNN_PROFILE)
  unsigned long long prof[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#define FTB_TICK(i) { const long long now_ = clock64(); prof[i] += now_ - tick_; tick_ = now_; }
  long long tick_ = clock64();
#else
#define FTB_TICK(i)
#endif
  auto gate_group = [&](int step) {
    if (CL) { FTB_TICK(7) }
    float hnew = 0.0f;
    bool dead = false;
    if (pv) {
      const char* xsb = xs0 + (step & (RNN_XS - 1)) * x_stage;
      float xs[G];
#pragma unroll
      for (int g = 0; g < G; ++g)
        xs[g] = x16 ? __half2float(*reinterpret_cast<const __half*>(xsb + g * U * 2))
                    : *reinterpret_cast<const float*>(xsb + g * U * 4);
      float hh[G];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float s = bias[g];
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) s += pw[ks * NBG * S + g * U];
        hh[g] = s;
      }
      if (G == 4) {
        const float ig = PR ? sigm_x(xs[0] + hh[0]) : sigm(xs[0] + hh[0]);
        const float fg = PR ? sigm_x(xs[1] + hh[1]) : sigm(xs[1] + hh[1]);
        const float gg = PR ? tanh_x(xs[2] + hh[2]) : tanh_fast(xs[2] + hh[2]);
        const float og = PR ? sigm_x(xs[G - 1] + hh[G - 1]) : sigm(xs[G - 1] + hh[G - 1]);
        c_state = fg * c_state + ig * gg;
        hnew = og * (PR ? tanh_x(c_state) : tanh_fast(c_state));
      } else {
        const float r = PR ? sigm_x(xs[0] + hh[0]) : sigm(xs[0] + hh[0]);
        const float z = PR ? sigm_x(xs[1] + hh[1]) : sigm(xs[1] + hh[1]);
        const float n = PR ? tanh_x(xs[2] + r * hh[2]) : tanh_fast(xs[2] + r * hh[2]);
        hnew = (1.0f - z) * n + z * h_state;
      }
      const bool live = (dir ? (L - 1 - step) : step) < len_b;
      if (!live) { hnew = 0.0f; c_state = 0.0f; }     // zero state until the backward pass reaches len_b - 1
      h_state = hnew;
      dead = !live;
    }
    const __half h16 = __float2half_rn(hnew);
    const unsigned mine = __half_as_ushort(h16);
    if (CL) { FTB_TICK(4) }
    if (CL) {
      // publish: all-gather the 8 units of this lane's octet (3 xor rounds) -> one 16-byte chunk of column pbl;
      // lane l pushes chunk l/8 into the tile of peer l%8 (and l%8 + 8) of the unit
      const unsigned o1 = __shfl_xor_sync(0xffffffffu, mine, 1);
      const unsigned w = (lane & 1) ? (o1 | (mine << 16)) : (mine | (o1 << 16));
      const unsigned o2 = __shfl_xor_sync(0xffffffffu, w, 2);
      const unsigned d0 = (lane & 2) ? o2 : w, d1 = (lane & 2) ? w : o2;
      const unsigned e0 = __shfl_xor_sync(0xffffffffu, d0, 4), e1 = __shfl_xor_sync(0xffffffffu, d1, 4);
      const uint4 v = (lane & 4) ? make_uint4(e0, e1, d0, d1) : make_uint4(d0, d1, e0, e1);
      uint4 vl = make_uint4(0u, 0u, 0u, 0u);
      if (PR) {                              // the same gather for the lo halves fp16(2048 (h - hi))
        const unsigned ml = __half_as_ushort(__float2half_rn((hnew - __half2float(h16)) * RNN_LO_SCALE));
        const unsigned p1 = __shfl_xor_sync(0xffffffffu, ml, 1);
        const unsigned wl = (lane & 1) ? (p1 | (ml << 16)) : (ml | (p1 << 16));
        const unsigned p2 = __shfl_xor_sync(0xffffffffu, wl, 2);
        const unsigned f0 = (lane & 2) ? p2 : wl, f1 = (lane & 2) ? wl : p2;
        const unsigned g0 = __shfl_xor_sync(0xffffffffu, f0, 4), g1 = __shfl_xor_sync(0xffffffffu, f1, 4);
        vl = (lane & 4) ? make_uint4(g0, g1, f0, f1) : make_uint4(f0, f1, g0, g1);
      }
      FTB_TICK(5)
      if (step + 1 < L) {
        const uint32_t poff = (step & 1) * (NG * SL * TILE * 2), boff = (step & 1) * (NG * 8);
#pragma unroll
        for (int d = 0; d < ND; ++d)
          if (rvalid & (1u << d)) {
            st_async_v4(rdata[d] + poff, v, rbar[d] + boff);
            if (PR) st_async_v4(rdata[d] + poff + T::HCL_TILE_BYTES, vl, rbar[d] + boff);
          }
      }
      if (pv) store_out(dead, h16, hnew);
      FTB_TICK(6)
    } else {
      // publish: fp16 pair (even lane stores its own + the odd neighbour's half) and the fp32 layer output
      const unsigned other = __shfl_down_sync(0xffffffffu, mine, 1);
      if (pv) {
        if ((lane & 1) == 0) st_relaxed_u32(hx_p, mine | (other << 16));
        store_out(dead, h16, hnew);
      }
      hx_p += h_step;
    }
    out_p += o_step;
  };

  unsigned retries = 0;
  const bool has_b = nbv > NBG;                     // group B holds at least one valid column

  // The two batch groups are independent recurrences that share the register-resident weights.  Their
  // per-step chains (publish -> L2 -> stage -> MMA -> gates -> publish) are interleaved so that the L2
  // round trip of one group is covered by the tensor-core and gate phases of the other.
  for (int i = tid; i < T::NG * KS * NBG * S; i += RNN_THREADS) p_s[i] = 0.0f;
  if (CL) {
    // "h tile full" barriers: one local arrival (the expect_tx that arms a phase) + TXB bytes from the unit's CTAs.
    // Tile / barrier [par] carries h_t for t = par (mod 2).  A phase is re-armed by the thread that
    // saw it complete, before this CTA's next publication (CTA barriers in between) and hence before any peer can
    // send the step after next; the two tiles of a group alternate, so a peer one step ahead never overwrites a
    // tile that is still being read.
    if (tid == 0) {
#pragma unroll
      for (int i = 0; i < 2; ++i) mbar_init(&bars[i], 1);
      fence_mbar_init();
#pragma unroll
      for (int i = 0; i < 2; ++i) mbar_expect_tx(&bars[i], T::TXB);
    }
    cluster_sync_all();                    // every peer's barriers exist before the first remote store
  }
  for (int step = 0; step < L; ++step) {
    FTB_TICK(0)
    if (CL) {
      const int par = (step - 1) & 1;
      if (step > 0 && mma_warp) {
        mbar_wait(&bars[par], ((step - 1) >> 1) & 1);    // h_{t-1} of all 16 columns has landed in this CTA's tile
        if (tid == 0 && step + 1 < L) mbar_expect_tx(&bars[par], T::TXB);
        FTB_TICK(1)
        mma_group(0, par);
      }
      cp_async_wait<RNN_XS - 2>();         // own x chunk of this step
      __syncthreads();                     // partial sums + everyone's x chunks visible; ring slot (step-1) free
      FTB_TICK(2)
      gate_group(step);                    // every thread owns one (unit, column) pair
      prefetch_x();                        // after the publication: off the step-to-step critical path
      FTB_TICK(3)
      continue;
    }
    // ---------------- group A --------------------------------------------------------------------
    if (step > 0 && mma_warp) {
      settle_h(0, -h_step, retries);       // A loads were issued in the previous iteration
      FTB_TICK(1)
      named_bar_sync(1 + kq, GS);
      FTB_TICK(8)
      if (has_b) issue_h(1, -h_step);      // group B of t-1 was published at the end of the previous iteration
      FTB_TICK(9)
      mma_group(0, 0);
      if (has_b) refresh_h(1, -h_step);
      FTB_TICK(10)
    }
    cp_async_wait<RNN_XS - 2>();           // own x chunk of this step (issued three steps ago)
    __syncthreads();                       // A partial sums + everyone's x chunks visible; ring slot (step-1) free
    FTB_TICK(2)
    if (pg == 0) gate_group(step);
    FTB_TICK(3)
    // ---------------- group B --------------------------------------------------------------------
    if (mma_warp) {
      if (step > 0 && has_b) settle_h(1, -h_step, retries);
      FTB_TICK(4)
      named_bar_sync(3 + kq, GS);          // also orders gate A's publication before the A re-fetch below
      if (step + 1 < L) issue_h(0, 0);     // group A of step t: published by the A-warps of all CTAs just now
      if (step > 0 && has_b) mma_group(1, 0);
      if (step + 1 < L) refresh_h(0, 0);
#pragma unroll
      for (int j = 0; j < IPT; ++j) hp[j] += h_step;
    }
    prefetch_x();                          // newest cp.async group: never waited for before the step after next
    __syncthreads();                       // B partial sums visible
    FTB_TICK(5)
    if (pg == 1) gate_group(step);
    FTB_TICK(6)
  }
  cp_async_wait<0>();
#ifdef FTB_RNN_PROFILE
  if (tid == 0 && (local_cta % 37) == 0)
    printf("rnn prof G=%d H=%d cta %d L %d: cycles/step top %llu waitA %llu mmaA+sync %llu gateA %llu waitB %llu mmaB+sync %llu gateB %llu p7 %llu barA %llu issueB %llu mmaA %llu | retries/step %.2f\n",
           G, H, local_cta, L, prof[0] / L, prof[1] / L, prof[2] / L, prof[3] / L, prof[4] / L, prof[5] / L, prof[6] / L, prof[7] / L, prof[8] / L, prof[9] / L, prof[10] / L,
           static_cast<float>(retries) / L);
#endif
}

template <bool CL>
__device__ __forceinline__ void rnn_dispatch(const RnnProblemDev& p, int local, unsigned char* smem) {
  if constexpr (CL) {
    if (p.precise) {                       // GRU only (rnn_plan_build): the duration predictors
      if (p.hidden == 128) rnn_cta<3, 128, true, true>(p, local, smem);
      else rnn_cta<3, 64, true, true>(p, local, smem);
      return;
    }
  }
  // sentences in this CTA's unit (rnn_cta derives the same numbers): units of <= 8 take the one-n-tile form
  const int unit_b0 = p.b_begin + ((local / p.slices) % p.bgroups) * RNN_NB;
  const bool one = CL && (p.b_end - unit_b0) <= 8;
  if (p.cell == FTB_CELL_LSTM) {
    if (p.hidden == 512 && p.x_idx) {
      if constexpr (CL) { if (one) { rnn_cta<4, 512, true, false, true, true>(p, local, smem); return; } }
      rnn_cta<4, 512, CL, false, true>(p, local, smem);
    } else if (p.hidden == 512) {
      if constexpr (CL) { if (one) { rnn_cta<4, 512, true, false, false, true>(p, local, smem); return; } }
      rnn_cta<4, 512, CL>(p, local, smem);
    }
    else if (p.hidden == 256) rnn_cta<4, 256, CL>(p, local, smem);
    else if (p.hidden == 128) rnn_cta<4, 128, CL>(p, local, smem);
    else rnn_cta<4, 64, CL>(p, local, smem);
  } else {
    if (p.hidden == 512) rnn_cta<3, 512, CL>(p, local, smem);
    else if (p.hidden == 256) {
      if constexpr (CL) { if (one) { rnn_cta<3, 256, true, false, false, true>(p, local, smem); return; } }
      rnn_cta<3, 256, CL>(p, local, smem);
    }
    else if (p.hidden == 128) rnn_cta<3, 128, CL>(p, local, smem);
    else rnn_cta<3, 64, CL>(p, local, smem);
  }
}

// L2 hand-off; cooperative launch (every CTA of the grid must be co-resident: CTAs poll each other's data)
__global__ void __launch_bounds__(RNN_THREADS, 1)
rnn_persistent_kernel(const RnnProblemDev* __restrict__ probs, int nprob) {
  extern __shared__ __align__(16) unsigned char rnn_smem[];
  int pi = 0;
  while (pi + 1 < nprob && static_cast<int>(blockIdx.x) >= probs[pi + 1].cta_base) ++pi;
  const RnnProblemDev p = probs[pi];
  const int local = blockIdx.x - p.cta_base;
  if (local >= p.ctas) return;
  rnn_dispatch<false>(p, local, rnn_smem);
}

// Cluster hand-off; launched with a cluster dimension that is a multiple of every part's slice count, every part
// starting on a cluster boundary.  Units (= SL consecutive CTAs) never straddle clusters and depend on nothing outside
// their cluster, so no grid-wide co-residency is needed.  Padding CTAs only take part in the two cluster barriers.
__global__ void __launch_bounds__(RNN_THREADS, 1)
rnn_cluster_kernel(const RnnProblemDev* __restrict__ probs, int nprob) {
  extern __shared__ __align__(16) unsigned char rnn_smem[];
  pdl_launch_dependents();
  int pi = 0;
  while (pi + 1 < nprob && static_cast<int>(blockIdx.x) >= probs[pi + 1].cta_base) ++pi;
  const RnnProblemDev p = probs[pi];
  const int local = blockIdx.x - p.cta_base;
  if (local < p.ctas) rnn_dispatch<true>(p, local, rnn_smem);
  else cluster_sync_all();                 // pairs with the barrier after the mbarrier set-up of the working CTAs
  cluster_sync_all();                      // no CTA leaves while a peer may still store into its shared memory
}

static int rnn_smem_bytes(int cell, int hidden, bool cluster, bool precise = false) {
  const bool lstm = cell == FTB_CELL_LSTM;
  if (precise) {
    switch (hidden) {
      case 128: return RnnTiling<3, 128, true, true>::SMEM_BYTES_CL;
      default: return RnnTiling<3, 64, true, true>::SMEM_BYTES_CL;
    }
  }
#define FTB_RNN_SMEM(GG, HH) (cluster ? RnnTiling<GG, HH, true>::SMEM_BYTES_CL : RnnTiling<GG, HH, false>::SMEM_BYTES)
  switch (hidden) {
    case 512: return lstm ? FTB_RNN_SMEM(4, 512) : FTB_RNN_SMEM(3, 512);
    case 256: return lstm ? FTB_RNN_SMEM(4, 256) : FTB_RNN_SMEM(3, 256);
    case 128: return lstm ? FTB_RNN_SMEM(4, 128) : FTB_RNN_SMEM(3, 128);
    default: return lstm ? FTB_RNN_SMEM(4, 64) : FTB_RNN_SMEM(3, 64);
  }
#undef FTB_RNN_SMEM
}

// ------------------------------------------------------------------------------------------
// Max co-resident clusters of `cs` CTAs of the cluster kernel (0 = this cluster size cannot be launched here).
static int rnn_cluster_capacity(int cs, int smem) {
  if (cudaFuncSetAttribute(rnn_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           rnn_smem_bytes(FTB_CELL_LSTM, 512, true)) != cudaSuccess) { cudaGetLastError(); return 0; }
  if (cs > 8 && cudaFuncSetAttribute(rnn_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(cs * 16);
  cfg.blockDim = dim3(RNN_THREADS);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, rnn_cluster_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// Lays the (problem, direction, batch group, slice) CTAs out over launches of at most `capacity` CTAs; every part starts
// on a multiple of `align` CTAs (cluster size; 1 for the L2 hand-off).
static int rnn_place(const ftb_rnn_problem* problems, const std::vector<int>& which, int capacity, int align, bool cluster,
                     int lane, std::vector<RnnProblemDev>* parts, std::vector<RnnLaunch>* launches) {
  RnnLaunch cur;
  cur.first = static_cast<int>(parts->size());
  cur.cluster = cluster ? align : 0;
  cur.lane = lane;
  int ngroups = 0;
  auto up = [align](int v) { return (v + align - 1) / align * align; };
  for (int i : which) {
    const ftb_rnn_problem& q = problems[i];
    const int slices = q.hidden / RNN_U;
    const int per_bgroup = 2 * slices;                 // CTAs per batch group (both directions)
    if (per_bgroup > capacity) return FTB_E_RESIDENT;
    int b = 0;
    while (b < q.batch) {
      int room = (capacity - up(cur.ctas)) / per_bgroup;    // batch groups that still fit in this launch
      if (room <= 0) {
        cur.ctas = up(cur.ctas);
        launches->push_back(cur);
        cur = RnnLaunch();
        cur.first = static_cast<int>(parts->size());
        cur.cluster = cluster ? align : 0;
        cur.lane = lane;
        room = capacity / per_bgroup;
      }
      const int want = (q.batch - b + RNN_NB - 1) / RNN_NB;
      const int take = want < room ? want : room;
      RnnProblemDev d{};
      d.xproj = q.xproj; d.w_hh = q.w_hh; d.b_hh = q.b_hh; d.out = q.out; d.hx = nullptr;
      d.lens = q.lens; d.pad_value = q.pad_value; d.out_f16 = q.out_f16; d.x_f16 = q.x_f16;
      d.x_idx = q.x_idx;
      d.precise = (cluster && q.precise) ? 1 : 0;
      d.cell = q.cell; d.hidden = q.hidden; d.batch = q.batch; d.seq_len = q.seq_len;
      d.seq_stride = q.seq_stride; d.out_ld = q.out_ld; d.out_col0 = q.out_col0;
      d.gates = q.cell == FTB_CELL_LSTM ? 4 : 3;
      d.units_per_cta = RNN_U; d.slices = slices; d.batch_per_cta = RNN_NB;
      d.bgroups = take;
      d.b_begin = b;
      d.b_end = (b + take * RNN_NB < q.batch) ? b + take * RNN_NB : q.batch;
      d.cta_base = up(cur.ctas);
      d.ctas = 2 * take * slices;
      const int sm = rnn_smem_bytes(q.cell, q.hidden, cluster, cluster && q.precise);
      if (sm > cur.smem) cur.smem = sm;
      d.group_base = ngroups;
      d.problem = i;
      ngroups += 2 * take;
      cur.ctas = d.cta_base + d.ctas;
      cur.count += 1;
      parts->push_back(d);
      b = d.b_end;
    }
  }
  if (cur.count) {
    cur.ctas = up(cur.ctas);
    launches->push_back(cur);
  }
  return ngroups;
}

int rnn_plan_build(const ftb_rnn_problem* problems, int nproblems, int num_sms, RnnPlan* plan) {
  if (!problems || nproblems < 1 || nproblems > FTB_MAX_RNN_PROBLEMS) return FTB_E_ARG;
  int max_slices = 0;
  bool any_precise = false;
  for (int i = 0; i < nproblems; ++i) {
    const ftb_rnn_problem& q = problems[i];
    if (!q.xproj || !q.w_hh || !q.b_hh || !q.out) return FTB_E_ARG;
    if (q.cell != FTB_CELL_GRU && q.cell != FTB_CELL_LSTM) return FTB_E_ARG;
    if (q.hidden != 64 && q.hidden != 128 && q.hidden != 256 && q.hidden != 512) return FTB_E_SHAPE;
    if (q.batch < 1 || q.seq_len < 1 || q.seq_stride < q.seq_len) return FTB_E_ARG;
    if ((q.out_ld % 4) != 0 || (q.out_col0 % 4) != 0 || (reinterpret_cast<uintptr_t>(q.out) & 15u)) return FTB_E_ALIGN;
    if (q.x_idx && (q.cell != FTB_CELL_LSTM || q.hidden != 512 || q.precise)) return FTB_E_SHAPE;   // instantiated for LSTM-512
    if (q.hidden / RNN_U > max_slices) max_slices = q.hidden / RNN_U;
    if (q.precise) {
      // split-precision recurrence: GRU, H <= 128, fp32 x-projection and output, cluster hand-off only
      if (q.cell != FTB_CELL_GRU || q.hidden > 128 || q.x_f16 || q.out_f16) return FTB_E_SHAPE;
      any_precise = true;
    }
  }
  // L2 hand-off layout (always possible on a device that holds one batch group)
  std::vector<RnnProblemDev> parts;
  std::vector<RnnLaunch> launches;
  std::vector<int> all(nproblems);
  for (int i = 0; i < nproblems; ++i) all[i] = i;
  int ngroups = rnn_place(problems, all, num_sms, 1, false, 0, &parts, &launches);
  if (ngroups < 0) return ngroups;
  // Cluster hand-off (FTB_RNN_HANDOFF=l2 disables it).  Problems are grouped by slice count = cluster size; every group
  // is its own launch sequence ("lane") and the lanes run concurrently (side streams forked from / joined to the
  // caller's stream), because a launch has ONE cluster size and large clusters pack badly into the GPCs (B200: 15
  // resident clusters of 8 CTAs, 7 of 16).  Taken when no lane needs more launches than the L2 layout.
  plan->cluster = 0;
  const char* mode = getenv("FTB_RNN_HANDOFF");
  if (!(mode && strcmp(mode, "l2") == 0)) {
    std::vector<RnnProblemDev> cparts;
    std::vector<RnnLaunch> claunches;
    int cg = 0, lanes = 0;
    size_t depth = 0;
    bool ok = true;
    for (int cs = 16; cs >= 2 && ok; cs >>= 1) {
      std::vector<int> which;
      int smem = 0;
      for (int i = 0; i < nproblems; ++i) {
        if (problems[i].hidden / RNN_U != cs) continue;
        which.push_back(i);
        const int sm = rnn_smem_bytes(problems[i].cell, problems[i].hidden, true, problems[i].precise != 0);
        if (sm > smem) smem = sm;
      }
      if (which.empty()) continue;
      const int cap = rnn_cluster_capacity(cs, smem) * cs;
      const size_t before = claunches.size();
      const int g = cap >= 2 * cs ? rnn_place(problems, which, cap, cs, true, lanes, &cparts, &claunches) : -1;
      if (getenv("FTB_RNN_DEBUG"))
        fprintf(stderr, "ftb: rnn plan: %zu problem(s) with cluster size %d: resident capacity %d CTAs -> %zu launch(es); "
                        "L2 layout %zu launch(es)\n", which.size(), cs, cap, claunches.size() - before, launches.size());
      if (g < 0) { ok = false; break; }
      cg += g;
      if (claunches.size() - before > depth) depth = claunches.size() - before;
      ++lanes;
    }
    if (ok && lanes > 0 && (depth <= launches.size() || any_precise || (mode && strcmp(mode, "cluster") == 0))) {
      parts.swap(cparts);
      launches.swap(claunches);
      ngroups = cg;
      plan->cluster = 1;
      plan->lanes = lanes;
      for (int i = 1; i < lanes; ++i) {
        cudaStream_t st = nullptr;
        cudaEvent_t ev = nullptr;
        cudaError_t se = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        if (se != cudaSuccess) return static_cast<int>(se);
        plan->side.push_back(st);
        se = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
        if (se != cudaSuccess) return static_cast<int>(se);
        plan->joined.push_back(ev);
      }
      if (lanes > 1) {
        const cudaError_t fe = cudaEventCreateWithFlags(&plan->forked, cudaEventDisableTiming);
        if (fe != cudaSuccess) return static_cast<int>(fe);
      }
    }
  }
  if (!plan->cluster && any_precise) return FTB_E_RESIDENT;      // no silent loss of precision on the L2 path
  if (!plan->cluster) {
    cudaError_t ae = cudaFuncSetAttribute(rnn_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          rnn_smem_bytes(FTB_CELL_LSTM, 512, false));
    if (ae != cudaSuccess) return static_cast<int>(ae);
    for (int i = 0; i < nproblems; ++i) {
      const ftb_rnn_problem& q = problems[i];
      const size_t hx_bytes = static_cast<size_t>(2) * q.seq_len * q.batch * q.hidden * sizeof(__half);
      void* hx = nullptr;
      cudaError_t he = cudaMalloc(&hx, hx_bytes);
      if (he != cudaSuccess) return static_cast<int>(he);
      plan->hx.push_back(hx);
      plan->hx_bytes.push_back(hx_bytes);
    }
    for (RnnProblemDev& d : parts) d.hx = static_cast<__half*>(plan->hx[d.problem]);
  }
  cudaError_t ce = cudaMalloc(&plan->dev, sizeof(RnnProblemDev) * parts.size());
  if (ce != cudaSuccess) return static_cast<int>(ce);
  ce = cudaMemcpy(plan->dev, parts.data(), sizeof(RnnProblemDev) * parts.size(), cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) return static_cast<int>(ce);
  plan->ngroups = ngroups;
  plan->launches = launches;
  plan->problems.assign(problems, problems + nproblems);
  plan->host_parts = parts;
  plan->cap_len = problems[0].seq_len;
  plan->cap_stride = problems[0].seq_stride;
  for (int i = 1; i < nproblems; ++i) {
    if (problems[i].seq_len < plan->cap_len) plan->cap_len = problems[i].seq_len;
    if (problems[i].seq_stride < plan->cap_stride) plan->cap_stride = problems[i].seq_stride;
  }
  return 0;
}

// Launch-time geometry (see gemm_plan_set_geometry): every problem of the plan runs `seq_len` steps over rows at
// stride `seq_stride`; buffers and the L2 exchange tensors keep the capacity they were built with.
int rnn_plan_set_geometry(RnnPlan& plan, int seq_len, int seq_stride, cudaStream_t stream) {
  if (seq_len < 1 || seq_stride < seq_len) return FTB_E_ARG;
  if (seq_len > plan.cap_len || seq_stride > plan.cap_stride) return FTB_E_LIMIT;
  for (RnnProblemDev& d : plan.host_parts) { d.seq_len = seq_len; d.seq_stride = seq_stride; }
  return static_cast<int>(cudaMemcpyAsync(plan.dev, plan.host_parts.data(), sizeof(RnnProblemDev) * plan.host_parts.size(),
                                          cudaMemcpyHostToDevice, stream));
}

void rnn_plan_free(RnnPlan& plan) {
  cudaFree(plan.dev);
  for (void* x : plan.hx) cudaFree(x);
  for (cudaStream_t st : plan.side) cudaStreamDestroy(st);
  for (cudaEvent_t ev : plan.joined) cudaEventDestroy(ev);
  if (plan.forked) cudaEventDestroy(plan.forked);
  plan = RnnPlan();
}

int rnn_plan_launch(RnnPlan& plan, cudaStream_t stream) {
  cudaError_t ce = cudaSuccess;
  if (plan.cluster) {
    // attributes were set by rnn_cluster_capacity() when the plan was built
    if (plan.lanes > 1) {
      ce = cudaEventRecord(plan.forked, stream);
      if (ce != cudaSuccess) return static_cast<int>(ce);
      for (cudaStream_t st : plan.side) {
        ce = cudaStreamWaitEvent(st, plan.forked, 0);
        if (ce != cudaSuccess) return static_cast<int>(ce);
      }
    }
    for (const RnnLaunch& l : plan.launches) {
      const RnnProblemDev* base = plan.dev + l.first;
      int count = l.count;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(l.ctas);
      cfg.blockDim = dim3(RNN_THREADS);
      cfg.dynamicSmemBytes = l.smem;
      cfg.stream = l.lane ? plan.side[l.lane - 1] : stream;
      cudaLaunchAttribute attr[2];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = l.cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[1].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = (l.lane == 0 && pdl_enabled()) ? 2 : 1;    // side lanes follow an event wait, not a kernel
      ce = cudaLaunchKernelEx(&cfg, rnn_cluster_kernel, base, count);
      if (ce != cudaSuccess) return static_cast<int>(ce);
    }
    for (size_t i = 0; i < plan.side.size(); ++i) {
      ce = cudaEventRecord(plan.joined[i], plan.side[i]);
      if (ce != cudaSuccess) return static_cast<int>(ce);
      ce = cudaStreamWaitEvent(stream, plan.joined[i], 0);
      if (ce != cudaSuccess) return static_cast<int>(ce);
    }
    return 0;
  }
  // (the shared-memory attribute of rnn_persistent_kernel was set for this device when the plan was built)
  // arm the in-band "not written yet" marker (0xFFFF halves) in every exchange slot the recurrences publish
  for (size_t i = 0; i < plan.hx.size(); ++i) {
    ce = cudaMemsetAsync(plan.hx[i], 0xFF, plan.hx_bytes[i], stream);
    if (ce != cudaSuccess) return static_cast<int>(ce);
  }
  for (const RnnLaunch& l : plan.launches) {
    const RnnProblemDev* base = plan.dev + l.first;
    int count = l.count;
    void* args[] = {&base, &count};
    // cooperative launch: the driver guarantees that all CTAs of the grid are co-resident,
    // which the data-polling hand-off of h between CTAs relies on.
    ce = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(rnn_persistent_kernel), dim3(l.ctas), dim3(RNN_THREADS),
                                     args, l.smem, stream);
    if (ce != cudaSuccess) return static_cast<int>(ce);
  }
  return 0;
}

}  // namespace ftb
