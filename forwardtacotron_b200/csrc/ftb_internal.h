// Internal declarations shared by the kernel translation units and the C-ABI layer.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <vector>

#include "../../include/ftb200.h"

namespace ftb {

// FTB_PDL=0 turns programmatic dependent launches off (every kernel then waits for its predecessor to drain).
inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("FTB_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

// ------------------------------------------------------------------ conv-GEMM (ftb_gemm.cu)
// Device-resident problem record.  CUtensorMap needs 64-byte alignment; the struct is
// padded to a multiple of 128 bytes and stored in a cudaMalloc'd array.
struct alignas(128) GemmProblemDev {
  CUtensorMap tm_a;      // A [rows, cin] fp32, box {32, 128}, SWIZZLE_128B
  CUtensorMap tm_w;      // W [taps*w_tap_rows, cin] fp32, box {32, bn}, SWIZZLE_128B
  const float* a;        // fp16 data when FTB_F_A_F16 (only the SIMT cross-check kernel dereferences a / w)
  const float* w;
  const float* bias0;
  const float* scale;
  const float* shift;
  const float* residual;
  float* out;
  float* out_t;
  float* out_full;
  long long rows;
  int lda, cin, taps, tap0, w_tap_rows, n, bn, epilogue, flags;
  int res_ld, out_ld, out_col0;
  int batch, seq_len, seq_stride;
  int kblocks;           // ceil(cin / kelems)
  int m_tiles, n_tiles;  // tile grid of this problem
  int m_step, m_shift;   // tile i covers rows [i*m_step + m_shift, +128)
  int tile_base;         // first global tile id
  int kelems;            // operand elements per 128-byte k-block: 32 (tf32) or 64 (fp16, FTB_F_A_F16)
  int plan_tiles;        // schedule slots of the whole plan at the CURRENT geometry (the kernels read record 0)
  long long cap_rows;    // rows the buffers / tensor maps were built for (ftb_gemm_plan_set_geometry may not exceed it)
};

struct GemmPlan {
  GemmProblemDev* dev = nullptr;
  int nprob = 0;
  int total_tiles = 0;
  long long simt_elems = 0;   // max rows*n over problems (SIMT cross-check grid)
  bool pair = false;          // CTA-pair plan: total_tiles / tile_base count tile pairs, tm_w boxes hold bn/2 rows
  int cap_tiles = 0;          // total_tiles at the geometry the plan was built with (sizes the grid; >= total_tiles)
  std::vector<GemmProblemDev>* host = nullptr;   // host copy of the records (re-uploaded by gemm_plan_set_geometry)
};

int gemm_plan_build(const ftb_gemm_problem* problems, int nproblems, GemmPlan* plan);
int gemm_plan_launch(const GemmPlan& plan, int impl, int num_sms, cudaStream_t stream);
int gemm_plan_set_geometry(GemmPlan& plan, int seq_len, int seq_stride, cudaStream_t stream);
void gemm_plan_free(GemmPlan& plan);

// 2-D row-major [rows, cols] fp32 or fp16 tensor map (leading dimension ld elements), box {128 bytes, box_rows}, 128B swizzle
int tensor_map_2d(CUtensorMap* m, const void* base, long long rows, int cols, int ld, int box_rows, bool f16);
// K8u (ftb_attn5.cu): split-precision attention on tcgen05, head width 128
int mha_tc5_launch(const void* qkv16, int e, int heads, const int64_t* key_tok, int batch, int seq_len, int seq_stride,
                   float* out, int round, cudaStream_t st);

// ------------------------------------------------------------------ fused CBHG tail (ftb_tail.cu)
struct alignas(128) TailProblemDev {
  CUtensorMap tm_a;      // proj2 output  [rows, cin]          fp16, box {64, 128}
  CUtensorMap tm_wpre;   // pre_highway   [256, cin]           fp16, box {64, 256}
  CUtensorMap tm_whw;    // highways      [nh * 512, 256]      fp16, box {64, 256}  (pack_highway layout, bn = 256)
  CUtensorMap tm_wih;    // W_ih          [ih_chunks*256, 256] fp16, box {64, 256}
  const float* b_hw;     // [nh][512] = b1 | b2
  const float* b_ih;     // [n_ih]
  void* out;             // x-projection [rows, out_ld] fp32 or fp16
  long long rows, cap_rows;
  int batch, seq_len, seq_stride;
  int m_tiles, n_highways, kb_in, n_ih, ih_chunks, out_ld, out_f16;
};

struct TailPlan {
  TailProblemDev* dev = nullptr;
  TailProblemDev host{};
  int cap_tiles = 0;
};

int tail_plan_build(const ftb_cbhg_tail_problem* q, int num_sms, TailPlan* plan);
int tail_plan_launch(const TailPlan& plan, int num_sms, cudaStream_t stream);
int tail_plan_set_geometry(TailPlan& plan, int seq_len, int seq_stride, cudaStream_t stream);
void tail_plan_free(TailPlan& plan);

// ------------------------------------------------------------------ recurrences (ftb_rnn.cu)
struct RnnProblemDev {
  const float* xproj;
  const float* w_hh;
  const float* b_hh;
  float* out;
  __half* hx;            // fp16 exchange tensor [2][seq_len][batch][hidden], sentinel-armed before each run
  const int* x_idx;      // optional [batch, seq_len] frame -> absolute row of xproj (ftb_rnn_problem.x_idx) or null
  const int* lens;       // optional [batch] valid lengths (packed-sequence semantics) or null
  float pad_value;       // value of out[b, t >= lens[b]]
  int out_f16;           // `out` holds fp16 (the operand copy of a following FTB_F_A_F16 problem)
  int x_f16;             // `xproj` holds fp16 (written by a FTB_F_OUT_F16 problem)
  int precise;           // split-precision recurrent product + exact gates (cluster hand-off; ftb_rnn_problem.precise)
  int cell, hidden, batch, seq_len, seq_stride, out_ld, out_col0;
  int gates;             // 3 or 4
  int units_per_cta;     // hidden units owned by one CTA
  int slices;            // hidden / units_per_cta
  int batch_per_cta;     // batch columns handled by one CTA (multiple of 8)
  int bgroups;           // ceil(batch / batch_per_cta)
  int cta_base;          // first CTA of this part inside its launch
  int ctas;              // 2 * bgroups * slices
  int group_base;        // first barrier-counter index (one per (dir, bgroup))
  int b_begin, b_end;    // batch columns covered by this part (a problem too large for one
                         // co-resident grid is split into parts that run in successive launches)
  int problem;           // index of the originating ftb_rnn_problem
};

struct RnnLaunch {
  int first = 0, count = 0;           // parts [first, first+count) of RnnPlan::dev
  int ctas = 0;
  int smem = 0;                       // dynamic shared memory of this launch (max over its parts)
  int cluster = 0;                    // cluster size of this launch (0: cooperative launch, L2 hand-off)
  int lane = 0;                       // 0: the caller's stream; k > 0: side stream k-1 of the plan (concurrent lanes)
};

struct RnnPlan {
  RnnProblemDev* dev = nullptr;       // all parts of all launches
  unsigned int* counters = nullptr;   // unused since the in-band sentinel protocol (kept for ABI of the struct)
  int ngroups = 0;
  int cluster = 0;                    // 1: thread-block-cluster hand-off (per-launch cluster size); 0: L2 hand-off
  int lanes = 1;                      // concurrent launch sequences (one per cluster size)
  std::vector<cudaStream_t> side;     // lanes - 1 side streams + their join events, forked from the caller's stream
  std::vector<cudaEvent_t> joined;
  cudaEvent_t forked = nullptr;
  std::vector<ftb_rnn_problem> problems;   // originals
  std::vector<void*> hx;                   // per problem: fp16 exchange tensor (owned by the plan)
  std::vector<size_t> hx_bytes;
  std::vector<RnnLaunch> launches;
  std::vector<RnnProblemDev> host_parts;   // host copy of `dev` (re-uploaded by rnn_plan_set_geometry)
  int cap_len = 0, cap_stride = 0;         // geometry the plan (exchange tensors included) was built for
};

int rnn_plan_build(const ftb_rnn_problem* problems, int nproblems, int num_sms, RnnPlan* plan);
int rnn_plan_launch(RnnPlan& plan, cudaStream_t stream);
int rnn_plan_set_geometry(RnnPlan& plan, int seq_len, int seq_stride, cudaStream_t stream);
void rnn_plan_free(RnnPlan& plan);

}  // namespace ftb

struct ftb_context {
  int device = 0;
  int num_sms = 0;
  int64_t launches = 0;
  int* err_word = nullptr;    // device word: bit 0 = a token / class id outside its embedding table (set by the embedding
                              // kernels, handed to the host through meta[3] of the next ftb_lr_scan and cleared there)
  int next_plan = 1;
  std::map<int, ftb::GemmPlan> gemm_plans;
  std::map<int, ftb::RnnPlan> rnn_plans;
  std::map<int, ftb::TailPlan> tail_plans;
  std::map<int, cudaGraphExec_t> graphs;
};
