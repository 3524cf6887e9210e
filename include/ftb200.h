/* ftb200.h -- C ABI of libftb200.so, the sm_100a kernel library behind
 * forwardtacotron_b200.models.*.generate().
 *
 * The reference (ziyaad30/ForwardTacotron) has no FFI layer of its own: its hot path
 * calls torch.nn modules.  Each entry point below therefore names the reference call
 * site(s) (file:line under the reference tree) whose arithmetic it replaces; the Python
 * side that binds them with ctypes is forwardtacotron_b200/runtime.py and the stub a
 * reference maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless noted;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - return 0 = OK, <0 = invalid argument (FTB_E_*), >0 = a cudaError_t value;
 *   - no entry point synchronises, allocates per call, or throws; plans own their
 *     descriptors (tensor maps, barrier counters) and are created once per shape;
 *   - one handle per device, calls on one handle are not thread safe.
 *
 * Activation layout ("padded-flat, channel-last"): a [batch, seq_len, C] fp32 tensor is
 * stored as rows r = b*seq_stride + t (seq_stride >= seq_len + 8), C contiguous; the
 * rows with t in [seq_len, seq_stride) are ZERO and are never written by any kernel, so
 * convolution taps that reach across a sentence edge read zero padding exactly like
 * nn.Conv1d(padding=k//2) (common_layers.py:50).
 */
#ifndef FTB200_H_
#define FTB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FTB_VERSION 100

#define FTB_E_ARG      (-1)   /* null pointer / bad size / bad enum                     */
#define FTB_E_ALIGN    (-2)   /* pointer or leading dimension not 16-byte aligned        */
#define FTB_E_SHAPE    (-3)   /* shape not supported by the sm_100a kernels              */
#define FTB_E_PLAN     (-4)   /* unknown plan id                                         */
#define FTB_E_DRIVER   (-5)   /* cuTensorMapEncodeTiled unavailable (no driver)          */
#define FTB_E_RESIDENT (-6)   /* persistent recurrence grid does not fit on the device   */
#define FTB_E_LIMIT    (-7)   /* size beyond the fixed capacity of a no-allocation entry  */

typedef struct ftb_context* ftb_handle;

int ftb_version(void);
const char* ftb_error_string(int code);
int ftb_create(int device, ftb_handle* out);
int ftb_destroy(ftb_handle h);
/* number of kernels launched through this handle since creation (bench.py's gpu_launches) */
int64_t ftb_launch_count(ftb_handle h);

/* Launch lists as CUDA graphs (plumbing; no counterpart in the reference, whose modules launch eagerly).  Everything the
 * entry points of this header enqueue on `stream` between ftb_graph_capture_begin and ftb_graph_capture_end -- the
 * cluster / cooperative launches of a recurrence plan and their side-stream forks included -- becomes one executable
 * graph; ftb_graph_launch replays it on any stream.  Capture mode is thread-local; `stream` must not be the legacy
 * default stream.  (torch.cuda.CUDAGraph does the same but prepends two one-element fill kernels to every replay for its
 * random-number generator state: ~8 us of stream time per replay, three replays per generate().) */
int ftb_graph_capture_begin(ftb_handle h, void* stream);
int ftb_graph_capture_end(ftb_handle h, void* stream, int* graph_out);
int ftb_graph_launch(ftb_handle h, int graph, void* stream);
int ftb_graph_destroy(ftb_handle h, int graph);

/* ------------------------------------------------------------------------------------
 * K1/K2  implicit-GEMM convolution / linear with fused epilogue   (tcgen05 + TMEM + TMA)
 *
 * Replaces: nn.Conv1d + F.relu + nn.BatchNorm1d(eval) in BatchNormConv.forward
 * (models/common_layers.py:54-57), the CBHG bank/cat/MaxPool1d (common_layers.py:97-105),
 * conv_project1/2 + residual (:109-114), pre_highway Linear (:118), HighwayNetwork.forward
 * (:35-40), the W_ih half of nn.GRU / nn.LSTM (common_layers.py:123,
 * forward_tacotron.py:37,224), self.lin / self.post_proj (forward_tacotron.py:226,230) and
 * the FFT-block Conv1d/Linear layers (common_layers.py:173-181).
 *
 *   acc[r, n] = sum_{j<taps} sum_{c<cin} A[r + tap0 + j, c] * W[j*w_tap_rows + n, c]
 *   (rows outside [0, rows) read as zero; tf32 operands, fp32 accumulation)
 *
 *   FTB_EPI_AFFINE : v = acc + bias0[n]; if RELU v = max(v,0); v = v*scale[n] + shift[n];
 *                    v += residual[r, n]
 *   FTB_EPI_MAXPOOL: y = AFFINE(acc); v[r] = (t == 0) ? y[r] : max(y[r-1], y[r])
 *                    (MaxPool1d(2,1,1)[:, :, :L], common_layers.py:105)
 *   FTB_EPI_HIGHWAY: n = 2*units; W rows of N-tile i are [W1 rows | W2 rows] of units
 *                    [i*bn/2, (i+1)*bn/2); bias0 = [b1 | b2];
 *                    g = sigmoid(x2); v = g*relu(x1) + (1-g)*residual[r, u]
 *   Only rows with (r % seq_stride) < seq_len are stored; null bias0/scale/shift mean 0/1/0.
 * ------------------------------------------------------------------------------------ */
enum { FTB_EPI_AFFINE = 0, FTB_EPI_MAXPOOL = 1, FTB_EPI_HIGHWAY = 2 };
/* FTB_F_A_F16  : `a` and `w` hold IEEE fp16 (same 11-bit significand as tf32; lda / cin counted in fp16 elements,
 *                multiples of 8); the MMAs are kind::f16 with fp32 accumulation -- twice the tensor rate and half
 *                the operand traffic of tf32.  Used where the producer of `a` is another K1/K2 problem.
 * FTB_F_OUT_F16: `out` is fp16 [rows, out_ld] (out_ld, out_col0, n multiples of 8), rounded to nearest: the operand
 *                copy for a following FTB_F_A_F16 problem.  Any epilogue; residual / out_t / out_full (always fp32)
 *                may accompany it except with MAXPOOL; ROUND_TF32 is ignored.
 * FTB_F_OUT_SPLIT: (with FTB_F_OUT_F16, AFFINE or MAXPOOL epilogue, no residual / out_t / out_full) `out` receives every
 *                value v as THREE fp16 column blocks of width seg = out_ld / 3: hi = fp16(v) at out_col0 + c, lo =
 *                fp16(v - hi) at seg + out_col0 + c, hi again at 2 seg + out_col0 + c (out_col0 + n <= seg, so several
 *                problems -- the CBHG bank -- fill one split buffer).  A consumer whose weights are packed as
 *                [W_hi | W_hi | W_lo] along K (cin = 3 seg, FTB_F_A_F16) then accumulates hi*W_hi + lo*W_hi + hi*W_lo: the
 *                product at ~2^-20 relative operand precision instead of 2^-11, on the same kind::f16 MMAs.  Used for
 *                the DURATION predictor, whose output is rounded to integers (forward_tacotron.py:174-177,
 *                common_layers.py:18-21): an operand error of 2^-11 moves durations across the .5 boundary.
 * FTB_F_TAPS_IN_K: (with FTB_F_A_F16, taps <= 8) the taps are folded into the reduction: `a` is DENSE (lda == cin) and at
 *                least 8 zero rows PRECEDE it in the same allocation (like the 8 gap rows that follow every sentence),
 *                so the `taps` rows r + tap0 .. r + tap0 + taps - 1 are one contiguous run of taps*cin elements, and `w`
 *                is packed [w_tap_rows][taps*cin] (column j*cin + c = tap j, channel c).  The kernel then reads A through an
 *                overlapping-row tensor map (row stride cin, row extent taps*cin) and runs ONE tap over K = taps*cin:
 *                ceil(taps*cin/64) k-blocks instead of taps*ceil(cin/64).  For channel counts that fill a 128-byte
 *                swizzle row badly -- the CBHG postnet's bank over the 80-channel mel (common_layers.py:97-99): 48
 *                k-blocks instead of 72 for its eight kernel sizes.  Same sums, same operands, same rounding.        */
enum { FTB_F_RELU = 1, FTB_F_ROUND_TF32 = 2, FTB_F_A_F16 = 4, FTB_F_OUT_F16 = 8, FTB_F_OUT_SPLIT = 16, FTB_F_TAPS_IN_K = 32 };
#define FTB_FOLD_LEAD_ROWS 8

typedef struct {
  const void* a;         /* [rows, lda] activations: fp32 tf32-rounded by their producer, or fp16 (A_F16) */
  int64_t rows;          /* batch * seq_stride                                               */
  int32_t lda;           /* floats, multiple of 4                                            */
  int32_t cin;           /* reduction extent per tap (<= lda)                                */
  const void* w;         /* packed [taps][w_tap_rows][cin] fp32 tf32-rounded, or fp16 (A_F16)   */
  int32_t taps, tap0;    /* tap j reads row r + tap0 + j   (tap0 = -(k/2))                   */
  int32_t w_tap_rows;    /* rows per tap in w (n rounded up to a multiple of bn)             */
  int32_t n;             /* output channels (HIGHWAY: 2*units)                               */
  int32_t bn;            /* N tile: multiple of 16, 16..256                                  */
  int32_t epilogue;      /* FTB_EPI_*                                                        */
  int32_t flags;         /* FTB_F_*                                                          */
  const float* bias0;    /* [n] or null                                                      */
  const float* scale;    /* [n] or null                                                      */
  const float* shift;    /* [n] or null                                                      */
  const float* residual; /* [rows, res_ld] or null                                           */
  int32_t res_ld;
  void* out;             /* [rows, out_ld] fp32 (fp16 with OUT_F16), written at columns out_col0 + n; may be null */
  int32_t out_ld, out_col0;
  float* out_t;          /* optional [batch, n, seq_len] transposed, never rounded; or null  */
  int32_t batch, seq_len, seq_stride;
  float* out_full;       /* optional [rows, out_ld]: the same values as `out` but never tf32-rounded (the
                            copy a later layer adds as its residual); requires `out`; or null        */
} ftb_gemm_problem;

/* A plan = up to FTB_MAX_PROBLEMS independent problems run by ONE persistent launch. */
#define FTB_MAX_PROBLEMS 32
int ftb_gemm_plan_create(ftb_handle h, const ftb_gemm_problem* problems, int nproblems, int* plan_out);
/* impl: 0 = tcgen05/TMA kernel (the product path); 1 = plain SIMT fp32 cross-check kernel
 * (test-only: same problems, no tensor cores, used by tests to bisect descriptor bugs).   */
int ftb_gemm_plan_run(ftb_handle h, int plan, int impl, void* stream);
int ftb_gemm_plan_destroy(ftb_handle h, int plan);
/* What the library chose for the plan: *cta_pairs = 1 when the launch runs as 2-CTA clusters (conv_gemm_tc_pair_kernel:
 * tcgen05.mma.cta_group::2, M = 256, half of the W tile staged per CTA, six-stage ring; picked for launches whose tiles
 * run >= 8 k-blocks on average, FTB_GEMM_PAIR=0/1 forces), *tiles = schedule slots (tiles, or tile pairs).  Either
 * pointer may be null. */
int ftb_gemm_plan_info(ftb_handle h, int plan, int* cta_pairs, int* tiles);
/* Launch-time geometry.  A plan is built for a CAPACITY (the rows / seq_len / seq_stride of its problems); afterwards every
 * problem can be re-pointed, in stream order, at `batch` sentences of `seq_len` <= capacity rows at stride `seq_stride`
 * (batch * seq_stride <= the built rows, else FTB_E_LIMIT) inside the same buffers: rows, tile grid and schedule are
 * recomputed and the device records re-uploaded; tensor maps, buffers and captured CUDA graphs stay valid (the kernels
 * read the geometry from the records, not from their arguments).  The caller keeps rows t in [seq_len, seq_stride) of
 * every sentence zero.  generate() of the reference sees a new regulated length M on almost every call
 * (common_layers.py:23 pads to the batch maximum); this is what lets one plan serve a whole bucket of lengths.  */
int ftb_gemm_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream);

/* ------------------------------------------------------------------------------------
 * K2f  fused row-local tail of a CBHG: pre_highway -> highways -> W_ih projection in ONE persistent kernel
 *
 * Replaces: `self.pre_highway(x)` (models/common_layers.py:118), `for h in self.highways: x = h(x)` (:119-120,
 * HighwayNetwork.forward :35-40) and the W_ih x + b_ih half of `self.rnn(x)` (:123).  Every layer maps row r to row r,
 * so a 128-row tile runs the whole chain on one SM: the activation tile stays in shared memory as an fp16 operand copy
 * plus an fp16 remainder (their sum is the residual term of the highway), weights stream through a TMA ring, and only
 * the x-projection is written to HBM.  Same arithmetic as the K1/K2 problems it replaces (fp16 operands, fp32
 * accumulation, fp32 epilogue).
 *   a      fp16 [rows, lda]: conv_project2 output + residual (the FTB_F_OUT_F16 operand copy), cin <= 256 columns
 *   w_pre  fp16 [256][cin]                 pre_highway.weight (no bias)
 *   w_hw   fp16 [n_highways][512][256]     per layer: [W1 rows | W2 rows] of units 0..127, then of units 128..255
 *   b_hw   fp32 [n_highways][512]          per layer: b1 (256) | b2 (256)
 *   w_ih   fp16 [n_ih rounded up to 256][256], b_ih fp32 [n_ih]: both directions, n_ih = 2 * 3 * H
 *   out    [rows, out_ld] fp32, or fp16 when out_f16 != 0 (ftb_rnn_problem.x_f16); only rows with t < seq_len
 * ------------------------------------------------------------------------------------ */
typedef struct {
  const void* a; int32_t lda, cin;
  int32_t channels;                 /* highway width: 256 */
  const void* w_pre;
  int32_t n_highways; const void* w_hw; const float* b_hw;
  const void* w_ih; const float* b_ih; int32_t n_ih;
  void* out; int32_t out_ld, out_f16;
  int32_t batch, seq_len, seq_stride;
} ftb_cbhg_tail_problem;
int ftb_cbhg_tail_plan_create(ftb_handle h, const ftb_cbhg_tail_problem* problem, int* plan_out);
int ftb_cbhg_tail_plan_run(ftb_handle h, int plan, void* stream);
int ftb_cbhg_tail_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream);
int ftb_cbhg_tail_plan_destroy(ftb_handle h, int plan);

/* ------------------------------------------------------------------------------------
 * K3/K4  persistent bidirectional GRU / LSTM recurrence
 *
 * Replaces: the time loop of nn.GRU (common_layers.py:123; forward_tacotron.py:37) and
 * nn.LSTM (forward_tacotron.py:224), batch_first, bidirectional, h0 = c0 = 0, run over the
 * full padded length (no packing in generate()).  xproj holds W_ih x + b_ih for
 * [forward | reverse] (K2 output); W_hh stays resident in registers (fp16 MMA fragments, fp32
 * accumulation) for all steps; h_t is handed between the CTAs of one (direction, 16-sentence group)
 * either through peer shared memory inside a thread-block cluster (st.async + mbarrier; the GRUs and
 * small-batch LSTMs) or through a plan-owned fp16 exchange tensor in L2 whose sentinel pre-fill doubles
 * as the per-step synchronisation (no counter barrier); the plan picks the protocol from the cluster
 * occupancy of the device (FTB_RNN_HANDOFF=l2|cluster overrides).  h_t is written to `out`, the layer
 * output, as fp32 or (out_f16) as the fp16 value it already is.  A plan may launch on plan-owned side streams that are forked from and joined to `stream`.
 *   GRU  (r,z,n): r = s(xr + Whr h + bhr); z = s(xz + Whz h + bhz);
 *                 n = tanh(xn + r*(Whn h + bhn)); h' = (1-z)*n + z*h
 *   LSTM (i,f,g,o): c' = s(f)*c + s(i)*tanh(g); h' = s(o)*tanh(c')
 * ------------------------------------------------------------------------------------ */
enum { FTB_CELL_GRU = 0, FTB_CELL_LSTM = 1 };

typedef struct {
  int32_t cell;          /* FTB_CELL_*                                                       */
  int32_t hidden;        /* H: 64, 128, 256 or 512                                           */
  int32_t batch, seq_len, seq_stride;
  const float* xproj;    /* [batch*seq_stride, 2*g*H], g = 3 (GRU) / 4 (LSTM)                */
  const float* w_hh;     /* [2][g*H][H] fp32 (weight_hh_l0, weight_hh_l0_reverse); rounded to fp16 in-kernel */
  const float* b_hh;     /* [2][g*H]                                                         */
  float* out;            /* [batch*seq_stride, out_ld]; fwd h at out_col0, bwd at out_col0+H; fp16-exact values */
  int32_t out_ld, out_col0;
  const int32_t* lens;   /* optional [batch] valid lengths: pack_padded_sequence / pad_packed_sequence semantics
                            (forward_tacotron.py:147-152) -- sentence b runs over t < lens[b] only (the backward
                            direction starts at lens[b]-1 from a zero state) and out[b, t >= lens[b]] = pad_value;
                            null = every sentence runs over all seq_len steps (the generate() path)             */
  float pad_value;
  int32_t out_f16;       /* != 0: `out` is fp16 [batch*seq_stride, out_ld] (h_t is fp16-exact, so this loses nothing):
                            the operand copy for a following FTB_F_A_F16 GEMM (lin / post_proj)                 */
  int32_t x_f16;         /* != 0: `xproj` is fp16 [batch*seq_stride, 2*g*H] (a FTB_F_OUT_F16 GEMM wrote it): half the
                            x-projection traffic for one more rounding to 11 bits              */
  int32_t precise;       /* != 0 (GRU, H <= 128, fp32 xproj / out): split-precision recurrence -- h_t and W_hh enter the
                            MMAs as fp16 hi + lo pairs (three products per k-step, ~2^-20 relative operand precision
                            instead of 2^-11), the gates use ex2/rcp instead of tanh.approx, `out` receives the
                            unrounded fp32 h_t.  For the duration predictors, whose output is rounded to integers
                            (forward_tacotron.py:174-177); needs the cluster hand-off (else FTB_E_RESIDENT)      */
  const int32_t* x_idx;  /* optional [batch, seq_len] (dense, row stride = seq_len): step t of sentence b reads row
                            x_idx[b, t] of `xproj` (an absolute row index, >= 0) instead of row b * seq_stride + t.
                            With it the recurrence consumes the LengthRegulator's output WITHOUT it ever being
                            written: W_ih x + b_ih is a per-row map, so it commutes with the regulator's row
                            repetition (common_layers.py:21) -- `xproj` is then computed once per PHONEME (half the
                            projection work at two frames per phoneme), x_idx is the frame -> row map ftb_lr_index
                            writes, and the padded frames (common_layers.py:23) point at one extra row that holds
                            b_ih, the projection of a zero row.  null = rows b * seq_stride + t as described above.
                            Instantiated for the LSTM of 512 hidden units (the BiLSTM of both reference configs,
                            forward_tacotron.py:224); other problems return FTB_E_SHAPE                          */
} ftb_rnn_problem;

#define FTB_MAX_RNN_PROBLEMS 8
int ftb_rnn_plan_create(ftb_handle h, const ftb_rnn_problem* problems, int nproblems, int* plan_out);
int ftb_rnn_plan_run(ftb_handle h, int plan, void* stream);
int ftb_rnn_plan_destroy(ftb_handle h, int plan);
/* How the plan was laid out on this device: *handoff = 1 thread-block-cluster hand-off, 0 L2 hand-off;
 * *kernel_launches = kernel launches per run; *lanes = concurrent launch sequences (1 = all on `stream`). */
int ftb_rnn_plan_info(ftb_handle h, int plan, int* handoff, int* kernel_launches, int* lanes);
/* Launch-time geometry of a recurrence plan (see ftb_gemm_plan_set_geometry): seq_len <= the built seq_len and
 * seq_stride <= the built seq_stride, else FTB_E_LIMIT; the L2 exchange tensors keep their built size. */
int ftb_rnn_plan_set_geometry(ftb_handle h, int plan, int seq_len, int seq_stride, void* stream);

/* ------------------------------------------------------------------------------------
 * K5/K6  duration rounding + LengthRegulator gather (+ fused pitch/energy projection add)
 *
 * Replaces: `if sum(dur.long()) <= 0: fill_(dur, 2.)` (forward_tacotron.py:176-177),
 * LengthRegulator.forward (common_layers.py:17-24) and the pitch/energy Conv1d(1->C,k=3)
 * projections + scaled adds (forward_tacotron.py:214-220).
 *
 * ftb_lr_scan : clamps dur in place (dur[dur<0]=0), applies the batch-global fallback,
 *               rep = trunc(dur + 0.5f) (fp32 add), per-sentence inclusive scan.
 *               Outputs cum [batch, T] int32, lens [batch] int32, and meta[0] = max len,
 *               meta[1] = fallback flag, meta[2] = total frames, meta[3] = the handle's error word (bit 0: an
 *               embedding kernel since the last scan saw an id outside its table -- nn.Embedding raises IndexError
 *               there; the kernels clamp the id to row 0 instead of reading out of bounds) (host reads meta once).
 * ftb_lr_gather: out[b*out_stride + m, :] = src(b, idx) for m < lens[b], where
 *               src(b,t) = x[b*x_stride + t, :C1] ++ extra[b, :C2]  (speaker embedding tail)
 *                          + ps*(wp (*) pitch)[t] + es*(we (*) energy)[t]   (3-tap, zero padded)
 *               rows m in [lens[b], M) are written as 0.0; also writes idx [batch, M] (-1 = pad).
 *               The gathered rows are bit-exact copies of src; src itself is fp32 FMA-free
 *               arithmetic in the reference's operation order.
 * ------------------------------------------------------------------------------------ */
int ftb_lr_scan(ftb_handle h, float* dur, int batch, int T, int apply_fallback,
                int32_t* cum, int32_t* lens, int32_t* meta, void* stream);

typedef struct {
  const float* x; int32_t x_stride /*rows per sentence*/, x_ld, c1;
  const float* extra; int32_t c2;                 /* [batch, c2] appended channels, or null  */
  const float* pitch; const float* energy;        /* [batch, T] or null (then no projection) */
  const float* wp; const float* bp; float ps;     /* [C,3], [C], strength                    */
  const float* we; const float* be; float es;
  const int32_t* cum; const int32_t* lens;
  int32_t batch, T, M;
  float* out; int32_t out_stride /*rows per sentence*/, out_ld;
  int32_t* idx;                                   /* [batch, M] or null                      */
  int32_t round_tf32;                             /* 0 as computed, 1 tf32-rounded, 2 = `out` is fp16 (out_ld in halves) */
} ftb_lr_gather_args;
int ftb_lr_gather(ftb_handle h, const ftb_lr_gather_args* args, void* stream);
/* Frame -> row map of LengthRegulator.forward (common_layers.py:21: repeat_interleave(x[i], rep)): with t = the token
 * whose frames include m (cum[b, t-1] <= m < cum[b, t]; cum = the inclusive cumsum ftb_lr_scan wrote),
 *   idx[b, m] = b * row_stride + t     for m < lens[b]
 *   idx[b, m] = pad_row                for the padded frames m >= lens[b] (pad_sequence, common_layers.py:23)
 * idx is dense [batch, M].  row_stride = 0, pad_row = -1 gives the bare frame -> phoneme map; row_stride = the token
 * stride of a padded-flat buffer gives absolute rows: the map ftb_lr_gather applies, without the copy -- the input of
 * ftb_rnn_problem.x_idx. */
int ftb_lr_index(ftb_handle h, const int32_t* cum, const int32_t* lens, int batch, int T, int M, int row_stride,
                 int pad_row, int32_t* idx, void* stream);

/* No-allocation LengthRegulator: ftb_lr_scan + ftb_lr_gather (fp32 output, no projection / speaker tail) as ONE call
 * whose intermediates (cum, lens, meta) live in a file-scope `static __device__` array of the library instead of
 * caller-allocated scratch tensors.
 *
 * This entry point has NO counterpart in the reference: it was added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  It calls no cudaMalloc / cudaFree / allocating library routine; the caller passes inputs and outputs only.
 *   - capacity: batch <= FTB_LR_STATIC_MAX_BATCH and batch*T <= FTB_LR_STATIC_MAX_TOKENS, else FTB_E_LIMIT;
 *     every GPU that loads the library holds FTB_LR_STATIC_BYTES of device memory for the array until exit;
 *   - the array is shared by all calls on a device and keeps the last call's contents (meta is re-zeroed per call,
 *     cum / lens are overwritten for the [batch, T] in use): calls that OVERLAP on the GPU (different streams without
 *     ordering) race -- serialise them;
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress); make it outside capture.
 * Semantics = ftb_lr_scan(dur in place, fallback) followed by ftb_lr_gather(round_tf32 = 0): out[b*out_stride + m, :c]
 * = x[b*x_stride + idx, :c] for m < min(lens[b], M), rows m in [lens[b], M) = 0.0; M is the caller's frame capacity
 * per sentence (frames beyond it are dropped).  idx [batch, M], lens_out [batch], meta_out [4] are optional outputs
 * (meta as in ftb_lr_scan: max len, fallback flag, total frames).                                                 */
#define FTB_LR_STATIC_MAX_TOKENS 65536
#define FTB_LR_STATIC_MAX_BATCH  1024
#define FTB_LR_STATIC_BYTES      (4 * (FTB_LR_STATIC_MAX_TOKENS + FTB_LR_STATIC_MAX_BATCH + 64))   /* 266 496 */
int ftb_lr_regulate_static(ftb_handle h, const float* x, int x_stride /*rows per sentence*/, int x_ld, int c,
                           float* dur, int batch, int T, int apply_fallback, int M,
                           float* out, int out_stride /*rows per sentence*/, int out_ld,
                           int32_t* idx, int32_t* lens_out, int32_t* meta_out, void* stream);

/* ------------------------------------------------------------------------------------
 * Small glue kernels
 * ------------------------------------------------------------------------------------ */
/* nn.Embedding (forward_tacotron.py:31,210) (+ concatenated conditioning embedding and speaker
 * vector, multi_forward_tacotron.py:40-42,82-86) into the padded-flat layout:
 * out[b*seq_stride+t, :] = table[tok[b,t], :e1] ++ table2[tok2[b,t], :e2] ++ extra[b, :e3].
 * round_tf32: 0 = fp32 as stored, 1 = fp32 rounded to tf32, 2 = `out` is an fp16 array (out_ld in fp16 elements):
 * the operand copy a FTB_F_A_F16 problem reads; 3 = `out` is an fp16 array of THREE column blocks of width out_ld/3,
 * [hi | lo | hi] with hi = fp16(v), lo = fp16(v - hi): the split-precision operand layout of FTB_F_OUT_SPLIT. */
int ftb_embed(ftb_handle h, const int64_t* tok, const float* table, int e1,
              const int64_t* tok2, const float* table2, int e2,
              const float* extra, int e3,
              int batch, int T, int seq_stride, float* out, int out_ld, int round_tf32,
              int rows1 /* rows of table */, int rows2 /* rows of table2 */, void* stream);

/* Predictor head: y[b,t,o] = (dot(x[b*seq_stride+t, :k], w[o,:]) + bias[o]) / alpha
 * (self.lin + `/ alpha`, forward_tacotron.py:38-39).  n_out <= 4.  y is dense [batch, T, n_out]. */
int ftb_rowdot(ftb_handle h, const float* x, int x_ld, int k, const float* w, const float* bias,
               int n_out, float alpha, int batch, int T, int seq_stride, float* y, void* stream);

/* Batched forms: up to FTB_MAX_GLUE_ITEMS independent embeddings / predictor heads over the same [batch, T] token grid
 * in ONE launch (grid.y = item).  The token-side program of generate() issues five embeddings (three predictors, the
 * prenet operand copy and its unrounded twin: forward_tacotron.py:31,210) and three heads (:38-39) back to back; each is
 * a few microseconds of work, so separate launches are pure launch latency. */
#define FTB_MAX_GLUE_ITEMS 8
typedef struct {
  const int64_t* tok; const float* table; int32_t e1;
  const int64_t* tok2; const float* table2; int32_t e2;
  const float* extra; int32_t e3;
  float* out; int32_t out_ld;
  int32_t round_tf32;      /* as ftb_embed */
  int32_t rows1, rows2;    /* rows of table / table2: ids outside [0, rows) set the error word (ftb_lr_scan meta[3]) */
} ftb_embed_item;
int ftb_embed_batch(ftb_handle h, const ftb_embed_item* items, int nitems, int batch, int T, int seq_stride, void* stream);

typedef struct {
  const float* x; int32_t x_ld, k;
  const float* w; const float* bias;
  int32_t n_out; float alpha;
  float* y;
} ftb_rowdot_item;
int ftb_rowdot_batch(ftb_handle h, const ftb_rowdot_item* items, int nitems, int batch, int T, int seq_stride, void* stream);

/* argmax over the last axis of [n, c] (c <= 8) -> int64 (multi_forward_tacotron.py:252, per (b,t)). */
int ftb_argmax_lastdim(ftb_handle h, const float* x, int64_t n, int c, int64_t* out, void* stream);

/* ------------------------------------------------------------------------------------
 * K7/K8  FFT-block pieces (FastPitch, common_layers.py:127-223)
 * ------------------------------------------------------------------------------------ */
/* y = LayerNorm(x (+ res)) * gamma + beta, eps = 1e-5, over the last axis e (<= 1024), only
 * valid rows; optional second LayerNorm chained (the ForwardTransformer's final norm).
 * `out` is tf32-rounded when round_tf32 != 0 (it feeds a GEMM); `out_full` (optional) receives
 * the unrounded values (the residual copy).  round_tf32 == 3 (here, in ftb_posenc and in ftb_mha; e % 8 == 0): `out`
 * is an fp16 array [rows, 3e] = [hi | lo | hi] with hi = fp16(v), lo = fp16(v - hi), the split-precision GEMM operand
 * (FTB_F_OUT_SPLIT) of the duration predictor's FFT blocks (fast_pitch.py:14-41).                                   */
int ftb_add_layernorm(ftb_handle h, const float* x, const float* res, int ld, int e,
                      const float* gamma, const float* beta,
                      const float* gamma2, const float* beta2,
                      int batch, int seq_len, int seq_stride, float* out, float* out_full, int round_tf32,
                      void* stream);

/* x[b,t,:] = emb[tok[b,t], :] + scale * pe[t, :]   (PositionalEncoding, common_layers.py:143-145);
 * tok == null: in-place x += scale*pe over an existing activation (post-LR postnet input). */
int ftb_posenc(ftb_handle h, const int64_t* tok, const float* table, const float* x_in, const float* pe,
               const float* scale_ptr, int e, int batch, int seq_len, int seq_stride,
               float* out, float* out_full, int round_tf32, int table_rows /* rows of `table` (ids are checked) */,
               void* stream);

/* 2..8-head self attention over qkv [rows, 3e] (packed in_proj output), softmax over keys with
 * optional key padding (tok[b,s] == 0 masked, common_layers.py:232-233); out [rows, e]. */
int ftb_mha(ftb_handle h, const float* qkv, int e, int heads, const int64_t* key_tok,
            int batch, int seq_len, int seq_stride, float* out, int round_tf32, void* stream);
/* round_tf32 of ftb_add_layernorm / ftb_posenc / ftb_mha / ftb_mha_split, as for ftb_embed: 0 = fp32 as computed, 1 = fp32
 * rounded to tf32, 2 = `out` is an fp16 array with the row stride of the fp32 form (ld resp. e halves; e % 8 == 0): the
 * operand copy a FTB_F_A_F16 problem reads, 3 = the split [hi | lo | hi] fp16 form (row stride 3e halves). */

/* The same attention on the TENSOR CORES with split-precision operands (head width 64 or 128, else FTB_E_SHAPE -> use
 * ftb_mha): qkv16 is the in_proj output as written by a FTB_F_OUT_SPLIT problem, fp16 [rows, 9e] = [hi | lo | hi] column
 * blocks of width 3e; both contractions run as three fp16 mma.sync products on hi / lo pairs (Qh Kh + Qh Kl + Ql Kh,
 * Ph Vh + Ph Vl + Pl Vh) with fp32 accumulation and an fp32 online softmax, i.e. ~2^-20 relative operand precision --
 * FastPitch's parity margin does not allow single-pass fp16 / tf32 scores.  out / round_tf32 as ftb_mha. */
int ftb_mha_split(ftb_handle h, const void* qkv16, int e, int heads, const int64_t* key_tok,
                  int batch, int seq_len, int seq_stride, float* out, int round_tf32, void* stream);

/* ------------------------------------------------------------------------------------
 * K9  mel -> waveform: the transforms of fast Griffin-Lim
 *
 * Replaces: the loop of librosa.core.griffinlim as called by DSP.griffinlim (utils/dsp.py:89-93: n_iter = 32,
 * hop_length, win_length = n_fft = 1024, librosa defaults momentum 0.99 / Hann window / centred frames).  librosa is a
 * third-party dependency that is absent from /root/reference; the entries restate its published algorithm.
 *   window    fp32 [n_fft] (periodic Hann); twiddles fp32 [n_fft/2][2] = exp(-2 pi i k / n_fft)
 *   ftb_stft  y [batch, y_stride] (valid samples [0, y_len)), centred frames, zero padding -> spec [batch, frames,
 *             n_fft/2 + 1][2] (interleaved complex)
 *   ftb_istft spec -> window * irfft per frame (frame_buf [batch, frames, n_fft], scratch) -> overlap-add, divided by
 *             the window sum-square, trimmed to hop * (frames - 1) samples -> y [batch, y_stride]
 *   ftb_gl_update  a = rebuilt - momentum/(1+momentum) * tprev; a /= |a| + tiny; next = mag * a; tprev <- rebuilt
 * n_fft: a power of two in [64, 2048].
 * ------------------------------------------------------------------------------------ */
int ftb_stft(ftb_handle h, const float* y, int y_len, int64_t y_stride, const float* window, const float* twiddles,
             int n_fft, int hop, int batch, int frames, float* spec, void* stream);
int ftb_istft(ftb_handle h, const float* spec, const float* window, const float* twiddles, int n_fft, int hop,
              int batch, int frames, float* frame_buf, float* y, int64_t y_stride, void* stream);
int ftb_gl_update(ftb_handle h, const float* mag, const float* rebuilt, float* tprev, float* next, int64_t count,
                  float momentum, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FTB200_H_ */
