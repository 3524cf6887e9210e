"""Launch programs for the mel-synthesis path: buffers + plans built once per shape.

A `Program` is an ordered list of native launches (GEMM plans, recurrence plans, glue
kernels) over a fixed set of device buffers in the padded-flat channel-last layout
(include/ftb200.h).  Independent layers that are ready at the same time are grouped into
ONE persistent GEMM launch (e.g. the three predictors' first convolutions and the 16 bank
convolutions), so the SMs see a few hundred tiles per launch instead of ~100.
"""
import os
from typing import Callable, Dict, List, Optional

import torch

from . import runtime as rt
from .pack import PackedGemm, PackedRnn, ceil_to


def pk_ceil8(n: int) -> int:
    return ceil_to(n, 8)


# fp16 x-projections for the frame-side recurrences (BiLSTM, postnet BiGRU): halves the fp32 `W_ih x` traffic of the two
# epilogue-bound projection GEMMs for one more rounding to 11 bits.  Measured on the goldens: mel rel-L2 7.01e-4 ->
# 7.13e-4 (ForwardTacotron worst case), 7.06e-4 -> 7.27e-4 (MultiForwardTacotron), tolerance 1e-3; generate() 2.65 ->
# 2.58 ms.  The token-side recurrences (predictors, prenet) keep fp32 x-projections.  FTB_X16=0 disables it.
X16 = os.environ.get('FTB_X16', '1') != '0'
F16_OPERANDS = _F16_OPERANDS = os.environ.get('FTB_F16_OPERANDS', '1') != '0'
# Split-precision duration predictor (conv stack + W_ih as [hi | lo | hi] fp16 operands, recurrence with hi/lo h and
# W_hh, exact gates): raw duration error ~1e-5 instead of ~3e-3, so the rounded durations (common_layers.py:21) match
# the reference's except within ~1e-4 of a .5 tie.  FTB_PRECISE_DUR=0 restores the single-fp16 path (diagnostic).
PRECISE_DUR = os.environ.get('FTB_PRECISE_DUR', '1') != '0'
# Debug guard: after every program of generate() / forward(), check the activation buffers for inf / NaN and for fp16
# operand copies near the end of fp16's range (Program.check_range; synchronises, off by default)
CHECK_RANGE = os.environ.get('FTB_CHECK_RANGE', '0') == '1'
# FFT-block attention on the tensor cores with split-precision operands (ftb_attn.cu); FTB_TC_ATTENTION=0: fp32 SIMT kernel
TC_ATTENTION = os.environ.get('FTB_TC_ATTENTION', '1') != '0'
# pre_highway -> highways -> W_ih of a CBHG as ONE persistent kernel (ftb_tail.cu) instead of six GEMM launches
FUSED_TAIL = os.environ.get('FTB_FUSED_TAIL', '1') != '0'
# taps folded into the reduction for convolutions over a channel count that fills a 128-byte swizzle row badly (the
# postnet bank over the 80-channel mel: 48 k-blocks instead of 72, FTB_F_TAPS_IN_K); FTB_FOLD_TAPS=0 disables
FOLD_TAPS = os.environ.get('FTB_FOLD_TAPS', '1') != '0'
# The BiLSTM's x-projection computed once per PHONEME and read through the LengthRegulator's frame -> phoneme map
# (ftb_rnn_problem.x_idx): the regulated sequence is never written and W_ih runs over T instead of M rows.
# FTB_XPROJ_TOKENS=0 restores the materialised LengthRegulator output + frame-level projection.
XPROJ_TOKENS = os.environ.get('FTB_XPROJ_TOKENS', '1') != '0'
# CUDA graphs through the library's own stream capture (ftb_graph_*); FTB_NATIVE_GRAPHS=0: torch.cuda.CUDAGraph
NATIVE_GRAPHS = os.environ.get('FTB_NATIVE_GRAPHS', '1') != '0'
# Independent glue launches of one level (the attention / LayerNorm calls of the four token-side transformers of FastPitch)
# on parallel streams, forked from and joined to the program's stream (parallel branches of its CUDA graph)
PARALLEL_GLUE = os.environ.get('FTB_PARALLEL_GLUE', '1') != '0'


def _tile_cost(q) -> float:
    """Cycles one 128 x bn tile keeps a CTA busy (ftb_gemm.cu): the MMAs (4 per 32-channel block and tap, bn/2 cycles
    each for kind::tf32) against the epilogue of its bn/32 chunks (two warp sets drain alternate tiles), plus the
    per-tile fixed cost."""
    kelems = 64 if (q.flags & rt.F_A_F16) else 32           # channels per 128-byte k-block
    kblocks = (q.taps * q.cin + kelems - 1) // kelems if (q.flags & rt.F_TAPS_IN_K) else q.taps * ((q.cin + kelems - 1) // kelems)
    mma = kblocks * 2.0 * q.bn
    epi = (q.bn / 32.0) * 1800.0 / 2.0
    return max(mma, epi) + 1500.0


def _order_tiles(pending):
    """Problems of one launch, heaviest tile first: the kernel's static snake schedule (wave i forwards, wave i+1
    backwards over the CTAs) then hands every CTA a similar mix.  Halving the tile width of a problem whose tile count
    quantises badly (204 tiles on 148 CTAs) was measured and rejected: a 128-wide tile re-reads A from shared memory
    twice as often and runs at ~64 % of the 256-wide MMA rate (profiles/r1m_gemm_epilogue.txt)."""
    return sorted(pending, key=lambda q: -_tile_cost(q))


def capacity_bucket(n: int) -> int:
    """Sequence-length capacity a program is built for: the next multiple of 64 (256 beyond 1024).  Only buffer sizes
    follow the bucket; rows, tiles and steps follow the actual length (Program.set_seq_len), so no work is padded."""
    step = 64 if n <= 1024 else 256
    return (n + step - 1) // step * step


class Geometry:
    """Row geometry of a [batch, seq_len, C] activation in the padded-flat layout.  Buffers are allocated for
    `cap` >= seq_len rows per sentence (+ the gap); seq_len / stride / rows are the CURRENT layout inside them."""

    def __init__(self, batch: int, seq_len: int):
        self.batch, self.cap = batch, seq_len
        self.cap_rows = batch * rt.seq_stride(seq_len)
        self.set_len(seq_len)

    def set_len(self, seq_len: int):
        assert 1 <= seq_len <= self.cap, (seq_len, self.cap)
        self.seq_len = seq_len
        self.stride = rt.seq_stride(seq_len)
        self.rows = self.batch * self.stride


class _Bufs(dict):
    """name -> tensor.  Padded-flat activations are stored at capacity; dense tensors with a sequence axis (shape
    marker -1 in Program.dense) are handed out as the view of the CURRENT length over the head of their storage."""

    def __init__(self, prog):
        super().__init__()
        self.prog, self.dyn = prog, {}

    def __getitem__(self, k):
        t = dict.__getitem__(self, k)
        shape = self.dyn.get(k)
        if shape is None:
            return t
        cur = tuple(self.prog.geo.seq_len if d == -1 else d for d in shape)
        n = 1
        for d in cur:
            n *= d
        return t.view(-1)[:n].view(cur)


class Program:
    def __init__(self, handle: rt.Handle, geo: Geometry):
        self.h, self.geo = handle, geo
        self.device = handle.device
        self.bufs: Dict[str, torch.Tensor] = _Bufs(self)
        self._flat: List[str] = []                    # names of the padded-flat activation buffers
        self.steps: List[Callable[[], None]] = []
        self._gemm_plans: List[int] = []
        self._rnn_plans: List[int] = []
        self._tail_plans: List[int] = []
        self._pending: List[rt.GemmProblem] = []
        self._glue: list = []                         # queued embeddings / predictor heads: one batched launch per kind
        self._keep: list = []                         # tensors the plans point into
        self._lead: Dict[int, int] = {}               # data_ptr of a padded-flat buffer -> zero rows in front of it
        self.info: List[dict] = []                    # per step: kind, label, algorithmic flops / bytes
        self._label = ''
        # CUDA graph of the launch list: captured on the second run() (the first one runs eagerly so one-time
        # cudaFuncSetAttribute calls happen outside capture), re-captured when a launch-time scalar changes
        self.use_graph = os.environ.get('FTB_CUDA_GRAPHS', '1') != '0'
        self._graphs: Dict[tuple, object] = {}        # (launch-time scalars, seq_len) -> captured graph
        self._baked_len = False                       # a glue kernel takes seq_len as a kernel ARGUMENT
        self._torch_steps = False                     # a step runs torch operators (see call())
        self._sides: list = []                        # (stream, fork event, join event) of the parallel glue steps
        self._eager_runs = 0

    # ---- buffers ---------------------------------------------------------------------------
    def buf(self, name: str, cols: int, dtype=torch.float32, lead: int = 0) -> torch.Tensor:
        """Zero-initialised [rows, cols] activation; gap rows are never written so they stay zero.  fp16 buffers are
        GEMM-to-GEMM operands (FTB_F_OUT_F16 producer, FTB_F_A_F16 consumer).  `lead`: zero rows kept IN FRONT of the
        buffer in the same allocation (what a FTB_F_TAPS_IN_K consumer reads for the taps left of the first sentence)."""
        if name not in self.bufs:
            full = torch.zeros((lead + self.geo.cap_rows, cols), dtype=dtype, device=self.device)
            self.bufs[name] = full[lead:]
            self._lead[self.bufs[name].data_ptr()] = lead
            self._flat.append(name)
        b = self.bufs[name]
        assert b.shape[1] == cols and b.dtype == dtype, (name, b.shape, cols, b.dtype)
        return b

    def dense(self, name: str, shape, dtype=torch.float32) -> torch.Tensor:
        """Dense tensor; a -1 in `shape` marks the sequence axis: storage is sized for the capacity and
        self.bufs[name] is always the contiguous view of the current length over its head."""
        if name not in self.bufs:
            full = tuple(self.geo.cap if d == -1 else d for d in shape)
            dict.__setitem__(self.bufs, name, torch.zeros(full, dtype=dtype, device=self.device))
            if -1 in shape:
                self.bufs.dyn[name] = tuple(shape)
        return self.bufs[name]

    def check_range(self, where: str = '') -> None:
        """Debug guard (FTB_CHECK_RANGE=1, called by the models after a program ran; it synchronises): the fp16 operand
        copies are stored without saturation, so a checkpoint whose activations leave fp16's range (|x| > 65504 -- unbounded
        post-ReLU BatchNorm outputs can) turns into inf silently.  Raises naming the first buffer that holds a non-finite
        value or has used more than ~90 % of the fp16 range (FTB_F16_OPERANDS=0 runs every layer on tf32 operands)."""
        for name in self._flat:
            b = dict.__getitem__(self.bufs, name)
            v = b[:self.geo.rows]
            if not bool(torch.isfinite(v).all()):
                raise RuntimeError(f'{where}: activation buffer {name!r} ({b.dtype}) holds inf / NaN -- activations outside '
                                   'the fp16 operand range?  FTB_F16_OPERANDS=0 keeps tf32 operands')
            if b.dtype == torch.float16 and float(v.abs().max()) > 6.0e4:
                raise RuntimeError(f'{where}: fp16 operand buffer {name!r} reaches |x| = {float(v.abs().max()):.3g} '
                                   '(limit 65504): FTB_F16_OPERANDS=0 keeps tf32 operands')

    def view(self, name: str) -> torch.Tensor:
        """[batch, seq_len, C] view of a padded-flat buffer (valid rows only) at the current geometry."""
        b = self.bufs[name]
        g = self.geo
        return b[:g.rows].view(g.batch, g.stride, b.shape[1])[:, :g.seq_len]

    def set_seq_len(self, seq_len: int):
        """Re-point the whole program at sentences of `seq_len` <= capacity rows (ftb_*_plan_set_geometry): plans keep
        their buffers, tensor maps and captured graphs; rows that become gap rows under the new stride are zeroed, the
        invariant every convolution tap relies on (include/ftb200.h)."""
        self.flush()
        g = self.geo
        if seq_len == g.seq_len:
            return
        g.set_len(seq_len)
        for name in self._flat:
            b = dict.__getitem__(self.bufs, name)
            b[:g.rows].view(g.batch, g.stride, b.shape[1])[:, seq_len:].zero_()
        for plan in self._gemm_plans:
            self.h.gemm_set_geometry(plan, seq_len, g.stride)
        for plan in self._rnn_plans:
            self.h.rnn_set_geometry(plan, seq_len, g.stride)
        for plan in self._tail_plans:
            self.h.tail_set_geometry(plan, seq_len, g.stride)
        valid = g.batch * seq_len
        for i in self.info:
            if 'flops_row' in i:
                i['flops'], i['bytes'] = i['flops_row'] * valid, i['bytes_row'] * valid + i['bytes_fixed']
                if 'flops_f16_row' in i:
                    i['flops_f16'] = i['flops_f16_row'] * valid
                if 'steps' in i:
                    i['steps'] = seq_len

    # ---- GEMM problems: queued, then flushed as one persistent launch ------------------------
    def gemm(self, a: torch.Tensor, w: PackedGemm, out: Optional[torch.Tensor], *, out_col0: int = 0,
             epilogue: int = rt.EPI_AFFINE, round_out: bool = True, residual: Optional[torch.Tensor] = None,
             out_t: Optional[torch.Tensor] = None, cin: Optional[int] = None,
             out_full: Optional[torch.Tensor] = None, split_in: bool = False, split_out: bool = False):
        """split_in: `a` is a [hi | lo | hi] fp16 operand (3 * ceil8(cin) columns) and the weight its w16s form;
        split_out: `out` (fp16, 3*n columns) receives the result in that layout (FTB_F_OUT_SPLIT)."""
        g = self.geo
        p = rt.GemmProblem()
        p.a, p.rows, p.lda = a.data_ptr(), g.rows, a.shape[1]
        a16 = a.dtype == torch.float16
        o16 = out is not None and out.dtype == torch.float16
        if split_in:
            assert a16 and a.shape[1] == 3 * pk_ceil8(w.cin), (a.shape, w.cin)
            wt = w.w16s
            p.cin = a.shape[1]
        else:
            p.cin = cin if cin is not None else w.cin
            assert p.cin == w.cin and p.cin <= a.shape[1], (p.cin, w.cin, a.shape)
            wt = w.w16 if a16 else w.w
        # taps folded into K: dense fp16 rows with zero lead rows in front, a channel count that wastes swizzle rows
        fold = (FOLD_TAPS and a16 and not split_in and 1 < w.taps <= rt.FOLD_LEAD_ROWS and a.shape[1] == w.cin
                and w.cin % 64 != 0 and self._lead.get(a.data_ptr(), 0) >= rt.FOLD_LEAD_ROWS
                and -(-(w.taps * w.cin) // 64) < w.taps * -(-w.cin // 64))
        if fold:
            wt = w.w16f
        assert not split_out or (o16 and out.shape[1] % 24 == 0 and out_col0 + w.n <= out.shape[1] // 3)
        p.w, p.taps, p.tap0, p.w_tap_rows = wt.data_ptr(), w.taps, w.tap0, w.n_pad
        p.n, p.bn, p.epilogue = w.n, w.bn, epilogue
        p.flags = ((rt.F_RELU if w.relu else 0) | (rt.F_ROUND if round_out and not o16 else 0) |
                   (rt.F_A_F16 if a16 else 0) | (rt.F_OUT_F16 if o16 else 0) | (rt.F_OUT_SPLIT if split_out else 0) |
                   (rt.F_TAPS_IN_K if fold else 0))
        assert not o16 or epilogue != rt.EPI_MAXPOOL or (residual is None and out_t is None and out_full is None)
        p.bias0, p.scale, p.shift = rt.ptr(w.bias0), rt.ptr(w.scale), rt.ptr(w.shift)
        assert residual is None or residual.dtype == torch.float32, 'the residual term reads the fp32 twin'
        assert out_full is None or out_full.dtype == torch.float32
        p.residual = rt.ptr(residual)
        p.res_ld = residual.shape[1] if residual is not None else 0
        p.out = rt.ptr(out)
        p.out_ld = out.shape[1] if out is not None else 0
        p.out_col0 = out_col0
        p.out_t = rt.ptr(out_t)
        p.out_full = rt.ptr(out_full)
        assert out_full is None or (out is not None and out_full.shape == out.shape)
        p.batch, p.seq_len, p.seq_stride = g.batch, g.seq_len, g.stride
        p._label = self._label
        p._alg_cin = w.cin                     # algorithmic reduction extent (a split operand carries 3x the columns)
        self._pending.append(p)
        self._keep += [a, w, wt, out, residual, out_t, out_full]

    # ---- embeddings / predictor heads: queued, flushed as ONE batched launch per kind ---------
    def embed(self, tok, table, out, tok2=None, table2=None, extra=None, round_tf32=True):
        """nn.Embedding (+ second table / per-sentence vector concat) into a padded-flat buffer; consecutive calls share
        one launch (ftb_embed_batch)."""
        if self._glue and self._glue[0][0] != 'embed':
            self._flush_glue()
        self._glue.append(('embed', (tok, table, out, tok2, table2, extra, round_tf32)))
        self._keep += [tok, table, out, tok2, table2, extra]

    def rowdot(self, x, w, bias, n_out, alpha, y):
        """Predictor head (lin + / alpha); `alpha` may be a callable evaluated at launch time.  Consecutive calls
        share one launch (ftb_rowdot_batch)."""
        if self._glue and self._glue[0][0] != 'rowdot':
            self._flush_glue()
        self._glue.append(('rowdot', (x, w, bias, n_out, alpha, y)))
        self._keep += [x, w, bias, y]

    def _flush_glue(self):
        g = self.geo
        if self._glue:
            self._baked_len = True
        while self._glue:
            kind = self._glue[0][0]
            items = [it for _, it in self._glue[:rt.MAX_GLUE_ITEMS]]
            self._glue = self._glue[rt.MAX_GLUE_ITEMS:]
            if kind == 'embed':
                self.steps.append(lambda items=items: self.h.embed_batch(items, g.batch, g.seq_len, g.stride))
            else:
                self.steps.append(lambda items=items: self.h.rowdot_batch(
                    [(x, w, b, n, (a() if callable(a) else a), y) for x, w, b, n, a, y in items],
                    g.batch, g.seq_len, g.stride))
            self.info.append({'kind': 'glue', 'label': f'{kind} x{len(items)}', 'flops': 0.0, 'bytes': 0.0})

    def flush(self):
        """Close the current group: everything queued since the last flush runs as one launch."""
        self._flush_glue()
        if not self._pending:
            return
        g = self.geo
        probs = _order_tiles(self._pending)
        self._pending = []
        valid = g.batch * g.seq_len
        for i in range(0, len(probs), 32):
            grp = probs[i:i + 32]
            plan = self.h.gemm_plan(grp)
            self._gemm_plans.append(plan)
            self.steps.append(lambda plan=plan: self.h.gemm_run(plan))
            # algorithmic work: valid rows only (gap rows and the maxpool halo are overhead)
            flops = sum(2.0 * valid * q._alg_cin * q.taps * q.n for q in grp)
            flops_f16 = sum(2.0 * valid * q._alg_cin * q.taps * q.n for q in grp if q.flags & rt.F_A_F16)
            nbytes = sum((2.0 if q.flags & rt.F_A_F16 else 4.0) * (valid * q.cin + q.taps * q.n * q.cin) +
                         (2.0 if q.flags & rt.F_OUT_F16 else 4.0) * valid * q.n for q in grp)
            labels = sorted({getattr(q, '_label', '') for q in grp})
            wbytes = sum((2.0 if q.flags & rt.F_A_F16 else 4.0) * q.taps * q.n * q.cin for q in grp)
            self.info.append({'kind': 'gemm', 'label': '+'.join(l for l in labels if l) or 'gemm',
                              'flops': flops, 'flops_f16': flops_f16, 'bytes': nbytes, 'problems': len(grp),
                              'flops_row': flops / valid, 'flops_f16_row': flops_f16 / valid,
                              'bytes_row': (nbytes - wbytes) / valid, 'bytes_fixed': wbytes,
                              **self.h.gemm_info(plan)})

    def rnn(self, specs, lens: Optional[torch.Tensor] = None, pad_value: float = 0.0):
        """specs: list of (PackedRnn, xproj buffer, out buffer, out_col0[, xmap]) run by one plan (cluster or cooperative
        launches).  `lens` (int32 [batch], device): packed-sequence semantics for every problem of the launch.
        xmap = x_idx int32 [batch, cap] (absolute rows of xproj): the problem reads its x-projection rows through the
        LengthRegulator's frame -> row map (ftb_rnn_problem.x_idx, ftb_lr_index)."""
        self.flush()
        g = self.geo
        probs = []
        specs = [tuple(sp) + (None,) * (5 - len(sp)) for sp in specs]
        for r, xproj, out, col0, xmap in specs:
            p = rt.RnnProblem()
            if xmap is not None:
                p.x_idx = xmap.data_ptr()
                self._keep.append(xmap)
            p.cell, p.hidden = r.cell, r.hidden
            p.batch, p.seq_len, p.seq_stride = g.batch, g.seq_len, g.stride
            p.xproj, p.w_hh, p.b_hh = xproj.data_ptr(), r.w_hh.data_ptr(), r.b_hh.data_ptr()
            p.out, p.out_ld, p.out_col0 = out.data_ptr(), out.shape[1], col0
            p.lens, p.pad_value = (lens.data_ptr() if lens is not None else None), float(pad_value)
            p.out_f16 = int(out.dtype == torch.float16)       # h_t is fp16-exact: an fp16 layer output loses nothing
            p.x_f16 = int(xproj.dtype == torch.float16)
            p.precise = int(getattr(r, 'precise', False))
            probs.append(p)
            self._keep += [r, xproj, out] + ([lens] if lens is not None else [])
        plan = self.h.rnn_plan(probs)
        self._rnn_plans.append(plan)
        self.steps.append(lambda plan=plan: self.h.rnn_run(plan))
        valid = g.batch * g.seq_len
        flops = sum(2.0 * 2 * valid * r.hidden * (4 if r.cell else 3) * r.hidden for r, *_ in specs)
        # x-projection read + h write per step, W_hh once
        nbytes = sum(4.0 * valid * (2 * (4 if r.cell else 3) * r.hidden + 2 * r.hidden) + 4.0 * r.w_hh.numel()
                     for r, *_ in specs)
        wbytes = sum(4.0 * r.w_hh.numel() for r, *_ in specs)
        self.info.append({'kind': 'rnn', 'label': self._label or 'rnn', 'flops': flops, 'bytes': nbytes,
                          'flops_row': flops / valid, 'bytes_row': (nbytes - wbytes) / valid, 'bytes_fixed': wbytes,
                          'steps': g.seq_len, 'problems': len(specs), **self.h.rnn_info(plan)})

    def cbhg_tail(self, a: torch.Tensor, c: 'CbhgPack', xp: torch.Tensor):
        """pre_highway -> highways -> W_ih projection of one CBHG as ONE fused launch (include/ftb200.h, K2f):
        a = conv_project2's fp16 operand copy [rows, proj2.n], xp = the x-projection buffer (fp32 or fp16)."""
        self.flush()
        g = self.geo
        w_pre = c.pre_highway.w16[0]                                             # [256, cin]
        w_hw = torch.cat([w.w16[0] for w in c.highways], dim=0).contiguous()      # [nh * 512, 256]
        b_hw = torch.cat([w.bias0 for w in c.highways]).contiguous()              # [nh * 512]
        w_ih, b_ih = c.rnn.ih.w16[0], c.rnn.ih.bias0
        p = rt.CbhgTailProblem()
        p.a, p.lda, p.cin, p.channels = a.data_ptr(), a.shape[1], c.pre_highway.cin, c.channels
        p.w_pre, p.n_highways, p.w_hw, p.b_hw = w_pre.data_ptr(), len(c.highways), w_hw.data_ptr(), b_hw.data_ptr()
        p.w_ih, p.b_ih, p.n_ih = w_ih.data_ptr(), b_ih.data_ptr(), c.rnn.ih.n
        p.out, p.out_ld, p.out_f16 = xp.data_ptr(), xp.shape[1], int(xp.dtype == torch.float16)
        p.batch, p.seq_len, p.seq_stride = g.batch, g.seq_len, g.stride
        plan = self.h.tail_plan(p)
        self._tail_plans.append(plan)
        self._keep += [a, w_pre, w_hw, b_hw, w_ih, b_ih, xp]
        self.steps.append(lambda plan=plan: self.h.tail_run(plan))
        valid = g.batch * g.seq_len
        ch = c.channels
        per_row = 2.0 * (c.pre_highway.cin * ch + len(c.highways) * 2 * ch * ch + ch * c.rnn.ih.n)
        wbytes = 2.0 * (w_pre.numel() + w_hw.numel() + w_ih.numel())
        brow = 2.0 * c.pre_highway.cin + xp.element_size() * c.rnn.ih.n
        self.info.append({'kind': 'gemm', 'label': self._label or 'cbhg_tail', 'flops': per_row * valid,
                          'flops_f16': per_row * valid, 'bytes': brow * valid + wbytes, 'problems': 2 + len(c.highways),
                          'flops_row': per_row, 'flops_f16_row': per_row, 'bytes_row': brow, 'bytes_fixed': wbytes,
                          'cta_pairs': False, 'slots': (g.rows + 127) // 128, 'fused_tail': True})

    def call(self, fn: Callable[[], None], label: str = 'glue', nbytes: float = 0.0, torch_ops: bool = False):
        """A step that is not a plan: a C-ABI call, or (torch_ops) torch operators -- those allocate through torch's
        caching allocator, so a program that holds one is captured with torch.cuda.graph instead of the library's own
        stream capture."""
        self.flush()
        self._baked_len = True
        self._torch_steps = self._torch_steps or torch_ops
        self.steps.append(fn)
        self.info.append({'kind': 'glue', 'label': label, 'flops': 0.0, 'bytes': nbytes, 'call': not torch_ops})

    # ---- levels: the steps queued between begin_level() and end_level() are independent of each other ----------
    def begin_level(self):
        self.flush()
        self._level_mark = len(self.steps)

    def end_level(self):
        """Close a level (run_levels).  When it consists of two or more C-ABI call steps -- one per branch: the attention
        or LayerNorm launches of independent transformers, each a few microseconds of work on a fraction of the SMs --
        they become ONE step that runs them on parallel streams."""
        self.flush()
        n0 = getattr(self, '_level_mark', len(self.steps))
        infos = self.info[n0:]
        if not (PARALLEL_GLUE and len(infos) >= 2 and all(i.get('call') for i in infos)):
            return
        fns = self.steps[n0:]
        del self.steps[n0:], self.info[n0:]
        while len(self._sides) < len(fns) - 1:
            self._sides.append((torch.cuda.Stream(self.device), torch.cuda.Event(), torch.cuda.Event()))
        sides = self._sides[:len(fns) - 1]

        def run(fns=fns, sides=sides):
            main = torch.cuda.current_stream(self.device)
            fork = sides[0][1]
            fork.record(main)
            for fn, (st, _, done) in zip(fns[1:], sides):
                st.wait_event(fork)
                with torch.cuda.stream(st):
                    fn()
                    done.record(st)
            fns[0]()
            for _, _, done in sides:
                main.wait_event(done)
        self.steps.append(run)
        self.info.append({'kind': 'glue', 'label': ' | '.join(i['label'] for i in infos), 'flops': 0.0,
                          'bytes': sum(i['bytes'] for i in infos), 'parallel': len(fns)})

    def label(self, text: str):
        self._label = text

    def run_eager(self):
        self.flush()
        for s in self.steps:
            s()

    def run(self):
        """Launch every step on the current stream: eagerly the first time, afterwards as ONE CUDA-graph
        replay (a 20-launch program otherwise leaves the SMs idle between its short kernels)."""
        # baked into kernel arguments of a captured graph: launch-time scalars, and (glue kernels only -- the GEMM and
        # recurrence kernels read their geometry from the device records) the sequence length
        if CHECK_RANGE:
            self.run_eager()
            return self.check_range('forwardtacotron_b200 (FTB_CHECK_RANGE=1)')
        key = (getattr(self, 'alpha', None), self.geo.seq_len if self._baked_len else None)
        if not self.use_graph:
            return self.run_eager()
        graph = self._graphs.get(key)
        if graph is not None:
            self.h.replayed_launches += self._launches_per_run     # kernels inside the replayed graph
            return self.h.graph_launch(graph) if isinstance(graph, int) else graph.replay()
        if self._eager_runs < 1:
            self._eager_runs += 1
            n0 = self.h.launch_count()
            self.run_eager()
            self._launches_per_run = self.h.launch_count() - n0
            return
        try:
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            if self._torch_steps or not NATIVE_GRAPHS:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side):
                    self.run_eager()
            else:
                # the library's own stream capture (ftb_graph_*): a torch.cuda.CUDAGraph replay prepends two one-element
                # fill kernels for its generator state (~8 us of stream time, three replays per generate())
                graph = self.h.graph_capture(self.run_eager, side)
            torch.cuda.current_stream(self.device).wait_stream(side)
        except Exception as exc:                              # capture unsupported: stay on eager launches
            import warnings
            warnings.warn(f'forwardtacotron_b200: CUDA-graph capture failed ({exc}); using eager launches')
            self.use_graph = False
            return self.run_eager()
        while len(self._graphs) >= 8:                         # a few distinct lengths / scalars stay captured
            old = self._graphs.pop(next(iter(self._graphs)))
            if isinstance(old, int):
                self.h.graph_destroy(old)
        self._graphs[key] = graph
        # (the capture pass already counted these launches once in the C counter)
        return self.h.graph_launch(graph) if isinstance(graph, int) else graph.replay()

    def run_timed(self):
        """Like run(), with a CUDA-event pair around every launch on the current stream.
        Returns [(info, milliseconds)] after a device synchronize (profiling aid for bench.py)."""
        evs = []
        self.flush()
        for s in self.steps:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            s()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize(self.device)
        return [(i, a.elapsed_time(b)) for i, (a, b) in zip(self.info, evs)]

    def close(self):
        xp = self.__dict__.pop('xprog', None)          # a token program's companion (models/_accel.token_x_program)
        if xp is not None:
            xp.close()
        for p in self._gemm_plans:
            self.h.gemm_destroy(p)
        for p in self._rnn_plans:
            self.h.rnn_destroy(p)
        for p in self._tail_plans:
            self.h.tail_destroy(p)
        self._tail_plans = []
        for g in self._graphs.values():
            if isinstance(g, int):
                self.h.graph_destroy(g)
        self._graphs = {}
        self._gemm_plans, self._rnn_plans, self.steps = [], [], []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# -------------------------------------------------------------------------------------------
# Layer groups.  Each `stage_*` function queues the GEMMs of ONE dependency level; the caller
# interleaves levels of independent branches and flushes between levels.
# -------------------------------------------------------------------------------------------

class CbhgPack:
    """Packed weights of one CBHG (common_layers.py:60-124)."""

    def __init__(self, bank: List[PackedGemm], proj1: PackedGemm, proj2: PackedGemm, pre_highway: PackedGemm,
                 highways: List[PackedGemm], rnn: PackedRnn, in_channels: int, channels: int):
        self.bank, self.proj1, self.proj2 = bank, proj1, proj2
        self.pre_highway, self.highways, self.rnn = pre_highway, highways, rnn
        self.in_channels, self.channels = in_channels, channels


def cbhg_levels(prog: Program, tag: str, c: CbhgPack, x: torch.Tensor, out: torch.Tensor, out_col0: int = 0,
                x_full: Optional[torch.Tensor] = None, xproj_f16: bool = False,
                x_split: Optional[torch.Tensor] = None):
    """Returns the list of per-level closures for one CBHG over input buffer x [rows, in_channels];
    level i may be queued only after level i-1 has been flushed.  The last entry runs the BiGRU."""
    K = len(c.bank)
    ch = c.channels
    # bank -> proj1 -> proj2 are GEMM-to-GEMM hand-offs with K = 3 * K*ch and 3 * proj1.n: fp16 operands (the same
    # 11-bit significand the tf32 path keeps) halve their traffic and double the MMA rate
    f16 = _F16_OPERANDS and (K * ch) % 8 == 0 and c.proj1.n % 8 == 0
    # x_split (the packed-sequence frame program): the whole conv chain bank -> proj1 -> proj2 runs split-precision,
    # because the |value| ~ 30 padding rows next to a packed sentence reach its valid frames through every conv tap
    chain_split = x_split is not None and c.proj1.w_src is not None and c.proj2.w_src is not None
    if chain_split:
        bank = prog.buf(f'{tag}.bank', 3 * K * ch, torch.float16)
        p1 = prog.buf(f'{tag}.proj1', 3 * c.proj1.n, torch.float16)
    else:
        bank = prog.buf(f'{tag}.bank', K * ch, torch.float16 if f16 else torch.float32)
        p1 = prog.buf(f'{tag}.proj1', c.proj1.n, torch.float16 if f16 else torch.float32)
    # proj2 -> pre_highway -> highways -> gru_ih: small-K layers paced by operand delivery from L2
    # (profiles/r1r_highway_ncu.txt): fp16 operand copies halve it; the residual terms keep their fp32 twins
    f16s = _F16_OPERANDS and c.proj2.n % 8 == 0 and ch % 8 == 0
    sdt = torch.float16 if f16s else torch.float32
    p2 = prog.buf(f'{tag}.proj2', c.proj2.n, sdt)
    xp = prog.buf(f'{tag}.xproj', c.rnn.ih.n, torch.float16 if (xproj_f16 and c.rnn.ih.n % 8 == 0) else torch.float32)
    levels = []

    def named(name, fn):
        def run():
            prog.label(f'{tag}.{name}')
            fn()
            prog.label('')
        return run

    def l_bank():   # conv bank k=1..K, each trimmed to L, ReLU->BN, + MaxPool1d(2,1,1)[:L] fused (:97-105)
        for k, w in enumerate(c.bank):
            if x_split is not None:          # [hi | lo | hi] copy of x (split-precision operand, PackedGemm.w16s)
                prog.gemm(x_split, w, bank, out_col0=k * ch, epilogue=rt.EPI_MAXPOOL, split_in=True,
                          split_out=chain_split)
            else:
                prog.gemm(x, w, bank, out_col0=k * ch, epilogue=rt.EPI_MAXPOOL)
    levels.append(named('bank', l_bank))
    levels.append(named('proj1', lambda: prog.gemm(bank, c.proj1, p1, split_in=chain_split,       # :109
                                                   split_out=chain_split)))
    levels.append(named('proj2', lambda: prog.gemm(p1, c.proj2, p2, split_in=chain_split,         # :111-114
                                                   residual=x_full if x_full is not None else x)))
    if (FUSED_TAIL and f16s and ch == 256 and c.pre_highway.cin % 8 == 0 and c.rnn.ih.n % 16 == 0
            and all(w.bn == 256 for w in c.highways)):
        # :118-123 (W_ih half) as one fused launch: the highway activations never leave the SM
        def l_tail():
            prog.label(f'{tag}.tail')
            prog.cbhg_tail(p2, c, xp)
            prog.label('')
        levels.append(l_tail)
        return levels, (c.rnn, xp, out, out_col0)
    hw = [prog.buf(f'{tag}.hw0', ch, sdt), prog.buf(f'{tag}.hw1', ch, sdt)]
    # unrounded twins of the highway activations: the residual term (1-g)*x uses full fp32 values, only the
    # GEMM operand copy is tf32-rounded
    hwf = [prog.buf(f'{tag}.hw0_full', ch), prog.buf(f'{tag}.hw1_full', ch)]
    levels.append(named('pre_highway', lambda: prog.gemm(p2, c.pre_highway, hw[0], out_full=hwf[0])))   # :118
    for i, w in enumerate(c.highways):                                                            # :119-120
        levels.append(named(f'highway{i}', lambda i=i, w=w: prog.gemm(
            hw[i & 1], w, hw[(i + 1) & 1], epilogue=rt.EPI_HIGHWAY, residual=hwf[i & 1],
            out_full=hwf[(i + 1) & 1])))
    last = hw[len(c.highways) & 1]
    levels.append(named('gru_ih', lambda: prog.gemm(last, c.rnn.ih, xp, round_out=False)))        # W_ih half of :123
    return levels, (c.rnn, xp, out, out_col0)


class PredictorPack:
    """Packed weights of one SeriesPredictor (forward_tacotron.py:14-39 and the multi variants)."""

    def __init__(self, convs: List[PackedGemm], rnn: PackedRnn, lin_w: torch.Tensor, lin_b: torch.Tensor,
                 precise: bool = False):
        self.convs, self.rnn, self.precise = convs, rnn, precise
        self.lin_w = lin_w.detach().float().contiguous()
        self.lin_b = lin_b.detach().float().contiguous()


def predictor_emb_buffer(prog: Program, tag: str, p: PredictorPack, width: int) -> torch.Tensor:
    """Input buffer of a predictor's first convolution: fp32 (tf32-rounded by the embedding kernel), or for a precise
    predictor the [hi | lo | hi] fp16 operand (embed mode 3)."""
    if p.precise:
        return prog.buf(f'{tag}.emb', 3 * pk_ceil8(width), torch.float16)
    return prog.buf(f'{tag}.emb', width)


def predictor_levels(prog: Program, tag: str, p: PredictorPack, emb: torch.Tensor):
    ch = p.convs[0].n
    if p.precise:
        return _precise_predictor_levels(prog, tag, p, emb)
    # conv0 -> conv1 -> conv2 -> gru_ih are GEMM-to-GEMM hand-offs: fp16 operand buffers (see cbhg_levels)
    cdt = torch.float16 if (_F16_OPERANDS and ch % 8 == 0) else torch.float32
    c = [prog.buf(f'{tag}.c0', ch, cdt), prog.buf(f'{tag}.c1', ch, cdt)]
    xp = prog.buf(f'{tag}.xproj', p.rnn.ih.n)
    h = prog.buf(f'{tag}.h', 2 * p.rnn.hidden)
    def named(name, fn):
        def run():
            prog.label(f'{tag}.{name}')
            fn()
            prog.label('')
        return run
    levels = [named('conv0', lambda: prog.gemm(emb, p.convs[0], c[0])),
              named('conv1', lambda: prog.gemm(c[0], p.convs[1], c[1])),
              named('conv2', lambda: prog.gemm(c[1], p.convs[2], c[0])),
              named('gru_ih', lambda: prog.gemm(c[0], p.rnn.ih, xp, round_out=False))]
    return levels, (p.rnn, xp, h, 0), h


def _precise_predictor_levels(prog: Program, tag: str, p: PredictorPack, emb: torch.Tensor):
    """The split-precision predictor: every activation travels as [hi | lo | hi] fp16 column blocks, every weight as
    [W_hi | W_hi | W_lo] (PackedGemm.w16s); the W_ih projection is written as unrounded fp32 and the recurrence runs with
    ftb_rnn_problem.precise.  3x the MMA work of the single-fp16 path on ~2.6 % of the model's FLOPs."""
    ch = p.convs[0].n
    assert ch % 8 == 0
    c = [prog.buf(f'{tag}.c0', 3 * ch, torch.float16), prog.buf(f'{tag}.c1', 3 * ch, torch.float16)]
    xp = prog.buf(f'{tag}.xproj', p.rnn.ih.n)
    h = prog.buf(f'{tag}.h', 2 * p.rnn.hidden)

    def named(name, fn):
        def run():
            prog.label(f'{tag}.{name}')
            fn()
            prog.label('')
        return run
    levels = [named('conv0', lambda: prog.gemm(emb, p.convs[0], c[0], split_in=True, split_out=True)),
              named('conv1', lambda: prog.gemm(c[0], p.convs[1], c[1], split_in=True, split_out=True)),
              named('conv2', lambda: prog.gemm(c[1], p.convs[2], c[0], split_in=True, split_out=True)),
              named('gru_ih', lambda: prog.gemm(c[0], p.rnn.ih, xp, round_out=False, split_in=True))]
    return levels, (p.rnn, xp, h, 0), h


def run_levels(prog: Program, branches: List[list]):
    """Queue level i of every branch, flush, then level i+1, ...  (branches are independent)."""
    depth = max(len(b) for b in branches)
    for i in range(depth):
        prog.begin_level()
        for b in branches:
            if i < len(b):
                b[i]()
        prog.end_level()
