"""ctypes binding of libftb200.so (include/ftb200.h) and the per-device handle.

This module is the ONLY place that touches the native library.  There is no CPU or
PyTorch fallback: if the library is missing, or the device is not sm_100, every entry
raises (north_star: "no CPU fallback").  PyTorch is used for device memory and streams only.
"""
import ctypes as C
import os
from typing import Dict, List, Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libftb200.so')

EPI_AFFINE, EPI_MAXPOOL, EPI_HIGHWAY = 0, 1, 2
F_RELU, F_ROUND, F_A_F16, F_OUT_F16, F_OUT_SPLIT, F_TAPS_IN_K = 1, 2, 4, 8, 16, 32
FOLD_LEAD_ROWS = 8
CELL_GRU, CELL_LSTM = 0, 1
SEQ_GAP = 8          # zero rows between sentences in the padded-flat layout (>= largest conv reach)

_fp = C.c_void_p


class GemmProblem(C.Structure):
    _fields_ = [('a', _fp), ('rows', C.c_int64), ('lda', C.c_int32), ('cin', C.c_int32),
                ('w', _fp), ('taps', C.c_int32), ('tap0', C.c_int32), ('w_tap_rows', C.c_int32),
                ('n', C.c_int32), ('bn', C.c_int32), ('epilogue', C.c_int32), ('flags', C.c_int32),
                ('bias0', _fp), ('scale', _fp), ('shift', _fp), ('residual', _fp), ('res_ld', C.c_int32),
                ('out', _fp), ('out_ld', C.c_int32), ('out_col0', C.c_int32),
                ('out_t', _fp), ('batch', C.c_int32), ('seq_len', C.c_int32), ('seq_stride', C.c_int32),
                ('out_full', _fp)]


class RnnProblem(C.Structure):
    _fields_ = [('cell', C.c_int32), ('hidden', C.c_int32), ('batch', C.c_int32), ('seq_len', C.c_int32),
                ('seq_stride', C.c_int32), ('xproj', _fp), ('w_hh', _fp), ('b_hh', _fp), ('out', _fp),
                ('out_ld', C.c_int32), ('out_col0', C.c_int32), ('lens', _fp), ('pad_value', C.c_float),
                ('out_f16', C.c_int32), ('x_f16', C.c_int32), ('precise', C.c_int32),
                ('x_idx', _fp)]


class CbhgTailProblem(C.Structure):
    _fields_ = [('a', _fp), ('lda', C.c_int32), ('cin', C.c_int32), ('channels', C.c_int32), ('w_pre', _fp),
                ('n_highways', C.c_int32), ('w_hw', _fp), ('b_hw', _fp), ('w_ih', _fp), ('b_ih', _fp),
                ('n_ih', C.c_int32), ('out', _fp), ('out_ld', C.c_int32), ('out_f16', C.c_int32),
                ('batch', C.c_int32), ('seq_len', C.c_int32), ('seq_stride', C.c_int32)]


class EmbedItem(C.Structure):
    _fields_ = [('tok', _fp), ('table', _fp), ('e1', C.c_int32),
                ('tok2', _fp), ('table2', _fp), ('e2', C.c_int32),
                ('extra', _fp), ('e3', C.c_int32),
                ('out', _fp), ('out_ld', C.c_int32), ('round_tf32', C.c_int32), ('rows1', C.c_int32), ('rows2', C.c_int32)]


class RowdotItem(C.Structure):
    _fields_ = [('x', _fp), ('x_ld', C.c_int32), ('k', C.c_int32), ('w', _fp), ('bias', _fp),
                ('n_out', C.c_int32), ('alpha', C.c_float), ('y', _fp)]


MAX_GLUE_ITEMS = 8


class LrGatherArgs(C.Structure):
    _fields_ = [('x', _fp), ('x_stride', C.c_int32), ('x_ld', C.c_int32), ('c1', C.c_int32),
                ('extra', _fp), ('c2', C.c_int32),
                ('pitch', _fp), ('energy', _fp),
                ('wp', _fp), ('bp', _fp), ('ps', C.c_float),
                ('we', _fp), ('be', _fp), ('es', C.c_float),
                ('cum', _fp), ('lens', _fp),
                ('batch', C.c_int32), ('T', C.c_int32), ('M', C.c_int32),
                ('out', _fp), ('out_stride', C.c_int32), ('out_ld', C.c_int32),
                ('idx', _fp), ('round_tf32', C.c_int32)]


# every symbol include/ftb200.h declares (tests check that the .so exports all of them)
SYMBOLS = ['ftb_version', 'ftb_error_string', 'ftb_create', 'ftb_destroy', 'ftb_launch_count',
           'ftb_graph_capture_begin', 'ftb_graph_capture_end', 'ftb_graph_launch', 'ftb_graph_destroy',
           'ftb_gemm_plan_create', 'ftb_gemm_plan_run', 'ftb_gemm_plan_destroy', 'ftb_gemm_plan_info',
           'ftb_gemm_plan_set_geometry', 'ftb_rnn_plan_set_geometry',
           'ftb_cbhg_tail_plan_create', 'ftb_cbhg_tail_plan_run', 'ftb_cbhg_tail_plan_set_geometry',
           'ftb_cbhg_tail_plan_destroy',
           'ftb_rnn_plan_create', 'ftb_rnn_plan_run', 'ftb_rnn_plan_destroy', 'ftb_rnn_plan_info',
           'ftb_lr_scan', 'ftb_lr_gather', 'ftb_lr_index', 'ftb_lr_regulate_static', 'ftb_embed', 'ftb_rowdot', 'ftb_embed_batch', 'ftb_rowdot_batch',
           'ftb_argmax_lastdim',
           'ftb_add_layernorm', 'ftb_posenc', 'ftb_mha', 'ftb_mha_split', 'ftb_stft', 'ftb_istft', 'ftb_gl_update']

_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f'{LIB_PATH} not found: build it with `python -m forwardtacotron_b200.build` '
                           '(there is no CPU fallback for the mel-synthesis path)')
    lib = C.CDLL(LIB_PATH)
    lib.ftb_error_string.restype = C.c_char_p
    lib.ftb_launch_count.restype = C.c_int64
    lib.ftb_launch_count.argtypes = [_fp]
    lib.ftb_create.argtypes = [C.c_int, C.POINTER(_fp)]
    lib.ftb_destroy.argtypes = [_fp]
    lib.ftb_gemm_plan_create.argtypes = [_fp, C.POINTER(GemmProblem), C.c_int, C.POINTER(C.c_int)]
    lib.ftb_gemm_plan_run.argtypes = [_fp, C.c_int, C.c_int, _fp]
    lib.ftb_gemm_plan_destroy.argtypes = [_fp, C.c_int]
    lib.ftb_gemm_plan_info.argtypes = [_fp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.ftb_gemm_plan_set_geometry.argtypes = [_fp, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_rnn_plan_set_geometry.argtypes = [_fp, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_cbhg_tail_plan_create.argtypes = [_fp, C.POINTER(CbhgTailProblem), C.POINTER(C.c_int)]
    lib.ftb_cbhg_tail_plan_run.argtypes = [_fp, C.c_int, _fp]
    lib.ftb_cbhg_tail_plan_set_geometry.argtypes = [_fp, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_cbhg_tail_plan_destroy.argtypes = [_fp, C.c_int]
    lib.ftb_rnn_plan_create.argtypes = [_fp, C.POINTER(RnnProblem), C.c_int, C.POINTER(C.c_int)]
    lib.ftb_rnn_plan_run.argtypes = [_fp, C.c_int, _fp]
    lib.ftb_rnn_plan_destroy.argtypes = [_fp, C.c_int]
    lib.ftb_embed_batch.argtypes = [_fp, C.POINTER(EmbedItem), C.c_int, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_rowdot_batch.argtypes = [_fp, C.POINTER(RowdotItem), C.c_int, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_rnn_plan_info.argtypes = [_fp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.ftb_lr_scan.argtypes = [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp, _fp, _fp, _fp]
    lib.ftb_lr_gather.argtypes = [_fp, C.POINTER(LrGatherArgs), _fp]
    lib.ftb_graph_capture_begin.argtypes = [_fp, _fp]
    lib.ftb_graph_capture_end.argtypes = [_fp, _fp, C.POINTER(C.c_int)]
    lib.ftb_graph_launch.argtypes = [_fp, C.c_int, _fp]
    lib.ftb_graph_destroy.argtypes = [_fp, C.c_int]
    lib.ftb_lr_index.argtypes = [_fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp]
    lib.ftb_lr_regulate_static.argtypes = [_fp, _fp, C.c_int, C.c_int, C.c_int, _fp, C.c_int, C.c_int, C.c_int, C.c_int,
                                           _fp, C.c_int, C.c_int, _fp, _fp, _fp, _fp]
    lib.ftb_embed.argtypes = [_fp, _fp, _fp, C.c_int, _fp, _fp, C.c_int, _fp, C.c_int,
                              C.c_int, C.c_int, C.c_int, _fp, C.c_int, C.c_int, C.c_int, C.c_int, _fp]
    lib.ftb_rowdot.argtypes = [_fp, _fp, C.c_int, C.c_int, _fp, _fp, C.c_int, C.c_float,
                               C.c_int, C.c_int, C.c_int, _fp, _fp]
    lib.ftb_argmax_lastdim.argtypes = [_fp, _fp, C.c_int64, C.c_int, _fp, _fp]
    lib.ftb_add_layernorm.argtypes = [_fp, _fp, _fp, C.c_int, C.c_int, _fp, _fp, _fp, _fp,
                                      C.c_int, C.c_int, C.c_int, _fp, _fp, C.c_int, _fp]
    lib.ftb_posenc.argtypes = [_fp, _fp, _fp, _fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int,
                               _fp, _fp, C.c_int, C.c_int, _fp]
    lib.ftb_mha.argtypes = [_fp, _fp, C.c_int, C.c_int, _fp, C.c_int, C.c_int, C.c_int, _fp, C.c_int, _fp]
    lib.ftb_mha_split.argtypes = [_fp, _fp, C.c_int, C.c_int, _fp, C.c_int, C.c_int, C.c_int, _fp, C.c_int, _fp]
    lib.ftb_stft.argtypes = [_fp, _fp, C.c_int, C.c_int64, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp]
    lib.ftb_istft.argtypes = [_fp, _fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, _fp, _fp, C.c_int64, _fp]
    lib.ftb_gl_update.argtypes = [_fp, _fp, _fp, _fp, _fp, C.c_int64, C.c_float, _fp]
    _lib = lib
    return lib


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def seq_stride(seq_len: int) -> int:
    """Rows per sentence in the padded-flat layout (include/ftb200.h)."""
    return seq_len + SEQ_GAP


class Handle:
    """One ftb_handle per CUDA device; raises RuntimeError on any non-zero status."""

    _per_device: Dict[int, 'Handle'] = {}

    def __init__(self, device: torch.device):
        if device.type != 'cuda':
            raise RuntimeError('forwardtacotron_b200 runs on CUDA (sm_100a) only; got device ' + str(device))
        self.lib = load_library()
        self.device = device
        self.index = device.index if device.index is not None else torch.cuda.current_device()
        h = _fp()
        self.check(self.lib.ftb_create(self.index, C.byref(h)))
        self.h = h
        self.num_sms = torch.cuda.get_device_properties(self.index).multi_processor_count
        self.replayed_launches = 0      # kernels launched through CUDA-graph replays (engine.Program)
        self.timing = None              # bench.py: a list -> (name, start event, end event) of the K5/K6 launches
        self.impl = int(os.environ.get('FTB_GEMM_IMPL', '0'))    # 1 = SIMT cross-check kernel (tests only)

    @classmethod
    def get(cls, device: torch.device) -> 'Handle':
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if idx not in cls._per_device:
            cls._per_device[idx] = Handle(torch.device('cuda', idx))
        return cls._per_device[idx]

    def check(self, code: int):
        if code != 0:
            msg = self.lib.ftb_error_string(code).decode()
            raise RuntimeError(f'libftb200: {msg} (code {code})')

    def stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def launch_count(self) -> int:
        return int(self.lib.ftb_launch_count(self.h)) + self.replayed_launches

    # ---- plans -----------------------------------------------------------------------------
    def gemm_plan(self, problems: List[GemmProblem]) -> int:
        arr = (GemmProblem * len(problems))(*problems)
        pid = C.c_int()
        self.check(self.lib.ftb_gemm_plan_create(self.h, arr, len(problems), C.byref(pid)))
        return pid.value

    def gemm_run(self, plan: int, impl: Optional[int] = None):
        self.check(self.lib.ftb_gemm_plan_run(self.h, plan, self.impl if impl is None else impl, self.stream()))

    def gemm_set_geometry(self, plan: int, seq_len: int, stride: int):
        self.check(self.lib.ftb_gemm_plan_set_geometry(self.h, plan, seq_len, stride, self.stream()))

    def rnn_set_geometry(self, plan: int, seq_len: int, stride: int):
        self.check(self.lib.ftb_rnn_plan_set_geometry(self.h, plan, seq_len, stride, self.stream()))

    def tail_plan(self, problem: CbhgTailProblem) -> int:
        pid = C.c_int()
        self.check(self.lib.ftb_cbhg_tail_plan_create(self.h, C.byref(problem), C.byref(pid)))
        return pid.value

    def tail_run(self, plan: int):
        self.check(self.lib.ftb_cbhg_tail_plan_run(self.h, plan, self.stream()))

    def tail_set_geometry(self, plan: int, seq_len: int, stride: int):
        self.check(self.lib.ftb_cbhg_tail_plan_set_geometry(self.h, plan, seq_len, stride, self.stream()))

    def tail_destroy(self, plan: int):
        self.check(self.lib.ftb_cbhg_tail_plan_destroy(self.h, plan))

    def gemm_destroy(self, plan: int):
        self.check(self.lib.ftb_gemm_plan_destroy(self.h, plan))

    def rnn_plan(self, problems: List[RnnProblem]) -> int:
        arr = (RnnProblem * len(problems))(*problems)
        pid = C.c_int()
        self.check(self.lib.ftb_rnn_plan_create(self.h, arr, len(problems), C.byref(pid)))
        return pid.value

    def rnn_run(self, plan: int):
        self.check(self.lib.ftb_rnn_plan_run(self.h, plan, self.stream()))

    def gemm_info(self, plan: int) -> dict:
        """What the library chose for the plan (include/ftb200.h: ftb_gemm_plan_info)."""
        pair, tiles = C.c_int(0), C.c_int(0)
        self.check(self.lib.ftb_gemm_plan_info(self.h, plan, C.byref(pair), C.byref(tiles)))
        return {'cta_pairs': bool(pair.value), 'slots': tiles.value}

    def rnn_info(self, plan: int) -> dict:
        """Layout the library chose for the plan on this device (include/ftb200.h: ftb_rnn_plan_info)."""
        handoff, launches, lanes = C.c_int(), C.c_int(), C.c_int()
        self.check(self.lib.ftb_rnn_plan_info(self.h, plan, C.byref(handoff), C.byref(launches), C.byref(lanes)))
        return {'handoff': 'cluster' if handoff.value else 'l2', 'kernel_launches': launches.value, 'lanes': lanes.value}

    def rnn_destroy(self, plan: int):
        self.check(self.lib.ftb_rnn_plan_destroy(self.h, plan))

    # ---- stateless kernels -------------------------------------------------------------------
    def _timed(self, name, fn):
        if self.timing is None:
            return fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        self.timing.append((name, a, b))

    def graph_capture(self, fn, stream: 'torch.cuda.Stream') -> int:
        """Run fn() with `stream` current and under capture; returns the id of the instantiated graph."""
        with torch.cuda.stream(stream):
            self.check(self.lib.ftb_graph_capture_begin(self.h, C.c_void_p(stream.cuda_stream)))
            try:
                fn()
            except Exception:
                gid = C.c_int(0)
                if self.lib.ftb_graph_capture_end(self.h, C.c_void_p(stream.cuda_stream), C.byref(gid)) == 0:
                    self.lib.ftb_graph_destroy(self.h, gid.value)
                raise
            gid = C.c_int(0)
            self.check(self.lib.ftb_graph_capture_end(self.h, C.c_void_p(stream.cuda_stream), C.byref(gid)))
        return gid.value

    def graph_launch(self, gid: int):
        self.check(self.lib.ftb_graph_launch(self.h, gid, self.stream()))

    def graph_destroy(self, gid: int):
        self.lib.ftb_graph_destroy(self.h, gid)

    def lr_scan(self, dur, batch, T, apply_fallback, cum, lens, meta):
        self._timed('lr_scan', lambda: self.check(self.lib.ftb_lr_scan(
            self.h, ptr(dur), batch, T, int(apply_fallback), ptr(cum), ptr(lens), ptr(meta), self.stream())))

    def lr_gather(self, args: LrGatherArgs):
        self._timed('lr_gather', lambda: self.check(self.lib.ftb_lr_gather(self.h, C.byref(args), self.stream())))

    def lr_index(self, cum, lens, batch, T, M, idx, row_stride=0, pad_row=-1):
        """Frame -> row map of the LengthRegulator (dense int32 [batch, M]): b * row_stride + phoneme, `pad_row` for the
        padded frames (defaults: the bare frame -> phoneme map, -1 = padded)."""
        self._timed('lr_index', lambda: self.check(self.lib.ftb_lr_index(
            self.h, ptr(cum), ptr(lens), batch, T, M, int(row_stride), int(pad_row), ptr(idx), self.stream())))

    def lr_regulate_static(self, x, x_stride, c, dur, batch, T, apply_fallback, M, out, out_stride,
                           idx=None, lens_out=None, meta_out=None):
        """Scan + gather with the intermediates in the library's static device scratch: no allocation on the call
        path (include/ftb200.h: ftb_lr_regulate_static).  x [.., x_ld] / out [.., out_ld] are 2-D row views."""
        self.check(self.lib.ftb_lr_regulate_static(
            self.h, ptr(x), x_stride, x.shape[-1], c, ptr(dur), batch, T, int(apply_fallback), M,
            ptr(out), out_stride, out.shape[-1], ptr(idx), ptr(lens_out), ptr(meta_out), self.stream()))

    def embed(self, tok, table, out, out_ld, batch, T, stride, tok2=None, table2=None, extra=None, round_tf32=True):
        e1 = table.shape[1]
        e2 = table2.shape[1] if table2 is not None else 0
        e3 = extra.shape[1] if extra is not None else 0
        # split fp16 [hi | lo | hi] / fp16 operand copy / tf32-rounded / as stored
        mode = 3 if int(round_tf32) == 3 else (2 if out.dtype == torch.float16 else int(round_tf32))
        self.check(self.lib.ftb_embed(self.h, ptr(tok), ptr(table), e1, ptr(tok2), ptr(table2), e2, ptr(extra), e3,
                                      batch, T, stride, ptr(out), out_ld, mode, table.shape[0],
                                      table2.shape[0] if table2 is not None else 0, self.stream()))

    def embed_batch(self, items, batch, T, stride):
        """items: [(tok, table, out, tok2, table2, extra, round_tf32)] -> ONE launch (include/ftb200.h: ftb_embed_batch)."""
        arr = (EmbedItem * len(items))()
        for a, (tok, table, out, tok2, table2, extra, round_tf32) in zip(arr, items):
            a.tok, a.table, a.e1 = ptr(tok), ptr(table), table.shape[1]
            a.tok2, a.table2, a.e2 = ptr(tok2), ptr(table2), (table2.shape[1] if table2 is not None else 0)
            a.extra, a.e3 = ptr(extra), (extra.shape[1] if extra is not None else 0)
            a.out, a.out_ld = ptr(out), out.shape[1]
            a.round_tf32 = 3 if int(round_tf32) == 3 else (2 if out.dtype == torch.float16 else int(round_tf32))
            a.rows1, a.rows2 = table.shape[0], (table2.shape[0] if table2 is not None else 0)
        self.check(self.lib.ftb_embed_batch(self.h, arr, len(items), batch, T, stride, self.stream()))

    def rowdot_batch(self, items, batch, T, stride):
        """items: [(x, w, bias, n_out, alpha, y)] -> ONE launch (include/ftb200.h: ftb_rowdot_batch)."""
        arr = (RowdotItem * len(items))()
        for a, (x, w, bias, n_out, alpha, y) in zip(arr, items):
            a.x, a.x_ld, a.k = ptr(x), x.shape[1], x.shape[1]
            a.w, a.bias, a.n_out, a.alpha, a.y = ptr(w), ptr(bias), n_out, float(alpha), ptr(y)
        self.check(self.lib.ftb_rowdot_batch(self.h, arr, len(items), batch, T, stride, self.stream()))

    def rowdot(self, x, x_ld, k, w, bias, n_out, alpha, batch, T, stride, y):
        self.check(self.lib.ftb_rowdot(self.h, ptr(x), x_ld, k, ptr(w), ptr(bias), n_out, float(alpha),
                                       batch, T, stride, ptr(y), self.stream()))

    def argmax(self, x, n, c, out):
        self.check(self.lib.ftb_argmax_lastdim(self.h, ptr(x), n, c, ptr(out), self.stream()))

    def add_layernorm(self, x, res, ld, e, g, b, g2, b2, batch, L, stride, out, round_tf32=True, out_full=None):
        self.check(self.lib.ftb_add_layernorm(self.h, ptr(x), ptr(res), ld, e, ptr(g), ptr(b), ptr(g2), ptr(b2),
                                              batch, L, stride, ptr(out), ptr(out_full), int(round_tf32),
                                              self.stream()))

    def posenc(self, tok, table, x_in, pe, scale, e, batch, L, stride, out, round_tf32=True, out_full=None):
        self.check(self.lib.ftb_posenc(self.h, ptr(tok), ptr(table), ptr(x_in), ptr(pe), ptr(scale), e, batch, L,
                                       stride, ptr(out), ptr(out_full), int(round_tf32),
                                       table.shape[0] if table is not None else 0, self.stream()))

    def stft(self, y, y_len, window, tw, n_fft, hop, frames, spec):
        self.check(self.lib.ftb_stft(self.h, ptr(y), y_len, y.shape[1], ptr(window), ptr(tw), n_fft, hop, y.shape[0],
                                     frames, ptr(spec), self.stream()))

    def istft(self, spec, window, tw, n_fft, hop, frames, frame_buf, y):
        self.check(self.lib.ftb_istft(self.h, ptr(spec), ptr(window), ptr(tw), n_fft, hop, y.shape[0], frames,
                                      ptr(frame_buf), ptr(y), y.shape[1], self.stream()))

    def gl_update(self, mag, rebuilt, tprev, nxt, momentum):
        self.check(self.lib.ftb_gl_update(self.h, ptr(mag), ptr(rebuilt), ptr(tprev), ptr(nxt), mag.numel(),
                                          float(momentum), self.stream()))

    def mha_split(self, qkv16, e, heads, key_tok, batch, L, stride, out, round_tf32=True):
        """Tensor-core attention on the split in_proj output (include/ftb200.h: ftb_mha_split)."""
        self.check(self.lib.ftb_mha_split(self.h, ptr(qkv16), e, heads, ptr(key_tok), batch, L, stride, ptr(out),
                                          int(round_tf32), self.stream()))

    def mha(self, qkv, e, heads, key_tok, batch, L, stride, out, round_tf32=True):
        self.check(self.lib.ftb_mha(self.h, ptr(qkv), e, heads, ptr(key_tok), batch, L, stride, ptr(out),
                                    int(round_tf32), self.stream()))
